#include "b200_synchro_monitor.h"
#include <arpa/inet.h>
#include <cstring>
#include <memory>
#include <netinet/in.h>
#include <sys/socket.h>
#include <unistd.h>

namespace gb200h
{
namespace
{
// ---- proto3 wire primitives: tag = (field << 3) | wire type; 0 varint, 1 fixed64, 2 length-delimited
void put_varint(std::string* o, uint64_t v)
{
    while (v >= 0x80)
        {
            o->push_back(static_cast<char>((v & 0x7F) | 0x80));
            v >>= 7;
        }
    o->push_back(static_cast<char>(v));
}
void put_tag(std::string* o, uint32_t field, uint32_t wire) { put_varint(o, (static_cast<uint64_t>(field) << 3) | wire); }
void put_u64(std::string* o, uint32_t field, uint64_t v)  // uint32 / uint64 / bool, and int32 / int64 through sign extension
{
    if (v == 0) return;
    put_tag(o, field, 0);
    put_varint(o, v);
}
void put_i64(std::string* o, uint32_t field, int64_t v) { put_u64(o, field, static_cast<uint64_t>(v)); }  // negative: ten bytes
void put_double(std::string* o, uint32_t field, double v)
{
    uint64_t bits;
    std::memcpy(&bits, &v, 8);
    if (bits == 0) return;  // only +0.0 is the default
    put_tag(o, field, 1);
    char b[8];
    for (int i = 0; i < 8; i++) b[i] = static_cast<char>(bits >> (8 * i));  // little-endian
    o->append(b, 8);
}
void put_string(std::string* o, uint32_t field, const std::string& s)
{
    if (s.empty()) return;
    put_tag(o, field, 2);
    put_varint(o, s.size());
    o->append(s);
}
}  // namespace

// field numbers of docs/protobuf/gnss_synchro.proto:8-43, values as serdes_gnss_synchro.h:78-114 sets them
void encode_gnss_synchro(const Gnss_Synchro& gs, std::string* o)
{
    put_string(o, 1, std::string(1, gs.System));            // a one-character string even for '\0' (:80-81)
    put_string(o, 2, std::string(gs.Signal, gs.Signal + 2));  // exactly two characters (:83-86)
    put_u64(o, 3, gs.PRN);
    put_i64(o, 4, gs.Channel_ID);
    put_double(o, 5, gs.Acq_delay_samples);
    put_double(o, 6, gs.Acq_doppler_hz);
    put_u64(o, 7, gs.Acq_samplestamp_samples);
    put_u64(o, 8, gs.Acq_doppler_step);
    put_u64(o, 9, gs.Flag_valid_acquisition ? 1 : 0);
    put_i64(o, 10, gs.fs);
    put_double(o, 11, gs.Prompt_I);
    put_double(o, 12, gs.Prompt_Q);
    put_double(o, 13, gs.CN0_dB_hz);
    put_double(o, 14, gs.Carrier_Doppler_hz);
    put_double(o, 15, gs.Carrier_phase_rads);
    put_double(o, 16, gs.Code_phase_samples);
    put_u64(o, 17, gs.Tracking_sample_counter);
    put_u64(o, 18, gs.Flag_valid_symbol_output ? 1 : 0);
    put_i64(o, 19, gs.correlation_length_ms);
    put_u64(o, 20, gs.Flag_valid_word ? 1 : 0);
    put_u64(o, 21, gs.TOW_at_current_symbol_ms);
    put_double(o, 22, gs.Pseudorange_m);
    put_double(o, 23, gs.RX_time);
    put_u64(o, 24, gs.Flag_valid_pseudorange ? 1 : 0);
    put_double(o, 25, gs.interp_TOW_ms);
    put_u64(o, 26, gs.Flag_PLL_180_deg_phase_locked ? 1 : 0);
}

std::string create_protobuffer(const std::vector<Gnss_Synchro>& vgs)
{
    std::string data, one;
    for (const auto& gs : vgs)
        {
            one.clear();
            encode_gnss_synchro(gs, &one);
            put_tag(&data, 1, 2);  // repeated GnssSynchro observable = 1: present even when the message body is empty
            put_varint(&data, one.size());
            data.append(one);
        }
    return data;
}

B200SynchroUdpSink::B200SynchroUdpSink(const std::vector<std::string>& addresses, uint16_t port)
{
    for (const auto& a : addresses)
        {
            in_addr ia{};
            if (inet_pton(AF_INET, a.c_str(), &ia) != 1) continue;  // the reference keeps going with the error code set
            endpoints_.push_back(Endpoint{ia.s_addr, htons(port)});
        }
    fd_ = ::socket(AF_INET, SOCK_DGRAM, 0);
}

B200SynchroUdpSink::~B200SynchroUdpSink()
{
    if (fd_ >= 0) ::close(fd_);
}

bool B200SynchroUdpSink::write_gnss_synchro(const std::vector<Gnss_Synchro>& stocks)
{
    if (fd_ < 0) return false;
    const std::string outbound = create_protobuffer(stocks);
    bool ok = true;
    for (const auto& e : endpoints_)
        {
            sockaddr_in sa{};
            sa.sin_family = AF_INET;
            sa.sin_addr.s_addr = e.addr_be;
            sa.sin_port = e.port_be;
            const ssize_t n = ::sendto(fd_, outbound.data(), outbound.size(), 0, reinterpret_cast<const sockaddr*>(&sa), sizeof sa);
            if (n < 0)
                ok = false;
            else
                sent_++;
        }
    return ok;
}

B200SynchroMonitor::B200SynchroMonitor(const ConfigurationInterface* configuration, const std::string& role, int n_channels)
    : count_(static_cast<size_t>(n_channels > 0 ? n_channels : 0), 0)
{
    if (!configuration->property(role + ".enable_monitor", false)) return;
    if (!configuration->property(role + ".enable_protobuf", true)) return;  // the Boost archive stream is not provided
    // gnss_flowgraph.cc:203-207: addresses separated by '_'
    const std::string all = configuration->property(role + ".client_addresses", std::string("127.0.0.1"));
    std::vector<std::string> addresses;
    size_t from = 0;
    while (from <= all.size())
        {
            const size_t to = all.find('_', from);
            const std::string a = all.substr(from, to == std::string::npos ? std::string::npos : to - from);
            if (!a.empty()) addresses.push_back(a);
            if (to == std::string::npos) break;
            from = to + 1;
        }
    const int default_port = role == "Monitor" ? 1234 : role == "AcquisitionMonitor" ? 1235 : 1236;  // :211,234,257
    const int port = configuration->property(role + ".udp_port", default_port);
    decimation_ = configuration->property(role + ".decimation_factor", 1);
    sink_ = std::make_unique<B200SynchroUdpSink>(addresses, static_cast<uint16_t>(port));
}

void B200SynchroMonitor::write(int channel, const Gnss_Synchro* items, int n)
{
    if (!sink_ || channel < 0 || channel >= static_cast<int>(count_.size())) return;
    std::vector<Gnss_Synchro> stocks(1);
    for (int i = 0; i < n; i++)
        {
            if (++count_[channel] >= decimation_)
                {
                    stocks[0] = items[i];
                    sink_->write_gnss_synchro(stocks);
                    count_[channel] = 0;
                }
        }
}
}  // namespace gb200h
