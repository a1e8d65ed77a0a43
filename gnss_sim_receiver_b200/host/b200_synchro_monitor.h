// Gnss_Synchro wire format out of the tracking path: the protobuf monitor stream.
//
// Reference shape: gnss_synchro_monitor (src/core/monitor/gnss_synchro_monitor.cc:41-100) decimates each channel's
// Gnss_Synchro stream and hands single-element vectors to Gnss_Synchro_Udp_Sink (gnss_synchro_udp_sink.cc:24-75), which
// serialises them with Serdes_Gnss_Synchro::createProtobuffer (serdes_gnss_synchro.h:72-119) into the `Observables` message
// of docs/protobuf/gnss_synchro.proto and sends one datagram per client address.  Configured by `Monitor.*`,
// `TrackingMonitor.*`, `AcquisitionMonitor.*` (gnss_flowgraph.cc:194-262: enable_monitor, enable_protobuf, client_addresses
// separated by '_', udp_port, decimation_factor).
//
// libprotobuf and Boost.Asio are not part of this tree: the proto3 wire encoding of that one message is written out by hand
// (fields in number order, zero-valued scalars and empty strings omitted, like the C++ runtime does) and the datagrams go out
// through POSIX sockets.  The Boost binary archive alternative (enable_protobuf=false) is not provided.
#ifndef GB200_SYNCHRO_MONITOR_H
#define GB200_SYNCHRO_MONITOR_H

#include "gnss_sdr_compat.h"
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

namespace gb200h
{
// serdes_gnss_synchro.h:72-119: `Observables{ repeated GnssSynchro observable = 1; }`
std::string create_protobuffer(const std::vector<Gnss_Synchro>& vgs);
// one GnssSynchro message body (without the enclosing field tag)
void encode_gnss_synchro(const Gnss_Synchro& gs, std::string* out);

class B200SynchroUdpSink
{
public:
    B200SynchroUdpSink(const std::vector<std::string>& addresses, uint16_t port);
    ~B200SynchroUdpSink();
    B200SynchroUdpSink(const B200SynchroUdpSink&) = delete;
    B200SynchroUdpSink& operator=(const B200SynchroUdpSink&) = delete;
    bool write_gnss_synchro(const std::vector<Gnss_Synchro>& stocks);  // gnss_synchro_udp_sink.cc:43-75
    uint64_t datagrams() const { return sent_; }

private:
    struct Endpoint
    {
        uint32_t addr_be;
        uint16_t port_be;
    };
    std::vector<Endpoint> endpoints_;
    int fd_{-1};
    uint64_t sent_{0};
};

class B200SynchroMonitor
{
public:
    // `role`: "Monitor", "TrackingMonitor" or "AcquisitionMonitor"; enabled() is false when <role>.enable_monitor is not set
    B200SynchroMonitor(const ConfigurationInterface* configuration, const std::string& role, int n_channels);
    bool enabled() const { return sink_ != nullptr; }
    // one channel's new items, oldest first: every decimation_factor-th goes out as its own datagram (:64-90; the count is
    // kept across calls here, the reference restarts it with every general_work)
    void write(int channel, const Gnss_Synchro* items, int n);
    uint64_t datagrams() const { return sink_ ? sink_->datagrams() : 0; }

private:
    std::unique_ptr<B200SynchroUdpSink> sink_;
    int decimation_{1};
    std::vector<int> count_;
};
}  // namespace gb200h
#endif
