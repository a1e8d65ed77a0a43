// Galileo E1, GPS L5 and Galileo E5a replica generators for the standalone host layer.
//
// Behaviour follows the reference's generators (cited per function); inside GNSS-SDR the adapters call the reference's own.
//   * Galileo memory codes (E1-B, E1-C, E5a-I, E5a-Q) are ICD tables, not LFSR outputs: they are read from
//     data/gnss_memory_codes.bin (one bit per chip, written by tools/make_code_tables.py), embedded below with .incbin;
//   * GPS L5 codes come from the two 13-stage registers of IS-GPS-705 (XA short-cycled at 8190, XB advanced per PRN).
// tests/test_codes_cpu.py checks every generator bit-exactly against the reference's own compiled generators for all PRNs.
#include "b200_gnss_codes.h"
#include <cmath>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

// the table file, placed in .rodata (path relative to this directory: the Makefile compiles from host/)
__asm__(
    ".section .rodata\n"
    ".balign 8\n"
    ".global gb200h_memory_codes_begin\n"
    "gb200h_memory_codes_begin:\n"
    ".incbin \"../data/gnss_memory_codes.bin\"\n"
    ".global gb200h_memory_codes_end\n"
    "gb200h_memory_codes_end:\n"
    ".byte 0\n"
    ".previous\n");
extern "C" const unsigned char gb200h_memory_codes_begin[];
extern "C" const unsigned char gb200h_memory_codes_end[];

namespace gb200h
{
namespace
{
constexpr int E1_LEN = 4092, L5_LEN = 10230, E5A_LEN = 10230;
enum Table
{
    E1B = 0,
    E1C = 1,
    E5AI = 2,
    E5AQ = 3,
    E5AQ_SECONDARY = 4
};

struct TableRef
{
    const unsigned char* base;
    uint32_t n_prn, n_chips, stride;
};

TableRef table(int which)
{
    const unsigned char* p = gb200h_memory_codes_begin;
    uint32_t version, n_tables;
    if (std::memcmp(p, "GB2C", 4) != 0) throw std::runtime_error("memory-code table: bad magic");
    std::memcpy(&version, p + 4, 4);
    std::memcpy(&n_tables, p + 8, 4);
    if (version != 1 || n_tables < 5) throw std::runtime_error("memory-code table: unknown version");
    const unsigned char* data = p + 12 + 8 * n_tables;
    TableRef r{};
    for (int t = 0; t <= which; t++)
        {
            std::memcpy(&r.n_prn, p + 12 + 8 * t, 4);
            std::memcpy(&r.n_chips, p + 16 + 8 * t, 4);
            r.stride = (r.n_chips + 7) / 8;
            r.base = data;
            data += static_cast<size_t>(r.n_prn) * r.stride;
        }
    if (data > gb200h_memory_codes_end) throw std::runtime_error("memory-code table: truncated");
    return r;
}

// chip i of PRN prn (1-based) as +-1
inline int chip(const TableRef& t, uint32_t prn, uint32_t i)
{
    const unsigned char b = t.base[static_cast<size_t>(prn - 1) * t.stride + (i >> 3)];
    return ((b >> (i & 7)) & 1) ? -1 : 1;
}

int e1_table_of(const char* signal_id)  // galileo_e1_signal_replica.cc:37,45: "1B" / "1C" anywhere in the id; -1 otherwise
{
    const std::string s(signal_id ? signal_id : "");
    if (s.size() >= 2 && s.rfind("1B") != std::string::npos) return E1B;
    if (s.size() >= 2 && s.rfind("1C") != std::string::npos) return E1C;
    return -1;
}

// galileo_e1_code_gen_int (:28-53): zeros when the PRN or the signal is not tabulated
void e1_chips(std::vector<int>& c, const char* signal_id, uint32_t prn)
{
    c.assign(E1_LEN, 0);
    const int which = e1_table_of(signal_id);
    if (prn < 1 || prn > 50 || which < 0) return;
    const TableRef t = table(which);
    for (int i = 0; i < E1_LEN; i++) c[i] = chip(t, prn, i);
}

inline int32_t aux_ceil2(float x) { return static_cast<int32_t>(static_cast<int64_t>(x + 1)); }  // gnss_signal_replica.cc:25

template <typename T>
void resample_any(const T* from, size_t n_from, T* dest, size_t n_dest, float fs_in, float fs_out)
{
    if (n_dest == 0) return;
    const float t_out = 1.0F / fs_out;
    for (size_t i = 0; i + 1 < n_dest; i++)
        {
            const float aux = (t_out * (static_cast<float>(i) + 1.0F)) * fs_in;
            dest[i] = from[static_cast<uint32_t>(aux_ceil2(aux) - 1)];
        }
    dest[n_dest - 1] = from[n_from - 1];
}

// GPS L5: 13-stage registers, s[12] is the output stage (gps_l5_signal_replica.cc:25-68)
struct Lfsr13
{
    unsigned s[13];
    Lfsr13()
    {
        for (auto& v : s) v = 1;
    }
    void push(unsigned in)
    {
        for (int i = 12; i > 0; i--) s[i] = s[i - 1];
        s[0] = in;
    }
};

// XB code advance per PRN 1..50 (IS-GPS-705 tables 3-Ia / 3-Ib, I5 and Q5)
constexpr int L5I_ADVANCE[50] = {266, 365, 804, 1138, 1509, 1559, 1756, 2084, 2170, 2303, 2527, 2687, 2930, 3471, 3940, 4132, 4332, 4924, 5343, 5443, 5641, 5816, 5898,
    5918, 5955, 6243, 6345, 6477, 6518, 6875, 7168, 7187, 7329, 7577, 7720, 7777, 8057, 5358, 3550, 3412, 819, 4608, 3698, 962, 3001, 4441, 4937, 3717, 4730, 7291};
constexpr int L5Q_ADVANCE[50] = {1701, 323, 5292, 2020, 5429, 7136, 1041, 5947, 4315, 148, 535, 1939, 5206, 5910, 3595, 5135, 6082, 6990, 3546, 1523, 4548, 4484, 1893,
    3961, 7106, 5299, 4660, 276, 4389, 3783, 1591, 1601, 749, 1387, 1661, 3210, 708, 4226, 5604, 6375, 3056, 1772, 3662, 4401, 5218, 2838, 6913, 1685, 1194, 6963};

// make_l5i / make_l5q (:127-161): 0/1 chips; XB is generated for one code length and indexed circularly over it
void l5_bits(std::vector<int>& bits, uint32_t prn, bool q)
{
    bits.assign(L5_LEN, 0);
    if (prn < 1 || prn > 50) return;  // :178: code_aux stays zero
    std::vector<unsigned> xa(L5_LEN), xb(L5_LEN);
    Lfsr13 a, b;
    for (int n = 0; n < L5_LEN; n++)
        {
            xa[n] = a.s[12];
            xb[n] = b.s[12];
            // XA: 1 + x^9 + x^10 + x^12 + x^13, reset to all ones after state 1111111111101 (short cycle of 8190)
            bool reset = a.s[11] == 0 && a.s[12] == 1;
            for (int i = 0; i < 11 && reset; i++) reset = a.s[i] == 1;
            if (reset)
                a = Lfsr13();
            else
                a.push(a.s[12] ^ a.s[11] ^ a.s[9] ^ a.s[8]);
            // XB: 1 + x + x^3 + x^4 + x^6 + x^7 + x^8 + x^12 + x^13
            b.push(b.s[12] ^ b.s[11] ^ b.s[7] ^ b.s[6] ^ b.s[5] ^ b.s[3] ^ b.s[2] ^ b.s[0]);
        }
    const int adv = q ? L5Q_ADVANCE[prn - 1] : L5I_ADVANCE[prn - 1];
    for (int n = 0; n < L5_LEN; n++) bits[n] = static_cast<int>(xa[n] ^ xb[(adv + n) % L5_LEN]);
}

int l5_sampled(std::complex<float>* dest, uint32_t prn, int32_t sampling_freq, bool q)
{
    // :193-222 (L5I: tc is the double quotient rounded to float, chips in the real part) / :262-291 (L5Q: float quotient,
    // chips in the IMAGINARY part); ts is a float quotient in both
    constexpr float tc_i = 1.0 / static_cast<float>(10230000);
    constexpr float tc_q = 1.0F / static_cast<float>(10230000);
    const float tc = q ? tc_q : tc_i;
    const auto samples_per_code = static_cast<int32_t>(static_cast<double>(sampling_freq) / (10230000.0 / static_cast<double>(L5_LEN)));
    const float ts = 1.0F / static_cast<float>(sampling_freq);
    std::vector<int> bits;
    l5_bits(bits, prn, q);
    for (int32_t i = 0; i < samples_per_code; i++)
        {
            const int32_t idx = i == samples_per_code - 1 ? L5_LEN - 1 : static_cast<int32_t>(std::ceil(ts * (i + 1.0F) / tc)) - 1;
            const float v = 1.0F - 2.0F * bits[idx];
            dest[i] = q ? std::complex<float>(0.0F, v) : std::complex<float>(v, 0.0F);
        }
    return samples_per_code;
}
}  // namespace

void resampler(const float* from, size_t n_from, float* dest, size_t n_dest, float fs_in, float fs_out) { resample_any(from, n_from, dest, n_dest, fs_in, fs_out); }
void resampler(const std::complex<float>* from, size_t n_from, std::complex<float>* dest, size_t n_dest, float fs_in, float fs_out)
{
    resample_any(from, n_from, dest, n_dest, fs_in, fs_out);
}

void galileo_e1_code_gen_sinboc11_float(float* dest, const char* signal_id, uint32_t prn)
{
    std::vector<int> c;
    e1_chips(c, signal_id, prn);
    for (int i = 0; i < E1_LEN; i++)
        {
            dest[2 * i] = static_cast<float>(c[i]);
            dest[2 * i + 1] = -dest[2 * i];
        }
}

int galileo_e1_code_gen_float_sampled(float* dest, const char* signal_id, bool cboc, uint32_t prn, int32_t sampling_freq, uint32_t chip_shift, bool secondary_flag)
{
    constexpr int32_t code_freq_basis = 1023000;
    const int32_t samples_per_chip = cboc ? 12 : 2;
    const uint32_t code_length = static_cast<uint32_t>(samples_per_chip) * E1_LEN;
    auto samples_per_code = static_cast<uint32_t>(static_cast<double>(sampling_freq) / (static_cast<double>(code_freq_basis) / E1_LEN));
    const uint32_t delay = ((static_cast<int32_t>(E1_LEN) - chip_shift) % static_cast<int32_t>(E1_LEN)) * samples_per_code / E1_LEN;  // unsigned arithmetic, as written (:150)
    std::vector<int> c;
    e1_chips(c, signal_id, prn);
    std::vector<float> signal(code_length);
    const int which = e1_table_of(signal_id);
    if (cboc)
        {
            // galileo_e1_gen_float (:113-139): alpha sinBOC(1,1) +- beta sinBOC(6,1) at 12 samples per chip
            const float alpha = std::sqrt(10.0F / 11.0F), beta = std::sqrt(1.0F / 11.0F);
            for (int i = 0; i < E1_LEN; i++)
                for (int j = 0; j < 12; j++)
                    {
                        const int s11 = j < 6 ? c[i] : -c[i];
                        const int s61 = (j & 1) ? -c[i] : c[i];
                        float v = 0.0F;  // neither "1B" nor "1C": the vector stays zero
                        if (which == E1B) v = alpha * static_cast<float>(s11) + beta * static_cast<float>(s61);
                        if (which == E1C) v = alpha * static_cast<float>(s11) - beta * static_cast<float>(s61);
                        signal[12 * i + j] = v;
                    }
        }
    else
        {
            for (int i = 0; i < E1_LEN; i++)
                {
                    signal[2 * i] = static_cast<float>(c[i]);
                    signal[2 * i + 1] = static_cast<float>(-c[i]);
                }
        }
    if (sampling_freq != samples_per_chip * code_freq_basis)
        {
            std::vector<float> resampled(samples_per_code);
            resampler(signal.data(), signal.size(), resampled.data(), resampled.size(), static_cast<float>(samples_per_chip * code_freq_basis), static_cast<float>(sampling_freq));
            signal = std::move(resampled);
        }
    if (which == E1C && secondary_flag)
        {
            static const char secondary[] = "0011100000001010110110010";  // E1-C secondary code CS25_1 (OS SIS ICD table 19)
            std::vector<float> with_secondary(25U * samples_per_code);
            for (uint32_t i = 0; i < 25U; i++)
                for (uint32_t k = 0; k < samples_per_code; k++) with_secondary[i * samples_per_code + k] = signal[k] * (secondary[i] == '0' ? 1.0F : -1.0F);
            samples_per_code *= 25U;
            signal = std::move(with_secondary);
        }
    for (uint32_t i = 0; i < samples_per_code; i++) dest[(i + delay) % samples_per_code] = signal[i];
    return static_cast<int>(samples_per_code);
}

int galileo_e1_code_gen_complex_sampled(std::complex<float>* dest, const char* signal_id, bool cboc, uint32_t prn, int32_t sampling_freq, uint32_t chip_shift, bool secondary_flag)
{
    auto samples_per_code = static_cast<uint32_t>(static_cast<double>(sampling_freq) / (1023000.0 / E1_LEN));
    if (e1_table_of(signal_id) == E1C && secondary_flag) samples_per_code *= 25U;
    std::vector<float> real_code(samples_per_code);
    galileo_e1_code_gen_float_sampled(real_code.data(), signal_id, cboc, prn, sampling_freq, chip_shift, secondary_flag);
    for (uint32_t i = 0; i < samples_per_code; i++) dest[i] = std::complex<float>(real_code[i], 0.0F);
    return static_cast<int>(samples_per_code);
}

void gps_l5i_code_gen_float(float* dest, uint32_t prn)
{
    std::vector<int> bits;
    l5_bits(bits, prn, false);
    for (int i = 0; i < L5_LEN; i++) dest[i] = 1.0F - 2.0F * static_cast<float>(bits[i]);
}

void gps_l5q_code_gen_float(float* dest, uint32_t prn)
{
    std::vector<int> bits;
    l5_bits(bits, prn, true);
    for (int i = 0; i < L5_LEN; i++) dest[i] = 1.0F - 2.0F * static_cast<float>(bits[i]);
}

int gps_l5i_code_gen_complex_sampled(std::complex<float>* dest, uint32_t prn, int32_t sampling_freq) { return l5_sampled(dest, prn, sampling_freq, false); }
int gps_l5q_code_gen_complex_sampled(std::complex<float>* dest, uint32_t prn, int32_t sampling_freq) { return l5_sampled(dest, prn, sampling_freq, true); }

void galileo_e5_a_code_gen_complex_primary(std::complex<float>* dest, int32_t prn, const char* signal_id)
{
    if (prn < 1 || prn > 50 || !signal_id || signal_id[0] != '5') return;  // :38-41: dest untouched
    const TableRef ti = table(E5AI), tq = table(E5AQ);
    for (int i = 0; i < E5A_LEN; i++)
        {
            if (signal_id[1] == 'Q')
                dest[i] = std::complex<float>(static_cast<float>(chip(tq, prn, i)), 0.0F);
            else if (signal_id[1] == 'I')
                dest[i] = std::complex<float>(static_cast<float>(chip(ti, prn, i)), 0.0F);
            else if (signal_id[1] == 'X')
                dest[i] = std::complex<float>(static_cast<float>(chip(ti, prn, i)), static_cast<float>(chip(tq, prn, i)));
        }
}

int galileo_e5_a_code_gen_complex_sampled(std::complex<float>* dest, uint32_t prn, const char* signal_id, int32_t sampling_freq, uint32_t chip_shift)
{
    constexpr uint32_t code_length = E5A_LEN;
    constexpr int32_t code_freq_basis = 10230000;
    const auto samples_per_code = static_cast<uint32_t>(static_cast<double>(sampling_freq) / (static_cast<double>(code_freq_basis) / static_cast<double>(code_length)));
    const uint32_t delay = ((code_length - chip_shift) % code_length) * samples_per_code / code_length;
    std::vector<std::complex<float>> code(code_length);
    galileo_e5_a_code_gen_complex_primary(code.data(), static_cast<int32_t>(prn), signal_id);
    if (sampling_freq != code_freq_basis)
        {
            std::vector<std::complex<float>> resampled(samples_per_code);
            resampler(code.data(), code.size(), resampled.data(), resampled.size(), static_cast<float>(code_freq_basis), static_cast<float>(sampling_freq));
            code = std::move(resampled);
        }
    for (uint32_t i = 0; i < samples_per_code; i++) dest[(i + delay) % samples_per_code] = code[i];
    return static_cast<int>(samples_per_code);
}

std::string galileo_e5a_q_secondary_code(uint32_t prn)
{
    const TableRef t = table(E5AQ_SECONDARY);
    std::string out(100, '0');
    if (prn < 1 || prn > t.n_prn) return out;
    for (uint32_t i = 0; i < 100; i++) out[i] = chip(t, prn, i) < 0 ? '1' : '0';
    return out;
}
}  // namespace gb200h
