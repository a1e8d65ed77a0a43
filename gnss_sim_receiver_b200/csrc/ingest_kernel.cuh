// Raw-sample ingest for sm_100a: the file item types and the 2-bit packed format of the reference's signal
// sources are turned into the ring's IQ type ON THE DEVICE, so only the raw bytes cross PCIe (a 2-bit "iq" stream is 1/16 of
// the gr_complex bytes the reference's flowgraph moves between its blocks).
//
// Replaces, in front of the tracking ring:
//   unpack_2bit_samples::work                 (src/algorithms/signal_source/gnuradio_blocks/unpack_2bit_samples.cc:121-208)
//   + interleaved_char_to_complex / char_to_float of Two_Bit_Packed_File_Signal_Source
//                                              (src/algorithms/signal_source/adapters/two_bit_packed_file_signal_source.cc:117-140)
//   Ibyte_To_Complex / Ishort_To_Complex       (src/algorithms/data_type_adapter/adapters/{ibyte,ishort}_to_complex.cc: plain
//                                               integer -> float conversion, no scaling)
// Layout: one thread per 4 output reals (= one packed byte) for 2-bit input, one thread per sample otherwise; consecutive
// threads write consecutive ring elements (coalesced); the first max_block samples of the ring are mirrored behind its end
// like every other ring write (trk.cu write_at).  HBM-bound, trivially: bytes out = ring element size x samples.
#pragma once
#include "gnss_b200.h"
#include <cstdint>

namespace gb200
{
struct IngestParams
{
    const unsigned char* src;  // staged raw bytes (device)
    char* ring;
    unsigned long long abs_start;  // absolute index of the first sample produced
    long long cap, mirror;         // ring_samples, max_block
    long long n_units;             // 2-bit: packed bytes; otherwise: samples
    int item_size;                 // 2-bit: bytes per file item
    int swap_items;                // 2-bit: reverse the bytes of every item (big-endian items on a little-endian machine)
    int order;                     // 2-bit: 0: s0 s1 s2 s3, 1: s3 s2 s1 s0, 2: s1 s0 s3 s2, 3: s2 s3 s0 s1 (unpack_2bit_samples.cc:146-205)
    int complex_pairs;             // 2-bit: 1 = consecutive reals are (I,Q); 0 = real stream, Q = 0
};

template <int DST>
__device__ __forceinline__ void ingest_store(const IngestParams& p, unsigned long long abs, float re, float im)
{
    const long long pos = (long long)(abs % (unsigned long long)p.cap);
#pragma unroll
    for (int rep = 0; rep < 2; rep++)
        {
            const long long at = rep ? p.cap + pos : pos;
            if (rep && pos >= p.mirror) break;
            if (DST == GB200_IQ_CF32)
                reinterpret_cast<float2*>(p.ring)[at] = make_float2(re, im);
            else if (DST == GB200_IQ_CS16)
                reinterpret_cast<short2*>(p.ring)[at] = make_short2((short)re, (short)im);
            else
                reinterpret_cast<char2*>(p.ring)[at] = make_char2((signed char)re, (signed char)im);
        }
}

// the four 2-bit fields of a byte, least significant first (gcc bit-field order on little-endian x86, which is what the
// reference's byte_2bit_struct resolves to), each a two's-complement value s in [-2, 1] mapped to 2 s + 1
__device__ __forceinline__ float ingest_field(unsigned b, int k)
{
    const int s = (int)(((b >> (2 * k)) & 3u) ^ 2u) - 2;
    return (float)(2 * s + 1);
}

template <int DST>
__global__ void __launch_bounds__(256) ingest_2bit_kernel(IngestParams p)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < p.n_units; i += stride)
        {
            long long j = i;
            if (p.swap_items)
                {
                    const long long item = i / p.item_size;
                    j = item * p.item_size + (p.item_size - 1 - (i - item * p.item_size));
                }
            const unsigned b = p.src[j];
            float v[4];
            // position of the field written first, second, ... for each order
            const int f0 = p.order == 0 ? 0 : p.order == 1 ? 3 : p.order == 2 ? 1 : 2;
            const int f1 = p.order == 0 ? 1 : p.order == 1 ? 2 : p.order == 2 ? 0 : 3;
            const int f2 = p.order == 0 ? 2 : p.order == 1 ? 1 : p.order == 2 ? 3 : 0;
            const int f3 = p.order == 0 ? 3 : p.order == 1 ? 0 : p.order == 2 ? 2 : 1;
            v[0] = ingest_field(b, f0);
            v[1] = ingest_field(b, f1);
            v[2] = ingest_field(b, f2);
            v[3] = ingest_field(b, f3);
            if (p.complex_pairs)
                {
                    const unsigned long long a = p.abs_start + 2ULL * (unsigned long long)i;
                    ingest_store<DST>(p, a, v[0], v[1]);
                    ingest_store<DST>(p, a + 1, v[2], v[3]);
                }
            else
                {
                    const unsigned long long a = p.abs_start + 4ULL * (unsigned long long)i;
#pragma unroll
                    for (int k = 0; k < 4; k++) ingest_store<DST>(p, a + k, v[k], 0.0f);
                }
        }
}

// LabSat 2/3 single-channel payload (labsat23_source.cc:364-433): little-endian int16 words, most significant bits first.
// BITS = 2: sample k of a word has I = bit 15-2k, Q = bit 14-2k, value 2 b - 1.  BITS = 4: I = (bit 15-4k, bit 13-4k),
// Q = (bit 14-4k, bit 12-4k) with (hi, lo) 00 -> +1, 01 -> +2, 10 -> -2, 11 -> -1.  One thread per OUTPUT sample: the
// 8 (4) threads of a word read the same 2 bytes (one L1 line per warp), the ring writes are fully coalesced.
__device__ __forceinline__ float labsat_level(unsigned hi, unsigned lo) { return hi ? (lo ? -1.0f : -2.0f) : (lo ? 2.0f : 1.0f); }

template <int BITS, int DST>
__global__ void __launch_bounds__(256) ingest_labsat_kernel(IngestParams p)
{
    constexpr int PER_WORD = 16 / (2 * (BITS / 2));  // 8 samples (BITS = 2) or 4 samples (BITS = 4)
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < p.n_units; i += stride)
        {
            const unsigned w = reinterpret_cast<const unsigned short*>(p.src)[i / PER_WORD];
            const int k = (int)(i % PER_WORD);
            float re, im;
            if (BITS == 2)
                {
                    re = (float)(2 * (int)((w >> (15 - 2 * k)) & 1u) - 1);
                    im = (float)(2 * (int)((w >> (14 - 2 * k)) & 1u) - 1);
                }
            else
                {
                    re = labsat_level((w >> (15 - 4 * k)) & 1u, (w >> (13 - 4 * k)) & 1u);
                    im = labsat_level((w >> (14 - 4 * k)) & 1u, (w >> (12 - 4 * k)) & 1u);
                }
            ingest_store<DST>(p, p.abs_start + (unsigned long long)i, re, im);
        }
}

// SRC: GB200_SRC_* of a whole-sample format (complex or real integer / float items)
template <int SRC, int DST>
__global__ void __launch_bounds__(256) ingest_convert_kernel(IngestParams p)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < p.n_units; i += stride)
        {
            float re, im = 0.0f;
            if (SRC == GB200_SRC_CF32)
                {
                    const float2 v = reinterpret_cast<const float2*>(p.src)[i];
                    re = v.x, im = v.y;
                }
            else if (SRC == GB200_SRC_CS16)
                {
                    const short2 v = reinterpret_cast<const short2*>(p.src)[i];
                    re = (float)v.x, im = (float)v.y;
                }
            else if (SRC == GB200_SRC_CS8)
                {
                    const char2 v = reinterpret_cast<const char2*>(p.src)[i];
                    re = (float)v.x, im = (float)v.y;
                }
            else if (SRC == GB200_SRC_S16)
                re = (float)reinterpret_cast<const short*>(p.src)[i];
            else
                re = (float)reinterpret_cast<const signed char*>(p.src)[i];
            ingest_store<DST>(p, p.abs_start + (unsigned long long)i, re, im);
        }
}
}  // namespace gb200
