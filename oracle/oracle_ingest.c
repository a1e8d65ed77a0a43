// TEST INFRASTRUCTURE — CPU restatement of the raw-sample conversions in front of the hot paths.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use anything under oracle/.
//
// Pinned (2-bit unpacker): the reference's unpack_2bit_samples.cc compiles unmodified into oracle/_ref against three stub declarations
// (oracle/shim/gr_block) and tests/test_ingest_cpu.py compares its work() with this file bit for bit over every option combination.
// The LabSat decoder stays on hand-worked known answers.  The restatement keeps the reference's formulation — a bit-field view of each byte, laid out by THIS
// compiler exactly as it lays out the reference's byte_2bit_struct — while the CUDA kernel uses shifts and masks, so the two
// sides of the parity test are independent formulations.  Known answers in tests/test_ingest_cpu.py.
#include "oracle.h"
#include <stdlib.h>
#include <string.h>

// four signed 2-bit fields in one byte (unpack_2bit_samples.cc:21-36)
typedef struct
{
    signed int f0 : 2;
    signed int f1 : 2;
    signed int f2 : 2;
    signed int f3 : 2;
} quad2_t;

typedef union
{
    int8_t byte;
    quad2_t q;
} byte_view_t;

static int host_is_big_endian(void)  // :39-48
{
    const uint32_t probe = 0x01020304u;
    return *(const unsigned char*)&probe == 1;
}

// unpack_2bit_samples::work (:121-208) with the constructor's endianness decisions (:99-115): out receives 4 * nbytes values
void orc_unpack_2bit(const int8_t* in, int64_t nbytes, int item_size, int big_endian_items, int big_endian_bytes, int reverse_interleaving, int8_t* out)
{
    const int swap_items = item_size > 1 && (host_is_big_endian() != (big_endian_items != 0));
    // systemBytesAreBigEndian() (:51-61) compares the first byte of a one-byte object with 1: false on every machine
    const int swap_bytes = big_endian_bytes != 0;
    int8_t* work = NULL;
    if (swap_items)  // swapEndianness (:64-82): reverse the bytes of every item
        {
            work = (int8_t*)malloc((size_t)(nbytes > 0 ? nbytes : 1));
            const int64_t items = nbytes / item_size;
            for (int64_t it = 0; it < items; it++)
                for (int k = 0; k < item_size; k++) work[it * item_size + k] = in[it * item_size + (item_size - 1 - k)];
            in = work;
        }
    int64_t n = 0;
    for (int64_t i = 0; i < nbytes; i++)
        {
            byte_view_t v;
            v.byte = in[i];
            const int8_t s0 = (int8_t)(2 * v.q.f0 + 1), s1 = (int8_t)(2 * v.q.f1 + 1), s2 = (int8_t)(2 * v.q.f2 + 1), s3 = (int8_t)(2 * v.q.f3 + 1);
            if (!reverse_interleaving)
                {
                    if (swap_bytes)
                        out[n] = s3, out[n + 1] = s2, out[n + 2] = s1, out[n + 3] = s0;  // :146-158
                    else
                        out[n] = s0, out[n + 1] = s1, out[n + 2] = s2, out[n + 3] = s3;  // :160-171
                }
            else
                {
                    if (swap_bytes)
                        out[n] = s2, out[n + 1] = s3, out[n + 2] = s0, out[n + 3] = s1;  // :176-188
                    else
                        out[n] = s1, out[n + 1] = s0, out[n + 2] = s3, out[n + 3] = s2;  // :190-204
                }
            n += 4;
        }
    free(work);
}

// interleaved_char_to_complex(false) / char_to_float + imaginary part 0 (two_bit_packed_file_signal_source.cc:117-140),
// Ibyte_To_Complex, Ishort_To_Complex: integer -> float without scaling.  kind: 0 cf32, 1 cs16, 2 cs8, 3 s16, 4 s8
// (GB200_SRC_* of include/gnss_b200.h); out: n_samples interleaved (re, im) floats
void orc_items_to_complex(const void* in, int64_t n_samples, int kind, float* out)
{
    for (int64_t i = 0; i < n_samples; i++)
        {
            float re, im = 0.0f;
            switch (kind)
                {
                case 0: re = ((const float*)in)[2 * i], im = ((const float*)in)[2 * i + 1]; break;
                case 1: re = (float)((const int16_t*)in)[2 * i], im = (float)((const int16_t*)in)[2 * i + 1]; break;
                case 2: re = (float)((const int8_t*)in)[2 * i], im = (float)((const int8_t*)in)[2 * i + 1]; break;
                case 3: re = (float)((const int16_t*)in)[i]; break;
                default: re = (float)((const int8_t*)in)[i]; break;
                }
            out[2 * i] = re;
            out[2 * i + 1] = im;
        }
}

// labsat23_source::decode_samples_one_channel (labsat23_source.cc:364-433) over a payload of native int16 words
// (general_work :960-972 for 2 bits per sample, :1016-1028 for 4): out receives 8 (bits = 2) or 4 (bits = 4) interleaved
// (re, im) float pairs per word.  Bit k of a word is tested the way std::bitset<16>::operator[] does.
static int word_bit(int16_t w, int k) { return (int)(((uint16_t)w >> k) & 1u); }

void orc_labsat_one_channel(const int16_t* in, int64_t n_words, int bits, float* out)
{
    int64_t o = 0;
    for (int64_t n = 0; n < n_words; n++)
        {
            const int16_t w = in[n];
            if (bits == 2)
                for (int i = 0; i < 8; i++)  // :370-376: (b, b') * 2 - (1, 1)
                    {
                        out[2 * o] = (float)word_bit(w, 15 - 2 * i) * 2.0f - 1.0f;
                        out[2 * o + 1] = (float)word_bit(w, 14 - 2 * i) * 2.0f - 1.0f;
                        o++;
                    }
            else
                for (int i = 0; i < 4; i++)
                    {
                        float re, im;
                        if (word_bit(w, 15 - 4 * i))  // :384-405
                            re = word_bit(w, 13 - 4 * i) ? -1.0f : -2.0f;
                        else
                            re = word_bit(w, 13 - 4 * i) ? 2.0f : 1.0f;
                        if (word_bit(w, 14 - 4 * i))  // :408-429
                            im = word_bit(w, 12 - 4 * i) ? -1.0f : -2.0f;
                        else
                            im = word_bit(w, 12 - 4 * i) ? 2.0f : 1.0f;
                        out[2 * o] = re;
                        out[2 * o + 1] = im;
                        o++;
                    }
        }
}
