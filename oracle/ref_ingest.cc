// TEST INFRASTRUCTURE — not part of the product.
// C door to the reference's own 2-bit unpacker, compiled UNMODIFIED from /root/reference
// (src/algorithms/signal_source/gnuradio_blocks/unpack_2bit_samples.{h,cc}) against the few declarations of oracle/shim/gr_block/.
#include "unpack_2bit_samples.h"
#include <cstdint>

extern "C" {
// n_bytes packed input bytes (a multiple of item_size) -> 4 * n_bytes samples in {-3, -1, 1, 3}; returns work()'s return value
int ref_unpack_2bit(const int8_t* in, long n_bytes, int item_size, int big_endian_items, int big_endian_bytes, int reverse_interleaving, int8_t* out)
{
    unpack_2bit_samples u(big_endian_bytes != 0, static_cast<size_t>(item_size), big_endian_items != 0, reverse_interleaving != 0);
    gr_vector_const_void_star iv{in};
    gr_vector_void_star ov{out};
    return u.work(static_cast<int>(4 * n_bytes), iv, ov);
}
}
