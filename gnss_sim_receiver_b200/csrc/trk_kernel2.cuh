// Tracking multicorrelator, run-based kernel (K1b) for sm_100a — used when the code advances by less than one table entry
// over a run of R samples (R * code_step <= 0.9, or 21 * code_step <= 0.92 for the 22-sample cf32 tier: e.g. GPS L1 at >= 12 MS/s,
// Galileo E1 at >= 25 MS/s).
//
// Same contract as trk_kernel.cuh (Cpu_Multicorrelator_Real_Codes::Carrier_wipeoff_multicorrelator_resampler,
// src/algorithms/tracking/libs/cpu_multicorrelator_real_codes.cc:103-126), different work decomposition:
//   * a CTA (4 warps) owns one command and walks its window tile by tile (128 runs of R samples).  Every WARP runs its own
//     pipeline over its 32 runs of a tile: its 32 * R contiguous samples are staged into shared memory by one TMA bulk copy
//     (cp.async.bulk + the warp's mbarrier); a lane pulls its run into registers, and as soon as the warp's runs are consumed
//     (after the wipe-off) lane 0 issues the copy of the NEXT tile into the same buffer, so the transfer overlaps the tap phase
//     and the next tile's set-up: registers are the second buffer, and there is no block-wide barrier inside the tile loop.
//     R is chosen so that a run is an ODD number of 16-byte chunks (cf32: 22, 18 or 10 samples, int16: 20 or 12, int8: 24 or 8):
//     the lanes' 128-bit reads of the plain contiguous rows are then conflict-free without padding;
//   * a lane owns one run: it wipes the carrier off once, keeps the R prefix sums of the wiped samples, and for every
//     tap finds the single chip boundary inside the run, so a tap costs O(1) per run instead of O(R):
//       out_t += c[idx] * P[k_t] + c[idx+1] * (S - P[k_t])
//     where k_t = number of samples of the run that still see table entry idx.  k_t comes from the reference's exact fp32
//     index law (volk_gnsssdr_32f_xn_resampler_32f_xn.h:73): an estimate from the fractional distance to the next integer,
//     confirmed by ONE exact evaluation of the law at the candidate sample (the estimate is within 0.02 sample of the fp32
//     crossing under the host's applicability check, so the boundary is the candidate or its successor);
//   * the arithmetic is packed: complex (re, im) pairs and PAIRS OF TAPS go through the sm_100 f32x2 instructions
//     (fma/add/mul.rn.f32x2 and add.rm.f32x2: SASS FFMA2 / FADD2 / FMUL2 / FADD2.RM), every lane of which rounds exactly like
//     the scalar instruction it replaces — the index law stays the reference's three separately rounded operations;
//   * the carrier phasor of a run is seeded exactly (sincospif of the 64-bit fixed-point phase) on the first tile and every 8th,
//     and advanced by one complex multiplication with exp(j 2 pi du TILE) in between;
//   * accumulators stay in registers across all tiles of a command: one block reduce per command, no cross-CTA partials.
#pragma once
#include "trk_kernel.cuh"
#include <type_traits>

namespace gb200
{
constexpr int TRK2_THREADS = 128;
constexpr int TRK2_WARPS = TRK2_THREADS / 32;
constexpr int TRK2_TAB = 160;     // code-pair window entries per warp per tile (0.9 * 32 entries of code advance + tap span <= 100 + margin)
constexpr int TRK2_RESEED = 8;    // tiles between exact phasor seeds

template <int IQ>
struct IqRow;  // bytes per element
template <>
struct IqRow<GB200_IQ_CF32>
{
    static constexpr int EB = 8;
};
template <>
struct IqRow<GB200_IQ_CS16>
{
    static constexpr int EB = 4;
};
template <>
struct IqRow<GB200_IQ_CS8>
{
    static constexpr int EB = 2;
};

// shared-memory accesses through 32-bit shared-window addresses computed once per kernel (a generic pointer costs a cvta, i.e. a read of
// the CTA's shared-window base, at every use inside the tile loop)
__device__ __forceinline__ uint4 ld_shared_v4(unsigned addr)
{
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void st_shared_f2(unsigned addr, float a, float b) { asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(addr), "f"(a), "f"(b) : "memory"); }
__device__ __forceinline__ void mbar_expect_tx_a(unsigned bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait_a(unsigned bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(bar),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s_a(unsigned dst, const void* gmem_src, uint32_t bytes, unsigned bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(gmem_src), "r"(bytes), "r"(bar) : "memory");
}

// one 16-byte chunk of a staged run -> 16/EB complex samples
template <int IQ>
__device__ __forceinline__ void trk2_unpack(const uint4 v, float2* x)
{
    if (IQ == GB200_IQ_CF32)
        {
            x[0] = make_float2(__uint_as_float(v.x), __uint_as_float(v.y));
            x[1] = make_float2(__uint_as_float(v.z), __uint_as_float(v.w));
        }
    else if (IQ == GB200_IQ_CS16)
        {
            const unsigned w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int i = 0; i < 4; i++) x[i] = make_float2((float)(short)(w[i] & 0xffffu), (float)(short)(w[i] >> 16));
        }
    else
        {
            const unsigned w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int i = 0; i < 4; i++)
                {
                    x[2 * i] = make_float2((float)(signed char)(w[i] & 0xffu), (float)(signed char)((w[i] >> 8) & 0xffu));
                    x[2 * i + 1] = make_float2((float)(signed char)((w[i] >> 16) & 0xffu), (float)(signed char)(w[i] >> 24));
                }
        }
}

// one step of the halving butterfly over CNT partial sums with partner lane ^ O, then the following steps (compile-time recursion
// keeps every array index static, so the sums stay in registers)
template <int CNT, int O>
__device__ __forceinline__ void trk2_halve(float* a, int lane)
{
    constexpr int HALF = (CNT + 1) / 2;
    const bool upper = (lane & O) != 0;
#pragma unroll
    for (int i = 0; i < HALF; i++)
        {
            const float hi = (i + HALF < CNT) ? a[i + HALF] : 0.0f;
            const float keep = upper ? hi : a[i];
            const float give = upper ? a[i] : hi;
            a[i] = keep + __shfl_xor_sync(0xffffffffu, give, O);
        }
    if constexpr (O > 1) trk2_halve<HALF, O / 2>(a, lane);
}

template <int IQ, int CH>
constexpr size_t trk2_smem_bytes()
{
    constexpr int R = CH * 16 / IqRow<IQ>::EB;
    return (size_t)TRK2_THREADS * R * IqRow<IQ>::EB + (size_t)(R + 2) * TRK2_THREADS * sizeof(float2) + (size_t)TRK2_WARPS * TRK2_TAB * sizeof(float2);
}

// resident CTAs per SM that the kernel's shared memory allows (228 KB per SM, 1 KB reserved per CTA), capped at 8: the register
// allocation is bounded accordingly (65536 / (128 * blocks))
template <int IQ, int CH>
constexpr int trk2_min_blocks()
{
    const size_t fit = (228 * 1024) / (trk2_smem_bytes<IQ, CH>() + 1024 + 512);
    return fit > 8 ? 8 : (int)fit;
}

// T taps, IQ sample type, CH = 16-byte chunks per run (odd)
template <int T, int IQ, int CH>
__global__ void __launch_bounds__(TRK2_THREADS, trk2_min_blocks<IQ, CH>()) trk_run_kernel(const TrkDevCmd* __restrict__ cmds, const unsigned char* __restrict__ ring, float* __restrict__ out)
{
    constexpr int EB = IqRow<IQ>::EB;
    constexpr int SPC = 16 / EB;        // samples per chunk
    constexpr int R = CH * SPC;         // samples per run
    constexpr int WROW = 32 * R;        // samples a warp stages per tile
    constexpr int WROW_BYTES = WROW * EB;
    constexpr int TILE = TRK2_THREADS * R;
    constexpr int TP = (T + 1) / 2;     // tap pairs
    constexpr float MAGIC = 12582912.0f;  // 1.5 * 2^23: floor(x) for |x| < 2^22 sits in the mantissa of RD(x + MAGIC)
    constexpr int MAGICBITS = 0x4B400000;

    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t s_bar[TRK2_WARPS];
    __shared__ float s_red[TRK2_WARPS][2 * T];
    __shared__ unsigned s_hide[TRK2_WARPS][2];

    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
    unsigned char* my_rows = smem + w * WROW_BYTES;
    float2* my_pref = reinterpret_cast<float2*>(smem + TRK2_THREADS * R * EB) + w * (R + 2) * 32;  // [R + 2][32]
    float2* my_tab = reinterpret_cast<float2*>(smem + TRK2_THREADS * R * EB + (R + 2) * TRK2_THREADS * sizeof(float2)) + w * TRK2_TAB;  // (c[j] - c[j+1], c[j+1])

    const unsigned rows_a = smem_u32(my_rows), row_a = rows_a + (unsigned)lane * (CH * 16), pref_addr = smem_u32(my_pref + lane), tab_a = smem_u32(my_tab), bar_a = smem_u32(&s_bar[w]);

    // ---- the command (every thread reads the same 176 bytes: broadcast)
    const TrkDevCmd& c = cmds[blockIdx.x];
    const long long ws = (long long)((static_cast<const unsigned char*>(c.iq) - ring) / EB);  // ring element index of the window's first sample
    const long long tile0 = ws & ~7LL;                                                           // 16-byte aligned for every sample type
    const int n = c.n, L = c.L;
    if (n <= 0) return;  // an empty slot of a device-formed command array (trk_loop.cuh): uniform for the CTA, before any barrier
    const int n_tiles = (int)((ws + n - tile0 + TILE - 1) / TILE);
    const int lead = (int)(tile0 - ws);  // <= 0: sample index of the first staged sample
    const unsigned long long u0 = c.u0, du = c.du;
    const float step = c.code_step, rem = c.rem_code, inv_step = __frcp_rn(c.code_step), shift_min = c.shift_min, shift_max = c.shift_max;
    const float* code = reinterpret_cast<const float*>(c.code);
    const float2 inc = trk_phasor(du);
    const float2 inc_tile = trk_phasor(du * (unsigned long long)TILE);
    f32x2 shp[TP];
#pragma unroll
    for (int p = 0; p < TP; p++) shp[p] = pk2(c.shifts[2 * p], c.shifts[2 * p + 1 < T ? 2 * p + 1 : T - 1]);

    if (lane == 0)
        {
            mbar_init(&s_bar[w], 1);
            fence_barrier_init();
            s_hide[w][0] = 0u - (unsigned)MAGICBITS;
            s_hide[w][1] = 0u;
        }
    my_pref[lane] = make_float2(0.0f, 0.0f);  // P[0]
    __syncwarp();
    // -bits(MAGIC) and a zero, read back through shared memory so that ptxas cannot see them: knowing the constant, it splits the biased
    // table and prefix addresses below into (variable part) + constant again, and without the XOR it re-associates tab_adj + 8 * bits into
    // (bits + variable) * 8 + base — either way an extra IADD3 per tap
    const unsigned hide = *reinterpret_cast<volatile unsigned*>(&s_hide[w][0]), zero = *reinterpret_cast<volatile unsigned*>(&s_hide[w][1]);
    const unsigned char* src = ring + ((size_t)tile0 + (size_t)w * WROW) * EB;
    if (lane == 0)
        {
            mbar_expect_tx_a(bar_a, WROW_BYTES);
            bulk_g2s_a(rows_a, src, WROW_BYTES, bar_a);
        }
    // entries of a warp's code window: the code advance over its 32 runs plus the tap span, constant per command
    const int span_c = min((int)(step * (float)WROW + (shift_max - shift_min)) + 5, TRK2_TAB);

    f32x2 acc[T];
#pragma unroll
    for (int t = 0; t < T; t++) acc[t] = 0ull;
    float2 ph0 = make_float2(1.0f, 0.0f);  // phasor of this lane's run start, carried from tile to tile

    // ---- code-pair window of the warp's 32 runs of a tile: raw table indices [idx_lo, idx_lo + span).  The window of tile k + 1 is
    // fetched (global -> registers) while tile k is computed and stored to shared memory at the top of the next iteration, so the
    // latency of those loads is never exposed; the table position modulo L advances incrementally (one integer division per command).
    int win_lo = 0, win_span = 0, win_k = 0;  // of the window being fetched: first raw index, entries, (first raw index) mod L
    float pre_c = 0.0f, pre_c1 = 0.0f;        // entry `lane` of that window: c[j], c[j + 1]
    auto fetch_window = [&](int tile, bool first) {
        const int n0w = lead + tile * TILE + w * WROW;
        if (!(n0w < n && n0w + WROW > 0))
            {
                win_span = 0;
                return;
            }
        const int lo = trk_raw_index(step, (float)max(n0w, -R), shift_min, rem);
        if (first || win_span == 0)
            win_k = math_mod(lo, L);
        else
            {
                win_k += lo - win_lo;  // the index law is monotone in n: the window only moves forward
                while (win_k >= L) win_k -= L;
            }
        win_lo = lo;
        win_span = span_c;
        if (lane < win_span)
            {
                int k = win_k + lane;
                while (k >= L) k -= L;
                const int k1 = (k + 1 == L) ? 0 : k + 1;
                pre_c = __ldg(code + k);
                pre_c1 = __ldg(code + k1);
            }
    };
    fetch_window(0, true);

    for (int tile = 0; tile < n_tiles; tile++)
        {
            const int n0w = lead + tile * TILE + w * WROW;  // sample index of the warp's first staged sample
            const int nr = n0w + lane * R;                  // ... and of this lane's run
            // ---- install the window fetched during the previous tile
            const unsigned tab_adj = (tab_a + (hide - (unsigned)win_lo) * 8u) ^ zero;  // address of entry j = tab_adj + 8 * bits(RD(x + MAGIC))
            {
                if (lane < win_span) st_shared_f2(tab_a + 8u * lane, pre_c - pre_c1, pre_c1);
                if (win_span > 32)  // wide tap spans only: the rest of the window, fetched here
                    {
                        int k = win_k + 32 + lane;
                        for (int j = 32 + lane; j < win_span; j += 32)
                            {
                                while (k >= L) k -= L;
                                const int k1 = (k + 1 == L) ? 0 : k + 1;
                                const float ck = __ldg(code + k), ck1 = __ldg(code + k1);
                                st_shared_f2(tab_a + 8u * j, ck - ck1, ck1);
                                k += 32;
                            }
                    }
            }
            if (tile + 1 < n_tiles) fetch_window(tile + 1, false);
            // phasor of the run start: exact seed, or one complex multiplication from the previous tile's
            if (tile % TRK2_RESEED == 0)
                ph0 = trk_phasor(u0 + du * (unsigned long long)(long long)nr);
            else
                ph0 = make_float2(ph0.x * inc_tile.x - ph0.y * inc_tile.y, ph0.x * inc_tile.y + ph0.y * inc_tile.x);

            mbar_wait_a(bar_a, tile & 1);
            uint4 v[CH];
#pragma unroll
            for (int ch = 0; ch < CH; ch++) v[ch] = ld_shared_v4(row_a + 16u * ch);
            __syncwarp();  // the warp's reads of its rows are ordered before lane 0's next copy into them; the window is installed
            if (lane == 0 && tile + 1 < n_tiles)
                {
                    fence_proxy_async();
                    mbar_expect_tx_a(bar_a, WROW_BYTES);
                    bulk_g2s_a(rows_a, src + (size_t)(tile + 1) * TILE * EB, WROW_BYTES, bar_a);
                }
            if (nr + R > 0 && nr < n)
                {
                    // ---- carrier wipe-off + prefix sums of the run (P[k] = sum of the first k wiped samples), packed (re, im)
                    f32x2 P = 0ull;
                    float phx = ph0.x, phy = ph0.y;
                    auto wipe = [&](auto full_tag) {
                        constexpr bool FULL = decltype(full_tag)::value;
                        const f32x2 incx = pk2(inc.x, inc.x), incy = pk2(inc.y, inc.y);
#pragma unroll
                        for (int ch = 0; ch < CH; ch++)
                            {
                                float2 x[SPC];
                                trk2_unpack<IQ>(v[ch], x);
#pragma unroll
                                for (int j = 0; j < SPC; j++)
                                    {
                                        const int i = ch * SPC + j;
                                        if (!FULL && (nr + i < 0 || nr + i >= n)) x[j] = make_float2(0.0f, 0.0f);
                                        const f32x2 A = pk2(phx, phy), B = pk2(-phy, phx);  // ph and j * ph
                                        P = fma2(pk2(x[j].x, x[j].x), A, P);              // P += x.re * ph + x.im * (j ph): the product is never formed on its own
                                        P = fma2(pk2(x[j].y, x[j].y), B, P);
                                        up2(fma2(incy, B, mul2(incx, A)), phx, phy);      // ph *= inc
                                        st_shared_b64(pref_addr + (unsigned)((i + 1) * 32 * sizeof(float2)), P);  // in order: P is not kept in registers
                                    }
                            }
                    };
                    if (nr >= 0 && nr + R <= n)
                        wipe(std::true_type{});
                    else
                        wipe(std::false_type{});
                    st_shared_b64(pref_addr + (unsigned)((R + 1) * 32 * sizeof(float2)), P);  // k = R + 1 reads the same total: no clamp in the tap loop

                    // ---- taps, two at a time: at most one table-entry boundary per run and tap
                    const float nf = (float)nr;
                    const float nfm = nf - MAGIC;  // exact
                    const float a0 = __fmul_rn(step, nf);
                    const f32x2 A0 = pk2(a0, a0), REM = pk2(rem, rem), INV = pk2(inv_step, inv_step), MG = pk2(MAGIC, MAGIC), NFM = pk2(nfm, nfm);
                    const float kq_max = MAGIC + (float)R;
                    const unsigned pref_adj = (pref_addr + 256u + hide * 256u) ^ zero;  // address of P[k] = pref_adj + 256 * (bits(MAGIC + k) - 1)
#pragma unroll
                    for (int p = 0; p < TP; p++)
                        {
                            const f32x2 SH = shp[p];
                            const f32x2 X0 = sub2(add2(A0, SH), REM);                 // the reference's law at the run start (rounded add, rounded sub)
                            const f32x2 M0 = add2_rd(X0, MG);                         // mantissa holds floor(x0)
                            const f32x2 DN = fma2(sub2(sub2(M0, MG), X0), INV, INV);  // samples until the next table entry
                            float kq0, kq1;
                            up2(add2(DN, MG), kq0, kq1);
                            kq0 = fminf(kq0, kq_max);                                 // MAGIC + min(rint(dn), R)
                            kq1 = fminf(kq1, kq_max);
                            float nq0, nq1;
                            up2(add2(pk2(kq0, kq1), NFM), nq0, nq1);                  // nf + kq, exact
                            // the law at the candidate: rounded mul, rounded add, rounded sub.  The product is formed by two SCALAR mul.rn.f32:
                            // ptxas (12.9) contracts mul.rn.f32x2 + add.rn.f32x2 into one FFMA2, which would change the reference's rounding
                            const f32x2 MQ = add2_rd(sub2(add2(pk2(__fmul_rn(step, nq0), __fmul_rn(step, nq1)), SH), REM), MG);
                            float m00, m01, mq0, mq1;
                            up2(M0, m00, m01);
                            up2(MQ, mq0, mq1);
#pragma unroll
                            for (int q = 0; q < 2; q++)
                                {
                                    const int t = 2 * p + q;
                                    if (t >= T) continue;
                                    const int m0i = __float_as_int(q ? m01 : m00), mqi = __float_as_int(q ? mq1 : mq0), kqi = __float_as_int(q ? kq1 : kq0);
                                    // samples of the run that still see entry floor(x0): kq when the candidate already sees the next entry
                                    // (mqi = m0i + 1), kq + 1 when it does not (mqi = m0i)  ->  k = kq + 1 - (mqi - m0i)
                                    const f32x2 Pk = ld_shared_b64(pref_adj + 256u * (unsigned)(kqi + m0i - mqi));
                                    const float2 cc = ld_shared_f2(tab_adj + 8u * (unsigned)m0i);
                                    acc[t] = fma2(pk2(cc.y, cc.y), P, fma2(pk2(cc.x, cc.x), Pk, acc[t]));  // c[j+1] * S + (c[j] - c[j+1]) * P[k]
                                }
                        }
                }
            __syncwarp();  // the code window is free again
        }

    // ---- one block reduce per command.  Warp level: a halving butterfly — at every step a lane keeps one half of its partial sums and
    // hands the other half to its partner, so 2T values take about 2T shuffles in total instead of 2T per step; after five steps lane l
    // holds the warp's sum of ONE output component (or a padding zero).
    constexpr int NV = 2 * T;
    float a[NV];
#pragma unroll
    for (int t = 0; t < T; t++) up2(acc[t], a[2 * t], a[2 * t + 1]);
    trk2_halve<NV, 16>(a, lane);
    {
        // which component the lane holds: walk the same splits (split points come from the PADDED counts the butterfly uses, validity
        // from the true counts: an upper half can be shorter than its lower half, or empty)
        int base = 0, pad = NV, cnt = NV;
        for (int o = 16; o > 0; o >>= 1)
            {
                const int half = (pad + 1) / 2;
                if (lane & o)
                    {
                        base += half;
                        cnt -= half;
                    }
                else
                    cnt = min(cnt, half);
                pad = half;
            }
        const bool ok = cnt > 0;
        if (ok) s_red[w][base] = a[0];
    }
    __syncthreads();
    if (tid < 2 * T)
        {
            float s = 0.0f;
#pragma unroll
            for (int k = 0; k < TRK2_WARPS; k++) s += s_red[k][tid];
            out[(size_t)blockIdx.x * (2 * TRK_OUT_STRIDE) + tid] = s;
        }
}
}  // namespace gb200
