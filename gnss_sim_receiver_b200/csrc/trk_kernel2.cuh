// Tracking multicorrelator, run-based kernel (K1b) for sm_100a — used when the code advances by less than one table entry
// over a run of R samples (R * code_step <= 0.9: e.g. GPS L1 at >= 12 MS/s, Galileo E1 at >= 25 MS/s).
//
// Same contract as trk_kernel.cuh (Cpu_Multicorrelator_Real_Codes::Carrier_wipeoff_multicorrelator_resampler,
// src/algorithms/tracking/libs/cpu_multicorrelator_real_codes.cc:103-126), different work decomposition:
//   * a CTA owns a GROUP of up to G commands whose windows overlap in the IQ ring and walks the ring tile by tile; each
//     tile (128 runs of R samples, contiguous) is staged ONCE into shared memory by one TMA bulk copy (cp.async.bulk +
//     mbarrier) and shared by the G commands and all their taps.  R is chosen so that a run is an ODD number of 16-byte
//     chunks (cf32: 18 or 10 samples, int16: 20 or 12, int8: 24 or 8): consecutive threads' runs then start in different
//     bank groups and the per-thread 128-bit reads of the plain contiguous tile are conflict-free without padding;
//   * a thread owns one run: it wipes the carrier off once, keeps the R prefix sums of the wiped samples, and for every
//     tap finds the single chip boundary inside the run, so a tap costs O(1) per run instead of O(R):
//       out_t += c[idx] * P[k_t] + c[idx+1] * (S - P[k_t])
//     where k_t = number of samples of the run that still see table entry idx.  k_t comes from the reference's exact fp32
//     index law (volk_gnsssdr_32f_xn_resampler_32f_xn.h:73): an estimate from the fractional distance to the next integer,
//     confirmed by ONE exact evaluation of the law at the candidate sample (the estimate is within 0.02 sample of the fp32
//     crossing under the host's applicability check, so the boundary is the candidate or its successor);
//   * accumulators stay in registers across all tiles of a command: one block reduce per command, no cross-CTA partials.
#pragma once
#include "trk_kernel.cuh"
#include <type_traits>

namespace gb200
{
constexpr int TRK2_THREADS = 128;
constexpr int TRK2_GMAX = 4;
constexpr int TRK2_TAB = 256;  // code-pair window entries per command per tile

struct TrkGroup
{
    long long tile0;  // ring element index of the first tile (multiple of 8 samples)
    int n_tiles;
    int n_cmds;
    int cmd[TRK2_GMAX];
    long long ws[TRK2_GMAX];  // ring element index of each command's first sample
};

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

template <int T>
struct Trk2Cmd  // per-command constants, staged once per CTA
{
    unsigned long long u0, du;
    long long ws;
    const float* code;
    float2 inc;
    float step, rem, inv_step, shift_min;
    int n, L, out_index, pad_;
    float sh[T];
};

// T taps, IQ sample type, CH = 16-byte chunks per run (odd), G = commands per group
template <int T, int IQ, int CH, int G>
__global__ void __launch_bounds__(TRK2_THREADS) trk_run_kernel(const TrkDevCmd* __restrict__ cmds, const TrkGroup* __restrict__ groups,
    const unsigned char* __restrict__ ring, float* __restrict__ out)
{
    constexpr int EB = IqRow<IQ>::EB;
    constexpr int SPC = 16 / EB;  // samples per chunk
    constexpr int R = CH * SPC;   // samples per run
    constexpr int TILE = TRK2_THREADS * R;
    constexpr int TILE_BYTES = TILE * EB;
    constexpr float MAGIC = 12582912.0f;
    constexpr int MAGICBITS = 0x4B400000;

    extern __shared__ __align__(128) unsigned char smem[];
    const uint4* s_rows = reinterpret_cast<const uint4*>(smem);
    float2* s_prefix = reinterpret_cast<float2*>(smem + TILE_BYTES);  // [R + 2][128]
    float2* s_tab = s_prefix + (R + 2) * TRK2_THREADS;               // [G][TRK2_TAB] (c[j] - c[j+1], c[j+1])
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ Trk2Cmd<T> s_cmd[G];
    __shared__ float s_red[TRK2_THREADS / 32][2 * T];

    // groups == nullptr: one command per CTA (the default plan) — the group is derived here instead of being built on the
    // host and copied (64 bytes per command)
    __shared__ TrkGroup s_grp;
    const int tid = threadIdx.x;
    if (tid == 0)
        {
            if (groups)
                s_grp = groups[blockIdx.x];
            else
                {
                    const TrkDevCmd& c = cmds[blockIdx.x];
                    const long long ws = (long long)((static_cast<const unsigned char*>(c.iq) - ring) / EB);
                    s_grp.tile0 = ws & ~7LL;
                    s_grp.n_tiles = (int)((ws + c.n - s_grp.tile0 + TILE - 1) / TILE);
                    s_grp.n_cmds = 1;
                    s_grp.cmd[0] = (int)blockIdx.x;
                    s_grp.ws[0] = ws;
                }
        }
    __syncthreads();
    const TrkGroup& grp = s_grp;
    const int n_tiles = grp.n_tiles;
    const int n_g = grp.n_cmds;

    if (tid == 0)
        {
            mbar_init(&s_bar, 1);
            fence_barrier_init();
        }
    if (tid < n_g)
        {
            const TrkDevCmd& c = cmds[grp.cmd[tid]];
            Trk2Cmd<T>& s = s_cmd[tid];
            s.u0 = c.u0;
            s.du = c.du;
            s.ws = grp.ws[tid];
            s.code = reinterpret_cast<const float*>(c.code);
            s.inc = trk_phasor(c.du);
            s.step = c.code_step;
            s.rem = c.rem_code;
            s.inv_step = __frcp_rn(c.code_step);
            s.shift_min = c.shift_min;
            s.n = c.n;
            s.L = c.L;
            s.out_index = grp.cmd[tid];
#pragma unroll
            for (int t = 0; t < T; t++) s.sh[t] = c.shifts[t];
        }
    s_prefix[tid] = make_float2(0.0f, 0.0f);  // P[0]
    __syncthreads();

    float acc_re[G][T], acc_im[G][T];
#pragma unroll
    for (int g = 0; g < G; g++)
#pragma unroll
        for (int t = 0; t < T; t++) acc_re[g][t] = acc_im[g][t] = 0.0f;

    for (int tile = 0; tile < n_tiles; tile++)
        {
            const long long tile_start = grp.tile0 + (long long)tile * TILE;
            if (tid == 0)
                {
                    mbar_expect_tx(&s_bar, TILE_BYTES);
                    bulk_g2s(smem, ring + (size_t)tile_start * EB, TILE_BYTES, &s_bar);
                }
            // ---- code-pair windows of this tile for every command of the group (overlaps the copy)
            int bias[G];
#pragma unroll
            for (int g = 0; g < G; g++)
                {
                    bias[g] = 0;
                    if (g < n_g)
                        {
                            const Trk2Cmd<T>& c = s_cmd[g];
                            const long long n0 = tile_start - c.ws;  // sample index of the tile's first sample
                            if (n0 < c.n && n0 + TILE > 0)
                                {
                                    const float nlo = (float)max(n0, (long long)-R);
                                    const int idx_lo = trk_raw_index(c.step, nlo, c.shift_min, c.rem);
                                    bias[g] = MAGICBITS + idx_lo;
                                    const int L = c.L;
                                    int k = math_mod(idx_lo + tid, L);
                                    const int adv = TRK2_THREADS % L;
                                    for (int j = tid; j < TRK2_TAB; j += TRK2_THREADS)
                                        {
                                            const int k1 = (k + 1 == L) ? 0 : k + 1;
                                            const float ck = __ldg(c.code + k), ck1 = __ldg(c.code + k1);
                                            s_tab[g * TRK2_TAB + j] = make_float2(ck - ck1, ck1);  // (c[j] - c[j+1], c[j+1]): the tap loop needs exactly these
                                            k += adv;
                                            if (k >= L) k -= L;
                                        }
                                }
                        }
                }
            mbar_wait(&s_bar, tile & 1);
            __syncthreads();
            const uint4* row = s_rows + tid * CH;
#pragma unroll
            for (int g = 0; g < G; g++)
                {
                    if (g >= n_g) continue;
                    const Trk2Cmd<T>& c = s_cmd[g];
                    const long long nr_ll = tile_start - c.ws + (long long)tid * R;
                    const int n = c.n;
                    if (nr_ll + R <= 0 || nr_ll >= n) continue;  // run outside the window
                    const int nr = (int)nr_ll;
                    // ---- carrier wipe-off + prefix sums of the run
                    float2 ph = trk_phasor(c.u0 + c.du * (unsigned long long)(long long)nr);
                    const float2 inc = c.inc;
                    float2 P = make_float2(0.0f, 0.0f);
                    auto wipe = [&](auto full_tag) {
                        constexpr bool FULL = decltype(full_tag)::value;
#pragma unroll
                        for (int ch = 0; ch < CH; ch++)
                            {
                                float2 x[SPC];
                                trk2_unpack<IQ>(row[ch], x);
#pragma unroll
                                for (int j = 0; j < SPC; j++)
                                    {
                                        const int i = ch * SPC + j;
                                        if (!FULL && (nr + i < 0 || nr + i >= n)) x[j] = make_float2(0.0f, 0.0f);
                                        P.x = fmaf(x[j].x, ph.x, fmaf(-x[j].y, ph.y, P.x));  // 4 FFMA: the product is never formed on its own
                                        P.y = fmaf(x[j].x, ph.y, fmaf(x[j].y, ph.x, P.y));
                                        const float pr = ph.x * inc.x - ph.y * inc.y;
                                        ph.y = ph.x * inc.y + ph.y * inc.x;
                                        ph.x = pr;
                                        s_prefix[(i + 1) * TRK2_THREADS + tid] = P;
                                    }
                            }
                    };
                    if (nr >= 0 && nr + R <= n)
                        wipe(std::true_type{});
                    else
                        wipe(std::false_type{});
                    s_prefix[(R + 1) * TRK2_THREADS + tid] = P;  // k = R + 1 reads the same total: no clamp in the tap loop
                    // ---- taps: at most one table-entry boundary per run
                    const float nf = (float)nr;
                    const float nfm = nf - MAGIC;  // exact
                    const float step = c.step, rem = c.rem, inv_step = c.inv_step;
                    const float a0 = __fmul_rn(step, nf);
                    const float2* tab = s_tab + g * TRK2_TAB;
                    const int tab_bias = bias[g];
#pragma unroll
                    for (int t = 0; t < T; t++)
                        {
                            const float sh = c.sh[t];
                            const float x0 = __fsub_rn(__fadd_rn(a0, sh), rem);                 // the reference's law at the run start
                            const float m0 = __fadd_rd(x0, MAGIC);                             // mantissa holds floor(x0)
                            const float fl = __fsub_rn(m0, MAGIC);
                            const float dn = fmaf(__fsub_rn(fl, x0), inv_step, inv_step);       // samples until the next table entry
                            const float kqm = fminf(__fadd_rn(dn, MAGIC), MAGIC + (float)R);    // MAGIC + min(rint(dn), R)
                            const float nfq = __fadd_rn(kqm, nfm);                             // nf + kq, exact
                            const float xq = __fsub_rn(__fadd_rn(__fmul_rn(step, nfq), sh), rem);  // the law at the candidate
                            const float mq = __fadd_rd(xq, MAGIC);
                            const int kb = __float_as_int(kqm) + (__float_as_int(mq) > __float_as_int(m0) ? 0 : 1);
                            const float2 cc = tab[__float_as_int(m0) - tab_bias];
                            const float2 Pk = s_prefix[(kb - MAGICBITS) * TRK2_THREADS + tid];
                            acc_re[g][t] = fmaf(cc.y, P.x, fmaf(cc.x, Pk.x, acc_re[g][t]));  // c[j+1] * S + (c[j] - c[j+1]) * P[k]
                            acc_im[g][t] = fmaf(cc.y, P.y, fmaf(cc.x, Pk.y, acc_im[g][t]));
                        }
                }
            __syncthreads();  // the tile and the code windows are free again
        }

    // ---- one block reduce per command
#pragma unroll
    for (int g = 0; g < G; g++)
        {
            if (g >= n_g) break;
#pragma unroll
            for (int t = 0; t < T; t++)
                {
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1)
                        {
                            acc_re[g][t] += __shfl_xor_sync(0xffffffffu, acc_re[g][t], o);
                            acc_im[g][t] += __shfl_xor_sync(0xffffffffu, acc_im[g][t], o);
                        }
                }
            if ((tid & 31) == 0)
                {
#pragma unroll
                    for (int t = 0; t < T; t++)
                        {
                            s_red[tid >> 5][2 * t] = acc_re[g][t];
                            s_red[tid >> 5][2 * t + 1] = acc_im[g][t];
                        }
                }
            __syncthreads();
            if (tid < 2 * T)
                {
                    float v = 0.0f;
#pragma unroll
                    for (int w = 0; w < TRK2_THREADS / 32; w++) v += s_red[w][tid];
                    out[(size_t)s_cmd[g].out_index * (2 * TRK_OUT_STRIDE) + tid] = v;
                }
            __syncthreads();
        }
}
}  // namespace gb200
