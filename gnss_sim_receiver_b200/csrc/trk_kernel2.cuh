// Tracking multicorrelator, run-based kernel (K1b) for sm_100a — used when the code advances by less than one table entry
// over a run of R samples (R * code_step <= 0.9: e.g. GPS L1 at >= 10 MS/s, Galileo E1 at >= 20 MS/s).
//
// Same contract as trk_kernel.cuh (Cpu_Multicorrelator_Real_Codes::Carrier_wipeoff_multicorrelator_resampler,
// src/algorithms/tracking/libs/cpu_multicorrelator_real_codes.cc:103-126), different work decomposition:
//   * a CTA owns a GROUP of up to G commands whose windows overlap in the IQ ring and walks the ring tile by tile; each
//     tile (128 rows of R samples) is staged ONCE into shared memory by TMA bulk copies (cp.async.bulk + mbarrier, two stages)
//     and shared by the G commands and all their taps;
//   * a thread owns one run of R consecutive samples: it wipes the carrier off once, keeps the R prefix sums of the wiped
//     samples, and for every tap finds the single chip boundary inside the run, so a tap costs O(1) per run instead of O(R):
//       out_t += c[idx] * P[k_t] + c[idx+1] * (S - P[k_t])
//     where k_t = number of samples of the run that still see table entry idx.  k_t comes from the reference's exact fp32
//     index law (volk_gnsssdr_32f_xn_resampler_32f_xn.h:73): an estimate from the fractional distance to the next integer,
//     confirmed by ONE exact evaluation of the law at the candidate sample (the estimate is within 0.02 sample of the fp32
//     crossing under the host's applicability check, so the boundary is the candidate or its successor);
//   * accumulators stay in registers across all tiles of a command: one block reduce per command, no cross-CTA partials.
#pragma once
#include "trk_kernel.cuh"

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

// sample i of a staged row -> float2
template <int IQ>
__device__ __forceinline__ float2 trk2_row_sample(const unsigned char* row, int i)
{
    if (IQ == GB200_IQ_CF32) return reinterpret_cast<const float2*>(row)[i];
    if (IQ == GB200_IQ_CS16)
        {
            const short2 v = reinterpret_cast<const short2*>(row)[i];
            return make_float2((float)v.x, (float)v.y);
        }
    const char2 v = reinterpret_cast<const char2*>(row)[i];
    return make_float2((float)v.x, (float)v.y);
}

template <int T, int IQ, int R, int G>
__global__ void __launch_bounds__(TRK2_THREADS) trk_run_kernel(const TrkDevCmd* __restrict__ cmds, const TrkGroup* __restrict__ groups,
    const unsigned char* __restrict__ ring, float* __restrict__ out)
{
    constexpr int EB = IqRow<IQ>::EB;
    constexpr int ROW_BYTES = R * EB;
    constexpr int ROW_STRIDE = ROW_BYTES + 16;  // padding keeps the per-thread 128-bit row reads conflict-free
    constexpr int TILE = TRK2_THREADS * R;
    constexpr int STAGE_BYTES = TRK2_THREADS * ROW_STRIDE;
    constexpr float MAGIC = 12582912.0f;

    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* s_rows = smem;                                                    // 2 stages
    float2* s_prefix = reinterpret_cast<float2*>(smem + 2 * STAGE_BYTES);             // [R + 1][128]
    float2* s_tab = s_prefix + (R + 1) * TRK2_THREADS;                               // [G][TRK2_TAB] (c[j], c[j+1])
    __shared__ __align__(8) uint64_t s_bar[2];
    __shared__ float s_red[TRK2_THREADS / 32][2 * T];

    const TrkGroup& grp = groups[blockIdx.x];
    const int tid = threadIdx.x;
    const int n_tiles = grp.n_tiles;

    if (tid == 0)
        {
            mbar_init(&s_bar[0], 1);
            mbar_init(&s_bar[1], 1);
            fence_barrier_init();
        }
    s_prefix[tid] = make_float2(0.0f, 0.0f);  // P[0]
    __syncthreads();

    auto issue_tile = [&](int tile, int stage) {
        if (tid == 0) mbar_expect_tx(&s_bar[stage], TRK2_THREADS * ROW_BYTES);
        __syncwarp();
        // every thread copies its own row: global is contiguous, shared is padded
        const unsigned char* src = ring + ((size_t)grp.tile0 + (size_t)tile * TILE + (size_t)tid * R) * EB;
        bulk_g2s(s_rows + stage * STAGE_BYTES + tid * ROW_STRIDE, src, ROW_BYTES, &s_bar[stage]);
    };
    // thread 0's expect_tx must precede the copies of other warps: order with a barrier on the first issues only
    if (n_tiles > 0)
        {
            if (tid == 0) mbar_expect_tx(&s_bar[0], TRK2_THREADS * ROW_BYTES);
            if (n_tiles > 1 && tid == 0) mbar_expect_tx(&s_bar[1], TRK2_THREADS * ROW_BYTES);
            __syncthreads();
            const unsigned char* src = ring + ((size_t)grp.tile0 + (size_t)tid * R) * EB;
            bulk_g2s(s_rows + tid * ROW_STRIDE, src, ROW_BYTES, &s_bar[0]);
            if (n_tiles > 1) bulk_g2s(s_rows + STAGE_BYTES + tid * ROW_STRIDE, src + (size_t)TILE * EB, ROW_BYTES, &s_bar[1]);
        }

    float acc_re[G][T], acc_im[G][T];
#pragma unroll
    for (int g = 0; g < G; g++)
#pragma unroll
        for (int t = 0; t < T; t++) acc_re[g][t] = acc_im[g][t] = 0.0f;

    for (int tile = 0; tile < n_tiles; tile++)
        {
            const int stage = tile & 1;
            // ---- code-pair windows of this tile for every command of the group
            int bias[G];
#pragma unroll
            for (int g = 0; g < G; g++)
                {
                    bias[g] = 0;
                    if (g < grp.n_cmds)
                        {
                            const TrkDevCmd& c = cmds[grp.cmd[g]];
                            const long long n0 = grp.tile0 + (long long)tile * TILE - grp.ws[g];  // sample index of the tile's first sample
                            if (n0 < c.n && n0 + TILE > 0)
                                {
                                    const float nlo = (float)max(n0, (long long)-R);
                                    const int idx_lo = trk_raw_index(c.code_step, nlo, c.shift_min, c.rem_code);
                                    bias[g] = 0x4B400000 + idx_lo;
                                    int k = math_mod(idx_lo + tid, c.L);
                                    const int adv = TRK2_THREADS % c.L;
                                    const float* code = reinterpret_cast<const float*>(c.code);
                                    for (int j = tid; j < TRK2_TAB; j += TRK2_THREADS)
                                        {
                                            const int k1 = (k + 1 == c.L) ? 0 : k + 1;
                                            s_tab[g * TRK2_TAB + j] = make_float2(__ldg(code + k), __ldg(code + k1));
                                            k += adv;
                                            if (k >= c.L) k -= c.L;
                                        }
                                }
                        }
                }
            mbar_wait(&s_bar[stage], (tile >> 1) & 1);
            __syncthreads();
            const unsigned char* row = s_rows + stage * STAGE_BYTES + tid * ROW_STRIDE;
#pragma unroll
            for (int g = 0; g < G; g++)
                {
                    if (g >= grp.n_cmds) continue;
                    const TrkDevCmd& c = cmds[grp.cmd[g]];
                    const long long n0 = grp.tile0 + (long long)tile * TILE - grp.ws[g];
                    const long long nr_ll = n0 + (long long)tid * R;
                    if (nr_ll + R <= 0 || nr_ll >= c.n) continue;  // run outside the window
                    const int nr = (int)nr_ll;
                    const int n = c.n;
                    const float step = c.code_step, rem = c.rem_code;
                    // ---- carrier wipe-off + prefix sums of the run
                    float2 ph = trk_phasor(c.u0 + c.du * (unsigned long long)(long long)nr);
                    const float2 inc = trk_phasor(c.du);
                    float2 P = make_float2(0.0f, 0.0f);
                    const bool full = nr >= 0 && nr + R <= n;
#pragma unroll
                    for (int i = 0; i < R; i++)
                        {
                            float2 x = trk2_row_sample<IQ>(row, i);
                            if (!full && (nr + i < 0 || nr + i >= n)) x = make_float2(0.0f, 0.0f);
                            P.x += x.x * ph.x - x.y * ph.y;
                            P.y += x.x * ph.y + x.y * ph.x;
                            const float pr = ph.x * inc.x - ph.y * inc.y;
                            ph.y = ph.x * inc.y + ph.y * inc.x;
                            ph.x = pr;
                            s_prefix[(i + 1) * TRK2_THREADS + tid] = P;
                        }
                    // ---- taps: one chip boundary per run at most
                    const float nf = (float)nr;
                    const float a0 = __fmul_rn(step, nf);
                    const float inv_step = __frcp_rn(step);
                    const float2* tab = s_tab + g * TRK2_TAB;
#pragma unroll
                    for (int t = 0; t < T; t++)
                        {
                            const float sh = c.shifts[t];
                            const float x0 = __fsub_rn(__fadd_rn(a0, sh), rem);
                            const float m0 = __fadd_rd(x0, MAGIC);  // mantissa holds floor(x0)
                            const float fl = __fsub_rn(m0, MAGIC);
                            const float frac = __fsub_rn(__fadd_rn(fl, 1.0f), x0);  // distance to the next table entry, (0, 1]
                            const float kq = fminf(__fsub_rn(__fadd_rn(frac * inv_step, MAGIC), MAGIC), (float)R);  // min(rint(estimate), R)
                            const float xq = __fsub_rn(__fadd_rn(__fmul_rn(step, nf + kq), sh), rem);  // the exact law at the candidate
                            const float mq = __fadd_rd(xq, MAGIC);
                            int k = (int)kq + (__float_as_int(mq) > __float_as_int(m0) ? 0 : 1);
                            k = min(k, R);
                            const float2 cc = tab[__float_as_int(m0) - bias[g]];
                            const float2 Pk = s_prefix[k * TRK2_THREADS + tid];
                            const float d = cc.x - cc.y;
                            acc_re[g][t] += cc.y * P.x + d * Pk.x;
                            acc_im[g][t] += cc.y * P.y + d * Pk.y;
                        }
                }
            __syncthreads();  // rows of this stage and the code windows are free again
            if (tile + 2 < n_tiles) issue_tile(tile + 2, stage);
        }

    // ---- one block reduce per command
#pragma unroll
    for (int g = 0; g < G; g++)
        {
            if (g >= grp.n_cmds) break;
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
                    out[(size_t)grp.cmd[g] * (2 * TRK_OUT_STRIDE) + tid] = v;
                }
            __syncthreads();
        }
}
}  // namespace gb200
