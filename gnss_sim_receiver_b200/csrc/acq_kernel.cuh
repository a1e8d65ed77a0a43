// PCPS acquisition kernels (K2) for sm_100a — see acq_fft.cuh for the FFT scheme.
//
// Reference path: pcps_acquisition::acquisition_core, src/algorithms/acquisition/gnuradio_blocks/pcps_acquisition.cc:600-697
//   forward  : volk_32fc_x2_multiply_32fc(in, wipeoff[bin]) :643 + d_fft_if->execute() :647   -> acq_fwd_kernel
//   inverse  : multiply by d_fft_codes :650 + d_ifft->execute() :653 + magnitude_squared / add :657-665
//              + per-bin volk_gnsssdr_32f_index_max_32u :505-514                                  -> acq_inv_kernel
//   statistic: max_to_input_power_statistic :496-528 / first_vs_second_peak_statistic :531-597   -> acq_stats_kernel
// The forward transform of (input x wipe-off[bin]) does not depend on the PRN, so it is computed once per Doppler bin
// and shared by every PRN (the reference recomputes it per channel).
#pragma once
#include "acq_fft.cuh"

namespace gb200
{
struct AcqCell  // reduction of one (slot, bin, part) over code phase
{
    float vmax;
    int imax;
    float sum;
    int pad_;
};

struct AcqParams
{
    FftPlan plan;
    int consumed;        // input samples, zero-padded to plan.n
    int n_bins;
    int neff;            // code phases searched
    int out_offset;      // first output sample searched (fft_size/2 with bit_transition, else 0)
    int n_parts;         // cells per (slot, bin): 1, or chunks of the global inverse stage
    int dwell_index;
    int keep_grid;
    int doppler_base;    // -doppler_max + doppler_center
    int doppler_step;
    int use_cfar;
    int samples_per_chip;
    int num_dwells;      // dwells accumulated so far, including this one
    float samples_per_code;
};

__device__ __forceinline__ float2 acq_phasor(unsigned long long u)
{
    const int a = (int)(u >> 32);
    float s, c;
    sincospif((float)a * 4.656612873077393e-10f, &s, &c);
    return make_float2(c, s);
}

template <int IQ>
__device__ __forceinline__ float2 acq_load_iq(const void* base, int s)
{
    if (IQ == GB200_IQ_CF32) return __ldg(reinterpret_cast<const float2*>(base) + s);
    if (IQ == GB200_IQ_CS16)
        {
            const short2 v = __ldg(reinterpret_cast<const short2*>(base) + s);
            return make_float2((float)v.x, (float)v.y);  // volk_gnsssdr_16ic_convert_32fc, pcps_acquisition.cc:610
        }
    const char2 v = __ldg(reinterpret_cast<const char2*>(base) + s);
    return make_float2((float)v.x, (float)v.y);
}

// sample n of (zero-padded input) x wipe-off of one Doppler bin.
//   phase_row != NULL (default, the reference's law): the wipe-off is (cosf, sinf) of the fp32 phase the reference
//     accumulates by repeated addition (update_local_carrier, pcps_acquisition.cc:232-245 ->
//     volk_gnsssdr_s32f_sincos_32fc_generic, kernels/volk_gnsssdr/volk_gnsssdr_s32f_sincos_32fc.h:390-401); the
//     accumulated phases are tabulated once per Doppler grid by acq_phase_table_kernel;
//   phase_row == NULL (exact mode): exp(-j*2*pi*f_bin*n/fs) from dphi = frac(-f_bin/fs) in 64-bit fixed-point turns.
template <int IQ>
__device__ __forceinline__ float2 acq_wiped_sample(const void* iq, int n, int consumed, unsigned long long dphi, const float* __restrict__ phase_row)
{
    if (n >= consumed) return make_float2(0.0f, 0.0f);
    const float2 x = acq_load_iq<IQ>(iq, n);
    if (phase_row)
        {
            float s, c;
            sincosf(__ldg(phase_row + n), &s, &c);
            return cmul(x, make_float2(c, s));
        }
    if (dphi == 0ull) return x;
    return cmul(x, acq_phasor(dphi * (unsigned long long)n));
}

// fp32 phase accumulation of the reference: tab[bin][n] = phase before sample n, accumulated by repeated fp32 addition (the order of
// volk_gnsssdr_s32f_sincos_32fc_generic, so the values are the reference's bit for bit).  The recurrence is serial per bin (25 000
// dependent additions: ~50 us of FADD latency whatever the layout), so the kernel is laid out for the store side: one warp per CTA owns
// ACQ_PT_BINS bins (80 bins -> 10 SMs instead of one), lanes 0..ACQ_PT_BINS-1 run the recurrences over a chunk of ACQ_PT_CHUNK samples
// into shared memory, then all 32 lanes write the chunk's rows as coalesced 128-byte stores.
constexpr int ACQ_PT_BINS = 8;
constexpr int ACQ_PT_CHUNK = 128;
__global__ void __launch_bounds__(32) acq_phase_table_kernel(const float* __restrict__ inc, int n_bins, int n, float* __restrict__ tab)
{
    __shared__ float tile[ACQ_PT_BINS][ACQ_PT_CHUNK + 1];
    const int lane = threadIdx.x;
    const int bin0 = blockIdx.x * ACQ_PT_BINS;
    const int rows = min(ACQ_PT_BINS, n_bins - bin0);
    const float d = lane < rows ? inc[bin0 + lane] : 0.0f;
    float ph = 0.0f;
    for (int i0 = 0; i0 < n; i0 += ACQ_PT_CHUNK)
        {
            if (lane < rows)
                {
#pragma unroll 16
                    for (int k = 0; k < ACQ_PT_CHUNK; k++)
                        {
                            tile[lane][k] = ph;
                            ph = __fadd_rn(ph, d);
                        }
                }
            __syncwarp();
            for (int r = 0; r < rows; r++)
                {
                    float* row = tab + (size_t)(bin0 + r) * n + i0;
#pragma unroll
                    for (int j = 0; j < ACQ_PT_CHUNK; j += 32)
                        if (i0 + j + lane < n) row[j + lane] = tile[r][j + lane];
                }
            __syncwarp();
        }
}

// Forward transform of one shared-memory block.  grid = (g_radix, batch).
//   SRC_IQ = true : block 0 only (g_radix == 1): input = raw IQ x wipe-off of bin blockIdx.y
//   SRC_IQ = false: input = cf32 buffer src[batch][N] (output of the global stage, or a zero-padded code replica)
// Output dst[batch][N] in digit-reversed order.  code_out: conjugate (code spectra, pcps_acquisition.cc:207) and store each
// block as [q][butterfly] of the first inverse stage (position id * RL + q -> q * (m / RL) + id), the order in which
// acq_inv_kernel's first stage reads it with coalesced loads.
template <int IQ, bool SRC_IQ>
__global__ void __launch_bounds__(ACQ_MAX_THREADS) acq_fwd_kernel(FftPlan pl, const void* __restrict__ src, int consumed, const unsigned long long* __restrict__ dphi, const float* __restrict__ phase_tab,
    const float2* __restrict__ gT, float2* __restrict__ dst, int code_out)
{
    extern __shared__ __align__(16) float2 sm[];
    float2* x = sm;
    float2* sT = sm + pl.m + pl.y_len;
    fft_load_tables(pl, gT, sT);
    const size_t base = (size_t)blockIdx.y * pl.n + (size_t)blockIdx.x * pl.m;
    if (SRC_IQ)
        {
            const unsigned long long d = dphi[blockIdx.y];
            const float* prow = phase_tab ? phase_tab + (size_t)blockIdx.y * consumed : nullptr;
            // (one phase-table word and one IQ item per sample, both from L2 / HBM: unrolled so that a thread has 8 of each in flight; worth 5 us of the
            // 47 us this launch of 80 CTAs takes — the rest is the latency of one transform per SM with nothing beside it)
#pragma unroll 8
            for (int i = threadIdx.x; i < pl.m; i += blockDim.x) x[i] = acq_wiped_sample<IQ>(src, i, consumed, d, prow);
        }
    else
        {
            const float2* s = reinterpret_cast<const float2*>(src) + base;
            for (int i = threadIdx.x; i < pl.m; i += blockDim.x) x[i] = s[i];
        }
    __syncthreads();
    fft_smem_forward(pl, x, sT);
    float2* d = dst + base;
    if (code_out)
        {
            const int rl = pl.perm_rl, nb = pl.m / rl;
            for (int i = threadIdx.x; i < pl.m; i += blockDim.x)
                {
                    const int id = i / rl, q = i - id * rl;
                    const float2 v = x[i];
                    d[q * nb + id] = make_float2(v.x, -v.y);
                }
            return;
        }
    for (int i = threadIdx.x; i < pl.m; i += blockDim.x) d[i] = x[i];
}

// First DIF stage through global memory (transforms longer than one SM's shared memory): radix R = plan.g_radix over
// stride m.  grid = (ceil(m / blockDim), batch).  dst[batch][k1*m + j2] = (sum_j1 in[j1*m + j2] W_R^{j1 k1}) * W_N^{j2 k1}
template <int IQ, bool SRC_IQ>
__global__ void __launch_bounds__(256) acq_gstage_fwd_kernel(FftPlan pl, const void* __restrict__ src, int consumed, const unsigned long long* __restrict__ dphi, const float* __restrict__ phase_tab,
    const float2* __restrict__ gA, const float2* __restrict__ gB, float2* __restrict__ dst)
{
    const int j2 = blockIdx.x * blockDim.x + threadIdx.x;
    if (j2 >= pl.m) return;
    const int R = pl.g_radix, unit = pl.n / R;
    float2 v[8];
    for (int j1 = 0; j1 < R; j1++)
        {
            const int n = j1 * pl.m + j2;
            if (SRC_IQ)
                v[j1] = acq_wiped_sample<IQ>(src, n, consumed, dphi[blockIdx.y], phase_tab ? phase_tab + (size_t)blockIdx.y * consumed : nullptr);
            else
                v[j1] = reinterpret_cast<const float2*>(src)[(size_t)blockIdx.y * pl.n + n];
        }
    float2* d = dst + (size_t)blockIdx.y * pl.n;
    for (int k1 = 0; k1 < R; k1++)
        {
            float2 acc = v[0];
            for (int j1 = 1; j1 < R; j1++) acc = cadd(acc, cmul(v[j1], fft_twiddle(gA, gB, ((j1 * k1) % R) * unit)));
            const long long e = (long long)j2 * k1;  // < N
            d[k1 * pl.m + j2] = k1 ? cmul(acc, fft_twiddle(gA, gB, (int)e)) : acc;
        }
}

// Folded, wiped input of every Doppler bin — pcps_quicksync_acquisition_cc.cc:304-318: the block times the bin's wipe-off
// (volk_32fc_x2_multiply_32fc: four products, one difference, one sum, each rounded), then chunk i of M samples added to the transform
// input for i = 0 ... K-1 in that order.  grid = (ceil(M / 256), bins); out[bin][out_stride]; phase: [bins][len] accumulated fp32 phases.
__global__ void acq_fold_kernel(const float2* __restrict__ iq, const float* __restrict__ phase, int len, int M, int K, float2* __restrict__ out, int out_stride)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= M) return;
    const float* prow = phase + (size_t)blockIdx.y * len;
    float2 acc = make_float2(0.0f, 0.0f);
    for (int i = 0; i < K; i++)
        {
            const int n = i * M + k;
            const float2 x = iq[n];
            float s, c;
            sincosf(__ldg(prow + n), &s, &c);
            acc.x = __fadd_rn(acc.x, __fsub_rn(__fmul_rn(x.x, c), __fmul_rn(x.y, s)));
            acc.y = __fadd_rn(acc.y, __fadd_rn(__fmul_rn(x.x, s), __fmul_rn(x.y, c)));
        }
    out[(size_t)blockIdx.y * out_stride + k] = acc;  // (a padded transform's tail behind M stays zero)
}

// Candidate correlations that resolve the folding ambiguity (:337-349): acc[i] = sum_{j < n_code} iq[d_i + j] * w_bin[d_i + j] * code[j],
// d_i = index + i * M.  grid = (n_candidates); the sum is a tree, not the reference's left-to-right loop (only the largest |acc|^2 is used).
__global__ void acq_fold_candidates_kernel(const float2* __restrict__ iq, const float* __restrict__ prow, int M, int index, const float2* __restrict__ code, int n_code,
    float2* __restrict__ acc_out)
{
    __shared__ float2 s_acc[32];
    const int d = index + (int)blockIdx.x * M;
    float2 acc = make_float2(0.0f, 0.0f);
    for (int j = threadIdx.x; j < n_code; j += blockDim.x)
        {
            const float2 x = iq[d + j];
            float s, c;
            sincosf(__ldg(prow + d + j), &s, &c);
            const float2 t = make_float2(__fsub_rn(__fmul_rn(x.x, c), __fmul_rn(x.y, s)), __fadd_rn(__fmul_rn(x.x, s), __fmul_rn(x.y, c)));
            const float2 cd = code[j];
            acc.x += __fsub_rn(__fmul_rn(t.x, cd.x), __fmul_rn(t.y, cd.y));
            acc.y += __fadd_rn(__fmul_rn(t.x, cd.y), __fmul_rn(t.y, cd.x));
        }
    for (int o = 16; o > 0; o >>= 1)
        {
            acc.x += __shfl_down_sync(0xffffffffu, acc.x, o);
            acc.y += __shfl_down_sync(0xffffffffu, acc.y, o);
        }
    if ((threadIdx.x & 31) == 0) s_acc[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0)
        {
            float2 t = make_float2(0.0f, 0.0f);
            for (int w = 0; w < (int)(blockDim.x + 31) / 32; w++) t.x += s_acc[w].x, t.y += s_acc[w].y;
            acc_out[blockIdx.x] = t;
        }
}

// block-wide reduction of (max value, first index of the maximum, sum); result valid on thread 0
__device__ __forceinline__ void acq_block_reduce(float& v, int& idx, float& sum)
{
    __shared__ float s_v[32];
    __shared__ int s_i[32];
    __shared__ float s_s[32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        {
            const float ov = __shfl_down_sync(0xffffffffu, v, o);
            const int oi = __shfl_down_sync(0xffffffffu, idx, o);
            sum += __shfl_down_sync(0xffffffffu, sum, o);
            if (ov > v || (ov == v && oi < idx))
                {
                    v = ov;
                    idx = oi;
                }
        }
    const int w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    if ((threadIdx.x & 31) == 0)
        {
            s_v[w] = v;
            s_i[w] = idx;
            s_s[w] = sum;
        }
    __syncthreads();
    if (threadIdx.x < 32)
        {
            v = threadIdx.x < nw ? s_v[threadIdx.x] : -1.0f;
            idx = threadIdx.x < nw ? s_i[threadIdx.x] : 0x7fffffff;
            sum = threadIdx.x < nw ? s_s[threadIdx.x] : 0.0f;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1)
                {
                    const float ov = __shfl_down_sync(0xffffffffu, v, o);
                    const int oi = __shfl_down_sync(0xffffffffu, idx, o);
                    sum += __shfl_down_sync(0xffffffffu, sum, o);
                    if (ov > v || (ov == v && oi < idx))
                        {
                            v = ov;
                            idx = oi;
                        }
                }
        }
}

// The multiplication by the conjugated code spectrum (volk_32fc_x2_multiply_32fc, pcps_acquisition.cc:650) as the input
// hook of the first inverse stage: point q of butterfly id, spectra stored [q][butterfly] (see acq_fwd_kernel)
struct AcqCodeProduct
{
    static constexpr bool on = true;
    const float2* c;
    int nb;
    __device__ __forceinline__ float2 operator()(int q, int id) const { return __ldg(c + q * nb + id); }
};

// |y|^2 of output sample n, dwell accumulation into the kept grid (pcps_acquisition.cc:657-665), running max / first index of
// the max / sum, as the output hook of the last inverse stage.  Samples arrive in no particular order: ties go to the lower
// index explicitly (volk_gnsssdr_32f_index_max_32u keeps the first maximum).
struct AcqSink
{
    static constexpr bool on = true;
    static constexpr bool batched = true;
    int fast;  // every output is searched and no grid is kept (the plain one-dwell CFAR sweep): no masking in the per-butterfly form below
    int out_offset, neff, accumulate;
    float* grid_row;
    float best, sum;
    int besti;
    __device__ __forceinline__ void operator()(float2 y, int n)
    {
        // branch-free: samples outside [out_offset, out_offset + neff) add nothing and can never win
        const int i = n - out_offset;
        const bool ok = (unsigned)i < (unsigned)neff;
        // volk_32fc_magnitude_squared_32f (pcps_acquisition.cc:659,663): two products and one sum, each rounded to fp32 (no FMA
        // contraction), so a near-tie between two code phases resolves as it does in the reference's generic kernel
        float g = __fadd_rn(__fmul_rn(y.x, y.x), __fmul_rn(y.y, y.y));
        if (grid_row)  // uniform
            {
                if (ok)
                    {
                        if (accumulate) g = __fadd_rn(grid_row[i], g);  // volk_32f_x2_add_32f, :664
                        grid_row[i] = g;
                    }
            }
        sum += ok ? g : 0.0f;
        const float gm = ok ? g : -2.0f;
        const bool better = gm > best || (gm == best && i < besti);
        best = better ? gm : best;
        besti = better ? i : besti;
    }
    // The R outputs of one butterfly, indices first + i * stride (ascending).  The per-sample form above is a serial compare / select chain R deep
    // (the final pass of the C3 kernel spent most of its stall samples waiting on it: profiles/r02g_acq_inv_stall_segments.txt, segment 10);
    // here the butterfly's maximum and its lowest index come out of a tournament (depth log2 R; the lower index sits on the left and only a
    // strictly larger right value replaces it, so ties go to the first index as in volk_gnsssdr_32f_index_max_32u), the sum out of a tree, and the
    // running (best, besti, sum) are touched once per butterfly.
    template <int R>
    __device__ __forceinline__ void batch(const float2* v, int first, int stride)
    {
        float mv[R], sv[R];
        int mi[R];
#pragma unroll
        for (int i = 0; i < R; i++)
            {
                float g = __fadd_rn(__fmul_rn(v[i].x, v[i].x), __fmul_rn(v[i].y, v[i].y));
                mi[i] = i;
                if (fast)  // uniform
                    {
                        mv[i] = g;
                        sv[i] = g;
                    }
                else
                    {
                        // samples outside [out_offset, out_offset + neff) add nothing and can never win; a kept grid takes the dwell's sum (:657-665)
                        const int idx = first + i * stride - out_offset;
                        const bool ok = (unsigned)idx < (unsigned)neff;
                        if (grid_row && ok)
                            {
                                if (accumulate) g = __fadd_rn(grid_row[idx], g);
                                grid_row[idx] = g;
                            }
                        mv[i] = ok ? g : -2.0f;
                        sv[i] = ok ? g : 0.0f;
                    }
            }
#pragma unroll
        for (int w = 1; w < R; w <<= 1)
#pragma unroll
            for (int i = 0; i + w < R; i += 2 * w)
                {
                    const bool t = mv[i + w] > mv[i];
                    mv[i] = t ? mv[i + w] : mv[i];
                    mi[i] = t ? mi[i + w] : mi[i];
                    sv[i] += sv[i + w];
                }
        sum += sv[0];
        const int n = first + mi[0] * stride - out_offset;
        const bool better = mv[0] > best || (mv[0] == best && n < besti);
        best = better ? mv[0] : best;
        besti = better ? n : besti;
    }
};
}  // namespace gb200

#include "acq_kernel2.cuh"
