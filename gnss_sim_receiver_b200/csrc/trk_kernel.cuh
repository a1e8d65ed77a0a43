// Tracking multicorrelator kernel (K1) for sm_100a.
//
// One launch evaluates, for every command (= one channel's code period), what the reference does with
//   volk_gnsssdr_32f_xn_resampler_32f_xn            (kernels/volk_gnsssdr/volk_gnsssdr_32f_xn_resampler_32f_xn.h:63-80)
//   volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn  (kernels/volk_gnsssdr/volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn.h:66-98)
// called from Cpu_Multicorrelator_Real_Codes::Carrier_wipeoff_multicorrelator_resampler
// (src/algorithms/tracking/libs/cpu_multicorrelator_real_codes.cc:103-126), without materialising the T resampled
// replicas: the chip index of every (tap, sample) is computed on the fly with the reference's exact fp32 law and
// looked up in a shared-memory window of the code table.
//
// grid = (n_chunks, n_cmds); CTA (x, y) correlates samples [x*chunk_len, (x+1)*chunk_len) of command y for all taps.
// Partial sums of the chunks are combined in fixed order by the last CTA of the command to finish (deterministic).
#pragma once
#include "common.cuh"

namespace gb200
{
constexpr int TRK_THREADS = 128;
constexpr int TRK_RESEED = 16;    // exact phasor re-seed period (in per-thread samples)
constexpr int TRK_PREFETCH = 8;   // samples a thread fetches ahead of the arithmetic (general kernel)
constexpr int TRK_OUT_STRIDE = GB200_MAX_TAPS;

// host-preprocessed command, one per Carrier_wipeoff_multicorrelator_resampler call
struct TrkDevCmd
{
    unsigned long long u0;  // carrier multiplier exp(+j*2*pi*(u0 + n*du)/2^64): 64-bit fixed-point turns
    unsigned long long du;
    const void* iq;         // first sample of the window in the device ring
    const void* code;       // channel's code table (float or float2), L entries
    int n;                  // samples
    int ntaps;
    int L;
    int flags;
    float rem_code;
    float code_step;
    float code_rate;
    float shift_min, shift_max;
    float shifts[GB200_MAX_TAPS];
    int pad_;
};

// Carrier angles of a batch of commands, on the device.  The host stores the two fp32-ROUNDED phasors the reference seeds its
// rotator with (cpu_multicorrelator_real_codes.cc:114-123: (cos phi, -sin phi) and exp(-j*step), computed with the host's
// cosf/sinf so that their rounding is the reference's) in the 8 bytes of u0 and du; this kernel replaces them in place by their
// angles as 64-bit fixed-point turns (float64 atan2: an angle error of 1e-16 rad is 1e-12 over a 25 000-sample block).  One
// thread per command, run once right after the command upload — 50 ns of host libm per command become a few microseconds here.
__global__ void __launch_bounds__(256) trk_phase_kernel(TrkDevCmd* cmds, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned long long* slot = &cmds[i].u0;  // u0 and du are adjacent
#pragma unroll
    for (int k = 0; k < 2; k++)
        {
            const unsigned long long raw = slot[k];
            const float re = __uint_as_float((unsigned)(raw & 0xffffffffu)), im = __uint_as_float((unsigned)(raw >> 32));
            const double turns = atan2((double)im, (double)re) * 0.15915494309189533576888376337251;  // 1 / (2 pi)
            double f = turns - floor(turns);
            if (f >= 1.0) f = 0.0;
            slot[k] = __double2ull_rz(f * 9007199254740992.0) << 11;  // 53 significant bits, shifted to the top
        }
}

// reference chip index BEFORE the modulo: floor(step*(float)n + shift - rem), every operation rounded to fp32
// separately (no FMA contraction), volk_gnsssdr_32f_xn_resampler_32f_xn.h:73.
__device__ __forceinline__ int trk_raw_index(float step, float nf, float shift, float rem)
{
    return (int)floorf(__fsub_rn(__fadd_rn(__fmul_rn(step, nf), shift), rem));
}

__device__ __forceinline__ int math_mod(int a, int L)
{
    int r = a % L;
    return r < 0 ? r + L : r;
}

template <int IQ>
__device__ __forceinline__ float2 trk_load_iq(const void* base, int s)
{
    if (IQ == GB200_IQ_CF32) return __ldg(reinterpret_cast<const float2*>(base) + s);
    if (IQ == GB200_IQ_CS16)
        {
            const short2 v = __ldg(reinterpret_cast<const short2*>(base) + s);
            return make_float2((float)v.x, (float)v.y);
        }
    const char2 v = __ldg(reinterpret_cast<const char2*>(base) + s);
    return make_float2((float)v.x, (float)v.y);
}

// phasor of the 64-bit fixed-point phase (turns): top 32 bits -> signed half-turn fraction -> sincospif
__device__ __forceinline__ float2 trk_phasor(unsigned long long u)
{
    const int a = (int)(u >> 32);
    float s, c;
    sincospif((float)a * 4.656612873077393e-10f, &s, &c);  // a * 2^-31 half-turns
    return make_float2(c, s);
}

// T = taps evaluated (commands with fewer taps are padded by the host with shift copies), CPLX = complex code table
template <int T, int IQ, bool CPLX>
__global__ void __launch_bounds__(TRK_THREADS) trk_correlate_kernel(const TrkDevCmd* __restrict__ cmds, int n_chunks,
    float* __restrict__ partials, unsigned* __restrict__ counters, float* __restrict__ out)
{
    extern __shared__ __align__(16) float s_tab[];
    __shared__ float s_red[TRK_THREADS / 32][2 * T];
    __shared__ int s_last;
    __shared__ unsigned s_adj;

    const TrkDevCmd& c = cmds[blockIdx.y];
    const int tid = threadIdx.x;
    const int n = c.n;
    if (n <= 0) return;  // an empty slot of a device-formed command array (trk_loop.cuh): uniform for the CTA, before any barrier
    const int chunk_len = (((n + n_chunks - 1) / n_chunks) + TRK_THREADS - 1) & ~(TRK_THREADS - 1);
    const int n0 = blockIdx.x * chunk_len;
    const int n1 = min(n, n0 + chunk_len);

    // accumulators as packed (re, im) pairs; taps are evaluated two at a time (f32x2 index law), see trk_kernel2.cuh for the packing rules
    constexpr int TP = (T + 1) / 2;
    f32x2 acc[T];
#pragma unroll
    for (int t = 0; t < T; t++) acc[t] = 0ull;

    if (n0 < n1)
        {
            const float step = c.code_step, rem = c.rem_code;
            const int L = c.L;
            // the index is monotone in n and in the shift: the chunk touches raw indices [idx_lo, idx_hi]
            const int idx_lo = trk_raw_index(step, (float)n0, c.shift_min, rem);
            const int idx_hi = trk_raw_index(step, (float)(n1 - 1), c.shift_max, rem);
            const int span = idx_hi - idx_lo + 1;
            if (tid == 0) s_adj = smem_u32(s_tab) - (unsigned)(0x4B400000 + idx_lo) * (CPLX ? 8u : 4u);
            {
                int k = math_mod(idx_lo + tid, L);
                const int adv = TRK_THREADS % L;
                for (int j = tid; j < span; j += TRK_THREADS)
                    {
                        if (CPLX)
                            reinterpret_cast<float2*>(s_tab)[j] = __ldg(reinterpret_cast<const float2*>(c.code) + k);
                        else
                            s_tab[j] = __ldg(reinterpret_cast<const float*>(c.code) + k);
                        k += adv;
                        if (k >= L) k -= L;
                    }
            }
            __syncthreads();

            f32x2 shp[TP];
#pragma unroll
            for (int p = 0; p < TP; p++) shp[p] = pk2(c.shifts[2 * p], c.shifts[2 * p + 1 < T ? 2 * p + 1 : T - 1]);
            // floor(x) for |x| < 2^22 as the low mantissa bits of RD(x + 1.5*2^23)
            const float MAGIC = 12582912.0f;
            const f32x2 MG = pk2(MAGIC, MAGIC), REM = pk2(rem, rem);
            // shared address of table entry j = bits(RD(x + MAGIC)) - bias:  tab_adj + ESZ * bits  (one multiply-add per tap)
            constexpr unsigned ESZ = CPLX ? 8u : 4u;
            // (read back through shared memory: given the constant, ptxas splits it off again and spends a LEA and an IADD3 per tap)
            const unsigned tab_adj = *reinterpret_cast<volatile unsigned*>(&s_adj);
            const unsigned long long u0 = c.u0, du = c.du;
            const void* iq = c.iq;
            const float2 stp = trk_phasor(du * (unsigned long long)TRK_THREADS);
            const f32x2 stpx = pk2(stp.x, stp.x), stpy = pk2(stp.y, stp.y);
            float nf = (float)(n0 + tid);
            // samples of a thread are TRK_THREADS apart; they are fetched TRK_PREFETCH at a time, one group ahead of the arithmetic, so the
            // L2 latency of a group is covered by the previous group's taps
            float2 xn[TRK_PREFETCH];
            auto fetch = [&](int first) {
#pragma unroll
                for (int k = 0; k < TRK_PREFETCH; k++)
                    {
                        const int s = first + k * TRK_THREADS;
                        xn[k] = s < n1 ? trk_load_iq<IQ>(iq, s) : make_float2(0.0f, 0.0f);
                    }
            };
            fetch(n0 + tid);
            int since_seed = TRK_RESEED;
            float2 ph = make_float2(1.0f, 0.0f);
            for (int base = n0 + tid; base < n1; base += TRK_THREADS * TRK_PREFETCH)
                {
                    float2 x[TRK_PREFETCH];
#pragma unroll
                    for (int k = 0; k < TRK_PREFETCH; k++) x[k] = xn[k];
                    if (base + TRK_THREADS * TRK_PREFETCH < n1) fetch(base + TRK_THREADS * TRK_PREFETCH);
                    if (since_seed >= TRK_RESEED)
                        {
                            ph = trk_phasor(u0 + du * (unsigned long long)base);  // exact re-seed every TRK_RESEED samples of the thread
                            since_seed = 0;
                        }
                    since_seed += TRK_PREFETCH;
#pragma unroll
                    for (int k = 0; k < TRK_PREFETCH; k++)
                        {
                            if (base + k * TRK_THREADS >= n1) break;  // (the table window only covers the chunk's own samples)
                            const f32x2 A = pk2(ph.x, ph.y), B = pk2(-ph.y, ph.x);               // ph and j * ph
                            const f32x2 W = fma2(pk2(x[k].y, x[k].y), B, mul2(pk2(x[k].x, x[k].x), A));  // wiped sample x * ph
                            up2(fma2(stpy, B, mul2(stpx, A)), ph.x, ph.y);                        // ph *= exp(j 2 pi du 128)
                            f32x2 JW = 0ull;
                            if (CPLX)
                                {
                                    float wr, wi;
                                    up2(W, wr, wi);
                                    JW = pk2(-wi, wr);
                                }
                            const float a = __fmul_rn(step, nf);  // the reference's law: rounded product, rounded add, rounded sub
                            nf += (float)TRK_THREADS;
                            const f32x2 A0 = pk2(a, a);
#pragma unroll
                            for (int p = 0; p < TP; p++)
                                {
                                    float m0, m1;
                                    up2(add2_rd(sub2(add2(A0, shp[p]), REM), MG), m0, m1);
#pragma unroll
                                    for (int q = 0; q < 2; q++)
                                        {
                                            const int t = 2 * p + q;
                                            if (t >= T) continue;
                                            const unsigned addr = tab_adj + ESZ * (unsigned)__float_as_int(q ? m1 : m0);
                                            if (CPLX)
                                                {
                                                    const float2 cc = ld_shared_f2(addr);
                                                    acc[t] = fma2(pk2(cc.y, cc.y), JW, fma2(pk2(cc.x, cc.x), W, acc[t]));
                                                }
                                            else
                                                {
                                                    const float cc = ld_shared_f1(addr);
                                                    acc[t] = fma2(pk2(cc, cc), W, acc[t]);
                                                }
                                        }
                                }
                        }
                }
        }
    float acc_re[T], acc_im[T];
#pragma unroll
    for (int t = 0; t < T; t++) up2(acc[t], acc_re[t], acc_im[t]);

    // block reduce: warp shuffle, then across the 4 warps through shared memory
#pragma unroll
    for (int t = 0; t < T; t++)
        {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1)
                {
                    acc_re[t] += __shfl_xor_sync(0xffffffffu, acc_re[t], o);
                    acc_im[t] += __shfl_xor_sync(0xffffffffu, acc_im[t], o);
                }
        }
    if ((tid & 31) == 0)
        {
#pragma unroll
            for (int t = 0; t < T; t++)
                {
                    s_red[tid >> 5][2 * t] = acc_re[t];
                    s_red[tid >> 5][2 * t + 1] = acc_im[t];
                }
        }
    __syncthreads();
    float v = 0.0f;
    if (tid < 2 * T)
        {
#pragma unroll
            for (int w = 0; w < TRK_THREADS / 32; w++) v += s_red[w][tid];
        }
    float* o = out + (size_t)blockIdx.y * (2 * TRK_OUT_STRIDE);
    if (n_chunks == 1)
        {
            if (tid < 2 * T) o[tid] = v;
            return;
        }
    float* p = partials + ((size_t)blockIdx.y * n_chunks) * (2 * TRK_OUT_STRIDE);
    if (tid < 2 * T) p[(size_t)blockIdx.x * (2 * TRK_OUT_STRIDE) + tid] = v;
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = (atomicAdd(&counters[blockIdx.y], 1u) == (unsigned)(n_chunks - 1));
    __syncthreads();
    if (s_last)
        {
            __threadfence();
            if (tid < 2 * T)
                {
                    float sum = 0.0f;
                    for (int k = 0; k < n_chunks; k++) sum += __ldcg(p + (size_t)k * (2 * TRK_OUT_STRIDE) + tid);
                    o[tid] = sum;
                }
            if (tid == 0) counters[blockIdx.y] = 0;
        }
}
}  // namespace gb200
