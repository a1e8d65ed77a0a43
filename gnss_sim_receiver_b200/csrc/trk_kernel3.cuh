// Tracking multicorrelator, high-dynamics variant (K1c) for sm_100a — Tracking_XX.high_dyn=true (dll_pll_conf.h:83).
//
// Replaces, per command, what Cpu_Multicorrelator_Real_Codes::Carrier_wipeoff_multicorrelator_resampler
// (src/algorithms/tracking/libs/cpu_multicorrelator_real_codes.cc:75-100) does with
//   volk_gnsssdr_32f_xn_high_dynamics_resampler_32f_xn          (kernels/volk_gnsssdr/...high_dynamics_resampler_32f_xn.h:67-91)
//   volk_gnsssdr_32fc_32f_high_dynamic_rotator_dot_prod_32fc_xn  (kernels/volk_gnsssdr/...high_dynamic_rotator_dot_prod_32fc_xn.h:68-109)
// without materialising the replicas:
//   * tap 0 follows the quadratic code phase  floor(step*(float)n + rate*(float)(n*n) + shift[0] - rem)  (fp32, every
//     operation rounded separately, n*n an unsigned product); tap t is tap 0 rotated by S_t SAMPLES (cumulative
//     round((shift[t] - shift[t-1]) / step)), circularly over the n samples of the window (:83-90) — the host computes S_t;
//   * carrier of sample n: phase_0 * inc^n * exp(j * (float)((n-1)^2) * arg(inc_rate)) for n >= 1 (the rate factor computed
//     in iteration n-1 is the one applied to sample n, :88-101); inc^n from the 64-bit fixed-point phase as in the other
//     kernels, the rate factor as (cosf, sinf) of the fp32 product exactly like the reference's closed form (:93-95).
// Commands without GB200_TRK_HIGH_DYN in the same call are evaluated with the plain law (trk_kernel.cuh) by the same kernel,
// so a call may mix both.  One CTA per command, fixed reduction order: deterministic.
#pragma once
#include "trk_kernel.cuh"

namespace gb200
{
constexpr int TRK3_THREADS = 256;

// For GB200_TRK_HIGH_DYN commands the host stores in TrkDevCmd: shifts[0] = first tap shift (chips), shifts[t >= 1] = the
// bit pattern of the integer S_t, pad_ = the bit pattern of arg(inc_rate) (float), shift_max = ln|inc| of the fp32-rounded
// carrier increment (the reference's un-normalised recursion gains |inc|^n; see prepare_cmd).
template <int IQ>
__global__ void __launch_bounds__(TRK3_THREADS) trk_high_dyn_kernel(const TrkDevCmd* __restrict__ cmds, float* __restrict__ out)
{
    __shared__ float s_red[TRK3_THREADS / 32][2 * GB200_MAX_TAPS];
    const TrkDevCmd& c = cmds[blockIdx.x];
    const int tid = threadIdx.x, n = c.n, T = c.ntaps, L = c.L;
    const bool hd = (c.flags & GB200_TRK_HIGH_DYN) != 0;
    const float step = c.code_step, rem = c.rem_code, rate = c.code_rate, sh0 = c.shifts[0];
    const float arga = __int_as_float(c.pad_);
    const float* code = reinterpret_cast<const float*>(c.code);
    float acc_re[GB200_MAX_TAPS], acc_im[GB200_MAX_TAPS];
#pragma unroll
    for (int t = 0; t < GB200_MAX_TAPS; t++) acc_re[t] = acc_im[t] = 0.0f;
    for (int s = tid; s < n; s += TRK3_THREADS)
        {
            const float2 x = trk_load_iq<IQ>(c.iq, s);
            float2 ph = trk_phasor(c.u0 + c.du * (unsigned long long)s);
            if (hd && s > 0)
                {
                    const unsigned m = (unsigned)(s - 1);
                    const float theta = __fmul_rn((float)(m * m), arga);
                    float sn, cs;
                    sincosf(theta, &sn, &cs);
                    const float gain = 1.0f + (float)s * c.shift_max;
                    ph = make_float2((ph.x * cs - ph.y * sn) * gain, (ph.x * sn + ph.y * cs) * gain);
                }
            const float wr = x.x * ph.x - x.y * ph.y;
            const float wi = x.x * ph.y + x.y * ph.x;
#pragma unroll
            for (int t = 0; t < GB200_MAX_TAPS; t++)
                {
                    if (t < T)
                        {
                            int idx;
                            if (hd)
                                {
                                    unsigned q = (unsigned)s + (t ? (unsigned)__float_as_int(c.shifts[t]) : 0u);
                                    if (q >= (unsigned)n) q -= (unsigned)n;
                                    const float a = __fadd_rn(__fmul_rn(step, (float)q), __fmul_rn(rate, (float)(q * q)));
                                    idx = (int)floorf(__fsub_rn(__fadd_rn(a, sh0), rem));
                                }
                            else
                                idx = trk_raw_index(step, (float)s, c.shifts[t], rem);
                            const float cc = __ldg(code + math_mod(idx, L));
                            acc_re[t] += wr * cc;
                            acc_im[t] += wi * cc;
                        }
                }
        }
#pragma unroll
    for (int t = 0; t < GB200_MAX_TAPS; t++)
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
            for (int t = 0; t < GB200_MAX_TAPS; t++)
                {
                    s_red[tid >> 5][2 * t] = acc_re[t];
                    s_red[tid >> 5][2 * t + 1] = acc_im[t];
                }
        }
    __syncthreads();
    if (tid < 2 * GB200_MAX_TAPS)
        {
            float v = 0.0f;
#pragma unroll
            for (int w = 0; w < TRK3_THREADS / 32; w++) v += s_red[w][tid];
            out[(size_t)blockIdx.x * (2 * TRK_OUT_STRIDE) + tid] = v;
        }
}
}  // namespace gb200
