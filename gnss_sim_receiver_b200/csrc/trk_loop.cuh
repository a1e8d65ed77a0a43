// Device-resident DLL/PLL loop closure: one thread per channel, state in HBM.
//
// What the reference does once per code period and channel inside dll_pll_veml_tracking::general_work
// (src/algorithms/tracking/gnuradio_blocks/dll_pll_veml_tracking.cc:1728-2094) AROUND its correlator call: pull-in alignment
// (:1757-1788), save_correlation_results (:1263-1373), cn0_and_tracking_lock_status (:972-1029), run_dll_pll (:1065-1152),
// update_tracking_vars (:1189-1260), the state machine 0 standby / 1 pull-in / 2 wide tracking + symbol sync / 3 extended
// integration / 4 narrow tracking (:1791-2094), and do_correlation_step's argument casts (:1037-1062) — plus what the host side of
// gb200_trk_enqueue does per command (csrc/trk_submit.inc: prepare_cmd) — so that a step of a closed loop over 10^4 channels is
//   correlate kernel -> this kernel -> correlate kernel -> ...
// with nothing but the new IQ block going to the device and the emitted Gnss_Synchro records coming back.
// Scope: real-code channels, high_dyn=false: data-component tracking, and pilot tracking (track_pilot) with a second one-tap correlator
// on the data code when the pilot's secondary code does not depend on the satellite (E1C / E1B, L5Q / L5I); the host block
// (host/b200_dll_pll_veml_tracking.cc, bit-exact scalar pieces) remains the path for the rest and the parity reference of this one.
// The scalar pieces restate tracking_discriminators.cc:68-149, tracking_loop_filter.cc:58-200, tracking_FLL_PLL_filter.cc:23-99,
// lock_detectors.cc:90-148, exponential_smoother.cc with the reference's operand types (float / double), like host/b200_tracking_libs.cc.
#pragma once
#include "trk_kernel.cuh"

namespace gb200
{
constexpr double LOOP_PI = 3.1415926535898;  // MATH_CONSTANTS.h:47-49 (IS-GPS-200 pi): the reference's loop uses this value, not M_PI
constexpr double LOOP_TWO_PI = 2.0 * LOOP_PI;
constexpr int LOOP_MAX_CN0 = 32;             // cn0_samples the per-channel prompt window can hold
constexpr int LOOP_MAX_SEC = 192;            // longest secondary code / bit-sync pattern (GPS L1 C/A preamble pattern: 160)

struct LoopFilterDll  // Tracking_loop_filter
{
    float in_coef[4], out_coef[3], x[4], y[4];
    int n_in, n_out, head;
};
struct LoopFilterPll  // Tracking_FLL_PLL_filter
{
    float w, x, w0p, w0p2, w0p3, w0f, w0f2, a2, a3, b3;
    int order;
};
struct LoopSmoother  // Exponential_Smoother
{
    float alpha, one_minus_alpha, old_value, min_value, offset, init_sum;
    int samples_for_initialization, init_counter, initializing;
};

// per-loop constants (one signal per loop object)
struct LoopConf
{
    gb200_loop_config c;
    unsigned sec_zero[LOOP_MAX_SEC / 32];  // bit i set: chip i of the secondary code / sync pattern is '0'
    unsigned data_sec_zero[2];             // the same for the data component's secondary code (track_pilot)
    int n_channels;                        // data command / data code slot of channel i: n_channels + i (track_pilot)
    float shifts_wide[GB200_MAX_TAPS], shifts_narrow[GB200_MAX_TAPS];
    int n_taps, idx_ve, idx_e, idx_p, idx_l, idx_vl;
    int enable_extended;
    // ring / engine geometry for building the device commands
    unsigned long long ring_base, code_base;
    unsigned long long ring_samples;
    int elem_bytes, max_code_len, max_block;
    float max_code_step_run;  // commands above this code step cannot go to the run-based kernel (0: general kernel in use)
};

struct LoopState
{
    int state;                 // d_state
    int pull_in_transitory, cloop, acc_carrier_phase_initialized, flag_pll_180;
    unsigned prn;
    unsigned long long nitems_read, acq_sample_stamp;
    double acq_code_phase_samples, acq_carrier_doppler_hz;
    double carrier_doppler_hz, carrier_phase_step_rad, code_freq_chips, code_phase_step_chips;
    double rem_code_phase_samples, rem_code_phase_chips, acc_carrier_phase_rad;
    double T_prn_samples, K_blk_samples, current_correlation_time_s;
    double carr_phase_error_hz, carr_freq_error_hz, carr_error_filt_hz, code_error_chips, code_error_filt_chips;
    double cn0_db_hz, carrier_lock_test;
    float rem_carr_phase_rad;  // a float member in the reference (dll_pll_veml_tracking.h:175)
    int current_prn_length_samples;
    int cn0_estimation_counter, carrier_lock_fail_counter, code_lock_fail_counter;
    int current_symbol, current_data_symbol, extend_count;
    int narrow;                // taps / spc in use
    float spc;
    float2 VE, E, P, L, VL, P_data, P_old;
    float2 prompt_buffer[LOOP_MAX_CN0];
    unsigned sign_bits[LOOP_MAX_SEC / 32];  // prompt sign history, oldest at bit 0
    int sign_count;
    int own_secondary;         // the channel carries its own secondary code (per-satellite codes: Galileo E5a-Q)
    unsigned own_sec_zero[LOOP_MAX_SEC / 32];
    LoopFilterDll dll;
    LoopFilterPll pll;
    LoopSmoother cn0_smoother, lock_smoother;
    int planned;               // a correlation command for this channel is in the command array
    int error;                 // sticky: command outside what the selected kernel serves
};

struct alignas(16) LoopCounters
{
    unsigned n_records, n_planned, n_ready_again, n_errors, n_loss_of_lock, pad_[3];  // pad_[0]: ticket of the publishing round (trk_loop_kernel)
};

// ------------------------------------------------------------------------------------------------ scalar pieces
__host__ __device__ __forceinline__ void loop_dll_update_coefficients(LoopFilterDll& f, float T, float bw, int order)
{
    // tracking_loop_filter.cc:95-200, include_last_integrator = false (the tracking block's setting)
    const float zeta = 1.0F / sqrtf(2.0F);
    float g1, g2, g3, wn;
    switch (order)
        {
        case 1:
            wn = bw * 4.0F;
            g1 = wn;
            f.in_coef[0] = g1;
            f.n_in = 1;
            f.n_out = 0;
            break;
        case 2:
            wn = bw * (8.0F * zeta) / (4.0F * zeta * zeta + 1.0F);
            g1 = wn * wn;
            g2 = wn * 2.0F * zeta;
            f.in_coef[0] = (float)(g1 * T / 2.0 + g2);
            f.in_coef[1] = (float)(g1 * T / 2.0 - g2);
            f.n_in = 2;
            f.out_coef[0] = 1.0F;
            f.n_out = 1;
            break;
        default:
            {
                wn = bw / 0.7845F;
                const float a3 = 1.1F, b3 = 2.4F;
                g1 = wn * wn * wn;
                g2 = a3 * wn * wn;
                g3 = b3 * wn;
                f.in_coef[0] = (float)(g3 + T / 2.0 * (g2 + T / 2.0 * g1));
                f.in_coef[1] = (float)(g1 * T * T / 2.0 - 2.0 * g3);
                f.in_coef[2] = (float)(g3 + T / 2.0 * (-g2 + T / 2.0 * g1));
                f.n_in = 3;
                f.out_coef[0] = 2.0F;
                f.out_coef[1] = -1.0F;
                f.n_out = 2;
                break;
            }
        }
}
__host__ __device__ __forceinline__ void loop_dll_initialize(LoopFilterDll& f, float initial_output)
{
    for (int i = 0; i < 4; i++)
        {
            f.x[i] = 0.0F;
            f.y[i] = initial_output;
        }
    f.head = 3;
}
__device__ __forceinline__ float loop_dll_apply(LoopFilterDll& f, float in)
{
    float result = 0.0F;
    for (int i = 0; i < f.n_out; i++) result = __fadd_rn(result, __fmul_rn(f.out_coef[i], f.y[(f.head + i) & 3]));
    f.head = (f.head + 3) & 3;
    f.x[f.head] = in;
    for (int i = 0; i < f.n_in; i++) result = __fadd_rn(result, __fmul_rn(f.in_coef[i], f.x[(f.head + i) & 3]));
    f.y[f.head] = result;
    return result;
}
__host__ __device__ __forceinline__ void loop_pll_set_params(LoopFilterPll& f, float fll_bw, float pll_bw, int order)
{
    f.order = order;
    f.a2 = 1.414F;
    if (order == 3)
        {
            f.b3 = 2.400F;
            f.a3 = 1.100F;
            f.w0p = pll_bw / 0.7845F;
            f.w0p2 = f.w0p * f.w0p;
            f.w0p3 = f.w0p2 * f.w0p;
            f.w0f = fll_bw / 0.53F;
            f.w0f2 = f.w0f * f.w0f;
        }
    else
        {
            f.w0p = pll_bw / 0.53F;
            f.w0p2 = f.w0p * f.w0p;
            f.w0f = fll_bw / 0.25F;
        }
}
__host__ __device__ __forceinline__ void loop_pll_initialize(LoopFilterPll& f, float doppler_hz)
{
    if (f.order == 3)
        {
            f.x = 2.0F * doppler_hz;
            f.w = 0;
        }
    else
        {
            f.w = doppler_hz;
            f.x = 0;
        }
}
// every product and sum rounded on its own (no FMA contraction), as the reference's x86 build evaluates it
__device__ __forceinline__ float loop_pll_get_carrier_error(LoopFilterPll& f, float fll, float pll, float Tc)
{
    if (f.order == 3)
        {
            f.w = __fadd_rn(f.w, __fmul_rn(Tc, __fadd_rn(__fmul_rn(f.w0p3, pll), __fmul_rn(f.w0f2, fll))));
            f.x = __fadd_rn(f.x, __fmul_rn(Tc, __fadd_rn(__fadd_rn(__fmul_rn(0.5F, f.w), __fmul_rn(__fmul_rn(f.a2, f.w0f), fll)), __fmul_rn(__fmul_rn(f.a3, f.w0p2), pll))));
            return __fadd_rn(__fmul_rn(0.5F, f.x), __fmul_rn(__fmul_rn(f.b3, f.w0p), pll));
        }
    const float w_new = __fadd_rn(__fadd_rn(f.w, __fmul_rn(__fmul_rn(pll, f.w0p2), Tc)), __fmul_rn(__fmul_rn(fll, f.w0f), Tc));
    const float out = __fadd_rn(__fmul_rn(0.5F, __fadd_rn(w_new, f.w)), __fmul_rn(__fmul_rn(f.a2, f.w0p), pll));
    f.w = w_new;
    return out;
}
__host__ __device__ __forceinline__ void loop_smoother_reset(LoopSmoother& s)
{
    s.initializing = 1;
    s.init_counter = 0;
    s.init_sum = 0.0F;
}
__device__ __forceinline__ float loop_smoother_smooth(LoopSmoother& s, float raw)
{
    if (!s.initializing)
        {
            s.old_value = __fadd_rn(__fmul_rn(s.alpha, raw), __fmul_rn(s.one_minus_alpha, s.old_value));
            return s.old_value;
        }
    s.init_sum = __fadd_rn(s.init_sum, raw);  // std::accumulate over the warm-up buffer, in arrival order
    if (++s.init_counter == s.samples_for_initialization)
        {
            s.old_value = s.init_sum / (float)s.init_counter;
            if (s.old_value < s.min_value + s.offset)
                {
                    s.init_counter = 0;
                    s.init_sum = 0.0F;
                }
            else
                s.initializing = 0;
        }
    return raw;
}
__device__ __forceinline__ double loop_phase_unwrap(double ph)
{
    if (ph >= LOOP_PI / 2.0) return ph - LOOP_PI;
    if (ph <= -LOOP_PI / 2.0) return ph + LOOP_PI;
    return ph;
}
__device__ __forceinline__ float loop_cabs(float2 a) { return hypotf(a.x, a.y); }  // std::abs(std::complex<float>)
__device__ __forceinline__ float loop_cn0_m2m4(const float2* p, int length, float T)
{
    float psig = 0.0F, m2 = 0.0F, m4 = 0.0F;
    const float n = (float)length;
    for (int i = 0; i < length; i++)
        {
            psig = __fadd_rn(psig, fabsf(p[i].x));
            const float pw = __fadd_rn(__fmul_rn(p[i].y, p[i].y), __fmul_rn(p[i].x, p[i].x));
            m2 = __fadd_rn(m2, pw);
            m4 = __fadd_rn(m4, __fmul_rn(pw, pw));
        }
    psig /= n;
    psig = __fmul_rn(psig, psig);
    m2 /= n;
    m4 /= n;
    const float root = sqrtf(__fsub_rn(__fmul_rn(__fmul_rn(2.0F, m2), m2), m4));
    const float snr = isnan(root) ? psig / __fsub_rn(m2, psig) : root / __fsub_rn(m2, root);
    return 10.0F * log10f(snr) - 10.0F * log10f(T);
}

// ------------------------------------------------------------------------------------------------ the per-period steps
__device__ __forceinline__ void loop_set_taps(const LoopConf& k, LoopState& s, bool narrow)
{
    s.narrow = narrow ? 1 : 0;
    if (narrow) s.spc = k.c.early_late_space_narrow_chips;  // set_narrow_taps also moves d_trk_parameters.spc
}

// dll_pll_veml_tracking.cc:972-1029; returns false on loss of lock
__device__ bool loop_cn0_and_lock(const LoopConf& k, LoopState& s, double coh_integration_time_s)
{
    const int ns = k.c.cn0_samples;
    if (s.cn0_estimation_counter < ns)
        {
            s.prompt_buffer[s.cn0_estimation_counter] = s.P;
            s.cn0_estimation_counter++;
            return true;
        }
    s.prompt_buffer[s.cn0_estimation_counter % ns] = s.P;
    s.cn0_estimation_counter++;
    const float cn0_raw = loop_cn0_m2m4(s.prompt_buffer, ns, (float)coh_integration_time_s);
    s.cn0_db_hz = (double)loop_smoother_smooth(s.cn0_smoother, cn0_raw);
    {
        // carrier_lock_detector over ONE prompt (:989): NBD / NBP
        const float si = s.prompt_buffer[0].x, sq = s.prompt_buffer[0].y;
        const float nbp = __fadd_rn(__fmul_rn(si, si), __fmul_rn(sq, sq)), nbd = __fsub_rn(__fmul_rn(si, si), __fmul_rn(sq, sq));
        s.carrier_lock_test = (double)loop_smoother_smooth(s.lock_smoother, nbd / nbp);
    }
    if (!s.pull_in_transitory)
        {
            if (s.carrier_lock_test < (double)k.c.carrier_lock_th)
                s.carrier_lock_fail_counter++;
            else if (s.carrier_lock_fail_counter > 0)
                s.carrier_lock_fail_counter--;
            if (s.cn0_db_hz < (double)k.c.cn0_min)
                s.code_lock_fail_counter++;
            else if (s.code_lock_fail_counter > 0)
                s.code_lock_fail_counter--;
        }
    if (s.carrier_lock_fail_counter > k.c.max_carrier_lock_fail || s.code_lock_fail_counter > k.c.max_code_lock_fail)
        {
            s.carrier_lock_fail_counter = 0;
            s.code_lock_fail_counter = 0;
            return false;
        }
    return true;
}

// :1065-1152 (enable_doppler_correction not served)
__device__ void loop_run_dll_pll(const LoopConf& k, LoopState& s)
{
    // PLL discriminator: Costas atan(Q / I) while data transitions are possible, four-quadrant once they are wiped (:1069-1078)
    if (s.cloop)
        s.carr_phase_error_hz = (s.P.x != 0.0F ? (double)atanf(s.P.y / s.P.x) : 0.0) / LOOP_TWO_PI;
    else
        s.carr_phase_error_hz = (double)atan2f(s.P.y, s.P.x) / LOOP_TWO_PI;
    const bool fll_pull_in = s.pull_in_transitory && k.c.enable_fll_pull_in;
    if (fll_pull_in || k.c.enable_fll_steady_state)
        {
            double d = (double)atanf(s.P.y / s.P.x) - (double)atanf(s.P_old.y / s.P_old.x);  // fll_diff_atan
            if (isnan(d)) d = 0;
            s.carr_freq_error_hz = loop_phase_unwrap(d) / (s.current_correlation_time_s - 0.0) / LOOP_TWO_PI;
            s.P_old = s.P;
            s.carr_error_filt_hz = (double)loop_pll_get_carrier_error(s.pll, (float)s.carr_freq_error_hz, fll_pull_in ? 0.0F : (float)s.carr_phase_error_hz, (float)s.current_correlation_time_s);
        }
    else
        s.carr_error_filt_hz = (double)loop_pll_get_carrier_error(s.pll, 0.0F, (float)s.carr_phase_error_hz, (float)s.current_correlation_time_s);
    s.carrier_doppler_hz = s.carr_error_filt_hz;
    // DLL
    if (k.c.veml)
        {
            const double early = sqrt((double)(s.VE.x * s.VE.x + s.VE.y * s.VE.y + s.E.x * s.E.x + s.E.y * s.E.y));
            const double late = sqrt((double)(s.L.x * s.L.x + s.L.y * s.L.y + s.VL.x * s.VL.x + s.VL.y * s.VL.y));
            const double sum = early + late;
            s.code_error_chips = sum == 0.0 ? 0.0 : (early - late) / sum;
        }
    else
        {
            const double pe = (double)loop_cabs(s.E), pl = (double)loop_cabs(s.L);
            const double sum = pe + pl;
            s.code_error_chips = sum == 0.0 ? 0.0 : (double)((k.c.y_intercept - k.c.slope * s.spc) / k.c.slope) * (pe - pl) / sum;
        }
    s.code_error_filt_chips = (double)loop_dll_apply(s.dll, (float)s.code_error_chips);
    s.code_freq_chips = k.c.code_chip_rate - s.code_error_filt_chips;
    if (k.c.carrier_aiding) s.code_freq_chips += s.carrier_doppler_hz * k.c.code_chip_rate / k.c.signal_carrier_freq;
}

// :1189-1260 (high_dyn = false)
__device__ void loop_update_tracking_vars(const LoopConf& k, LoopState& s)
{
    const double T_chip = 1.0 / s.code_freq_chips;
    const double T_prn = T_chip * (double)k.c.code_length_chips;
    s.T_prn_samples = T_prn * k.c.fs_in;
    s.K_blk_samples = s.T_prn_samples + s.rem_code_phase_samples;
    s.current_prn_length_samples = (int)floor(s.K_blk_samples);
    const double n = (double)s.current_prn_length_samples;
    s.carrier_phase_step_rad = LOOP_TWO_PI * s.carrier_doppler_hz / k.c.fs_in;
    const double advance = s.carrier_phase_step_rad * n;
    s.rem_carr_phase_rad += (float)advance;
    s.rem_carr_phase_rad = (float)fmod((double)s.rem_carr_phase_rad, LOOP_TWO_PI);
    s.acc_carrier_phase_rad -= advance;
    s.code_phase_step_chips = s.code_freq_chips / k.c.fs_in;
    s.rem_code_phase_samples = s.K_blk_samples - n;
    s.rem_code_phase_chips = s.code_freq_chips * s.rem_code_phase_samples / k.c.fs_in;
}

// :1263-1373.  dprompt: the data correlator's prompt (track_pilot), else unused
__device__ void loop_save_correlation_results(const LoopConf& k, LoopState& s, const float2* taps, float2 dprompt)
{
    const unsigned* sec_zero = s.own_secondary ? s.own_sec_zero : k.sec_zero;
    const bool flip = k.c.secondary && !((sec_zero[s.current_symbol >> 5] >> (s.current_symbol & 31)) & 1u);
    const float sign = flip ? -1.0F : 1.0F;
    if (k.c.veml)
        {
            s.VE.x += sign * taps[k.idx_ve].x;
            s.VE.y += sign * taps[k.idx_ve].y;
            s.VL.x += sign * taps[k.idx_vl].x;
            s.VL.y += sign * taps[k.idx_vl].y;
        }
    s.E.x += sign * taps[k.idx_e].x;
    s.E.y += sign * taps[k.idx_e].y;
    s.P.x += sign * taps[k.idx_p].x;
    s.P.y += sign * taps[k.idx_p].y;
    s.L.x += sign * taps[k.idx_l].x;
    s.L.y += sign * taps[k.idx_l].y;
    if (k.c.secondary) s.current_symbol = (s.current_symbol + 1) % k.c.secondary_code_length;
    // the symbol comes from the data component's own correlator when the pilot is tracked (:1305-1362)
    const float2 dp = k.c.track_pilot ? dprompt : taps[k.idx_p];
    if (k.c.symbols_per_bit > 1)
        {
            if (k.c.data_secondary_code_length > 0)
                {
                    const bool zero = (k.data_sec_zero[s.current_data_symbol >> 5] >> (s.current_data_symbol & 31)) & 1u;
                    const float ds = zero ? 1.0F : -1.0F;
                    s.P_data.x += ds * dp.x;
                    s.P_data.y += ds * dp.y;
                    s.current_data_symbol = (s.current_data_symbol + 1) % k.c.data_secondary_code_length;
                }
            else
                {
                    s.P_data.x += dp.x;
                    s.P_data.y += dp.y;
                    s.current_data_symbol = (s.current_data_symbol + 1) % k.c.symbols_per_bit;
                }
        }
    else
        s.P_data = dp;
    s.cloop = k.c.track_pilot ? 0 : 1;  // tracking the pilot: no Costas loop (:1364-1372)
}

__device__ void loop_clear_tracking_vars(LoopState& s)
{
    s.P_old = make_float2(0.0F, 0.0F);
    s.carr_phase_error_hz = s.carr_freq_error_hz = s.carr_error_filt_hz = 0.0;
    s.code_error_chips = s.code_error_filt_chips = 0.0;
    s.current_symbol = 0;
    s.current_data_symbol = 0;
    s.sign_count = 0;
}

// push the sign of a prompt into the history (d_Prompt_circular_buffer, oldest first) and test it against the pattern (:925-969)
__device__ bool loop_push_and_acquire_secondary(const LoopConf& k, LoopState& s, float prompt_re)
{
    const int len = k.c.secondary_code_length;
    const unsigned neg = prompt_re < 0.0F ? 1u : 0u;
    constexpr int W = LOOP_MAX_SEC / 32;
    if (s.sign_count == len)
        {
            // pop_front: shift the whole history down by one bit
            for (int i = 0; i < W; i++) s.sign_bits[i] = (s.sign_bits[i] >> 1) | (i + 1 < W ? (s.sign_bits[i + 1] << 31) : 0u);
            s.sign_count--;
        }
    {
        const int b = s.sign_count;
        s.sign_bits[b >> 5] = (s.sign_bits[b >> 5] & ~(1u << (b & 31))) | (neg << (b & 31));
        s.sign_count++;
    }
    if (s.sign_count != len) return false;
    int matches = 0;
    for (int i = 0; i < W; i++)
        {
            const int lo = i * 32;
            if (lo >= len) break;
            const unsigned mask = (len - lo >= 32) ? 0xffffffffu : ((1u << (len - lo)) - 1u);
            matches += __popc(~(s.sign_bits[i] ^ (s.own_secondary ? s.own_sec_zero[i] : k.sec_zero[i])) & mask);  // negative == zero_chip
        }
    if (matches == len || matches == 0)
        {
            s.flag_pll_180 = matches == 0;
            return true;
        }
    return false;
}

// one record per emitted Gnss_Synchro (layout of gb200_loop_record)
__device__ void loop_emit(const LoopConf& k, const LoopState& s, int ch, bool valid, gb200_loop_record* records, unsigned max_records, LoopCounters* cnt)
{
    const unsigned i = atomicAdd(&cnt->n_records, 1u);
    if (i >= max_records) return;
    gb200_loop_record r;
    r.channel = ch;
    r.prn = s.prn;
    r.flag_valid_symbol_output = valid ? 1 : 0;
    r.correlation_length_ms = (int)(k.c.code_period_s * 1000.0 + 0.5);
    r.prompt_i = valid ? (double)(k.c.interchange_iq ? s.P_data.y : s.P_data.x) : 0.0;
    r.prompt_q = valid ? (double)(k.c.interchange_iq ? s.P_data.x : s.P_data.y) : 0.0;
    r.cn0_db_hz = valid ? s.cn0_db_hz : 0.0;
    r.carrier_doppler_hz = valid ? s.carrier_doppler_hz : 0.0;
    r.carrier_phase_rads = valid ? s.acc_carrier_phase_rad : 0.0;
    r.code_phase_samples = valid ? s.rem_code_phase_samples : 0.0;
    r.tracking_sample_counter = s.nitems_read + (unsigned long long)s.current_prn_length_samples;  // nitems_read after consume_each
    records[i] = r;
}

// second half of general_work (:1791-2094) for a channel whose command was correlated
__device__ void loop_finish(const LoopConf& k, LoopState& s, int ch, const float2* taps, float2 dprompt, gb200_loop_record* records, unsigned max_records, LoopCounters* cnt)
{
    bool emit = false, loss = false, reset_acc = false;
    switch (s.state)
        {
        case 2:
            {
                if (k.c.veml)
                    {
                        s.VE = taps[k.idx_ve];
                        s.VL = taps[k.idx_vl];
                    }
                s.E = taps[k.idx_e];
                s.P = taps[k.idx_p];
                s.L = taps[k.idx_l];
                s.spc = k.c.early_late_space_chips;
                if ((unsigned long long)k.c.bit_synchronization_time_limit_s < (s.nitems_read - s.acq_sample_stamp) / (unsigned long long)(int)k.c.fs_in) s.carrier_lock_fail_counter = 300000;
                if (!loop_cn0_and_lock(k, s, k.c.code_period_s))
                    {
                        loop_clear_tracking_vars(s);
                        s.state = 0;
                        loss = true;
                        break;
                    }
                loop_run_dll_pll(k, s);
                loop_update_tracking_vars(k, s);
                bool next_state = false;
                if (!s.pull_in_transitory)
                    {
                        if (k.c.secondary || k.c.symbols_per_bit > 1)
                            next_state = loop_push_and_acquire_secondary(k, s, taps[k.idx_p].x);
                        else
                            next_state = true;
                    }
                if (next_state)
                    {
                        s.VE = s.E = s.P = s.L = s.VL = s.P_data = make_float2(0.0F, 0.0F);
                        s.sign_count = 0;
                        s.current_symbol = 0;
                        s.current_data_symbol = 0;
                        if (k.enable_extended)
                            {
                                s.extend_count = 0;
                                s.current_correlation_time_s = (double)((float)k.c.extend_correlation_symbols * (float)k.c.code_period_s);
                                s.state = 3;
                                loop_dll_update_coefficients(s.dll, (float)s.current_correlation_time_s, k.c.dll_bw_narrow_hz, k.c.dll_filter_order);
                                loop_pll_set_params(s.pll, k.c.fll_bw_hz, k.c.pll_bw_narrow_hz, k.c.pll_filter_order);
                                loop_set_taps(k, s, true);
                            }
                        else
                            s.state = 4;
                    }
                break;
            }
        case 3:
            {
                loop_save_correlation_results(k, s, taps, dprompt);
                loop_update_tracking_vars(k, s);
                if (s.current_data_symbol == 0) emit = true;
                s.extend_count++;
                if (s.extend_count == k.c.extend_correlation_symbols - 1)
                    {
                        s.extend_count = 0;
                        s.state = 4;
                    }
                break;
            }
        case 4:
            {
                loop_save_correlation_results(k, s, taps, dprompt);
                if (!loop_cn0_and_lock(k, s, k.c.code_period_s * (double)k.c.extend_correlation_symbols))
                    {
                        loop_clear_tracking_vars(s);
                        s.state = 0;
                        loss = true;
                        break;
                    }
                loop_run_dll_pll(k, s);
                loop_update_tracking_vars(k, s);
                if (!s.acc_carrier_phase_initialized)  // check_carrier_phase_coherent_initialization (:1155-1162)
                    {
                        s.acc_carrier_phase_rad = -(double)s.rem_carr_phase_rad;
                        s.acc_carrier_phase_initialized = 1;
                    }
                if (s.current_data_symbol == 0) emit = true;
                reset_acc = true;
                if (k.enable_extended) s.state = 3;
                break;
            }
        default: break;
        }
    if (emit || loss)
        {
            loop_emit(k, s, ch, !loss, records, max_records, cnt);  // fill_synchro (:1791-1800) before the accumulators restart
            if (loss) atomicAdd(&cnt->n_loss_of_lock, 1u);
        }
    if (emit) s.P_data = make_float2(0.0F, 0.0F);
    if (reset_acc) s.VE = s.E = s.P = s.L = s.VL = make_float2(0.0F, 0.0F);
}

// Device restatement of trk_submit.inc: prepare_cmd for a command formed on the device
__device__ void loop_write_cmd(const LoopConf& k, const LoopState& s, int ch, TrkDevCmd& d)
{
    const float spc = (float)k.c.code_samples_per_chip;
    const float rem_carr = s.rem_carr_phase_rad;
    const float phase_step = (float)s.carrier_phase_step_rad;
    // the two fp32-rounded phasors the reference seeds its rotator with (cpu_multicorrelator_real_codes.cc:114-123), then their angles
    const float ph_re = cosf(rem_carr), ph_im = -sinf(rem_carr);
    const float inc_re = cosf(-phase_step), inc_im = sinf(-phase_step);
    auto fix = [](float im, float re) {
        const double turns = atan2((double)im, (double)re) * 0.15915494309189533576888376337251;
        double f = turns - floor(turns);
        if (f >= 1.0) f = 0.0;
        return __double2ull_rz(f * 9007199254740992.0) << 11;
    };
    d.u0 = fix(ph_im, ph_re);
    d.du = fix(inc_im, inc_re);
    d.iq = reinterpret_cast<const void*>(k.ring_base + (s.nitems_read % k.ring_samples) * (unsigned long long)k.elem_bytes);
    d.code = reinterpret_cast<const void*>(k.code_base + (unsigned long long)ch * k.max_code_len * 2ull * sizeof(float));
    d.n = k.c.vector_length;
    d.ntaps = k.n_taps;
    d.L = k.c.code_length_chips * k.c.code_samples_per_chip;
    d.flags = 0;
    d.rem_code = (float)s.rem_code_phase_chips * spc;
    d.code_step = (float)s.code_phase_step_chips * spc;
    d.code_rate = 0.0F;
    const float* sh = s.narrow ? k.shifts_narrow : k.shifts_wide;
    float lo = sh[0], hi = sh[0];
    for (int t = 0; t < GB200_MAX_TAPS; t++)
        {
            const float v = sh[t < k.n_taps ? t : k.n_taps - 1];
            d.shifts[t] = v;
            lo = fminf(lo, v);
            hi = fmaxf(hi, v);
        }
    d.shift_min = lo;
    d.shift_max = hi;
    d.pad_ = 0;
}

// One round of loop closure for every channel: consume the correlator outputs of the channels whose command was in the last launch
// (finish), then form the next command of every channel whose window is in the ring (plan).  LOOP_CTA threads per CTA, one channel each.
// The CTA's channel states travel HBM -> shared memory -> HBM with coalesced copies, once per round: the loop closure is a long chain of
// dependent reads and writes of its 768-byte state, which cost an HBM / L2 round trip each when done in place (29 us per round for
// 10 240 channels against 9 us from shared memory).  The shared copies sit at a stride of LOOP_STATE_STRIDE bytes (194 words: a stride
// of the struct's own 192 words would put the same field of all 32 lanes in one bank).
constexpr int LOOP_CTA = 32;
constexpr int LOOP_STATE_WORDS = (int)(sizeof(LoopState) / 4);
constexpr int LOOP_STATE_STRIDE = (int)sizeof(LoopState) + 8;
static_assert(sizeof(LoopState) % 8 == 0, "LoopState is copied as words and holds doubles");
__global__ void __launch_bounds__(LOOP_CTA) trk_loop_kernel(const LoopConf* __restrict__ conf, LoopState* __restrict__ states, int n_channels, unsigned long long head,
    const float* __restrict__ outs, TrkDevCmd* __restrict__ cmds, gb200_loop_record* __restrict__ records, unsigned max_records, LoopCounters* __restrict__ cnt,
    LoopCounters* __restrict__ publish)
{
    __shared__ LoopConf k;
    __shared__ __align__(16) unsigned char sm_states[LOOP_CTA * LOOP_STATE_STRIDE];
    {
        const unsigned* src = reinterpret_cast<const unsigned*>(conf);
        unsigned* dst = reinterpret_cast<unsigned*>(&k);
        for (int i = threadIdx.x; i < (int)(sizeof(LoopConf) / 4); i += blockDim.x) dst[i] = src[i];
    }
    const int ch0 = blockIdx.x * LOOP_CTA;
    const int n_here = min(LOOP_CTA, n_channels - ch0);
    {
        const unsigned* src = reinterpret_cast<const unsigned*>(states + ch0);
        for (int i = threadIdx.x; i < n_here * LOOP_STATE_WORDS; i += LOOP_CTA)
            reinterpret_cast<unsigned*>(sm_states + (i / LOOP_STATE_WORDS) * LOOP_STATE_STRIDE)[i % LOOP_STATE_WORDS] = src[i];
    }
    __syncthreads();
    const int ch = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = ch < n_channels;
    bool planned_now = false, ready_again = false, bad = false;
    if (valid)
        {
            LoopState& s = *reinterpret_cast<LoopState*>(sm_states + threadIdx.x * LOOP_STATE_STRIDE);
            if (s.planned)
                {
                    float2 taps[GB200_MAX_TAPS];
                    const float2* o = reinterpret_cast<const float2*>(outs) + (size_t)ch * TRK_OUT_STRIDE;
                    for (int t = 0; t < k.n_taps; t++) taps[t] = o[t];
                    const float2 dprompt = k.c.track_pilot ? reinterpret_cast<const float2*>(outs)[(size_t)(k.n_channels + ch) * TRK_OUT_STRIDE] : make_float2(0.0F, 0.0F);
                    loop_finish(k, s, ch, taps, dprompt, records, max_records, cnt);
                    s.nitems_read += (unsigned long long)s.current_prn_length_samples;  // consume_each(d_current_prn_length_samples), :2092
                    s.planned = 0;
                }
            const unsigned long long need = 2ull * (unsigned long long)k.c.vector_length;  // forecast(): 2 * vector_length items (:607-614)
            for (int it = 0; it < 2; it++)
                {
                    if (s.state == 0)
                        {
                            s.nitems_read = head;  // standby consumes whatever arrives (:1746-1752)
                            break;
                        }
                    if (s.nitems_read + need > head) break;
                    if (s.pull_in_transitory && (unsigned long long)k.c.pull_in_time_s < (s.nitems_read - s.acq_sample_stamp) / (unsigned long long)(int)k.c.fs_in)
                        {
                            s.pull_in_transitory = 0;  // :1738-1747
                            s.carrier_lock_fail_counter = 0;
                            s.code_lock_fail_counter = 0;
                        }
                    if (s.state == 1)
                        {
                            // pull-in: skip samples until the incoming code is aligned with the local replica (:1757-1788)
                            const long long acq_trk_diff_samples = (long long)s.nitems_read - (long long)s.acq_sample_stamp;
                            const double delta_trk_to_acq_prn_start_samples = (double)acq_trk_diff_samples - s.acq_code_phase_samples;
                            s.code_freq_chips = k.c.code_chip_rate;
                            s.code_phase_step_chips = s.code_freq_chips / k.c.fs_in;
                            const double T_prn_mod_samples = (1.0 / s.code_freq_chips) * (double)k.c.code_length_chips * k.c.fs_in;
                            s.acq_code_phase_samples = T_prn_mod_samples - fmod(delta_trk_to_acq_prn_start_samples, T_prn_mod_samples);
                            s.current_prn_length_samples = (int)round(T_prn_mod_samples);
                            const int samples_offset = (int)round(s.acq_code_phase_samples);
                            s.acc_carrier_phase_rad -= s.carrier_phase_step_rad * (double)samples_offset;
                            s.state = 2;
                            loop_smoother_reset(s.cn0_smoother);
                            loop_smoother_reset(s.lock_smoother);
                            s.nitems_read += (unsigned long long)samples_offset;
                            continue;
                        }
                    TrkDevCmd d;
                    loop_write_cmd(k, s, ch, d);
                    if (k.max_code_step_run > 0.0F && !(d.code_step <= k.max_code_step_run && d.code_step >= 1e-3F))
                        {
                            s.error = 1;  // outside what the selected correlator kernel serves: reported by gb200_loop_step, never silently mis-correlated
                            bad = true;
                            break;
                        }
                    cmds[ch] = d;
                    if (k.c.track_pilot)
                        {
                            // the data correlator: same window, carrier and code NCO, the data code, one tap at the prompt shift
                            // (do_correlation_step :1051-1062; d_prompt_data_shift :498,510)
                            const float ps = d.shifts[k.idx_p];
                            d.code = reinterpret_cast<const void*>(k.code_base + (unsigned long long)(k.n_channels + ch) * k.max_code_len * 2ull * sizeof(float));
                            d.ntaps = 1;
                            for (int t = 0; t < GB200_MAX_TAPS; t++) d.shifts[t] = ps;
                            d.shift_min = d.shift_max = ps;
                            cmds[k.n_channels + ch] = d;
                        }
                    s.planned = 1;
                    planned_now = true;
                    ready_again = head - s.nitems_read >= need + (unsigned long long)k.c.vector_length;
                    break;
                }
            if (!s.planned)
                {
                    cmds[ch].n = 0;
                    if (k.c.track_pilot) cmds[k.n_channels + ch].n = 0;
                }
        }
    __syncthreads();
    {
        unsigned* dst = reinterpret_cast<unsigned*>(states + ch0);
        for (int i = threadIdx.x; i < n_here * LOOP_STATE_WORDS; i += LOOP_CTA)
            dst[i] = reinterpret_cast<const unsigned*>(sm_states + (i / LOOP_STATE_WORDS) * LOOP_STATE_STRIDE)[i % LOOP_STATE_WORDS];
    }
    // counters: one atomic per warp (10^4 single atomics on one address cost more than the loop closure itself)
    const unsigned m_planned = __ballot_sync(0xffffffffu, planned_now), m_again = __ballot_sync(0xffffffffu, ready_again), m_bad = __ballot_sync(0xffffffffu, bad);
    if ((threadIdx.x & 31) == 0)
        {
            if (m_planned) atomicAdd(&cnt->n_planned, (unsigned)__popc(m_planned));
            if (m_again) atomicAdd(&cnt->n_ready_again, (unsigned)__popc(m_again));
            if (m_bad) atomicAdd(&cnt->n_errors, (unsigned)__popc(m_bad));
        }
    // The last round of a step publishes its totals itself (publish = the step's host-mapped counter block): the CTA that finishes last
    // moves them over and leaves the device counters zero for the slot's next step — no memset before and no copy after the kernels on the
    // stream, a period is the correlator kernel and this one.
    if (publish)
        {
            __shared__ unsigned s_last;
            __syncthreads();
            if (threadIdx.x == 0)
                {
                    __threadfence();
                    s_last = atomicAdd(&cnt->pad_[0], 1u) == gridDim.x - 1 ? 1u : 0u;
                }
            __syncthreads();
            if (s_last && threadIdx.x == 0)
                {
                    // every other CTA's atomics are in L2 (their fence came before their ticket): two vector loads past L1, two vector stores to
                    // the mapped block, two to clear; the host reads after the kernel's completion event, which orders the stores
                    __threadfence();
                    uint4* c = reinterpret_cast<uint4*>(cnt);
                    const uint4 a = __ldcg(c), b = __ldcg(c + 1);
                    uint4* o = reinterpret_cast<uint4*>(publish);
                    o[0] = a;
                    o[1] = make_uint4(b.x, 0u, 0u, 0u);
                    __stcg(c, make_uint4(0u, 0u, 0u, 0u));
                    __stcg(c + 1, make_uint4(0u, 0u, 0u, 0u));
                }
        }
}
}  // namespace gb200
