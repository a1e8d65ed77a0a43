// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// extern "C" doors onto the reference's own C++ (compiled unmodified from
// /root/reference by oracle/Makefile): Cpu_Multicorrelator_Real_Codes,
// Cpu_Multicorrelator, the PRN/replica generators, discriminators, loop
// filters, lock detectors, exponential smoother. Used to pin oracle/ (the C
// restatement) and to generate tests/golden/*. Lives in oracle/_ref/libgnssref.so.
#include "cpu_multicorrelator.h"
#include "cpu_multicorrelator_16sc.h"
#include "tracking_dump_reader.h"
#include "cpu_multicorrelator_real_codes.h"
#include "exponential_smoother.h"
#include "galileo_e1_signal_replica.h"
#include "galileo_e5_signal_replica.h"
#include "gps_l5_signal_replica.h"
#include "gps_sdr_signal_replica.h"
#include "lock_detectors.h"
#include "tracking_FLL_PLL_filter.h"
#include "tracking_discriminators.h"
#include "tracking_loop_filter.h"
#include <array>
#include <chrono>
#include <complex>
#include <cstring>
#include <thread>
#include <vector>

extern "C" void ref_set_arch(int arch);

using cf = std::complex<float>;

extern "C" {

// cpu_multicorrelator_real_codes.cc:36-160 — one call = resample + rotate + T dot products
int ref_mc_real_codes(int arch, int high_dyn, const float* sig_ri, int n, const float* code, int L,
    const float* shifts, int ntaps, float rem_carr, float phase_step, float phase_rate,
    float rem_code, float code_step, float code_rate, float* out_ri)
{
    ref_set_arch(arch);
    Cpu_Multicorrelator_Real_Codes mc;
    mc.init(2 * n, ntaps);
    mc.set_high_dynamics_resampler(high_dyn != 0);
    std::vector<float> sh(shifts, shifts + ntaps);
    mc.set_local_code_and_taps(L, code, sh.data());
    std::vector<cf> out(ntaps);
    mc.set_input_output_vectors(out.data(), reinterpret_cast<const cf*>(sig_ri));
    mc.Carrier_wipeoff_multicorrelator_resampler(rem_carr, phase_step, phase_rate, rem_code, code_step, code_rate, n);
    std::memcpy(out_ri, out.data(), sizeof(cf) * ntaps);
    mc.free();
    ref_set_arch(0);
    return 0;
}

// cpu_multicorrelator.cc:73-100 — complex-code sibling
int ref_mc_complex_codes(int arch, const float* sig_ri, int n, const float* code_ri, int L,
    const float* shifts, int ntaps, float rem_carr, float phase_step,
    float rem_code, float code_step, float* out_ri)
{
    ref_set_arch(arch);
    Cpu_Multicorrelator mc;
    mc.init(2 * n, ntaps);
    std::vector<float> sh(shifts, shifts + ntaps);
    mc.set_local_code_and_taps(L, reinterpret_cast<const cf*>(code_ri), sh.data());
    std::vector<cf> out(ntaps);
    mc.set_input_output_vectors(out.data(), reinterpret_cast<const cf*>(sig_ri));
    mc.Carrier_wipeoff_multicorrelator_resampler(rem_carr, phase_step, rem_code, code_step, n);
    std::memcpy(out_ri, out.data(), sizeof(cf) * ntaps);
    mc.free();
    ref_set_arch(0);
    return 0;
}

// cpu_multicorrelator_16sc.cc:64-91 — the int16 path: rotate in fp32, rintf to int16, SATURATING int16 multiply-accumulate
// (volk_gnsssdr_16ic_x2_rotator_dot_prod_16ic_xn.h:66-102); sig / code / out are interleaved int16 (re, im)
int ref_mc_16sc(const int16_t* sig_ri, int n, const int16_t* code_ri, int L, const float* shifts, int ntaps, float rem_carr, float phase_step,
    float rem_code, float code_step, int16_t* out_ri)
{
    Cpu_Multicorrelator_16sc mc;
    mc.init(2 * n, ntaps);
    std::vector<float> sh(shifts, shifts + ntaps);
    mc.set_local_code_and_taps(L, reinterpret_cast<const lv_16sc_t*>(code_ri), sh.data());
    std::vector<lv_16sc_t> out(ntaps);
    mc.set_input_output_vectors(out.data(), reinterpret_cast<const lv_16sc_t*>(sig_ri));
    mc.Carrier_wipeoff_multicorrelator_resampler(rem_carr, phase_step, rem_code, code_step, n);
    std::memcpy(out_ri, out.data(), sizeof(lv_16sc_t) * ntaps);
    mc.free();
    return 0;
}

// Throughput probe shaped like CpuMulticorrelatorRealCodesTest
// (src/tests/unit-tests/signal-processing-blocks/tracking/cpu_multicorrelator_real_codes_test.cc:54-160):
// one std::thread per "channel", each owning a Cpu_Multicorrelator_Real_Codes and
// correlating the same input block `iters` times. Returns wall seconds.
double ref_mc_real_codes_bench(int arch, int nthreads, int n, int ntaps, int L, int iters,
    const float* sig_ri, const float* code, const float* shifts,
    float phase_step, float code_step, float rem_code, float* out_ri_last)
{
    ref_set_arch(arch);
    std::vector<Cpu_Multicorrelator_Real_Codes*> mcs(nthreads);
    std::vector<std::vector<cf>> outs(nthreads, std::vector<cf>(ntaps));
    std::vector<std::vector<float>> sh(nthreads, std::vector<float>(shifts, shifts + ntaps));
    for (int t = 0; t < nthreads; t++)
        {
            mcs[t] = new Cpu_Multicorrelator_Real_Codes();
            mcs[t]->init(2 * n, ntaps);
            mcs[t]->set_high_dynamics_resampler(false);
            mcs[t]->set_local_code_and_taps(L, code, sh[t].data());
            mcs[t]->set_input_output_vectors(outs[t].data(), reinterpret_cast<const cf*>(sig_ri));
        }
    auto body = [&](int t) {
        for (int i = 0; i < iters; i++)
            {
                mcs[t]->Carrier_wipeoff_multicorrelator_resampler(0.0F, phase_step, 0.0F, rem_code, code_step, 0.0F, n);
            }
    };
    auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; t++) th.emplace_back(body, t);
    for (auto& x : th) x.join();
    auto t1 = std::chrono::steady_clock::now();
    if (out_ri_last) std::memcpy(out_ri_last, outs[0].data(), sizeof(cf) * ntaps);
    for (int t = 0; t < nthreads; t++) { mcs[t]->free(); delete mcs[t]; }
    ref_set_arch(0);
    return std::chrono::duration<double>(t1 - t0).count();
}

// ---- PRN / replica generators (a17) ----
void ref_gps_l1_ca_code_gen_float(float* dest, int prn, unsigned chip_shift)
{
    gps_l1_ca_code_gen_float(own::span<float>(dest, 1023), prn, chip_shift);
}
void ref_gps_l1_ca_code_gen_complex_sampled(float* dest_ri, int n, unsigned prn, int fs, unsigned chip_shift)
{
    gps_l1_ca_code_gen_complex_sampled(own::span<cf>(reinterpret_cast<cf*>(dest_ri), n), prn, fs, chip_shift);
}
void ref_galileo_e1_code_gen_sinboc11_float(float* dest, int n, const char* sig, unsigned prn)
{
    std::array<char, 3> id{{sig[0], sig[1], '\0'}};
    galileo_e1_code_gen_sinboc11_float(own::span<float>(dest, n), id, prn);
}
void ref_galileo_e1_code_gen_complex_sampled(float* dest_ri, int n, const char* sig, int cboc, unsigned prn, int fs, unsigned chip_shift, int secondary)
{
    std::array<char, 3> id{{sig[0], sig[1], '\0'}};
    galileo_e1_code_gen_complex_sampled(own::span<cf>(reinterpret_cast<cf*>(dest_ri), n), id, cboc != 0, prn, fs, chip_shift, secondary != 0);
}
void ref_gps_l5i_code_gen_float(float* dest, unsigned prn) { gps_l5i_code_gen_float(own::span<float>(dest, 10230), prn); }
void ref_gps_l5q_code_gen_float(float* dest, unsigned prn) { gps_l5q_code_gen_float(own::span<float>(dest, 10230), prn); }
void ref_gps_l5i_code_gen_complex_sampled(float* dest_ri, int n, unsigned prn, int fs)
{
    gps_l5i_code_gen_complex_sampled(own::span<cf>(reinterpret_cast<cf*>(dest_ri), n), prn, fs);
}
void ref_gps_l5q_code_gen_complex_sampled(float* dest_ri, int n, unsigned prn, int fs)
{
    gps_l5q_code_gen_complex_sampled(own::span<cf>(reinterpret_cast<cf*>(dest_ri), n), prn, fs);
}
void ref_galileo_e5_a_code_gen_complex_primary(float* dest_ri, int prn, const char* sig)
{
    std::array<char, 3> id{{sig[0], sig[1], '\0'}};
    galileo_e5_a_code_gen_complex_primary(own::span<cf>(reinterpret_cast<cf*>(dest_ri), 10230), prn, id);
}
void ref_galileo_e5_a_code_gen_complex_sampled(float* dest_ri, int n, unsigned prn, const char* sig, int fs, unsigned chip_shift)
{
    std::array<char, 3> id{{sig[0], sig[1], '\0'}};
    galileo_e5_a_code_gen_complex_sampled(own::span<cf>(reinterpret_cast<cf*>(dest_ri), n), prn, id, fs, chip_shift);
}

// ---- discriminators (tracking_discriminators.cc:53-149) ----
double ref_pll_cloop_two_quadrant_atan(float re, float im) { return pll_cloop_two_quadrant_atan(cf(re, im)); }
double ref_pll_four_quadrant_atan(float re, float im) { return pll_four_quadrant_atan(cf(re, im)); }
double ref_fll_four_quadrant_atan(float r1, float i1, float r2, float i2, double t1, double t2) { return fll_four_quadrant_atan(cf(r1, i1), cf(r2, i2), t1, t2); }
double ref_fll_diff_atan(float r1, float i1, float r2, float i2, double t1, double t2) { return fll_diff_atan(cf(r1, i1), cf(r2, i2), t1, t2); }
double ref_dll_nc_e_minus_l_normalized(float er, float ei, float lr, float li, float spc, float slope, float y_intercept)
{
    return dll_nc_e_minus_l_normalized(cf(er, ei), cf(lr, li), spc, slope, y_intercept);
}
double ref_dll_nc_vemlp_normalized(float ver, float vei, float er, float ei, float lr, float li, float vlr, float vli)
{
    return dll_nc_vemlp_normalized(cf(ver, vei), cf(er, ei), cf(lr, li), cf(vlr, vli));
}
double ref_sinboc_slope_abs(double x) { return CalculateSlopeAbs(&SinBocCorrelationFunction<1, 1>, x); }
double ref_sinboc_yintercept_abs(double x) { return GetYInterceptAbs(&SinBocCorrelationFunction<1, 1>, x); }

// ---- lock detectors (lock_detectors.cc:54-148) ----
float ref_cn0_m2m4_estimator(const float* prompt_ri, int length, float T) { return cn0_m2m4_estimator(reinterpret_cast<const cf*>(prompt_ri), length, T); }
float ref_cn0_svn_estimator(const float* prompt_ri, int length, float T) { return cn0_svn_estimator(reinterpret_cast<const cf*>(prompt_ri), length, T); }
float ref_carrier_lock_detector(const float* prompt_ri, int length) { return carrier_lock_detector(reinterpret_cast<const cf*>(prompt_ri), length); }

// ---- loop filters: run a whole input sequence through a fresh filter ----
void ref_tracking_loop_filter_run(float update_interval, float bw, int order, int include_last, float init, const float* in, float* out, int n)
{
    Tracking_loop_filter f(update_interval, bw, order, include_last != 0);
    f.initialize(init);
    for (int i = 0; i < n; i++) out[i] = f.apply(in[i]);
}
void ref_fll_pll_filter_run(float fll_bw, float pll_bw, int order, float init_doppler, const float* fll_in, const float* pll_in, float T, float* out, int n)
{
    Tracking_FLL_PLL_filter f;
    f.set_params(fll_bw, pll_bw, order);
    f.initialize(init_doppler);
    for (int i = 0; i < n; i++) out[i] = f.get_carrier_error(fll_in[i], pll_in[i], T);
}
void ref_exponential_smoother_run(float alpha, int init_samples, float min_value, float offset, const float* in, float* out, int n)
{
    Exponential_Smoother s;
    s.set_alpha(alpha);
    s.set_samples_for_initialization(init_samples);
    s.set_min_value(min_value);
    s.set_offset(offset);
    for (int i = 0; i < n; i++) out[i] = s.smooth(in[i]);
}

// ---- the reference's own reader of the tracking dump (src/tests/unit-tests/signal-processing-blocks/libs/tracking_dump_reader.cc) ----
// out[e][10] per epoch: abs_E, abs_P, abs_L, prompt_I, prompt_Q, PRN_start_sample_count, carrier_doppler_hz, CN0_SNV_dB_Hz, carrier_lock_test, PRN
long ref_trk_dump_read(const char* path, double* out, long max_epochs)
{
    Tracking_Dump_Reader r;
    if (!r.open_obs_file(std::string(path))) return -1;
    const long n = static_cast<long>(r.num_epochs());
    long e = 0;
    for (; e < n && e < max_epochs; e++)
        {
            if (!r.read_binary_obs()) break;
            double* o = out + 10 * e;
            o[0] = r.abs_E, o[1] = r.abs_P, o[2] = r.abs_L, o[3] = r.prompt_I, o[4] = r.prompt_Q;
            o[5] = static_cast<double>(r.PRN_start_sample_count), o[6] = r.carrier_doppler_hz, o[7] = r.CN0_SNV_dB_Hz, o[8] = r.carrier_lock_test;
            o[9] = static_cast<double>(r.PRN);
        }
    return n == e ? n : -2 - e;
}

}  // extern "C"
