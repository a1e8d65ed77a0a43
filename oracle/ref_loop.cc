// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// Closed-loop oracle: one DLL/PLL (VEML) tracking channel built from the reference's OWN objects, compiled unmodified from
// /root/reference by oracle/Makefile — Cpu_Multicorrelator_Real_Codes (volk_gnsssdr protokernels), Tracking_loop_filter,
// Tracking_FLL_PLL_filter, Exponential_Smoother, the discriminators and lock detectors.  Only the glue between them is
// written here, following dll_pll_veml_tracking.cc (start_tracking :643-883, general_work :1728-2094, do_correlation_step
// :1037-1062, run_dll_pll :1065-1152, update_tracking_vars :1189-1260, save_correlation_results :1263-1373,
// cn0_and_tracking_lock_status :972-1029, acquire_secondary :926-969); that block itself needs GNU Radio and cannot be
// compiled here.  Shape: one monolithic loop over a host sample buffer per channel (the product splits it into
// plan/finish around a batched GPU launch).  Lives in oracle/_ref/libgnssref.so; used by tests/test_closed_loop_gpu.py.
#include "cpu_multicorrelator_real_codes.h"
#include "exponential_smoother.h"
#include "lock_detectors.h"
#include "tracking_FLL_PLL_filter.h"
#include "tracking_discriminators.h"
#include "tracking_loop_filter.h"
#include "MATH_CONSTANTS.h"
#include <cmath>
#include <complex>
#include <cstdint>
#include <cstring>
#include <deque>
#include <string>
#include <vector>

extern "C" void ref_set_arch(int arch);
using cf = std::complex<float>;

extern "C" {
struct RefLoopParams
{
    double fs_in, code_period, chip_rate, carrier_freq, carrier_lock_th;
    float pll_bw, dll_bw, fll_bw, pll_bw_narrow, dll_bw_narrow, el, vel, el_narrow, vel_narrow, cn0_smoother_alpha, lock_smoother_alpha;
    int32_t vector_length, veml, secondary, track_pilot, symbols_per_bit, samples_per_chip, code_length_chips, pll_order, dll_order, extend;
    int32_t cn0_samples, cn0_min, max_code_lock_fail, max_carrier_lock_fail, cn0_smoother_samples, lock_smoother_samples;
    int32_t carrier_aiding, enable_fll_pull_in, enable_fll_steady, interchange_iq, arch, correlation_length_ms;
    uint32_t pull_in_time_s, bit_sync_time_limit_s, secondary_code_length, data_secondary_code_length;
    char secondary_code[256], data_secondary_code[256];
};
struct RefLoopRecord  // same layout as gb200h_trk_record
{
    int32_t channel;
    uint32_t prn;
    int32_t valid, correlation_length_ms;
    double prompt_i, prompt_q, cn0_db_hz, carrier_doppler_hz, carrier_phase_rads, code_phase_samples;
    uint64_t tracking_sample_counter;
};
}

namespace
{
struct Loop
{
    RefLoopParams p;
    uint32_t prn;
    Cpu_Multicorrelator_Real_Codes mc, mc_data;
    Tracking_loop_filter dll;
    Tracking_FLL_PLL_filter pll;
    Exponential_Smoother cn0_s, lock_s;
    std::vector<float> code, data_code, shifts;
    std::vector<cf> outs, prompt_buf;
    std::deque<cf> circ;
    std::vector<cf> pending;  // unread samples, pending[0] has absolute index `read`
    uint64_t read, acq_stamp;
    cf prompt_data{0, 0}, VE{0, 0}, E{0, 0}, P{0, 0}, P_old{0, 0}, L{0, 0}, VL{0, 0}, Pd{0, 0};
    double acq_code_phase, doppler, phase_step, code_freq, code_step, rem_code_samples = 0, rem_code_chips = 0, acc_phase = 0;
    double corr_time, cn0 = 0, lock_test = 1, spc, slope = 1, yint = 1, code_err_filt = 0;
    float rem_carr = 0;
    int state = 1, n_taps, prn_len, ext_count = 0, cn0_count = 0, carr_fail = 0, code_fail = 0, iE, iP, iL;
    uint32_t cur_sym = 0, cur_data_sym = 0;
    bool pull_in = true, cloop = true, acc_init = false, flag180 = false;

    void taps(float el, float vel)
    {
        const float s = static_cast<float>(p.samples_per_chip);
        shifts.resize(p.veml ? 5 : 3);  // the multicorrelator borrows shifts.data(): written in place, never reallocated
        if (p.veml)
            {
                shifts[0] = -vel * s;
                shifts[1] = -el * s;
                shifts[2] = 0.0F;
                shifts[3] = el * s;
                shifts[4] = vel * s;
            }
        else
            {
                shifts[0] = -el * s;
                shifts[1] = 0.0F;
                shifts[2] = el * s;
            }
    }
};
}  // namespace

extern "C" {

void* ref_loop_new(const RefLoopParams* params, const float* tracking_code, const float* data_code, int n_entries, unsigned prn,
    double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp, uint64_t nitems_read)
{
    auto* k = new Loop;
    k->p = *params;
    const RefLoopParams& p = k->p;
    k->prn = prn;
    k->n_taps = p.veml ? 5 : 3;
    k->iE = p.veml ? 1 : 0;
    k->iP = p.veml ? 2 : 1;
    k->iL = p.veml ? 3 : 2;
    k->code.assign(tracking_code, tracking_code + n_entries);
    if (p.track_pilot) k->data_code.assign(data_code, data_code + n_entries);
    k->outs.assign(k->n_taps, cf(0, 0));
    k->prompt_buf.assign(p.cn0_samples > 0 ? p.cn0_samples : 1, cf(0, 0));
    k->taps(p.el, p.vel);
    // constructor (:462-547)
    k->dll = Tracking_loop_filter(static_cast<float>(p.code_period), p.dll_bw, p.dll_order, false);
    k->pll.set_params(p.fll_bw, p.pll_bw, p.pll_order);
    k->mc.init(2 * p.vector_length, k->n_taps);
    k->mc.set_high_dynamics_resampler(false);  // :536 with Tracking_XX.high_dyn=false (the class default is true)
    if (p.track_pilot)
        {
            k->mc_data.init(2 * p.vector_length, 1);
            k->mc_data.set_high_dynamics_resampler(false);  // :530
        }
    k->cn0_s.set_alpha(p.cn0_smoother_alpha);
    k->cn0_s.set_samples_for_initialization(p.cn0_smoother_samples / static_cast<int>(p.code_period * 1000.0));
    k->lock_s.set_alpha(p.lock_smoother_alpha);
    k->lock_s.set_min_value(-1.0);
    k->lock_s.set_offset(0.0);
    k->lock_s.set_samples_for_initialization(p.lock_smoother_samples);
    k->spc = p.el;
    if (p.veml)
        {
            k->slope = static_cast<float>(-CalculateSlopeAbs(&SinBocCorrelationFunction<1, 1>, p.el));
            k->yint = static_cast<float>(GetYInterceptAbs(&SinBocCorrelationFunction<1, 1>, p.el));
        }
    // start_tracking (:643-883)
    k->acq_code_phase = acq_delay_samples;
    k->acq_stamp = acq_samplestamp;
    k->doppler = acq_doppler_hz;
    k->phase_step = TWO_PI * k->doppler / p.fs_in;
    k->mc.set_local_code_and_taps(p.samples_per_chip * p.code_length_chips, k->code.data(), k->shifts.data());
    if (p.track_pilot) k->mc_data.set_local_code_and_taps(p.samples_per_chip * p.code_length_chips, k->data_code.data(), &k->shifts[k->iP]);
    k->corr_time = p.code_period;
    k->code_freq = p.chip_rate;
    k->code_step = 0.0;
    k->pll.initialize(static_cast<float>(acq_doppler_hz));
    k->dll.initialize();
    k->read = nitems_read;
    k->prn_len = p.vector_length;
    return k;
}

void ref_loop_free(void* h)
{
    auto* k = static_cast<Loop*>(h);
    if (!k) return;
    k->mc.free();
    if (k->p.track_pilot) k->mc_data.free();
    delete k;
}

void ref_loop_state(void* h, double* v)
{
    auto* k = static_cast<Loop*>(h);
    v[0] = k->state;
    v[1] = k->doppler;
    v[2] = k->cn0;
    v[3] = k->lock_test;
    v[4] = k->code_freq;
    v[5] = k->rem_code_samples;
    v[6] = static_cast<double>(k->read);
    v[7] = k->prn_len;
}
}  // extern "C"

namespace
{
bool lock_status(Loop& k, double coh_time)  // :972-1029
{
    const RefLoopParams& p = k.p;
    if (k.cn0_count < p.cn0_samples)
        {
            k.prompt_buf[k.cn0_count++] = k.P;
            return true;
        }
    k.prompt_buf[k.cn0_count % p.cn0_samples] = k.P;
    k.cn0_count++;
    const float raw = cn0_m2m4_estimator(k.prompt_buf.data(), p.cn0_samples, static_cast<float>(coh_time));
    k.cn0 = k.cn0_s.smooth(raw);
    k.lock_test = k.lock_s.smooth(carrier_lock_detector(k.prompt_buf.data(), 1));
    if (!k.pull_in)
        {
            if (k.lock_test < p.carrier_lock_th)
                k.carr_fail++;
            else if (k.carr_fail > 0)
                k.carr_fail--;
            if (k.cn0 < p.cn0_min)
                k.code_fail++;
            else if (k.code_fail > 0)
                k.code_fail--;
        }
    if (k.carr_fail > p.max_carrier_lock_fail || k.code_fail > p.max_code_lock_fail)
        {
            k.carr_fail = k.code_fail = 0;
            return false;
        }
    return true;
}

void dll_pll(Loop& k)  // :1065-1152 (without the experimental Doppler correction)
{
    const RefLoopParams& p = k.p;
    const double ph_err = (k.cloop ? pll_cloop_two_quadrant_atan(k.P) : pll_four_quadrant_atan(k.P)) / TWO_PI;
    double filt;
    if ((k.pull_in && p.enable_fll_pull_in) || p.enable_fll_steady)
        {
            const double f_err = fll_diff_atan(k.P_old, k.P, 0, k.corr_time) / TWO_PI;
            k.P_old = k.P;
            if (k.pull_in && p.enable_fll_pull_in)
                filt = k.pll.get_carrier_error(static_cast<float>(f_err), 0.0F, static_cast<float>(k.corr_time));
            else
                filt = k.pll.get_carrier_error(static_cast<float>(f_err), static_cast<float>(ph_err), static_cast<float>(k.corr_time));
        }
    else
        {
            filt = k.pll.get_carrier_error(0, static_cast<float>(ph_err), static_cast<float>(k.corr_time));
        }
    k.doppler = filt;
    double code_err;
    if (p.veml)
        code_err = dll_nc_vemlp_normalized(k.VE, k.E, k.L, k.VL);
    else
        code_err = dll_nc_e_minus_l_normalized(k.E, k.L, static_cast<float>(k.spc), static_cast<float>(k.slope), static_cast<float>(k.yint));
    k.code_err_filt = k.dll.apply(static_cast<float>(code_err));
    k.code_freq = p.chip_rate - k.code_err_filt;
    if (p.carrier_aiding) k.code_freq += k.doppler * p.chip_rate / p.carrier_freq;
}

void update_vars(Loop& k)  // :1189-1260, high_dyn off
{
    const RefLoopParams& p = k.p;
    const double T_prn_samples = (1.0 / k.code_freq) * static_cast<double>(p.code_length_chips) * p.fs_in;
    const double K_blk = T_prn_samples + k.rem_code_samples;
    k.prn_len = static_cast<int32_t>(std::floor(K_blk));
    k.phase_step = TWO_PI * k.doppler / p.fs_in;
    k.rem_carr += static_cast<float>(k.phase_step * static_cast<double>(k.prn_len));
    k.rem_carr = fmod(k.rem_carr, TWO_PI);
    k.acc_phase -= k.phase_step * static_cast<double>(k.prn_len);
    k.code_step = k.code_freq / p.fs_in;
    k.rem_code_samples = K_blk - static_cast<double>(k.prn_len);
    k.rem_code_chips = k.code_freq * k.rem_code_samples / p.fs_in;
}

void save_results(Loop& k)  // :1263-1373
{
    const RefLoopParams& p = k.p;
    const bool neg = p.secondary && p.secondary_code[k.cur_sym] != '0';
    if (neg)
        {
            if (p.veml)
                {
                    k.VE -= k.outs[0];
                    k.VL -= k.outs[4];
                }
            k.E -= k.outs[k.iE];
            k.P -= k.outs[k.iP];
            k.L -= k.outs[k.iL];
        }
    else
        {
            if (p.veml)
                {
                    k.VE += k.outs[0];
                    k.VL += k.outs[4];
                }
            k.E += k.outs[k.iE];
            k.P += k.outs[k.iP];
            k.L += k.outs[k.iL];
        }
    if (p.secondary)
        {
            k.cur_sym++;
            k.cur_sym %= p.secondary_code_length;
        }
    const cf dp = p.track_pilot ? k.prompt_data : k.outs[k.iP];
    if (p.symbols_per_bit > 1)
        {
            if (p.data_secondary_code_length > 0)
                {
                    if (p.data_secondary_code[k.cur_data_sym] == '0')
                        k.Pd += dp;
                    else
                        k.Pd -= dp;
                    k.cur_data_sym++;
                    k.cur_data_sym %= p.data_secondary_code_length;
                }
            else
                {
                    k.Pd += dp;
                    k.cur_data_sym++;
                    k.cur_data_sym %= p.symbols_per_bit;
                }
        }
    else
        {
            k.Pd = dp;
        }
    k.cloop = !p.track_pilot;
}

bool sync_secondary(Loop& k)  // :926-969
{
    const RefLoopParams& p = k.p;
    int32_t corr = 0;
    for (uint32_t i = 0; i < p.secondary_code_length; i++)
        {
            if (k.circ[i].real() < 0.0)
                corr += (p.secondary_code[i] == '0') ? 1 : -1;
            else
                corr += (p.secondary_code[i] == '0') ? -1 : 1;
        }
    if (std::abs(corr) == static_cast<int32_t>(p.secondary_code_length))
        {
            k.flag180 = corr < 0;
            return true;
        }
    return false;
}

void correlate(Loop& k, const cf* in)  // :1037-1062
{
    const RefLoopParams& p = k.p;
    const float s = static_cast<float>(p.samples_per_chip);
    k.mc.set_input_output_vectors(k.outs.data(), in);
    k.mc.Carrier_wipeoff_multicorrelator_resampler(k.rem_carr, static_cast<float>(k.phase_step), 0.0F, static_cast<float>(k.rem_code_chips) * s,
        static_cast<float>(k.code_step) * s, 0.0F, p.vector_length);
    if (p.track_pilot)
        {
            k.mc_data.set_input_output_vectors(&k.prompt_data, in);
            k.mc_data.Carrier_wipeoff_multicorrelator_resampler(k.rem_carr, static_cast<float>(k.phase_step), 0.0F, static_cast<float>(k.rem_code_chips) * s,
                static_cast<float>(k.code_step) * s, 0.0F, p.vector_length);
        }
}

void fill(const Loop& k, RefLoopRecord* r, bool valid)
{
    const RefLoopParams& p = k.p;
    r->channel = 0;
    r->prn = k.prn;
    r->valid = valid;
    r->correlation_length_ms = p.correlation_length_ms;
    r->prompt_i = p.interchange_iq ? k.Pd.imag() : k.Pd.real();
    r->prompt_q = p.interchange_iq ? k.Pd.real() : k.Pd.imag();
    r->cn0_db_hz = k.cn0;
    r->carrier_doppler_hz = k.doppler;
    r->carrier_phase_rads = k.acc_phase;
    r->code_phase_samples = k.rem_code_samples;
}
}  // namespace

extern "C" int ref_loop_feed(void* h, const float* sig_ri, int64_t n, RefLoopRecord* out, int max_out)
{
    auto* kp = static_cast<Loop*>(h);
    Loop& k = *kp;
    const RefLoopParams& p = k.p;
    ref_set_arch(p.arch);
    const cf* sig = reinterpret_cast<const cf*>(sig_ri);
    k.pending.insert(k.pending.end(), sig, sig + n);
    size_t pos = 0;  // consumed prefix of pending
    int n_out = 0;
    for (;;)
        {
            if (k.state == 0)  // standby: consume everything
                {
                    k.read += k.pending.size() - pos;
                    pos = k.pending.size();
                    break;
                }
            if (k.pending.size() - pos < static_cast<size_t>(2 * p.vector_length)) break;  // forecast (:607-614)
            const cf* in = k.pending.data() + pos;
            if (k.pull_in && p.pull_in_time_s < (k.read - k.acq_stamp) / static_cast<int>(p.fs_in))
                {
                    k.pull_in = false;
                    k.carr_fail = k.code_fail = 0;
                }
            bool produced = false, lost = false, narrow_done = false;
            if (k.state == 1)  // pull-in alignment (:1757-1789)
                {
                    const int64_t diff = static_cast<int64_t>(k.read) - static_cast<int64_t>(k.acq_stamp);
                    const double delta = static_cast<double>(diff) - k.acq_code_phase;
                    k.code_freq = p.chip_rate;
                    k.code_step = k.code_freq / p.fs_in;
                    const double T_prn = (1.0 / k.code_freq) * static_cast<double>(p.code_length_chips) * p.fs_in;
                    k.acq_code_phase = T_prn - std::fmod(delta, T_prn);
                    k.prn_len = round(T_prn);
                    const int32_t offset = round(k.acq_code_phase);
                    k.acc_phase -= k.phase_step * static_cast<double>(offset);
                    k.state = 2;
                    k.cn0_s.reset();
                    k.lock_s.reset();
                    pos += offset;
                    k.read += offset;
                    continue;
                }
            correlate(k, in);
            if (k.state == 2)  // :1790-1914
                {
                    if (p.veml)
                        {
                            k.VE = k.outs[0];
                            k.VL = k.outs[4];
                        }
                    k.E = k.outs[k.iE];
                    k.P = k.outs[k.iP];
                    k.L = k.outs[k.iL];
                    k.spc = p.el;
                    if (p.bit_sync_time_limit_s < (k.read - k.acq_stamp) / static_cast<int>(p.fs_in)) k.carr_fail = 300000;
                    if (!lock_status(k, p.code_period))
                        {
                            k.state = 0;
                            lost = true;
                        }
                    else
                        {
                            dll_pll(k);
                            update_vars(k);
                            bool next = false;
                            if (!k.pull_in)
                                {
                                    if (p.secondary || p.symbols_per_bit > 1)
                                        {
                                            k.circ.push_back(k.outs[k.iP]);
                                            if (k.circ.size() > p.secondary_code_length) k.circ.pop_front();
                                            if (k.circ.size() == p.secondary_code_length) next = sync_secondary(k);
                                        }
                                    else
                                        next = true;
                                }
                            if (next)
                                {
                                    k.VE = k.E = k.P = k.Pd = k.L = k.VL = cf(0, 0);
                                    k.circ.clear();
                                    k.cur_sym = k.cur_data_sym = 0;
                                    if (p.extend > 1)
                                        {
                                            k.ext_count = 0;
                                            k.corr_time = static_cast<float>(p.extend) * static_cast<float>(p.code_period);
                                            k.state = 3;
                                            k.dll.set_update_interval(static_cast<float>(k.corr_time));
                                            k.dll.set_noise_bandwidth(p.dll_bw_narrow);
                                            k.pll.set_params(p.fll_bw, p.pll_bw_narrow, p.pll_order);
                                            k.taps(p.el_narrow, p.vel_narrow);
                                            k.spc = p.el_narrow;
                                        }
                                    else
                                        k.state = 4;
                                }
                        }
                }
            else if (k.state == 3)  // :1915-1956
                {
                    save_results(k);
                    update_vars(k);
                    if (k.cur_data_sym == 0) produced = true;
                    k.ext_count++;
                    if (k.ext_count == p.extend - 1)
                        {
                            k.ext_count = 0;
                            k.state = 4;
                        }
                }
            else  // state 4 (:1957-2020)
                {
                    save_results(k);
                    if (!lock_status(k, p.code_period * static_cast<double>(p.extend)))
                        {
                            k.state = 0;
                            lost = true;
                        }
                    else
                        {
                            dll_pll(k);
                            update_vars(k);
                            if (!k.acc_init)
                                {
                                    k.acc_phase = -k.rem_carr;
                                    k.acc_init = true;
                                }
                            if (k.cur_data_sym == 0) produced = true;
                            narrow_done = true;
                        }
                }
            if ((produced || lost) && n_out < max_out)
                {
                    fill(k, &out[n_out], !lost);
                    out[n_out].tracking_sample_counter = k.read + static_cast<uint64_t>(k.prn_len);
                    out[n_out].carrier_phase_rads = k.acc_phase;
                    n_out++;
                }
            if (produced) k.Pd = cf(0, 0);
            if (narrow_done)  // reset of the extended accumulators after a narrow-tracking update (:2011-2019)
                {
                    k.VE = k.E = k.P = k.L = k.VL = cf(0, 0);
                    if (p.extend > 1) k.state = 3;
                }
            pos += k.prn_len;
            k.read += k.prn_len;
        }
    k.pending.erase(k.pending.begin(), k.pending.begin() + pos);
    ref_set_arch(0);
    return n_out;
}
