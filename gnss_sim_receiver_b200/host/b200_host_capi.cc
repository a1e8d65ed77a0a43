#include "b200_host_capi.h"
#include "b200_synchro_monitor.h"
#include "b200_adapters.h"
#include "b200_channel_manager.h"
#include "b200_mat_writer.h"
#include "b200_configuration.h"
#include "b200_gnss_codes.h"
#include "b200_gr_blocks.h"
#include "b200_tracking_bank.h"
#include <algorithm>
#include <cstring>
#include <sstream>
#include <thread>

using namespace gb200h;

namespace
{
thread_local std::string t_err, t_ret;
int fail(const std::exception& e)
{
    t_err = e.what();
    return -1;
}
cf32 c(float re, float im) { return cf32(re, im); }
}  // namespace

extern "C" {

const char* gb200h_last_error(void) { return t_err.c_str(); }

const char* gb200h_conf_get(const char* conf_text, const char* name, const char* def)
{
    B200Configuration cfg;
    cfg.load_string(conf_text);
    t_ret = cfg.property(std::string(name), std::string(def));
    return t_ret.c_str();
}

const char* gb200h_dll_pll_conf(const char* conf_text, const char* role)
{
    B200Configuration cfg;
    cfg.load_string(conf_text);
    Dll_Pll_Conf p;
    p.SetFromConfiguration(&cfg, role);
    std::ostringstream o;
    o.precision(9);
    o << "item_type=" << p.item_type << ";fs_in=" << p.fs_in << ";pll_bw_hz=" << p.pll_bw_hz << ";dll_bw_hz=" << p.dll_bw_hz << ";pll_bw_narrow_hz=" << p.pll_bw_narrow_hz
      << ";dll_bw_narrow_hz=" << p.dll_bw_narrow_hz << ";fll_bw_hz=" << p.fll_bw_hz << ";early_late_space_chips=" << p.early_late_space_chips
      << ";very_early_late_space_chips=" << p.very_early_late_space_chips << ";early_late_space_narrow_chips=" << p.early_late_space_narrow_chips
      << ";very_early_late_space_narrow_chips=" << p.very_early_late_space_narrow_chips << ";pll_filter_order=" << p.pll_filter_order << ";dll_filter_order=" << p.dll_filter_order
      << ";fll_filter_order=" << p.fll_filter_order << ";extend_correlation_symbols=" << p.extend_correlation_symbols << ";cn0_samples=" << p.cn0_samples << ";cn0_min=" << p.cn0_min
      << ";max_code_lock_fail=" << p.max_code_lock_fail << ";max_carrier_lock_fail=" << p.max_carrier_lock_fail << ";carrier_lock_th=" << p.carrier_lock_th
      << ";pull_in_time_s=" << p.pull_in_time_s << ";bit_synchronization_time_limit_s=" << p.bit_synchronization_time_limit_s << ";track_pilot=" << p.track_pilot
      << ";high_dyn=" << p.high_dyn << ";carrier_aiding=" << p.carrier_aiding << ";enable_fll_pull_in=" << p.enable_fll_pull_in << ";smoother_length=" << p.smoother_length
      << ";cn0_smoother_samples=" << p.cn0_smoother_samples << ";cn0_smoother_alpha=" << p.cn0_smoother_alpha << ";dump=" << p.dump << ";device=" << p.device << ";";
    t_ret = o.str();
    return t_ret.c_str();
}

const char* gb200h_acq_conf(const char* conf_text, const char* role, double chip_rate, double opt_freq, int ms_per_code)
{
    B200Configuration cfg;
    cfg.load_string(conf_text);
    Acq_Conf p;
    p.ms_per_code = static_cast<uint32_t>(ms_per_code);
    try
        {
            p.SetFromConfiguration(&cfg, role, chip_rate, opt_freq);
        }
    catch (const std::invalid_argument& e)
        {
            t_ret = std::string("invalid_argument=") + e.what() + ";";
            return t_ret.c_str();
        }
    std::ostringstream o;
    o.precision(9);
    o << "item_type=" << p.item_type << ";it_size=" << p.it_size << ";fs_in=" << p.fs_in << ";resampled_fs=" << p.resampled_fs << ";doppler_max=" << p.doppler_max
      << ";doppler_step=" << p.doppler_step << ";sampled_ms=" << p.sampled_ms << ";max_dwells=" << p.max_dwells << ";pfa=" << p.pfa << ";pfa2=" << p.pfa2
      << ";bit_transition_flag=" << p.bit_transition_flag << ";use_CFAR_algorithm_flag=" << p.use_CFAR_algorithm_flag << ";samples_per_ms=" << p.samples_per_ms
      << ";samples_per_chip=" << p.samples_per_chip << ";samples_per_code=" << p.samples_per_code << ";resampler_ratio=" << p.resampler_ratio << ";blocking=" << p.blocking
      << ";make_2_steps=" << p.make_2_steps << ";";
    t_ret = o.str();
    return t_ret.c_str();
}

const char* gb200h_factory(const char* conf_text, const char* role)
{
    B200Configuration cfg;
    cfg.load_string(conf_text);
    t_ret.clear();
    try
        {
            // construct the tracking adapters only: the acquisition adapter needs a GPU at construction
            const std::string impl = cfg.property(std::string(role) + ".implementation", std::string(""));
            if (auto t = GetB200TrkBlock(&cfg, role, 1, 1))
                t_ret = t->implementation() + ";item_size=" + std::to_string(t->item_size()) + ";role=" + t->role() + ";";
            else if (impl == "GPS_L1_CA_PCPS_Acquisition_B200" || impl == "Galileo_E1_PCPS_Ambiguous_Acquisition_B200" || impl == "GPS_L5i_PCPS_Acquisition_B200" ||
                     impl == "Galileo_E5a_Pcps_Acquisition_B200")
                t_ret = impl + ";acquisition;";
        }
    catch (const std::exception& e)
        {
            t_ret = std::string("error=") + e.what();
        }
    return t_ret.c_str();
}

// Two tracking adapters built from one configuration and started the way Channel does (set_gnss_synchro, start_tracking): which
// correlator engine each one attached to.  "same=<0|1>;stream_a=..;stream_b=..;slot_a=..;slot_b=..;max_block=..;" or "error=...".
const char* gb200h_engine_probe(const char* conf_text, const char* role_a, unsigned prn_a, const char* role_b, unsigned prn_b)
{
    B200Configuration cfg;
    cfg.load_string(conf_text);
    t_ret.clear();
    try
        {
            B200CorrelatorEngine::reset_instances();
            auto a = GetB200TrkBlock(&cfg, role_a, 1, 1);
            auto b = GetB200TrkBlock(&cfg, role_b, 1, 1);
            if (!a || !b) throw std::runtime_error("not a B200 tracking implementation");
            Gnss_Synchro sa{}, sb{};
            auto* ta = static_cast<B200DllPllTracking*>(a.get());
            auto* tb = static_cast<B200DllPllTracking*>(b.get());
            sa.System = ta->conf().system;
            std::memcpy(sa.Signal, ta->conf().signal, 3);
            sa.PRN = prn_a;
            sb.System = tb->conf().system;
            std::memcpy(sb.Signal, tb->conf().signal, 3);
            sb.PRN = prn_b;
            a->set_gnss_synchro(&sa);
            b->set_gnss_synchro(&sb);
            a->set_channel(0);
            b->set_channel(1);
            a->start_tracking();
            b->start_tracking();
            auto ea = ta->block()->engine(), eb = tb->block()->engine();
            if (!ea || !eb) throw std::runtime_error("a block did not attach to an engine");
            std::ostringstream o;
            o << "same=" << (ea == eb ? 1 : 0) << ";stream_a=" << ea->config().stream_id << ";stream_b=" << eb->config().stream_id << ";slot_a=" << ta->block()->engine_channel()
              << ";slot_b=" << tb->block()->engine_channel() << ";max_block=" << ea->config().max_block << ";max_block_b=" << eb->config().max_block << ";";
            t_ret = o.str();
            a.reset();
            b.reset();
            B200CorrelatorEngine::reset_instances();
        }
    catch (const std::exception& e)
        {
            B200CorrelatorEngine::reset_instances();
            t_ret = std::string("error=") + e.what();
        }
    return t_ret.c_str();
}

double gb200h_pll_cloop_two_quadrant_atan(float re, float im) { return pll_cloop_two_quadrant_atan(c(re, im)); }
double gb200h_pll_four_quadrant_atan(float re, float im) { return pll_four_quadrant_atan(c(re, im)); }
double gb200h_fll_four_quadrant_atan(float r1, float i1, float r2, float i2, double t1, double t2) { return fll_four_quadrant_atan(c(r1, i1), c(r2, i2), t1, t2); }
double gb200h_fll_diff_atan(float r1, float i1, float r2, float i2, double t1, double t2) { return fll_diff_atan(c(r1, i1), c(r2, i2), t1, t2); }
double gb200h_dll_nc_e_minus_l_normalized(float er, float ei, float lr, float li, float spc, float slope, float y_intercept)
{
    return dll_nc_e_minus_l_normalized(c(er, ei), c(lr, li), spc, slope, y_intercept);
}
double gb200h_dll_nc_vemlp_normalized(float ver, float vei, float er, float ei, float lr, float li, float vlr, float vli)
{
    return dll_nc_vemlp_normalized(c(ver, vei), c(er, ei), c(lr, li), c(vlr, vli));
}
int gb200h_e5a_q_secondary_code(unsigned prn, char* out)
{
    if (!out || prn < 1 || prn > 50) return 0;
    const std::string s = gb200h::galileo_e5a_q_secondary_code(prn);
    std::memcpy(out, s.c_str(), s.size() + 1);
    return static_cast<int>(s.size());
}

double gb200h_sinboc_slope_abs(double x) { return sinboc11_slope_abs(x); }
double gb200h_sinboc_yintercept_abs(double x) { return sinboc11_y_intercept_abs(x); }
float gb200h_cn0_m2m4_estimator(const float* p, int length, float T) { return cn0_m2m4_estimator(reinterpret_cast<const cf32*>(p), length, T); }
float gb200h_cn0_svn_estimator(const float* p, int length, float T) { return cn0_svn_estimator(reinterpret_cast<const cf32*>(p), length, T); }
float gb200h_carrier_lock_detector(const float* p, int length) { return carrier_lock_detector(reinterpret_cast<const cf32*>(p), length); }

void gb200h_tracking_loop_filter_run(float update_interval, float bw, int order, int include_last, float init, const float* in, float* out, int n)
{
    Tracking_loop_filter f(update_interval, bw, order, include_last != 0);
    f.initialize(init);
    for (int i = 0; i < n; i++) out[i] = f.apply(in[i]);
}

void gb200h_fll_pll_filter_run(float fll_bw, float pll_bw, int order, float init_doppler, const float* fll_in, const float* pll_in, float T, float* out, int n)
{
    Tracking_FLL_PLL_filter f;
    f.set_params(fll_bw, pll_bw, order);
    f.initialize(init_doppler);
    for (int i = 0; i < n; i++) out[i] = f.get_carrier_error(fll_in[i], pll_in[i], T);
}

void gb200h_exponential_smoother_run(float alpha, int init_samples, float min_value, float offset, const float* in, float* out, int n)
{
    Exponential_Smoother s;
    s.set_alpha(alpha);
    s.set_samples_for_initialization(init_samples);
    s.set_min_value(min_value);
    s.set_offset(offset);
    for (int i = 0; i < n; i++) out[i] = s.smooth(in[i]);
}

int gb200h_gps_l1_ca_code_gen_float(float* dest, unsigned prn, unsigned chip_shift)
{
    try
        {
            gps_l1_ca_code_gen_float(dest, prn, chip_shift);
            return 0;
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

int gb200h_gps_l1_ca_code_gen_complex_sampled(float* dest_ri, int n, unsigned prn, int fs, unsigned chip_shift)
{
    try
        {
            gps_l1_ca_code_gen_complex_sampled(reinterpret_cast<std::complex<float>*>(dest_ri), n, prn, fs, chip_shift);
            return 0;
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

// one door for the other signals' generators (a17); what: 0 E1 sinboc11 float(8184), 1 E1 complex sampled, 2 L5I float, 3 L5Q float,
// 4 L5I complex sampled, 5 L5Q complex sampled, 6 E5a complex primary, 7 E5a complex sampled.  Returns the samples written or < 0.
int gb200h_code_gen(int what, float* dest, const char* signal_id, unsigned prn, int fs, unsigned chip_shift, int cboc, int secondary)
{
    try
        {
            auto* c = reinterpret_cast<std::complex<float>*>(dest);
            switch (what)
                {
                case 0: galileo_e1_code_gen_sinboc11_float(dest, signal_id, prn); return 8184;
                case 1: return galileo_e1_code_gen_complex_sampled(c, signal_id, cboc != 0, prn, fs, chip_shift, secondary != 0);
                case 2: gps_l5i_code_gen_float(dest, prn); return 10230;
                case 3: gps_l5q_code_gen_float(dest, prn); return 10230;
                case 4: return gps_l5i_code_gen_complex_sampled(c, prn, fs);
                case 5: return gps_l5q_code_gen_complex_sampled(c, prn, fs);
                case 6: galileo_e5_a_code_gen_complex_primary(c, static_cast<int32_t>(prn), signal_id); return 10230;
                case 7: return galileo_e5_a_code_gen_complex_sampled(c, prn, signal_id, fs, chip_shift);
                default: throw std::invalid_argument("gb200h_code_gen: unknown generator");
                }
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

}  // extern "C"

// ------------------------------------------------------------------ tracking bank
struct gb200h_bank
{
    B200Configuration cfg;
    std::string role;
    std::shared_ptr<B200CorrelatorEngine> engine;
    std::unique_ptr<B200TrackingBank> bank;
    std::vector<std::unique_ptr<Gnss_Synchro>> synchros;
    std::vector<B200TrackingBank::Output> outs;
    std::unique_ptr<B200SynchroMonitor> monitor;  // TrackingMonitor.* of the bank's configuration
};

namespace
{
void to_record(int ch, const Gnss_Synchro& s, gb200h_trk_record* r)
{
    r->channel = ch;
    r->prn = s.PRN;
    r->valid = s.Flag_valid_symbol_output;
    r->correlation_length_ms = s.correlation_length_ms;
    r->prompt_i = s.Prompt_I;
    r->prompt_q = s.Prompt_Q;
    r->cn0_db_hz = s.CN0_dB_hz;
    r->carrier_doppler_hz = s.Carrier_Doppler_hz;
    r->carrier_phase_rads = s.Carrier_phase_rads;
    r->code_phase_samples = s.Code_phase_samples;
    r->tracking_sample_counter = s.Tracking_sample_counter;
}
}  // namespace

extern "C" {

int gb200h_bank_create(const char* conf_text, const char* role, int n_channels, int max_engine_channels, gb200h_bank** out)
{
    try
        {
            auto b = std::make_unique<gb200h_bank>();
            b->cfg.load_string(conf_text);
            b->role = role;
            for (int i = 0; i < n_channels; i++)
                {
                    auto adapter = GetB200TrkBlock(&b->cfg, role, 1, 1);
                    if (!adapter) throw std::runtime_error(std::string(role) + ".implementation is not a B200 tracking block");
                    auto* a = static_cast<B200DllPllTracking*>(adapter.get());
                    if (a->item_size() == 0) throw std::runtime_error("unsupported tracking item type");
                    if (!b->engine)
                        {
                            EngineConfig ec;
                            ec.device = a->conf().device;
                            ec.iq_dtype = item_type_iq(a->conf().item_type);
                            ec.max_channels = std::max(max_engine_channels, 4 * n_channels);  // bank slots + facade slots, pilot + data
                            ec.max_code_len = 2 * a->block()->code_table_entries();
                            ec.max_block = static_cast<int>(2 * a->conf().vector_length);
                            ec.ring_samples = std::max<int64_t>(int64_t{1} << 21, 64LL * a->conf().vector_length);
                            b->engine = std::make_shared<B200CorrelatorEngine>(ec);  // a private engine per bank
                            b->bank = std::make_unique<B200TrackingBank>(b->engine);
                        }
                    b->synchros.push_back(std::make_unique<Gnss_Synchro>());
                    auto blk = std::make_unique<b200_dll_pll_veml_tracking>(a->conf());
                    blk->set_gnss_synchro(b->synchros.back().get());
                    blk->set_channel(static_cast<uint32_t>(i));
                    b->bank->add_channel(std::move(blk));
                }
            b->monitor = std::make_unique<B200SynchroMonitor>(&b->cfg, "TrackingMonitor", n_channels);
            *out = b.release();
            return 0;
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

void gb200h_bank_destroy(gb200h_bank* b) { delete b; }

int gb200h_bank_set_codes(gb200h_bank* b, int ch, const float* tracking_code, const float* data_code, int n_entries)
{
    try
        {
            if (tracking_code) b->bank->block(ch).set_tracking_code(tracking_code, n_entries);
            if (data_code) b->bank->block(ch).set_data_code(data_code, n_entries);
            return 0;
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

int gb200h_bank_start(gb200h_bank* b, int ch, unsigned prn, double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp, uint64_t nitems_read)
{
    try
        {
            Gnss_Synchro& s = *b->synchros[ch];
            const Dll_Pll_Conf& p = b->bank->block(ch).parameters();
            s.System = p.system;
            std::memcpy(s.Signal, p.signal, 3);
            s.PRN = prn;
            s.Channel_ID = ch;
            s.Acq_delay_samples = acq_delay_samples;
            s.Acq_doppler_hz = acq_doppler_hz;
            s.Acq_samplestamp_samples = acq_samplestamp;
            s.Flag_valid_acquisition = true;
            b->bank->start_tracking(ch, nitems_read);
            return 0;
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

int gb200h_bank_push(gb200h_bank* b, const void* iq, int64_t n_samples)
{
    try
        {
            b->bank->push(iq, n_samples);
            return 0;
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

int64_t gb200h_bank_push_raw(gb200h_bank* b, const void* raw, int64_t n_bytes, int kind, int sample_type, int item_size, int big_endian_items, int big_endian_bytes)
{
    try
        {
            const gb200_ingest_format fmt{kind, sample_type, item_size, big_endian_items, big_endian_bytes};
            return b->bank->push_raw(raw, n_bytes, fmt);
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

int gb200h_bank_step(gb200h_bank* b, gb200h_trk_record* records, int max_records)
{
    try
        {
            b->outs.clear();
            b->bank->step(&b->outs);
            if (static_cast<int>(b->outs.size()) > max_records) throw std::runtime_error("gb200h_bank_step: " + std::to_string(b->outs.size()) + " records do not fit the caller's buffer");
            const int n = static_cast<int>(b->outs.size());
            for (int i = 0; i < n; i++) to_record(b->outs[i].channel, b->outs[i].synchro, &records[i]);
            if (b->monitor->enabled())
                for (int i = 0; i < n; i++) b->monitor->write(b->outs[i].channel, &b->outs[i].synchro, 1);
            return n;
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

int gb200h_bank_channel_state(gb200h_bank* b, int ch, double* v)
{
    if (!b || ch < 0 || ch >= b->bank->size()) return -1;
    const auto& k = b->bank->block(ch);
    v[0] = k.state();
    v[1] = k.carrier_doppler_hz();
    v[2] = k.cn0_db_hz();
    v[3] = k.carrier_lock_test();
    v[4] = k.code_freq_chips();
    v[5] = k.rem_code_phase_samples();
    v[6] = static_cast<double>(b->bank->nitems_read(ch));
    v[7] = k.current_prn_length_samples();
    return 0;
}

uint64_t gb200h_bank_launches(gb200h_bank* b) { return b->engine->launches(); }

// Thread-per-block shape: each thread owns a share of the channels and calls the blocking general_work on the common
// host stream, exactly what the GNU Radio scheduler does with one thread per tracking block.
int gb200h_bank_run_threaded(gb200h_bank* b, const void* iq, int64_t n_samples, uint64_t first_index, int n_threads, gb200h_trk_record* records, int max_records)
{
    try
        {
            const int nch = b->bank->size();
            const int eb = b->engine->config().iq_dtype == GB200_IQ_CF32 ? 8 : b->engine->config().iq_dtype == GB200_IQ_CS16 ? 4 : 2;
            n_threads = std::max(1, std::min(n_threads, nch));
            for (int i = 0; i < nch; i++)
                {
                    b->bank->block(i).set_engine(b->engine);
                    b->bank->block(i).upload_codes();
                }
            b->engine->set_expected_batch(n_threads);
            std::vector<std::vector<B200TrackingBank::Output>> per(n_threads);
            std::vector<std::string> errs(n_threads);
            std::vector<std::thread> th;
            for (int t = 0; t < n_threads; t++)
                th.emplace_back([&, t] {
                    try
                        {
                            // every thread walks the stream block by block; its channels keep their own read pointers
                            std::vector<uint64_t> rd;
                            for (int i = t; i < nch; i += n_threads) rd.push_back(b->bank->nitems_read(i));
                            bool progress = true;
                            while (progress)
                                {
                                    progress = false;
                                    int k = 0;
                                    for (int i = t; i < nch; i += n_threads, k++)
                                        {
                                            auto& blk = b->bank->block(i);
                                            const int64_t avail = static_cast<int64_t>(first_index) + n_samples - static_cast<int64_t>(rd[k]);
                                            if (avail < blk.required_samples() || rd[k] < first_index) continue;
                                            Gnss_Synchro s{};
                                            int consumed = 0;
                                            const char* in = static_cast<const char*>(iq) + (rd[k] - first_index) * eb;
                                            const int produced = blk.general_work(in, static_cast<int>(std::min<int64_t>(avail, blk.required_samples())), rd[k], &s, &consumed);
                                            rd[k] += static_cast<uint64_t>(consumed);
                                            if (produced) per[t].push_back({i, s});
                                            progress = true;
                                        }
                                }
                        }
                    catch (const std::exception& e)
                        {
                            errs[t] = e.what();
                        }
                    // a finished thread no longer takes part in the batches
                    static std::mutex m;
                    std::lock_guard<std::mutex> lk(m);
                    b->engine->set_expected_batch(0);
                });
            for (auto& x : th) x.join();
            for (auto& e : errs)
                if (!e.empty()) throw std::runtime_error(e);
            int n = 0;
            for (auto& v : per)
                for (auto& o : v)
                    if (n < max_records) to_record(o.channel, o.synchro, &records[n++]);
            return n;
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

}  // extern "C"

extern "C" {
// One tracking adapter started like Channel starts it and run behind its gr::block face (b200_tracking_gr): the stream goes through
// forecast / general_work / consume_each of the wrapper, Gnss_Synchro items come out of its output buffer.  `telemetry_fault_at` >= 0:
// a "telemetry_to_trk" message with event 1 is posted once nitems_read passes that sample.  Returns the records written; events[0..3] =
// number of messages on "events", the last one's value, items consumed, the block's final state.
int gb200h_gr_tracking_run(const char* conf_text, const char* role, unsigned prn, double acq_delay_samples, double acq_doppler_hz, const void* iq, int64_t n_samples,
    int64_t telemetry_fault_at, gb200h_trk_record* records, int max_records, int64_t* events)
{
    try
        {
            B200Configuration cfg;
            cfg.load_string(conf_text);
            B200CorrelatorEngine::reset_instances();
            auto adapter = GetB200TrkBlock(&cfg, role, 1, 1);
            if (!adapter) throw std::runtime_error(std::string(role) + ".implementation is not a B200 tracking block");
            auto* a = static_cast<B200DllPllTracking*>(adapter.get());
            Gnss_Synchro syn{};
            syn.System = a->conf().system;
            std::memcpy(syn.Signal, a->conf().signal, 3);
            syn.PRN = prn;
            syn.Acq_delay_samples = acq_delay_samples;
            syn.Acq_doppler_hz = acq_doppler_hz;
            syn.Acq_samplestamp_samples = 0;
            a->set_gnss_synchro(&syn);
            a->set_channel(0);
            a->start_tracking();
            auto blk = make_b200_tracking_gr(a->block(), a->item_size(), a->conf().vector_length);
            const size_t isz = a->item_size();
            std::vector<char> out;
            int64_t done = 0;
            bool posted = telemetry_fault_at < 0;
            const int64_t piece = 64 * static_cast<int64_t>(a->conf().vector_length);
            while (done < n_samples)
                {
                    // the scheduler hands the block what its input buffer holds: here pieces of the stream, each piece's unconsumed tail carried over
                    const int64_t n = std::min<int64_t>(piece, n_samples - done);
                    const int64_t c = blk->stub_feed(static_cast<const char*>(iq) + done * isz, n, &out);
                    if (c == 0) break;  // what is left does not meet the block's forecast
                    done += c;
                    if (!posted && done >= telemetry_fault_at)
                        {
                            blk->stub_post("telemetry_to_trk", pmt::from_long(1));
                            posted = true;
                        }
                }
            const int n_rec = static_cast<int>(out.size() / sizeof(Gnss_Synchro));
            for (int i = 0; i < n_rec && i < max_records; i++) to_record(0, reinterpret_cast<const Gnss_Synchro*>(out.data())[i], &records[i]);
            auto& pub = blk->stub_published("events");
            if (events)
                {
                    events[0] = static_cast<int64_t>(pub.size());
                    events[1] = pub.empty() ? 0 : pmt::to_long(pub.back());
                    events[2] = done;
                    events[3] = a->block()->state();
                }
            adapter.reset();
            B200CorrelatorEngine::reset_instances();
            return std::min(n_rec, max_records);
        }
    catch (const std::exception& e)
        {
            B200CorrelatorEngine::reset_instances();
            return fail(e);
        }
}

}  // extern "C"

// ------------------------------------------------------------------ acquisition block
struct gb200h_acq
{
    B200Configuration cfg;
    std::unique_ptr<B200PcpsAcquisition> adapter;
    Gnss_Synchro synchro;
    int last_event{0};
};

struct gb200h_acqcaf
{
    B200Configuration cfg;
    std::unique_ptr<GalileoE5aNoncoherentIQAcquisitionCafB200> adapter;
    Gnss_Synchro synchro;
    int last_event{0};
};

struct gb200h_acqqs
{
    B200Configuration cfg;
    std::unique_ptr<GpsL1CaPcpsQuickSyncAcquisitionB200> adapter;
    Gnss_Synchro synchro;
    int last_event{0};
};

struct gb200h_acqfd
{
    B200Configuration cfg;
    std::unique_ptr<GpsL1CaPcpsAcquisitionFineDopplerB200> adapter;
    Gnss_Synchro synchro;
    int last_event{0};
};

extern "C" {

int gb200h_acqcaf_create(const char* conf_text, const char* role, unsigned prn, const char* signal, gb200h_acqcaf** out)
{
    try
        {
            auto a = std::make_unique<gb200h_acqcaf>();
            a->cfg.load_string(conf_text);
            auto blk = GetB200AcqBlock(&a->cfg, role, 1, 0);
            auto* caf = dynamic_cast<GalileoE5aNoncoherentIQAcquisitionCafB200*>(blk.get());
            if (!caf) throw std::runtime_error(std::string(role) + ".implementation is not Galileo_E5a_Noncoherent_IQ_Acquisition_CAF_B200");
            blk.release();
            a->adapter.reset(caf);
            a->synchro.System = 'E';
            std::memset(a->synchro.Signal, 0, 3);
            a->synchro.Signal[0] = signal[0];
            a->synchro.Signal[1] = signal[0] ? signal[1] : '\0';
            a->synchro.PRN = prn;
            a->adapter->set_gnss_synchro(&a->synchro);
            a->adapter->set_channel(0);
            gb200h_acqcaf* raw = a.get();
            a->adapter->block()->set_event_sink([raw](int ev) { raw->last_event = ev; });
            *out = a.release();
            return 0;
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

void gb200h_acqcaf_destroy(gb200h_acqcaf* a) { delete a; }

int gb200h_acqcaf_arm(gb200h_acqcaf* a, unsigned doppler_max, unsigned doppler_step, float threshold)
{
    try
        {
            auto& ad = *a->adapter;
            ad.set_doppler_max(doppler_max);
            ad.set_doppler_step(doppler_step);
            ad.set_threshold(threshold);
            ad.init();
            ad.set_local_code();
            a->last_event = 0;
            ad.set_state(1);
            ad.reset();
            return 0;
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

int64_t gb200h_acqcaf_work(gb200h_acqcaf* a, const void* items, int64_t n_items, int chunk)
{
    try
        {
            auto& blk = *a->adapter->block();
            const auto* in = static_cast<const std::complex<float>*>(items);
            int64_t pos = 0;
            int idle = 0;
            while (pos < n_items && idle < 3)
                {
                    const int before = blk.state();
                    const int c = blk.general_work(in + pos, static_cast<int>(std::min<int64_t>(n_items - pos, chunk)));
                    pos += c;
                    if (blk.state() == 0 && !blk.active()) break;
                    idle = (c == 0 && blk.state() == before) ? idle + 1 : 0;
                }
            return pos;
        }
    catch (const std::exception& e)
        {
            fail(e);
            return -1;
        }
}

int gb200h_acqcaf_get_status(gb200h_acqcaf* a, gb200h_acqcaf_status* s)
{
    const auto& blk = *a->adapter->block();
    std::memset(s, 0, sizeof(*s));
    s->acq_delay_samples = a->synchro.Acq_delay_samples;
    s->acq_doppler_hz = a->synchro.Acq_doppler_hz;
    s->acq_samplestamp_samples = a->synchro.Acq_samplestamp_samples;
    s->acq_doppler_step = a->synchro.Acq_doppler_step;
    s->sample_counter = blk.sample_counter();
    s->test_statistics = blk.test_statistics();
    s->threshold = blk.threshold();
    s->mag = blk.mag_f();
    s->input_power = blk.input_power();
    s->state = blk.state();
    s->active = blk.active() ? 1 : 0;
    s->last_event = a->last_event;
    s->num_doppler_bins = blk.num_doppler_bins();
    s->fft_size = blk.fft_size();
    s->well_count = blk.well_count();
    return 0;
}

int gb200h_acqcaf_caf_vector(gb200h_acqcaf* a, float* dst, int max_bins)
{
    const auto& v = a->adapter->block()->caf_vector();
    const int n = std::min<int>(max_bins, static_cast<int>(v.size()));
    if (n > 0 && dst) std::copy(v.begin(), v.begin() + n, dst);
    return static_cast<int>(v.size());
}

int gb200h_acqqs_create(const char* conf_text, const char* role, unsigned prn, gb200h_acqqs** out) { return gb200h_acqqs_create2(conf_text, role, prn, 'G', "1C", out); }

int gb200h_acqqs_create2(const char* conf_text, const char* role, unsigned prn, char system, const char* signal, gb200h_acqqs** out)
{
    try
        {
            auto a = std::make_unique<gb200h_acqqs>();
            a->cfg.load_string(conf_text);
            auto blk = GetB200AcqBlock(&a->cfg, role, 1, 0);
            auto* qs = dynamic_cast<GpsL1CaPcpsQuickSyncAcquisitionB200*>(blk.get());
            if (!qs) throw std::runtime_error(std::string(role) + ".implementation is not a B200 QuickSync acquisition adapter");
            blk.release();
            a->adapter.reset(qs);
            a->synchro.System = system;
            std::memset(a->synchro.Signal, 0, 3);
            a->synchro.Signal[0] = signal[0];
            a->synchro.Signal[1] = signal[0] ? signal[1] : '\0';
            a->synchro.PRN = prn;
            a->adapter->set_gnss_synchro(&a->synchro);
            a->adapter->set_channel(0);
            gb200h_acqqs* raw = a.get();
            a->adapter->block()->set_event_sink([raw](int ev) { raw->last_event = ev; });
            *out = a.release();
            return 0;
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

void gb200h_acqqs_destroy(gb200h_acqqs* a) { delete a; }

int gb200h_acqqs_arm(gb200h_acqqs* a, unsigned doppler_max, unsigned doppler_step, float threshold)
{
    try
        {
            auto& ad = *a->adapter;
            ad.set_doppler_max(doppler_max);
            ad.set_doppler_step(doppler_step);
            ad.set_threshold(threshold);
            ad.init();
            ad.set_local_code();
            a->last_event = 0;
            ad.set_state(1);
            return 0;
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

int64_t gb200h_acqqs_work(gb200h_acqqs* a, const void* samples, int64_t n_samples, int max_items)
{
    try
        {
            auto& blk = *a->adapter->block();
            const auto* in = static_cast<const std::complex<float>*>(samples);
            const int64_t len = blk.item_length(), n_items = n_samples / len;
            int64_t pos = 0;
            while (pos < n_items)
                {
                    pos += blk.general_work(in + pos * len, static_cast<int>(std::min<int64_t>(n_items - pos, max_items)));
                    if (blk.state() == 0 && !blk.active()) break;
                }
            return pos;
        }
    catch (const std::exception& e)
        {
            fail(e);
            return -1;
        }
}

int gb200h_acqqs_get_status(gb200h_acqqs* a, gb200h_acqqs_status* s)
{
    const auto& blk = *a->adapter->block();
    std::memset(s, 0, sizeof(*s));
    s->acq_delay_samples = a->synchro.Acq_delay_samples;
    s->acq_doppler_hz = a->synchro.Acq_doppler_hz;
    s->acq_samplestamp_samples = a->synchro.Acq_samplestamp_samples;
    s->acq_doppler_step = a->synchro.Acq_doppler_step;
    s->sample_counter = blk.sample_counter();
    s->test_statistics = blk.test_statistics();
    s->threshold = blk.threshold();
    s->mag = blk.mag_f();
    s->input_power = blk.input_power();
    s->state = blk.state();
    s->active = blk.active() ? 1 : 0;
    s->last_event = a->last_event;
    s->num_doppler_bins = static_cast<int32_t>(blk.num_doppler_bins());
    s->fft_size = static_cast<int32_t>(blk.fft_size());
    s->well_count = static_cast<int32_t>(blk.well_count());
    s->item_length = blk.item_length();
    s->folding_factor = static_cast<int32_t>(a->adapter->folding_factor());
    s->sampled_ms = static_cast<int32_t>(a->adapter->sampled_ms());
    return 0;
}

int gb200h_acqfd_create(const char* conf_text, const char* role, unsigned prn, gb200h_acqfd** out)
{
    try
        {
            auto a = std::make_unique<gb200h_acqfd>();
            a->cfg.load_string(conf_text);
            auto blk = GetB200AcqBlock(&a->cfg, role, 1, 0);
            auto* fd = dynamic_cast<GpsL1CaPcpsAcquisitionFineDopplerB200*>(blk.get());
            if (!fd) throw std::runtime_error(std::string(role) + ".implementation is not GPS_L1_CA_PCPS_Acquisition_Fine_Doppler_B200");
            blk.release();
            a->adapter.reset(fd);
            a->synchro.System = 'G';
            std::memset(a->synchro.Signal, 0, 3);
            a->synchro.Signal[0] = '1';
            a->synchro.Signal[1] = 'C';
            a->synchro.PRN = prn;
            a->adapter->set_gnss_synchro(&a->synchro);
            a->adapter->set_channel(0);
            gb200h_acqfd* raw = a.get();
            a->adapter->block()->set_event_sink([raw](int ev) { raw->last_event = ev; });
            *out = a.release();
            return 0;
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

void gb200h_acqfd_destroy(gb200h_acqfd* a) { delete a; }

int gb200h_acqfd_arm(gb200h_acqfd* a, unsigned doppler_max, unsigned doppler_step, float threshold)
{
    try
        {
            auto& ad = *a->adapter;
            ad.set_doppler_max(doppler_max);
            ad.set_doppler_step(doppler_step);
            ad.set_threshold(threshold);
            ad.init();
            ad.set_local_code();
            ad.block()->start();
            a->last_event = 0;
            ad.set_state(1);
            return 0;
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

int64_t gb200h_acqfd_work(gb200h_acqfd* a, const void* items, int64_t n_items, int noutput_items)
{
    try
        {
            auto& blk = *a->adapter->block();
            const auto* in = static_cast<const std::complex<float>*>(items);
            const int64_t need = std::max(blk.fft_size(), noutput_items);
            int64_t pos = 0;
            while (n_items - pos >= need)
                {
                    pos += blk.general_work(noutput_items, in + pos, static_cast<int>(std::min<int64_t>(n_items - pos, 1 << 30)));
                    if (blk.state() == 0 && !blk.active()) break;
                }
            return pos;
        }
    catch (const std::exception& e)
        {
            fail(e);
            return -1;
        }
}

extern "C++" {
namespace
{
// one strategy block through its gr::block face: feed, read the "events" port, hand the event sink back to the door
template <class Gr, class Door>
int64_t work_through_gr(Door* a, const void* items, int64_t n_items, int* last_event_from_port)
{
    auto blk = std::make_shared<Gr>(a->adapter->block());  // re-points the block's event sink at the port
    const int64_t done = blk->stub_feed(items, n_items, nullptr);
    auto& pub = blk->stub_published("events");
    if (last_event_from_port) *last_event_from_port = pub.empty() ? 0 : static_cast<int>(pmt::to_long(pub.back()));
    if (!pub.empty()) a->last_event = static_cast<int>(pmt::to_long(pub.back()));
    a->adapter->block()->set_event_sink([a](int ev) { a->last_event = ev; });
    return done;
}
}  // namespace
}  // extern "C++"

int64_t gb200h_acqfd_work_gr(gb200h_acqfd* a, const void* items, int64_t n_items, int* last_event_from_port)
{
    try
        {
            return work_through_gr<b200_fine_doppler_acquisition_gr>(a, items, n_items, last_event_from_port);
        }
    catch (const std::exception& e)
        {
            fail(e);
            return -1;
        }
}

int64_t gb200h_acqcaf_work_gr(gb200h_acqcaf* a, const void* items, int64_t n_items, int* last_event_from_port)
{
    try
        {
            return work_through_gr<b200_e5a_caf_acquisition_gr>(a, items, n_items, last_event_from_port);
        }
    catch (const std::exception& e)
        {
            fail(e);
            return -1;
        }
}

int64_t gb200h_acqqs_work_gr(gb200h_acqqs* a, const void* samples, int64_t n_samples, int* last_event_from_port)
{
    try
        {
            return work_through_gr<b200_quicksync_acquisition_gr>(a, samples, n_samples / a->adapter->block()->item_length(), last_event_from_port);
        }
    catch (const std::exception& e)
        {
            fail(e);
            return -1;
        }
}

int gb200h_acqfd_get_status(gb200h_acqfd* a, gb200h_acqfd_status* s)
{
    const auto& blk = *a->adapter->block();
    std::memset(s, 0, sizeof(*s));
    s->acq_delay_samples = a->synchro.Acq_delay_samples;
    s->acq_doppler_hz = a->synchro.Acq_doppler_hz;
    s->acq_samplestamp_samples = a->synchro.Acq_samplestamp_samples;
    s->acq_doppler_step = a->synchro.Acq_doppler_step;
    s->sample_counter = blk.sample_counter();
    s->fine_index = blk.fine_index();
    s->test_statistics = blk.test_statistics();
    s->threshold = blk.threshold();
    s->state = blk.state();
    s->active = blk.active() ? 1 : 0;
    s->last_event = a->last_event;
    s->positive_acq = blk.positive_acq();
    s->num_doppler_points = blk.num_doppler_points();
    s->fft_size = blk.fft_size();
    return 0;
}

int gb200h_acq_create(const char* conf_text, const char* role, unsigned prn, gb200h_acq** out) { return gb200h_acq_create2(conf_text, role, prn, 'G', "1C", out); }

int gb200h_acq_create2(const char* conf_text, const char* role, unsigned prn, char system, const char* signal, gb200h_acq** out)
{
    try
        {
            auto a = std::make_unique<gb200h_acq>();
            a->cfg.load_string(conf_text);
            auto blk = GetB200AcqBlock(&a->cfg, role, 1, 0);
            if (!blk) throw std::runtime_error(std::string(role) + ".implementation is not a B200 acquisition block");
            a->adapter.reset(static_cast<B200PcpsAcquisition*>(blk.release()));
            a->synchro.System = system;
            std::memset(a->synchro.Signal, 0, 3);
            std::strncpy(a->synchro.Signal, signal, 2);
            a->synchro.PRN = prn;
            a->adapter->set_gnss_synchro(&a->synchro);
            a->adapter->set_channel(0);
            gb200h_acq* raw = a.get();
            a->adapter->block()->set_event_sink([raw](int ev) { raw->last_event = ev; });
            *out = a.release();
            return 0;
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

void gb200h_acq_destroy(gb200h_acq* a) { delete a; }

int gb200h_acq_arm(gb200h_acq* a, float threshold_if_no_pfa)
{
    try
        {
            auto& ad = *a->adapter;
            ad.set_doppler_max(static_cast<unsigned>(ad.conf().doppler_max));
            ad.set_doppler_step(static_cast<unsigned>(ad.conf().doppler_step));
            ad.set_threshold(threshold_if_no_pfa);
            ad.init();
            ad.set_local_code();
            ad.block()->start();
            a->last_event = 0;
            ad.set_state(1);
            return 0;
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

int64_t gb200h_acq_work(gb200h_acq* a, const void* items, int64_t n_items, int chunk)
{
    try
        {
            auto& blk = *a->adapter->block();
            const size_t isz = a->adapter->item_size();
            int64_t done = 0;
            int idle_calls = 0;
            while (done < n_items && idle_calls < 8)  // a two-step search makes 4 calls in a row that consume nothing
                {
                    const int n = static_cast<int>(std::min<int64_t>(chunk, n_items - done));
                    const int c = blk.general_work(static_cast<const char*>(items) + done * isz, n);
                    done += c;
                    if (!blk.active() && !blk.step_two()) break;  // a two-step search re-activates itself for the fine grid
                    idle_calls = (c == 0) ? idle_calls + 1 : 0;
                }
            return done;
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

int gb200h_acq_get_status(gb200h_acq* a, gb200h_acq_status* s)
{
    auto& blk = *a->adapter->block();
    s->acq_delay_samples = a->synchro.Acq_delay_samples;
    s->acq_doppler_hz = a->synchro.Acq_doppler_hz;
    s->acq_samplestamp_samples = a->synchro.Acq_samplestamp_samples;
    s->test_statistics = blk.test_statistics();
    s->threshold = blk.threshold();
    s->mag = static_cast<float>(blk.mag());
    s->input_power = blk.input_power();
    s->state = blk.state();
    s->active = blk.active();
    s->last_event = a->last_event;
    s->indext = blk.last_result().indext;
    s->doppler_index = blk.last_result().doppler_index;
    return 0;
}

// The acquisition block behind its gr::block face (b200_gr_blocks.h): items go through general_work(noutput, ninput_items, input_items, ...)
// and consume_each, events come out of the "events" message port.  Returns the items consumed; *last_event_from_port = the last PMT published.
int64_t gb200h_acq_work_gr(gb200h_acq* a, const void* items, int64_t n_items, int* last_event_from_port)
{
    try
        {
            auto blk = make_b200_acquisition_gr(a->adapter->block(), a->adapter->item_size());  // re-points the block's event sink at the port
            const int64_t done = blk->stub_feed(items, n_items, nullptr);
            auto& pub = blk->stub_published("events");
            if (last_event_from_port) *last_event_from_port = pub.empty() ? 0 : static_cast<int>(pmt::to_long(pub.back()));
            if (!pub.empty()) a->last_event = static_cast<int>(pmt::to_long(pub.back()));
            gb200h_acq* raw = a;
            a->adapter->block()->set_event_sink([raw](int ev) { raw->last_event = ev; });
            return done;
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

int64_t gb200h_acq_local_code(gb200h_acq* a, float* dst, int64_t max_samples)
{
    const auto& c = a->adapter->local_code();
    const int64_t n = std::min<int64_t>(max_samples, static_cast<int64_t>(c.size()));
    if (dst && n > 0) std::memcpy(dst, c.data(), static_cast<size_t>(n) * sizeof(std::complex<float>));
    return static_cast<int64_t>(c.size());
}

int gb200h_mat_write_sample(const char* filename, const float* grid, int rows, int cols, int32_t i32, uint32_t u32, uint64_t u64, float f32)
{
    MatFile5 m(filename);
    if (!m.ok()) return -1;
    m.write("acq_grid", MatFile5::SINGLE, static_cast<uint32_t>(rows), static_cast<uint32_t>(cols), grid);
    m.scalar("doppler_max", MatFile5::INT32, i32);
    m.scalar("PRN", MatFile5::UINT32, u32);
    m.scalar("sample_counter", MatFile5::UINT64, u64);
    m.scalar("threshold", MatFile5::SINGLE, f32);
    return 0;
}

struct gb200h_rx
{
    B200Configuration cfg;
    std::unique_ptr<B200ChannelManager> mgr;
    std::vector<B200TrackingBank::Output> outs;
};

int gb200h_rx_create(const char* conf_text, const char* acq_role, const char* trk_role, int n_channels, const unsigned* prns, int n_prns, gb200h_rx** out)
{
    try
        {
            auto r = std::make_unique<gb200h_rx>();
            r->cfg.load_string(conf_text);
            r->mgr = std::make_unique<B200ChannelManager>(&r->cfg, acq_role, trk_role, n_channels, std::vector<unsigned>(prns, prns + n_prns));
            *out = r.release();
            return 0;
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

void gb200h_rx_destroy(gb200h_rx* r) { delete r; }

int gb200h_rx_push(gb200h_rx* r, const void* iq, int64_t n_samples, gb200h_trk_record* records, int max_records)
{
    try
        {
            r->outs.clear();
            r->mgr->push(iq, n_samples, &r->outs);
            if (static_cast<int>(r->outs.size()) > max_records) throw std::runtime_error("gb200h_rx_push: " + std::to_string(r->outs.size()) + " records do not fit the caller's buffer");
            const int n = static_cast<int>(r->outs.size());
            for (int i = 0; i < n; i++) to_record(r->outs[i].channel, r->outs[i].synchro, &records[i]);
            return n;
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

int gb200h_rx_channel_info(gb200h_rx* r, int ch, double* v)
{
    if (!r || ch < 0 || ch >= r->mgr->size()) return -1;
    const auto c = r->mgr->info(ch);
    v[0] = c.state;
    v[1] = c.prn;
    v[2] = c.acquisitions;
    v[3] = c.losses;
    v[4] = c.acq_doppler_hz;
    v[5] = c.acq_delay_samples;
    v[6] = static_cast<double>(c.acq_samplestamp_samples);
    v[7] = c.test_statistics;
    v[8] = r->mgr->bank().block(ch).state();
    v[9] = r->mgr->bank().block(ch).cn0_db_hz();
    return 0;
}

uint64_t gb200h_rx_sweeps(gb200h_rx* r) { return r->mgr->sweeps(); }
float gb200h_rx_threshold(gb200h_rx* r) { return r->mgr->threshold(); }

// ---- Gnss_Synchro monitor stream
struct gb200h_monitor
{
    B200Configuration cfg;
    std::unique_ptr<B200SynchroMonitor> mon;
};

namespace
{
Gnss_Synchro from_c(const gb200h_synchro& v)
{
    Gnss_Synchro s;
    s.System = v.system;
    std::memcpy(s.Signal, v.signal, 3);
    s.PRN = v.prn;
    s.Channel_ID = v.channel_id;
    s.Acq_delay_samples = v.acq_delay_samples;
    s.Acq_doppler_hz = v.acq_doppler_hz;
    s.Acq_samplestamp_samples = v.acq_samplestamp_samples;
    s.Acq_doppler_step = v.acq_doppler_step;
    s.Flag_valid_acquisition = v.flag_valid_acquisition != 0;
    s.fs = v.fs;
    s.Prompt_I = v.prompt_i;
    s.Prompt_Q = v.prompt_q;
    s.CN0_dB_hz = v.cn0_db_hz;
    s.Carrier_Doppler_hz = v.carrier_doppler_hz;
    s.Carrier_phase_rads = v.carrier_phase_rads;
    s.Code_phase_samples = v.code_phase_samples;
    s.Tracking_sample_counter = v.tracking_sample_counter;
    s.Flag_valid_symbol_output = v.flag_valid_symbol_output != 0;
    s.correlation_length_ms = v.correlation_length_ms;
    s.Flag_valid_word = v.flag_valid_word != 0;
    s.TOW_at_current_symbol_ms = v.tow_at_current_symbol_ms;
    s.Pseudorange_m = v.pseudorange_m;
    s.RX_time = v.rx_time;
    s.Flag_valid_pseudorange = v.flag_valid_pseudorange != 0;
    s.Flag_PLL_180_deg_phase_locked = v.flag_pll_180_deg_phase_locked != 0;
    s.interp_TOW_ms = v.interp_tow_ms;
    return s;
}
}  // namespace

int gb200h_synchro_serialize(const gb200h_synchro* items, int n, char* buf, int cap)
{
    try
        {
            std::vector<Gnss_Synchro> v;
            for (int i = 0; i < n; i++) v.push_back(from_c(items[i]));
            const std::string data = create_protobuffer(v);
            if (static_cast<int>(data.size()) > cap) throw std::runtime_error("gb200h_synchro_serialize: buffer too small");
            std::memcpy(buf, data.data(), data.size());
            return static_cast<int>(data.size());
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

int gb200h_monitor_create(const char* conf_text, const char* role, int n_channels, gb200h_monitor** out)
{
    try
        {
            auto m = std::make_unique<gb200h_monitor>();
            m->cfg.load_string(conf_text);
            m->mon = std::make_unique<B200SynchroMonitor>(&m->cfg, role, n_channels);
            *out = m.release();
            return 0;
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

void gb200h_monitor_destroy(gb200h_monitor* m) { delete m; }
int gb200h_monitor_enabled(gb200h_monitor* m) { return m && m->mon->enabled() ? 1 : 0; }

int64_t gb200h_monitor_write(gb200h_monitor* m, int channel, const gb200h_synchro* items, int n)
{
    try
        {
            std::vector<Gnss_Synchro> v;
            for (int i = 0; i < n; i++) v.push_back(from_c(items[i]));
            m->mon->write(channel, v.data(), n);
            return static_cast<int64_t>(m->mon->datagrams());
        }
    catch (const std::exception& e)
        {
            return fail(e);
        }
}

int64_t gb200h_bank_monitor_datagrams(gb200h_bank* b) { return b && b->monitor ? static_cast<int64_t>(b->monitor->datagrams()) : 0; }

}  // extern "C"
