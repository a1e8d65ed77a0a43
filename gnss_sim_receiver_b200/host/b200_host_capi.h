/* C door onto the C++ host layer (libgnss_b200_host.so) for tests/, bench.py and non-C++ embedders.
 * Plain pointers and sizes; every function returns 0 / a count, or -1 with the message in gb200h_last_error().
 * The GPU entry points (bank, acquisition block) fail when libgnss_b200.so cannot create its engines: no CPU fallback. */
#ifndef GB200_HOST_CAPI_H
#define GB200_HOST_CAPI_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

const char* gb200h_last_error(void);

/* ---- CPU-only pieces (checked against the reference's own objects in tests/test_host_cpu.py) */
/* value of `name` in a .conf text, or `def` — B200Configuration::property(string) */
const char* gb200h_conf_get(const char* conf_text, const char* name, const char* def);
/* Dll_Pll_Conf / Acq_Conf after SetFromConfiguration, as "key=value;" text */
const char* gb200h_dll_pll_conf(const char* conf_text, const char* role);
const char* gb200h_acq_conf(const char* conf_text, const char* role, double chip_rate, double opt_freq, int ms_per_code);
/* two tracking adapters of one configuration, started like Channel starts them: which per-(device, item type, stream) engine each
 * attached to ("same=..;stream_a=..;stream_b=..;slot_a=..;slot_b=..;max_block=..;" or "error=...").  Needs a GPU. */
const char* gb200h_engine_probe(const char* conf_text, const char* role_a, unsigned prn_a, const char* role_b, unsigned prn_b);
/* implementation()/item_size of the adapter GetB200Block selects for `role`, "" when the string is not ours */
const char* gb200h_factory(const char* conf_text, const char* role);
double gb200h_pll_cloop_two_quadrant_atan(float re, float im);
double gb200h_pll_four_quadrant_atan(float re, float im);
double gb200h_fll_four_quadrant_atan(float r1, float i1, float r2, float i2, double t1, double t2);
double gb200h_fll_diff_atan(float r1, float i1, float r2, float i2, double t1, double t2);
double gb200h_dll_nc_e_minus_l_normalized(float er, float ei, float lr, float li, float spc, float slope, float y_intercept);
double gb200h_dll_nc_vemlp_normalized(float ver, float vei, float er, float ei, float lr, float li, float vlr, float vli);
/* GALILEO_E5A_Q_SECONDARY_CODE[prn - 1] as 100 '0' / '1' characters + NUL into out[101]; returns 100, or 0 for a PRN outside 1..50 */
int gb200h_e5a_q_secondary_code(unsigned prn, char* out);
double gb200h_sinboc_slope_abs(double x);
double gb200h_sinboc_yintercept_abs(double x);
float gb200h_cn0_m2m4_estimator(const float* prompt_ri, int length, float T);
float gb200h_cn0_svn_estimator(const float* prompt_ri, int length, float T);
float gb200h_carrier_lock_detector(const float* prompt_ri, int length);
void gb200h_tracking_loop_filter_run(float update_interval, float bw, int order, int include_last, float init, const float* in, float* out, int n);
void gb200h_fll_pll_filter_run(float fll_bw, float pll_bw, int order, float init_doppler, const float* fll_in, const float* pll_in, float T, float* out, int n);
void gb200h_exponential_smoother_run(float alpha, int init_samples, float min_value, float offset, const float* in, float* out, int n);
int gb200h_gps_l1_ca_code_gen_float(float* dest, unsigned prn, unsigned chip_shift);
int gb200h_gps_l1_ca_code_gen_complex_sampled(float* dest_ri, int n, unsigned prn, int fs, unsigned chip_shift);
/* Galileo E1 / GPS L5 / Galileo E5a generators (b200_gnss_codes.h); `what` 0..7, see b200_host_capi.cc; returns samples written */
int gb200h_code_gen(int what, float* dest, const char* signal_id, unsigned prn, int fs, unsigned chip_shift, int cboc, int secondary);

/* ---- tracking bank: N dll_pll_veml tracking blocks of one signal over one GPU engine */
typedef struct gb200h_bank gb200h_bank;
typedef struct
{
    int32_t channel;
    uint32_t prn;
    int32_t valid;        /* Flag_valid_symbol_output; 0 = loss-of-lock notification */
    int32_t correlation_length_ms;
    double prompt_i, prompt_q, cn0_db_hz, carrier_doppler_hz, carrier_phase_rads, code_phase_samples;
    uint64_t tracking_sample_counter;
} gb200h_trk_record;
/* `role` names the tracking block in conf_text (e.g. "Tracking_1C"); its .implementation selects the adapter */
int gb200h_bank_create(const char* conf_text, const char* role, int n_channels, int max_engine_channels, gb200h_bank** out);
void gb200h_bank_destroy(gb200h_bank* b);
/* optional code tables for signals generated outside this tree (Galileo E1, L5, E5a); before bank_start */
int gb200h_bank_set_codes(gb200h_bank* b, int ch, const float* tracking_code, const float* data_code, int n_entries);
int gb200h_bank_start(gb200h_bank* b, int ch, unsigned prn, double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp, uint64_t nitems_read);
int gb200h_bank_push(gb200h_bank* b, const void* iq, int64_t n_samples);
/* raw file items (gb200_ingest_format fields), converted on the device; returns the number of samples appended or < 0 */
int64_t gb200h_bank_push_raw(gb200h_bank* b, const void* raw, int64_t n_bytes, int kind, int sample_type, int item_size, int big_endian_items, int big_endian_bytes);
/* runs rounds until no channel can advance; returns the number of records written (at most max_records) */
int gb200h_bank_step(gb200h_bank* b, gb200h_trk_record* records, int max_records);
/* v[0..7] = state, carrier_doppler_hz, cn0_db_hz, carrier_lock_test, code_freq_chips, rem_code_phase_samples, nitems_read, current_prn_length */
int gb200h_bank_channel_state(gb200h_bank* b, int ch, double* v);
uint64_t gb200h_bank_launches(gb200h_bank* b);
/* per-channel blocking form from `n_threads` host threads (the GNU Radio thread-per-block shape): every thread runs
 * general_work over the same host stream for its share of channels; returns the records like bank_step */
/* one tracking adapter behind its gr::block face (b200_tracking_gr): forecast / general_work / consume_each, "events" and "telemetry_to_trk"
 * message ports.  events[4] = messages published on "events", the last one's value, items consumed, the block's final d_state. */
int gb200h_gr_tracking_run(const char* conf_text, const char* role, unsigned prn, double acq_delay_samples, double acq_doppler_hz, const void* iq, int64_t n_samples,
    int64_t telemetry_fault_at, gb200h_trk_record* records, int max_records, int64_t* events);
int gb200h_bank_run_threaded(gb200h_bank* b, const void* iq, int64_t n_samples, uint64_t first_index, int n_threads, gb200h_trk_record* records, int max_records);

/* ---- acquisition block (one channel) */
typedef struct gb200h_acq gb200h_acq;
typedef struct
{
    double acq_delay_samples, acq_doppler_hz;
    uint64_t acq_samplestamp_samples;
    float test_statistics, threshold, mag, input_power;
    int32_t state, active, last_event; /* last_event: 0 none, 1 positive, 2 negative */
    uint32_t indext;
    int32_t doppler_index;
} gb200h_acq_status;
int gb200h_acq_create(const char* conf_text, const char* role, unsigned prn, gb200h_acq** out); /* System 'G', Signal "1C" */
/* same for any of the four acquisition adapters: `system` / `signal` fill the channel's Gnss_Synchro ('E' "1B", 'G' "L5", 'E' "5X" ...) */
int gb200h_acq_create2(const char* conf_text, const char* role, unsigned prn, char system, const char* signal, gb200h_acq** out);
/* the same through the block's gr::block face (host/b200_gr_blocks.h; GNU Radio stub outside GNSS-SDR): general_work + consume_each, events
 * read back from the "events" message port */
int64_t gb200h_acq_work_gr(gb200h_acq* a, const void* items, int64_t n_items, int* last_event_from_port);
/* the replica the adapter handed to set_local_code (complex float pairs); returns its length in samples */
int64_t gb200h_acq_local_code(gb200h_acq* a, float* dst, int64_t max_samples);
void gb200h_acq_destroy(gb200h_acq* a);
/* adapter calls, in the order Channel/ChannelFsm make them: set_doppler_*, set_threshold (pfa = 0), init, set_local_code, start, set_state(1) */
int gb200h_acq_arm(gb200h_acq* a, float threshold_if_no_pfa);
/* general_work over a host buffer of the configured item type: loops until every item is consumed or the block went idle; returns items consumed */
int64_t gb200h_acq_work(gb200h_acq* a, const void* items, int64_t n_items, int chunk);
int gb200h_acq_get_status(gb200h_acq* a, gb200h_acq_status* s);

/* ---- fine-Doppler acquisition block (GPS_L1_CA_PCPS_Acquisition_Fine_Doppler_B200; b200_pcps_acquisition_fine_doppler.h) */
typedef struct gb200h_acqfd gb200h_acqfd;
typedef struct
{
    double acq_delay_samples, acq_doppler_hz;
    uint64_t acq_samplestamp_samples, sample_counter;
    int64_t fine_index; /* index of the padded-spectrum maximum of the last estimate_Doppler, -1 before */
    float test_statistics, threshold;
    int32_t state, active, last_event, positive_acq, num_doppler_points, fft_size;
    uint32_t acq_doppler_step;
    int32_t pad_;
} gb200h_acqfd_status;
int gb200h_acqfd_create(const char* conf_text, const char* role, unsigned prn, gb200h_acqfd** out);
void gb200h_acqfd_destroy(gb200h_acqfd* a);
/* the adapter calls a channel makes: set_doppler_max, set_doppler_step, set_threshold, init, set_local_code, start, set_state(1) */
int gb200h_acqfd_arm(gb200h_acqfd* a, unsigned doppler_max, unsigned doppler_step, float threshold);
/* plays the scheduler over a host buffer of gr_complex items: general_work(noutput_items, ...) while at least max(fft_size, noutput_items)
 * items are left, until the block has raised its event and gone idle; returns the items consumed, -1 on error */
int64_t gb200h_acqfd_work(gb200h_acqfd* a, const void* items, int64_t n_items, int noutput_items);
int gb200h_acqfd_get_status(gb200h_acqfd* a, gb200h_acqfd_status* s);
/* the same items through the block's gr::block face (b200_fine_doppler_acquisition_gr; GNU Radio stub outside GNSS-SDR: one output item per call): items consumed, event from the "events" port */
int64_t gb200h_acqfd_work_gr(gb200h_acqfd* a, const void* items, int64_t n_items, int* last_event_from_port);

/* ---- Galileo E5a non-coherent I/Q acquisition with CAF (Galileo_E5a_Noncoherent_IQ_Acquisition_CAF_B200; b200_e5a_noncoherent_iq_acquisition_caf.h) */
typedef struct gb200h_acqcaf gb200h_acqcaf;
typedef struct
{
    double acq_delay_samples, acq_doppler_hz;
    uint64_t acq_samplestamp_samples, sample_counter;
    float test_statistics, threshold, mag, input_power;
    int32_t state, active, last_event, num_doppler_bins, fft_size, well_count;
    uint32_t acq_doppler_step;
    int32_t pad_;
} gb200h_acqcaf_status;
/* signal: the channel's Gnss_Synchro signal ("5X": I and Q replicas, else one "5X"-composed replica, adapters/...caf.cc:218-233) */
int gb200h_acqcaf_create(const char* conf_text, const char* role, unsigned prn, const char* signal, gb200h_acqcaf** out);
void gb200h_acqcaf_destroy(gb200h_acqcaf* a);
/* set_doppler_max, set_doppler_step, set_threshold (a configured pfa wins), init, set_local_code, set_state(1), set_active(true) */
int gb200h_acqcaf_arm(gb200h_acqcaf* a, unsigned doppler_max, unsigned doppler_step, float threshold);
/* plays the scheduler: general_work with at most `chunk` items per call until the block has raised its event and gone idle or the buffer
 * is used up; returns the items consumed, -1 on error */
int64_t gb200h_acqcaf_work(gb200h_acqcaf* a, const void* items, int64_t n_items, int chunk);
int gb200h_acqcaf_get_status(gb200h_acqcaf* a, gb200h_acqcaf_status* s);
int64_t gb200h_acqcaf_work_gr(gb200h_acqcaf* a, const void* items, int64_t n_items, int* last_event_from_port); /* b200_e5a_caf_acquisition_gr */
/* the filtered CAF vector of the last search (num_doppler_bins floats; 0 returned when CAF_window_hz = 0) */
int gb200h_acqcaf_caf_vector(gb200h_acqcaf* a, float* dst, int max_bins);

/* ---- QuickSync acquisition (GPS_L1_CA_PCPS_QuickSync_Acquisition_B200; b200_pcps_quicksync_acquisition.h) */
typedef struct gb200h_acqqs gb200h_acqqs;
typedef struct
{
    double acq_delay_samples, acq_doppler_hz;
    uint64_t acq_samplestamp_samples, sample_counter;
    float test_statistics, threshold, mag, input_power;
    int32_t state, active, last_event, num_doppler_bins, fft_size, well_count, item_length, folding_factor, sampled_ms;
    uint32_t acq_doppler_step;
} gb200h_acqqs_status;
int gb200h_acqqs_create(const char* conf_text, const char* role, unsigned prn, gb200h_acqqs** out);
/* same for either QuickSync adapter: `system` / `signal` fill the channel's Gnss_Synchro ('G' "1C", 'E' "1B" / "1C") */
int gb200h_acqqs_create2(const char* conf_text, const char* role, unsigned prn, char system, const char* signal, gb200h_acqqs** out);
void gb200h_acqqs_destroy(gb200h_acqqs* a);
/* set_doppler_max, set_doppler_step, set_threshold (a configured pfa wins), init, set_local_code, set_state(1) */
int gb200h_acqqs_arm(gb200h_acqqs* a, unsigned doppler_max, unsigned doppler_step, float threshold);
/* plays stream_to_vector + the scheduler: the stream is cut into vector items of item_length samples and general_work is called with at most
 * `max_items` of them until the block has raised its event and gone idle or the items are used up; returns the ITEMS consumed, -1 on error */
int64_t gb200h_acqqs_work(gb200h_acqqs* a, const void* samples, int64_t n_samples, int max_items);
int gb200h_acqqs_get_status(gb200h_acqqs* a, gb200h_acqqs_status* s);
int64_t gb200h_acqqs_work_gr(gb200h_acqqs* a, const void* samples, int64_t n_samples, int* last_event_from_port); /* b200_quicksync_acquisition_gr: vector ITEMS consumed */

/* MAT level-5 writer used by the acquisition dump (b200_mat_writer.h): one file with a single matrix and one scalar of each class */
int gb200h_mat_write_sample(const char* filename, const float* grid, int rows, int cols, int32_t i32, uint32_t u32, uint64_t u64, float f32);

/* ---- channel manager (B200ChannelManager): batched acquisition sweep feeding a lock-step tracking bank */
typedef struct gb200h_rx gb200h_rx;
int gb200h_rx_create(const char* conf_text, const char* acq_role, const char* trk_role, int n_channels, const unsigned* prns, int n_prns, gb200h_rx** out);
void gb200h_rx_destroy(gb200h_rx* r);
/* feed host samples of the configured item type; returns the number of tracking records written (<= max_records) */
int gb200h_rx_push(gb200h_rx* r, const void* iq, int64_t n_samples, gb200h_trk_record* records, int max_records);
/* v[0..9] = state (0 standby, 1 acquisition, 2 tracking), prn, acquisitions, losses, acq_doppler_hz, acq_delay_samples,
 * acq_samplestamp_samples, test_statistics, tracking block state, tracking CN0 (dB-Hz) */
int gb200h_rx_channel_info(gb200h_rx* r, int ch, double* v);
uint64_t gb200h_rx_sweeps(gb200h_rx* r);
float gb200h_rx_threshold(gb200h_rx* r);

/* ---- Gnss_Synchro monitor stream (b200_synchro_monitor.h; serdes_gnss_synchro.h:72-119, gnss_synchro_udp_sink.cc:43-75) ---- */
typedef struct gb200h_synchro
{
    char system;
    char signal[3];
    uint32_t prn;
    int32_t channel_id;
    double acq_delay_samples, acq_doppler_hz;
    uint64_t acq_samplestamp_samples;
    uint32_t acq_doppler_step;
    int32_t flag_valid_acquisition;
    int64_t fs;
    double prompt_i, prompt_q, cn0_db_hz, carrier_doppler_hz, carrier_phase_rads, code_phase_samples;
    uint64_t tracking_sample_counter;
    int32_t flag_valid_symbol_output, correlation_length_ms, flag_valid_word;
    uint32_t tow_at_current_symbol_ms;
    double pseudorange_m, rx_time;
    int32_t flag_valid_pseudorange, flag_pll_180_deg_phase_locked;
    double interp_tow_ms;
} gb200h_synchro;
/* Serdes_Gnss_Synchro::createProtobuffer: the `Observables` message of n items; returns its length (<= cap) or < 0 */
int gb200h_synchro_serialize(const gb200h_synchro* items, int n, char* buf, int cap);
typedef struct gb200h_monitor gb200h_monitor;
/* role: "Monitor", "TrackingMonitor", "AcquisitionMonitor" (gnss_flowgraph.cc:194-262) */
int gb200h_monitor_create(const char* conf_text, const char* role, int n_channels, gb200h_monitor** out);
void gb200h_monitor_destroy(gb200h_monitor* m);
int gb200h_monitor_enabled(gb200h_monitor* m);
/* one channel's new items; returns the datagrams sent so far or < 0 */
int64_t gb200h_monitor_write(gb200h_monitor* m, int channel, const gb200h_synchro* items, int n);
/* datagrams the bank's own TrackingMonitor (TrackingMonitor.enable_monitor=true in the bank's .conf) has sent */
int64_t gb200h_bank_monitor_datagrams(gb200h_bank* b);

#ifdef __cplusplus
}
#endif
#endif
