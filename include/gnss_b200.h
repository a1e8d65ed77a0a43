/*
 * gnss_b200.h — C ABI of the B200-native (sm_100a) GNSS hot paths.
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++/torch types.
 * Every entry point names the reference interface it replaces (paths relative
 * to the root of ShingoNishimoto/gnss_sim_receiver).  Host code (the C++
 * adapters under gnss_sim_receiver_b200/host/, or GNSS-SDR itself through the
 * patch in INTEGRATION.md) calls only these functions.
 *
 * Conventions
 *   - every function returns 0 (GB200_OK) or a negative GB200_E* code;
 *     gb200_last_error() gives the message of the last failure on this thread;
 *   - there is NO CPU fallback: without a CUDA device every *_create fails
 *     with GB200_ENODEVICE;
 *   - complex samples are interleaved (re, im);
 *   - all host pointers may be pageable; pinned memory only makes copies faster.
 */
#ifndef GNSS_B200_H
#define GNSS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GB200_OK 0
#define GB200_EINVAL (-1)    /* bad argument */
#define GB200_ENODEVICE (-2) /* no usable CUDA device / wrong architecture */
#define GB200_ECUDA (-3)     /* CUDA runtime error (message in gb200_last_error) */
#define GB200_ENOMEM (-4)
#define GB200_ESTATE (-5)    /* call out of order (e.g. correlate before set_code) */
#define GB200_ERANGE (-6)    /* sample window not (or no longer) in the IQ ring */

#define GB200_MAX_TAPS 16

/* IQ sample types accepted on the device ring / acquisition input.
 * Replaces the item types of the reference's blocks: gr_complex, cshort
 * (lv_16sc_t, converted by volk_gnsssdr_16ic_convert_32fc at
 * pcps_acquisition.cc:610) and cbyte (lv_8sc_t). */
#define GB200_IQ_CF32 0 /* 2 x float   */
#define GB200_IQ_CS16 1 /* 2 x int16_t */
#define GB200_IQ_CS8 2  /* 2 x int8_t  */

typedef struct
{
    float re, im;
} gb200_cf32;

const char* gb200_last_error(void);
/* number of CUDA kernels this library has launched in this process (all handles) */
uint64_t gb200_launch_count(void);
/* number of visible CUDA devices (0 when there is none; never fails) */
int gb200_device_count(void);
const char* gb200_version(void);
/* Page-locked host memory for buffers the caller keeps across calls (correlator outputs, IQ staging): the place of
 * volk_gnsssdr_malloc in the reference (src/algorithms/tracking/libs/cpu_multicorrelator_real_codes.cc:44-49, d_correlator_outs in
 * src/algorithms/tracking/gnuradio_blocks/dll_pll_veml_tracking.cc:520-522).  gb200_trk_correlate / gb200_trk_collect copy device
 * results straight into an `out` that lives in such memory; any other pointer goes through the engine's own staging buffer. */
int gb200_host_alloc(void** out, size_t bytes);
int gb200_host_free(void* p);

/* ======================================================================
 * Tracking multicorrelator engine
 *
 * Replaces, for ALL channels resident on one GPU at once,
 *   Cpu_Multicorrelator_Real_Codes   src/algorithms/tracking/libs/cpu_multicorrelator_real_codes.{h:37-61,cc:36-160}
 *   Cpu_Multicorrelator              src/algorithms/tracking/libs/cpu_multicorrelator.{h,cc:73-100}
 *   cuda_multicorrelator (legacy)    src/algorithms/tracking/libs/cuda_multicorrelator.{h:111-161,cu:32-346}
 * i.e. volk_gnsssdr_32f_xn_resampler_32f_xn + volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn
 * (and the 32fc_x2 complex-code sibling).  One engine owns a device-resident IQ
 * ring shared by every channel; one gb200_trk_correlate call evaluates one
 * Carrier_wipeoff_multicorrelator_resampler per command, all in one launch.
 * ====================================================================== */
typedef struct gb200_trk gb200_trk;

typedef struct
{
    int32_t device;       /* CUDA ordinal */
    int32_t max_channels; /* channels addressable by gb200_trk_set_code / cmd.ch */
    int32_t max_code_len; /* longest code table (entries), e.g. 1023, 8184, 10230 */
    int32_t iq_dtype;     /* GB200_IQ_* of the ring */
    int64_t ring_samples; /* ring capacity in complex samples (>= 2 * max_block) */
    int32_t max_block;    /* longest correlation window (samples); init()'s max_signal_length_samples */
    int32_t max_cmds;     /* most commands in one correlate call */
} gb200_trk_config;

/* One Carrier_wipeoff_multicorrelator_resampler call
 * (cpu_multicorrelator_real_codes.cc:103-126): same argument meaning, all float. */
typedef struct
{
    int32_t ch;             /* channel whose code table is used */
    int32_t n_samples;      /* signal_length_samples */
    int32_t n_taps;         /* d_n_correlators, 1..GB200_MAX_TAPS */
    int32_t flags;          /* GB200_TRK_HIGH_DYN: use the high-dynamics resampler/rotator pair */
    uint64_t sample_offset; /* absolute index (since engine creation) of the window's first sample: d_sig_in */
    float rem_carrier_phase_rad;
    float phase_step_rad;
    float phase_rate_step_rad; /* used only with GB200_TRK_HIGH_DYN */
    float rem_code_phase_chips;
    float code_phase_step_chips;
    float code_phase_rate_step_chips; /* used only with GB200_TRK_HIGH_DYN */
    float shifts_chips[GB200_MAX_TAPS]; /* d_shifts_chips, copied on every call (the caller may change them) */
} gb200_trk_cmd;

#define GB200_TRK_HIGH_DYN 1

/* init() — cpu_multicorrelator_real_codes.cc:36-50 */
int gb200_trk_create(const gb200_trk_config* cfg, gb200_trk** out);
/* free() — cpu_multicorrelator_real_codes.cc:140-153 */
int gb200_trk_destroy(gb200_trk* h);
/* set_local_code_and_taps() code part — cpu_multicorrelator_real_codes.cc:53-63.
 * The table is copied to the device (the reference borrows the pointer). */
int gb200_trk_set_code(gb200_trk* h, int ch, const float* code, int code_len);
/* Cpu_Multicorrelator::set_local_code_and_taps — cpu_multicorrelator.cc:48-58 (complex code table) */
int gb200_trk_set_code_complex(gb200_trk* h, int ch, const gb200_cf32* code, int code_len);
/* set_input_output_vectors() input part — cpu_multicorrelator_real_codes.cc:66-72.  Appends n_samples
 * of cfg.iq_dtype to the ring (ONE host->device copy per GPU, shared by all channels);
 * *first_index (may be NULL) receives the absolute index of the first appended sample. */
int gb200_trk_push_iq(gb200_trk* h, const void* iq_host, int64_t n_samples, uint64_t* first_index);
/* same, from a device pointer on cfg.device (device->device copy on the engine's stream) */
int gb200_trk_push_iq_device(gb200_trk* h, const void* iq_dev, int64_t n_samples, uint64_t* first_index);
/* Advance the ring by n_samples without copying anything (the skipped samples read as undefined): used when no
 * channel needs a stretch of the stream, e.g. all channels idle (state 0 of dll_pll_veml_tracking.cc:1746-1752 just
 * consumes its input). */
int gb200_trk_skip(gb200_trk* h, int64_t n_samples);
/* Back-fill a stretch that is already behind the head (e.g. one that gb200_trk_skip jumped over and a slower channel turns
 * out to need): samples of absolute indices [abs_index, abs_index + n_samples), which must still be inside the ring. */
int gb200_trk_write_iq(gb200_trk* h, uint64_t abs_index, const void* iq_host, int64_t n_samples);
/* absolute index one past the newest sample */
uint64_t gb200_trk_head(const gb200_trk* h);

/* ---- raw-sample ingest: the item types of the reference's file sources, converted on the device ----
 * File_Signal_Source item types (src/algorithms/signal_source/adapters/file_signal_source.cc, file_source_base.cc:
 * "gr_complex", "ishort", "ibyte", "short", "byte") and the 2-bit packed stream of Two_Bit_Packed_File_Signal_Source
 * (adapters/two_bit_packed_file_signal_source.cc:33-36,79-91; gnuradio_blocks/unpack_2bit_samples.cc:121-208). */
#define GB200_SRC_CF32 0 /* gr_complex                                        */
#define GB200_SRC_CS16 1 /* ishort: interleaved I,Q int16 (Ishort_To_Complex)  */
#define GB200_SRC_CS8 2  /* ibyte: interleaved I,Q int8 (Ibyte_To_Complex)     */
#define GB200_SRC_S16 3  /* short: real int16, Q = 0                           */
#define GB200_SRC_S8 4   /* byte: real int8, Q = 0                             */
#define GB200_SRC_2BIT 5 /* four 2-bit samples per byte, values -3,-1,+1,+3    */
#define GB200_SRC_LABSAT2 6 /* LabSat 2/3 single-channel payload, 1-bit I + 1-bit Q: 8 samples per int16 word, +-1  */
#define GB200_SRC_LABSAT4 7 /* LabSat 2/3 single-channel payload, 2-bit I + 2-bit Q: 4 samples per int16 word, +-1,+-2 */
/* (labsat23_source::decode_samples_one_channel, src/algorithms/signal_source/gnuradio_blocks/labsat23_source.cc:364-433, fed by
 * general_work's native int16 reads :960-972,1016-1028; the file header (:141-361) is parsed by the caller, who pushes the
 * payload bytes that follow it) */
#define GB200_2BIT_REAL 0 /* sample_type=real: four real samples per byte      */
#define GB200_2BIT_IQ 1   /* sample_type=iq: two complex samples per byte      */
#define GB200_2BIT_QI 2   /* sample_type=qi: same with reverse_interleaving    */
typedef struct gb200_ingest_format
{
    int32_t kind;             /* GB200_SRC_* */
    int32_t sample_type;      /* GB200_SRC_2BIT only: GB200_2BIT_* (two_bit_packed_file_signal_source.cc:79-91) */
    int32_t item_size;        /* GB200_SRC_2BIT only: bytes per file item, 1 2 4 or 8 (:55-77 byte/short/int/long) */
    int32_t big_endian_items; /* GB200_SRC_2BIT only: the bytes of an item are stored most significant first (:34) */
    int32_t big_endian_bytes; /* GB200_SRC_2BIT only: the first sample sits in the top bits of a byte (:35) */
} gb200_ingest_format;
/* Appends the samples held in n_bytes of raw host data to the ring, converting to cfg.iq_dtype in a kernel (the raw bytes
 * are what crosses PCIe).  Integer sources convert without scaling like the reference's adapters; a source wider than the
 * ring type is refused (GB200_EINVAL).  *first_index / *n_samples (may be NULL) receive the absolute index of the first
 * appended sample and how many were appended. */
int gb200_trk_push_raw(gb200_trk* h, const void* raw_host, int64_t n_bytes, const gb200_ingest_format* fmt, uint64_t* first_index, int64_t* n_samples);
/* Read back samples [abs_index, abs_index + n_samples) of the ring in cfg.iq_dtype (diagnostics and tests). */
int gb200_trk_read_iq(gb200_trk* h, uint64_t abs_index, void* iq_host, int64_t n_samples);

/* Carrier_wipeoff_multicorrelator_resampler() for n_cmds commands:
 * out[i * GB200_MAX_TAPS + t] = correlator output of command i, tap t (d_corr_out).
 * Blocking: copies the commands in, launches, copies the results back. */
int gb200_trk_correlate(gb200_trk* h, const gb200_trk_cmd* cmds, int n_cmds, gb200_cf32* out);
/* Split form: enqueue (async copy-in + launch on the engine's stream) ... collect (copy-out + sync). */
int gb200_trk_enqueue(gb200_trk* h, const gb200_trk_cmd* cmds, int n_cmds);
int gb200_trk_collect(gb200_trk* h, gb200_cf32* out);
/* Re-launch the most recently enqueued command set `iters` times with inputs and commands
 * resident on the device; *ms_per_iter = CUDA-event time / iters on the engine's stream. */
int gb200_trk_replay(gb200_trk* h, int iters, float* ms_per_iter);
/* Allocation-free correlate.  NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
 * Same arithmetic as gb200_trk_correlate's general kernel (trk_correlate_kernel: real codes, gr_complex IQ), but with no engine
 * handle: the caller passes DEVICE pointers for the inputs and the output, and everything the kernel needs besides them (the device
 * copy of the commands, the per-chunk partial sums and the per-command arrival counters) lives in ONE file-scope
 * `static __device__` array of GB200_NOALLOC_SCRATCH_BYTES inside the library, held by every GPU that loads the library until exit.
 * The call performs no cudaMalloc / cudaFree / cudaHostAlloc and no other allocating library routine.
 *   iq_dev     n_iq gr_complex samples in device memory; cmds[i].sample_offset indexes it from 0 (window must lie inside [0, n_iq))
 *   codes_dev  n_codes tables of code_len floats each, device memory, table c at codes_dev + c * code_len; cmds[i].ch selects
 *   cmds       HOST array, same fields and meaning as for gb200_trk_correlate (GB200_TRK_HIGH_DYN is rejected)
 *   out_dev    device memory for n_cmds * GB200_MAX_TAPS gb200_cf32; out_dev[i * GB200_MAX_TAPS + t], t < padded tap count
 *              (1, 2, 3, 5, 8, 11 or 16 — the next size up from the largest n_taps), is written, the rest of a slot is left untouched
 *   stream     a cudaStream_t passed as void* (NULL = the legacy default stream); the work is queued on it and the call returns
 *              after synchronising that stream, i.e. with out_dev complete
 * Limits (GB200_EINVAL with a message beyond them): n_cmds <= GB200_NOALLOC_MAX_CMDS, n_samples <= GB200_NOALLOC_MAX_SAMPLES.
 * All calls on a device share the one scratch array, which keeps the previous call's contents; the entry point clears what must
 * start at zero (the arrival counters) on every call.  Calls that OVERLAP on the same GPU (two host threads, or two streams) race on
 * it and produce garbage: serialise them.  A process that never calls this pays only the array's device memory. */
#define GB200_NOALLOC_MAX_CMDS 64
#define GB200_NOALLOC_MAX_CHUNKS 64
#define GB200_NOALLOC_MAX_SAMPLES (GB200_NOALLOC_MAX_CHUNKS * 16384)
#define GB200_NOALLOC_SCRATCH_BYTES (GB200_NOALLOC_MAX_CMDS * 256 + GB200_NOALLOC_MAX_CMDS * 4 + GB200_NOALLOC_MAX_CMDS * GB200_NOALLOC_MAX_CHUNKS * 2 * GB200_MAX_TAPS * 4)
int gb200_trk_correlate_noalloc(int device, const gb200_cf32* iq_dev, int64_t n_iq, const float* codes_dev, int code_len, int n_codes,
    const gb200_trk_cmd* cmds, int n_cmds, gb200_cf32* out_dev, void* stream);
/* name of the kernel variant the last enqueue selected (diagnostics / bench reporting), e.g.
 * "trk_correlate_kernel<T=11,chunks=2>" or "trk_run_kernel<T=11,R=16,G=4>" */
const char* gb200_trk_last_kernel(const gb200_trk* h);

/* ======================================================================
 * Device-resident DLL/PLL loop closure over a tracking engine
 *
 * Replaces, for every channel of a GPU at once, what dll_pll_veml_tracking::general_work does AROUND the correlator call once per
 * code period (src/algorithms/tracking/gnuradio_blocks/dll_pll_veml_tracking.cc): start_tracking :643-883, the pull-in alignment
 * :1757-1788, do_correlation_step's argument casts :1037-1062, save_correlation_results :1263-1373,
 * cn0_and_tracking_lock_status :972-1029, run_dll_pll :1065-1152, update_tracking_vars :1189-1260 and the state machine :1791-2094
 * (0 standby, 1 pull-in, 2 wide tracking + symbol synchronisation, 3 extended integration, 4 narrow tracking), with the channel
 * state resident in HBM: one gb200_loop_step = upload the new IQ block, [correlate kernel -> loop kernel] for every channel whose
 * next window is in the ring, download the Gnss_Synchro records emitted in this step.  No per-channel host work.
 * Served: real-code channels with high_dyn = false — data-component tracking (GPS L1 C/A, GPS L5I, Galileo E1B, Galileo E5aI) and, with
 * track_pilot = 1, pilot tracking (Galileo E1C with E1B as data component, GPS L5Q with L5I, Galileo E5aQ with E5aI): the pilot's taps close the loop, a second one-tap correlator on the data code at the prompt shift supplies the symbols
 * (d_correlator_data_cpu, :526-531,1051-1062; save_correlation_results :1305-1372).  The host block
 * (gnss_sim_receiver_b200/host/b200_dll_pll_veml_tracking.cc) serves the rest (high_dyn).  Galileo E5a-Q, whose secondary code depends on the
 * satellite, takes it per channel through gb200_loop_set_secondary_code.
 * With track_pilot the engine needs max_channels and max_cmds >= 2 * n_channels: channel i's tracking code is the engine's code slot
 * i, its data code slot n_channels + i.
 * ====================================================================== */
typedef struct gb200_loop gb200_loop;

typedef struct
{
    /* the signal table of dll_pll_veml_tracking.cc:171-447 */
    double fs_in;                /* Dll_Pll_Conf::fs_in */
    double code_period_s;        /* d_code_period */
    double code_chip_rate;       /* d_code_chip_rate */
    double signal_carrier_freq;  /* d_signal_carrier_freq */
    int32_t code_length_chips;   /* d_code_length_chips */
    int32_t code_samples_per_chip; /* d_code_samples_per_chip: table entries per chip (2 for sinBOC(1,1)) */
    int32_t vector_length;       /* Dll_Pll_Conf::vector_length: samples correlated per code period */
    int32_t veml;                /* d_veml: 5 taps (VE, E, P, L, VL) instead of 3 */
    int32_t symbols_per_bit;     /* d_symbols_per_bit */
    int32_t secondary;           /* d_secondary: the tracked component carries a secondary code */
    int32_t secondary_code_length; /* d_secondary_code_length (GPS L1 C/A: the 160-symbol preamble pattern used for bit synchronisation) */
    int32_t interchange_iq;      /* d_interchange_iq */
    char secondary_code[192];    /* d_secondary_code_string: '0' / '1' */
    /* Dll_Pll_Conf (src/algorithms/tracking/libs/dll_pll_conf.h:30-88) */
    float pll_bw_hz, dll_bw_hz, fll_bw_hz, pll_bw_narrow_hz, dll_bw_narrow_hz;
    float early_late_space_chips, very_early_late_space_chips, early_late_space_narrow_chips, very_early_late_space_narrow_chips;
    float slope, y_intercept;    /* of |R| at the early-late spacing (1, 1 for BPSK) */
    float carrier_lock_th, cn0_smoother_alpha, carrier_lock_test_smoother_alpha;
    int32_t pll_filter_order, dll_filter_order, extend_correlation_symbols;
    int32_t enable_fll_pull_in, enable_fll_steady_state, carrier_aiding;
    int32_t cn0_samples, cn0_min, max_code_lock_fail, max_carrier_lock_fail;
    int32_t cn0_smoother_samples, carrier_lock_test_smoother_samples;
    uint32_t pull_in_time_s, bit_synchronization_time_limit_s;
    /* extended multicorrelator: total taps per channel, 0 or 3 (5 with veml) = the loop's own E, P, L (VE, VL) only.  With more taps
     * (odd, up to GB200_MAX_TAPS; veml = 0) they are spread uniformly over [-early_late_space, +early_late_space] (the narrow spacing once
     * the loop narrows): E is the first, P the middle, L the last; the others are correlated in the same launch and stay readable on
     * the device through gb200_loop_read_taps (multipath / signal-quality monitors). */
    int32_t n_correlators;
    int32_t track_pilot;         /* Dll_Pll_Conf::track_pilot (see "Served" above) */
    int32_t data_secondary_code_length; /* d_data_secondary_code_length: secondary code of the DATA component, removed from its prompt (track_pilot only) */
    char data_secondary_code[36];       /* d_data_secondary_code_string: '0' / '1' */
} gb200_loop_config;

/* one emitted Gnss_Synchro (what general_work writes to its output, :2063-2090), or a loss of lock (flag_valid_symbol_output = 0) */
typedef struct
{
    int32_t channel;
    uint32_t prn;
    int32_t flag_valid_symbol_output;
    int32_t correlation_length_ms;
    double prompt_i, prompt_q, cn0_db_hz, carrier_doppler_hz, carrier_phase_rads, code_phase_samples;
    uint64_t tracking_sample_counter;
} gb200_loop_record;

/* diagnostics: the loop variables of one channel (d_state, d_carrier_doppler_hz, d_CN0_SNV_dB_Hz, ...) */
typedef struct
{
    int32_t state, pull_in_transitory, current_prn_length_samples, narrow;
    double carrier_doppler_hz, cn0_db_hz, carrier_lock_test, code_freq_chips, rem_code_phase_samples, acc_carrier_phase_rad;
    uint64_t nitems_read;
} gb200_loop_channel_state;

/* `engine`: the tracking engine whose ring, code tables and kernels the loop drives; channel ch of the loop uses the engine's
 * code-table slot ch (gb200_trk_set_code).  The engine must outlive the loop; while a loop is attached, call gb200_trk_push_iq
 * only through gb200_loop_step. */
int gb200_loop_create(gb200_trk* engine, const gb200_loop_config* cfg, int n_channels, gb200_loop** out);
int gb200_loop_destroy(gb200_loop* l);
/* start_tracking() of channel ch (:643-883): acquisition results as left in Gnss_Synchro; `nitems_read` = absolute index of the
 * first sample the channel has not consumed yet (where its block would start reading) */
int gb200_loop_start(gb200_loop* l, int ch, uint32_t prn, double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp_samples, uint64_t nitems_read);
/* stop_tracking(): back to standby */
int gb200_loop_stop(gb200_loop* l, int ch);
/* A secondary code of the TRACKED component that depends on the satellite (Galileo E5a-Q: GALILEO_E5A_Q_SECONDARY_CODE[prn - 1],
 * dll_pll_veml_tracking.cc:733-735 d_secondary_code_string = &GALILEO_E5A_Q_SECONDARY_CODE[PRN - 1]): '0' / '1' characters, as long as
 * gb200_loop_config::secondary_code_length; call before gb200_loop_start of that channel.  Channels without one use the configuration's. */
int gb200_loop_set_secondary_code(gb200_loop* l, int ch, const char* code, int length);
/* One step.  Appends n_samples of the engine's item type from host memory (NULL / 0: none), runs every round that is due and
 * copies the records emitted during this step to `records` (at most max_records; *n_records receives how many were emitted). */
int gb200_loop_step(gb200_loop* l, const void* iq_host, int64_t n_samples, gb200_loop_record* records, int max_records, int* n_records);
/* Pipelined form: gb200_loop_submit enqueues the upload, the rounds and the counters' download and returns at once; gb200_loop_collect waits
 * for the OLDEST submitted step and hands its records over.  Up to two steps may be in flight (the host reads step k while the device runs
 * step k + 1); at most 7 code periods of samples per submit.  `iq_host` must stay valid — and should be page-locked — until collected. */
int gb200_loop_submit(gb200_loop* l, const void* iq_host, int64_t n_samples);
int gb200_loop_collect(gb200_loop* l, gb200_loop_record* records, int max_records, int* n_records);
int gb200_loop_channel(gb200_loop* l, int ch, gb200_loop_channel_state* out);
/* correlator outputs of the last correlated round of channel ch: n_correlators complex values (out[t] = tap t) */
int gb200_loop_read_taps(gb200_loop* l, int ch, gb200_cf32* out);
/* CUDA-event time of the kernels of the last step (correlate + loop closure, without the copies), and how many correlation
 * commands they served */
int gb200_loop_last_step(const gb200_loop* l, float* device_ms, int* n_correlated, int* n_rounds);
/* name of the correlator kernel the loop drives, e.g. "trk_run_kernel<T=3,R=18>" */
const char* gb200_loop_kernel(const gb200_loop* l);

/* ======================================================================
 * PCPS acquisition engine
 *
 * Replaces pcps_acquisition (src/algorithms/acquisition/gnuradio_blocks/pcps_acquisition.{h:92-290,cc:175-899})
 * for all PRNs x Doppler bins of one signal at once:
 *   set_local_code :175-208, update_grid_doppler_wipeoffs :232-245,295-302, acquisition_core :600-697,
 *   max_to_input_power_statistic :496-528, first_vs_second_peak_statistic :531-597, calculate_threshold :884-899.
 * The FFTs (FFTW through gr::fft in the reference) are hand-written mixed-radix kernels; no cuFFT.
 * ====================================================================== */
typedef struct gb200_acq gb200_acq;

typedef struct
{
    int32_t device;
    int32_t iq_dtype;         /* GB200_IQ_*: acq_parameters.it_size / item_type (acq_conf.cc:24-31) */
    int32_t fft_size;         /* d_fft_size           pcps_acquisition.cc:84-93  (any size, like FFTW behind gr::fft: sizes with a prime factor above 61 run on a padded transform that gives the same circular correlation) */
    int32_t consumed_samples; /* d_consumed_samples   pcps_acquisition.cc:84-91  (input zero-padded to fft_size, :612-619) */
    int32_t bit_transition;   /* acq_parameters.bit_transition_flag */
    int32_t use_cfar;         /* d_use_CFAR_algorithm_flag (pfa > 0), pcps_acquisition.cc:81 */
    int32_t samples_per_chip; /* acq_conf.cc:110 */
    int32_t max_dwells;       /* acq_parameters.max_dwells */
    int32_t max_prns;         /* code spectra slots */
    float samples_per_code;   /* acq_conf.cc:116-117 (float) */
    double fs_in;             /* acq_parameters.fs_in */
} gb200_acq_config;

/* what acquisition_core leaves in Gnss_Synchro / the block members (pcps_acquisition.cc:505-527, 691-696) */
typedef struct
{
    uint32_t indext;         /* code-phase index of the maximum, first index wins ties */
    int32_t doppler_index;   /* Doppler bin of the maximum, first bin wins ties */
    int32_t doppler_hz;      /* -doppler_max + doppler_center + doppler_step * doppler_index */
    float peak;              /* grid maximum (d_mag) */
    float input_power;       /* CFAR: sum(grid[opposite bin]) / N_eff / 2 / dwells;  2-peak: second peak */
    float test_statistics;   /* d_test_statistics */
    float acq_delay_samples; /* fmod((float)indext, samples_per_code) */
    float acq_doppler_hz;
} gb200_acq_result;

int gb200_acq_create(const gb200_acq_config* cfg, gb200_acq** out);
int gb200_acq_destroy(gb200_acq* h);
/* set_local_code() for one slot — pcps_acquisition.cc:175-208: `code` holds the sampled local replica
 * (consumed_samples complex values, or fft_size/2 with bit_transition); the device zero-pads, transforms
 * and conjugates it. */
int gb200_acq_set_local_code(gb200_acq* h, int slot, const gb200_cf32* code, int n_code);
/* set_doppler_max/step/center + update_grid_doppler_wipeoffs — pcps_acquisition.cc:248-302 */
int gb200_acq_set_doppler_grid(gb200_acq* h, int doppler_max, int doppler_step, int doppler_center);
/* Explicit list of wipe-off frequencies, one per Doppler bin, as update_grid_doppler_wipeoffs_step2 forms them for the
 * second (fine) step of a two-step acquisition (pcps_acquisition.cc:304-312): floats, not necessarily on an integer grid.
 * Results then carry doppler_index only (doppler_hz and acq_doppler_hz are 0: the caller maps the index). */
int gb200_acq_set_doppler_bins(gb200_acq* h, const float* doppler_hz, int n_bins);
/* Doppler wipe-off law.  0 (default): the reference's — (cosf, sinf) of an fp32 phase accumulated by repeated addition
 * (update_local_carrier :232-245 -> volk_gnsssdr_s32f_sincos_32fc_generic), so the magnitude grid matches the reference's
 * to ~1e-6; 1: exact phase exp(-j*2*pi*f*n/fs) (about 1e-4 of the peak away from the reference's own table, and closer
 * to the float64 truth). */
int gb200_acq_set_wipeoff_mode(gb200_acq* h, int exact);
int gb200_acq_num_doppler_bins(const gb200_acq* h);
int gb200_acq_neff(const gb200_acq* h);
/* One dwell of acquisition_core for every slot in `slots` (NULL = 0..n_slots-1): `iq_host` holds
 * consumed_samples samples of cfg.iq_dtype.  dwell_index 0 overwrites the magnitude grid, > 0 accumulates
 * (pcps_acquisition.cc:657-665); `out[i]` is the statistic after this dwell for slots[i]. */
int gb200_acq_run(gb200_acq* h, const void* iq_host, int dwell_index, const int32_t* slots, int n_slots, gb200_acq_result* out);
/* same with the input already on the device */
int gb200_acq_run_device(gb200_acq* h, const void* iq_dev, int dwell_index, const int32_t* slots, int n_slots, gb200_acq_result* out);
/* Re-run the last gb200_acq_run_device sweep `iters` times, everything resident; event-timed. */
int gb200_acq_replay(gb200_acq* h, int iters, float* ms_per_iter);
/* calculate_threshold — pcps_acquisition.cc:884-899 */
float gb200_acq_threshold(const gb200_acq* h, float pfa);
/* d_magnitude_grid of one slot after the last run: bins x N_eff floats (dump support, pcps_acquisition.cc:393-493).
 * Only available when the engine keeps a grid (max_dwells > 1 or keep_grid set through gb200_acq_keep_grid). */
int gb200_acq_keep_grid(gb200_acq* h, int keep);
int gb200_acq_read_grid(gb200_acq* h, int slot, float* grid_host);
/* Per-(slot, Doppler bin) reductions of the last run, [n_slots][n_bins] each: the maximum over code phase (what
 * volk_gnsssdr_32f_index_max_32u finds in one grid row, pcps_acquisition.cc:505-514), its first index, and the row sum (the CFAR
 * input power of the opposite bin, :518-519).  This is all a rank has to publish when ONE sweep is tiled over several GPUs by
 * Doppler bin (gb200_acq_set_doppler_bins per rank): the merge keeps the reference's first-bin / first-index tie rule. */
int gb200_acq_read_cells(gb200_acq* h, int n_slots, float* vmax, int32_t* imax, float* sum);
/* forward transform of one fft_size block through the engine's FFT (test hook: natural-order output) */
int gb200_acq_fft(gb200_acq* h, const gb200_cf32* in_host, gb200_cf32* out_host, int inverse);
/* Folded input — pcps_quicksync_acquisition_cc (pcps_quicksync_acquisition_cc.cc:296-318): every run takes a block of block_len samples,
 * wipes the WHOLE block off per Doppler bin (phase accumulated over block_len samples) and adds its first `chunks` pieces of fft_size samples,
 * in order, into the one fft_size-point transform input of that bin; the rest of the run (code product, inverse transform, |.|^2, maxima,
 * statistics) is unchanged.  gr_complex engines with consumed_samples == fft_size only; chunks = 0 switches it off.  With folding on,
 * gb200_acq_run / _run_device read block_len samples. */
int gb200_acq_set_input_folding(gb200_acq* h, int chunks, int block_len);
/* The correlations that resolve the folding ambiguity (:337-349), on the block of the last folded run:
 * acc_out[i] = sum_{j < n_code} x[d_i + j] * wipeoff_bin[d_i + j] * code[j],  d_i = index + i * fft_size,  i < n_candidates. */
int gb200_acq_folded_candidates(gb200_acq* h, int bin, int index, const gb200_cf32* code_host, int n_code, int n_candidates, gb200_cf32* acc_out);
/* Non-coherent combination of kept grid rows, per Doppler bin — galileo_e5a_noncoherentIQ_acquisition_caf_cc's "d_magnitudeIA[i] += d_magnitudeQA[i]"
 * then volk_gnsssdr_32f_index_max_32u (galileo_e5a_noncoherent_iq_acquisition_caf_cc.cc:464-467,475-478,494-497,505-508,525-531): for every bin
 * of the last run, row slot_a[bin] + row slot_b[bin] of the kept grid (slot_b[bin] < 0: row slot_a[bin] alone), its maximum and the first
 * index of it.  slot_a / slot_b / vmax / imax: n_bins entries each, host memory.  Needs gb200_acq_keep_grid(h, 1). */
int gb200_acq_grid_pair_max(gb200_acq* h, const int32_t* slot_a, const int32_t* slot_b, float* vmax, int32_t* imax);
/* n single values of the kept grid: value_out[k] = grid[slot[k]][bin[k]][index[k]] (the CAF block reads d_magnitudeIB at the index of
 * another component's maximum, :435) */
int gb200_acq_grid_probe(gb200_acq* h, int n, const int32_t* slot, const int32_t* bin, const int32_t* index, float* value_out);
/* Fine Doppler estimate — pcps_acquisition_fine_doppler_cc::estimate_Doppler (pcps_acquisition_fine_doppler_cc.cc:338-402): the n_x
 * buffered samples x (10 code periods) times the delay-aligned replica rep (one period, repeated: :357-366), transformed zero-padded to
 * pad_factor * n_x points (:340-345,:368), |.|^2 (:371), index of the FIRST maximum (volk_gnsssdr_32f_index_max_32u, :374) in
 * [0, pad_factor * n_x) and its value.  The handle's fft_size must divide n_x (fold = n_x / fft_size <= 64): the padded transform is
 * evaluated as pad_factor * fold transforms of fft_size points on the engine's own FFT (csrc/acq_peak.inc), nothing of the padded
 * length is stored.  Host buffers in, blocking; the handle needs a Doppler grid (any) and a factorable fft_size. */
int gb200_acq_spectrum_peak(gb200_acq* h, const gb200_cf32* x_host, int64_t n_x, const gb200_cf32* rep_host, int rep_len, int pad_factor,
    int64_t* index_out, float* value_out);

#ifdef __cplusplus
}
#endif
#endif /* GNSS_B200_H */
