/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.
 *
 * CPU restatement (plain C, scalar) of the two hot paths of
 * ShingoNishimoto/gnss_sim_receiver that the CUDA library replaces.  Every
 * function cites the reference file:line it follows (paths relative to the
 * reference root).  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load liboracle.so; the product
 * (gnss_sim_receiver_b200/, include/) never links or calls it.
 *
 * Pinning: the tracking side is checked bit-for-bit against the reference's
 * own code compiled into oracle/_ref/libgnssref.so (tests/test_oracle_pin.py)
 * and against golden vectors generated from it (tests/golden/).  The
 * acquisition side restates pcps_acquisition.cc, whose FFT (FFTW via GNU
 * Radio) and element-wise kernels (upstream VOLK) are NOT in the reference
 * tree.  Round 1: "parity unpinned".  Since round 2 the reference's own
 * pcps_acquisition.cc + acq_conf.cc are compiled UNMODIFIED into oracle/_ref
 * against stand-in declarations (oracle/shim/acq, oracle/ref_acq.cc) with this
 * file's FFT behind gr::fft: the restatement's grids equal the reference
 * block's bit for bit and the frozen vectors are reproduced by the reference
 * block (tests/test_acq_refblock_cpu.py).  FFTW's own rounding stays
 * unpinned: index assertions must hold with the fp32 and the float64 FFT.
 */
#ifndef GNSS_ORACLE_H
#define GNSS_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ---------------- tracking multicorrelator (oracle_trk.c) ---------------- */
/* volk_gnsssdr_32f_xn_resampler_32f_xn_generic — kernels/volk_gnsssdr/volk_gnsssdr_32f_xn_resampler_32f_xn.h:63-80 */
void orc_resampler_32f(float** result, const float* code, float rem_code_phase_chips, float code_phase_step_chips,
    const float* shifts_chips, unsigned code_length, int ntaps, unsigned n);
/* index table only (int32 per tap per sample), same law */
void orc_resampler_index(int32_t* idx, float rem_code_phase_chips, float code_phase_step_chips,
    const float* shifts_chips, unsigned code_length, int ntaps, unsigned n);
/* volk_gnsssdr_32f_xn_high_dynamics_resampler_32f_xn_generic — ..._high_dynamics_resampler_32f_xn.h:67-91 */
void orc_resampler_high_dyn_32f(float** result, const float* code, float rem, float step, float rate,
    const float* shifts_chips, unsigned code_length, int ntaps, unsigned n);
/* volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn_generic — ..._rotator_dot_prod_32fc_xn.h:66-98 */
void orc_rotator_dot_prod_32fc_32f(float* result_ri, const float* in_ri, float inc_re, float inc_im, float* phase_ri,
    const float* const* in_a, int ntaps, unsigned n);
/* volk_gnsssdr_32fc_x2_rotator_dot_prod_32fc_xn_generic — ..._32fc_x2_rotator_dot_prod_32fc_xn.h:67-101 */
void orc_rotator_dot_prod_32fc_x2(float* result_ri, const float* in_ri, float inc_re, float inc_im, float* phase_ri,
    const float* const* in_a_ri, int ntaps, unsigned n);
/* Cpu_Multicorrelator_Real_Codes::Carrier_wipeoff_multicorrelator_resampler —
 * src/algorithms/tracking/libs/cpu_multicorrelator_real_codes.cc:75-126 (fp32, generic protokernels) */
int orc_mc_real_codes(int high_dyn, const float* sig_ri, int n, const float* code, int L, const float* shifts, int ntaps,
    float rem_carr, float phase_step, float phase_rate, float rem_code, float code_step, float code_rate, float* out_ri);
/* Cpu_Multicorrelator::Carrier_wipeoff_multicorrelator_resampler — cpu_multicorrelator.cc:73-100 (complex codes) */
int orc_mc_complex_codes(const float* sig_ri, int n, const float* code_ri, int L, const float* shifts, int ntaps,
    float rem_carr, float phase_step, float rem_code, float code_step, float* out_ri);
/* Same call, float64 "truth": identical fp32 index law, carrier exp(-j(rem + n*step)) evaluated in double from the
 * float arguments, double accumulation.  mimic_inc=1 uses the angle of the fp32-rounded phase increment
 * exp(-j*step) the reference actually multiplies by (cpu_multicorrelator_real_codes.cc:114-123). */
int orc_mc_real_codes_f64(const float* sig_ri, int n, const float* code, int L, const float* shifts, int ntaps,
    float rem_carr, float phase_step, float rem_code, float code_step, int mimic_inc, double* out_ri);
/* Batched open-loop correlation used as the scalar CPU baseline ("port"): nch channels over one IQ ring. */
int orc_mc_batch(const float* ring_ri, int64_t ring_len, int nch, int ntaps, const int64_t* offsets, const int32_t* lens,
    const float* codes, int L, const float* shifts, const float* rem_carr, const float* phase_step, const float* rem_code,
    const float* code_step, float* out_ri, int nthreads);

/* ---------------- PRN / replica generators (oracle_codes.c) ---------------- */
/* gps_l1_ca_code_gen_int — src/algorithms/libs/gps_sdr_signal_replica.cc:25-110 */
void orc_gps_l1_ca_code_gen_int(int32_t* dest1023, int prn, unsigned chip_shift);
void orc_gps_l1_ca_code_gen_float(float* dest1023, int prn, unsigned chip_shift);
/* gps_l1_ca_code_gen_complex_sampled — gps_sdr_signal_replica.cc:145-187 */
void orc_gps_l1_ca_code_gen_complex_sampled(float* dest_ri, int n, unsigned prn, int fs, unsigned chip_shift);
/* resampler() — src/algorithms/libs/gnss_signal_replica.cc:257-291 */
void orc_code_resampler_32f(const float* from, int from_len, float* dest, int dest_len, float fs_in, float fs_out);

/* ---------------- FFT + PCPS acquisition (oracle_fft.c, oracle_acq.c) ---------------- */
/* Unnormalised complex DFT of arbitrary length (mixed radix, Bluestein-free: n must factor into 2,3,5,7).
 * Stand-in for FFTW behind gr::fft::fft_complex_fwd/rev (src/algorithms/libs/gnss_sdr_fft.h:27-61). */
int orc_fft_c32(const float* in_ri, float* out_ri, int n, int inverse);
int orc_fft_c64(const double* in_ri, double* out_ri, int n, int inverse);

typedef struct
{
    int32_t fft_size;           /* d_fft_size          pcps_acquisition.cc:84-93 */
    int32_t consumed;           /* d_consumed_samples  pcps_acquisition.cc:84-91 */
    int32_t num_doppler_bins;   /* pcps_acquisition.cc:261 */
    int32_t bit_transition;     /* acq_parameters.bit_transition_flag */
    int32_t use_cfar;           /* d_use_CFAR_algorithm_flag pcps_acquisition.cc:81 (pfa > 0) */
    int32_t samples_per_chip;   /* acq_conf.cc:110 */
    int32_t doppler_max, doppler_step, doppler_center;
    int32_t max_dwells;
    float samples_per_code;     /* acq_conf.cc:116-117 (float) */
    double fs_in;
    int32_t fft_f64;            /* 0: fp32 FFT, 1: float64 FFT (cross-check oracle) */
    int32_t sincos_arch;        /* reserved */
} orc_acq_conf;

typedef struct
{
    uint32_t indext;            /* pcps_acquisition.cc:505-514 */
    int32_t doppler_index;      /* bin number */
    int32_t doppler_hz;         /* pcps_acquisition.cc:521-522 */
    float peak;                 /* grid maximum (d_mag) */
    float input_power;          /* CFAR: mean of opposite bin / 2 / dwells ; 2-peak: second peak */
    float test_statistics;      /* pcps_acquisition.cc:527 / :596 */
    float acq_delay_samples;    /* pcps_acquisition.cc:693 */
    float acq_doppler_hz;       /* pcps_acquisition.cc:694 */
} orc_acq_result;

/* pcps_acquisition::set_local_code — pcps_acquisition.cc:175-208: code (consumed or fft_size complex) -> conj(FFT) */
int orc_acq_set_local_code(const orc_acq_conf* c, const float* code_ri, float* fft_codes_ri);
/* update_local_carrier / update_grid_doppler_wipeoffs — pcps_acquisition.cc:232-245,295-302; sincos restated from
 * volk_gnsssdr_s32f_sincos_32fc_generic (kernels/volk_gnsssdr/volk_gnsssdr_s32f_sincos_32fc.h:390-401) */
void orc_acq_wipeoff(float* out_ri, int n, float freq, double fs);
/* acquisition_core hot loop — pcps_acquisition.cc:600-672: one dwell, accumulates into grid[bins][N_eff] (dwell 0 overwrites) */
int orc_acq_dwell(const orc_acq_conf* c, const float* in_ri, const float* fft_codes_ri, float* grid, int dwell_index, int nthreads);
/* max_to_input_power_statistic :496-528 / first_vs_second_peak_statistic :531-597 + :691-696 */
int orc_acq_statistic(const orc_acq_conf* c, const float* grid, int num_dwells, orc_acq_result* r);
/* calculate_threshold — pcps_acquisition.cc:884-899 (Boost gamma_p_inv restated by bisection/Newton on the regularised gamma) */
float orc_acq_threshold(const orc_acq_conf* c, float pfa);
int orc_acq_neff(const orc_acq_conf* c);

/* ---- raw-sample conversions in front of the paths (oracle_ingest.c) ---- */
/* unpack_2bit_samples::work — unpack_2bit_samples.cc:121-208; out receives 4 * nbytes values in {-3,-1,1,3} */
void orc_unpack_2bit(const int8_t* in, int64_t nbytes, int item_size, int big_endian_items, int big_endian_bytes, int reverse_interleaving, int8_t* out);
/* integer / float items -> interleaved (re, im) floats, no scaling; kind = GB200_SRC_* 0..4 */
void orc_items_to_complex(const void* in, int64_t n_samples, int kind, float* out);
// labsat23_source::decode_samples_one_channel (labsat23_source.cc:364-433) over n_words int16 words, bits = 2 or 4
void orc_labsat_one_channel(const int16_t* in, int64_t n_words, int bits, float* out);

#ifdef __cplusplus
}
#endif
#endif
