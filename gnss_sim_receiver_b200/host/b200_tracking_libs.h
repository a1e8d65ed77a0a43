// Scalar loop-closure pieces of the DLL/PLL tracking block: O(1) work per channel per code period, host side.
// Same names, argument meaning and arithmetic types (float/double) as the reference's tracking/libs:
//   tracking_discriminators.cc:27-149, tracking_loop_filter.cc:26-265, tracking_FLL_PLL_filter.cc:23-99,
//   lock_detectors.cc:54-148, exponential_smoother.cc:23-115.
// tests/test_host_cpu.py checks every function against the reference's own objects (built by the test suite) and against the
// expected values of the reference's unit tests (discriminator_test.cc, tracking_loop_filter_test.cc).
#ifndef GB200_TRACKING_LIBS_H
#define GB200_TRACKING_LIBS_H

#include <complex>
#include <vector>

namespace gb200h
{
using cf32 = std::complex<float>;

// ---- discriminators (tracking_discriminators.cc)
double phase_unwrap(double phase_rad);
double fll_four_quadrant_atan(cf32 prompt_s1, cf32 prompt_s2, double t1, double t2);
double fll_diff_atan(cf32 prompt_s1, cf32 prompt_s2, double t1, double t2);
double pll_four_quadrant_atan(cf32 prompt_s1);
double pll_cloop_two_quadrant_atan(cf32 prompt_s1);
double dll_nc_e_minus_l_normalized(cf32 early_s1, cf32 late_s1, float spc = 0.5, float slope = 1.0, float y_intercept = 1.0);
double dll_nc_vemlp_normalized(cf32 very_early_s1, cf32 early_s1, cf32 late_s1, cf32 very_late_s1);
// sinBOC(1,1) correlation function, slope and y-intercept of |R| at offset x (tracking_discriminators.h:117-186)
double sinboc11_correlation(double offset_in_chips);
double sinboc11_slope_abs(double x);
double sinboc11_y_intercept_abs(double x);

// ---- lock detectors (lock_detectors.cc)
float cn0_svn_estimator(const cf32* Prompt_buffer, int length, float coh_integration_time_s);
float cn0_m2m4_estimator(const cf32* Prompt_buffer, int length, float coh_integration_time_s);
float carrier_lock_detector(const cf32* Prompt_buffer, int length);

// ---- DLL loop filter (tracking_loop_filter.cc): 1st/2nd/3rd order, bilinear integrators, 4-deep history ring
class Tracking_loop_filter
{
public:
    Tracking_loop_filter();
    Tracking_loop_filter(float update_interval, float noise_bandwidth, int loop_order = 2, bool include_last_integrator = false);
    float get_noise_bandwidth() const { return bw_; }
    float get_update_interval() const { return T_; }
    bool get_include_last_integrator() const { return last_integrator_; }
    int get_order() const { return order_; }
    void set_noise_bandwidth(float noise_bandwidth);
    void set_update_interval(float update_interval);
    void set_include_last_integrator(bool include_last_integrator);
    void set_order(int loop_order);
    void initialize(float initial_output = 0.0);
    float apply(float current_input);

private:
    static constexpr int HISTORY = 4;
    void update_coefficients();
    std::vector<float> in_coef_, out_coef_;
    float x_[HISTORY]{}, y_[HISTORY]{};
    float bw_, T_;
    int order_, head_{0};
    bool last_integrator_;
};

// ---- carrier loop filter: 2nd-order PLL + 1st-order FLL assist, or 3rd + 2nd (tracking_FLL_PLL_filter.cc)
class Tracking_FLL_PLL_filter
{
public:
    void set_params(float fll_bw_hz, float pll_bw_hz, int order);
    void initialize(float d_acq_carrier_doppler_hz);
    float get_carrier_error(float FLL_discriminator, float PLL_discriminator, float correlation_time_s);

private:
    float w_{0}, x_{0}, w0p_{0}, w0p2_{0}, w0p3_{0}, w0f_{0}, w0f2_{0}, a2_{0}, a3_{0}, b3_{0};
    int order_{0};
};

// ---- CN0 / lock-test smoother (exponential_smoother.cc): mean of the first N samples, then a one-pole filter
class Exponential_Smoother
{
public:
    void set_alpha(float alpha);
    void set_samples_for_initialization(int num_samples);
    void reset();
    void set_min_value(float value) { min_value_ = value; }
    void set_offset(float offset) { offset_ = offset; }
    float smooth(float raw);
    double smooth(double raw) { return static_cast<double>(smooth(static_cast<float>(raw))); }

private:
    std::vector<float> init_buffer_;
    float alpha_{0.001F}, one_minus_alpha_{0.999F}, old_value_{0.0F}, min_value_{25.0F}, offset_{12.0F};
    int samples_for_initialization_{200}, init_counter_{0};
    bool initializing_{true};
};
}  // namespace gb200h
#endif
