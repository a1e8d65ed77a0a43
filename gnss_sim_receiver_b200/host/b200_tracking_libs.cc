#include "b200_tracking_libs.h"
#include <cmath>
#include <numeric>

namespace gb200h
{
namespace
{
// pi as the reference defines it for GNSS use (IS-GPS-200 value; MATH_CONSTANTS.h:47-49): NOT M_PI, and the loop must agree with it
constexpr double PI_D = 3.1415926535898;
constexpr double HALF_PI_D = PI_D / 2.0;
}  // namespace

// ------------------------------------------------------------------ discriminators
double phase_unwrap(double phase_rad)
{
    if (phase_rad >= HALF_PI_D) return phase_rad - PI_D;
    if (phase_rad <= -HALF_PI_D) return phase_rad + PI_D;
    return phase_rad;
}

// The reference calls gr::fast_atan2f (GNU Radio's table-driven approximation, |err| < 1e-3 rad... documented as ~0.01 deg);
// libm atan2f is used here.
double fll_four_quadrant_atan(cf32 a, cf32 b, double t1, double t2)
{
    const float dot = a.real() * b.real() + a.imag() * b.imag();
    const float cross = a.real() * b.imag() - b.real() * a.imag();
    return std::atan2(cross, dot) / (t2 - t1);
}

double fll_diff_atan(cf32 a, cf32 b, double t1, double t2)
{
    // float divisions and float atan, the difference promoted to double: the reference's operand types
    double d = std::atan(b.imag() / b.real()) - std::atan(a.imag() / a.real());
    if (std::isnan(d)) d = 0;
    return phase_unwrap(d) / (t2 - t1);
}

double pll_four_quadrant_atan(cf32 p) { return std::atan2(p.imag(), p.real()); }

double pll_cloop_two_quadrant_atan(cf32 p)
{
    if (p.real() != 0.0) return static_cast<double>(std::atan(p.imag() / p.real()));
    return 0.0;
}

double dll_nc_e_minus_l_normalized(cf32 e, cf32 l, float spc, float slope, float y_intercept)
{
    const double pe = std::abs(e), pl = std::abs(l);
    const double sum = pe + pl;
    if (sum == 0.0) return 0.0;
    return ((y_intercept - slope * spc) / slope) * (pe - pl) / sum;
}

double dll_nc_vemlp_normalized(cf32 ve, cf32 e, cf32 l, cf32 vl)
{
    const double early = std::sqrt(ve.real() * ve.real() + ve.imag() * ve.imag() + e.real() * e.real() + e.imag() * e.imag());
    const double late = std::sqrt(l.real() * l.real() + l.imag() * l.imag() + vl.real() * vl.real() + vl.imag() * vl.imag());
    const double sum = early + late;
    if (sum == 0.0) return 0.0;
    return (early - late) / sum;
}

// Sousa & Nunes (NAVIGATION, 2013) closed form for sinBOC(1,1): 2p = 2
double sinboc11_correlation(double offset_in_chips)
{
    const double tau = std::abs(offset_in_chips);
    if (tau > 1.0) return 0.0;
    const int k = static_cast<int>(std::ceil(2 * tau));
    const double sgn = (k & 1) ? -1.0 : 1.0;
    return sgn * (2.0 * (k * k - k * 2 - k) / 2 + 1.0 + (2 * 2 - 2 * k + 1) * tau);
}

double sinboc11_slope_abs(double x)
{
    constexpr double dx = 1e-6;
    return (std::abs(sinboc11_correlation(x + dx / 2.0)) - std::abs(sinboc11_correlation(x - dx / 2.0))) / dx;
}

double sinboc11_y_intercept_abs(double x) { return std::abs(sinboc11_correlation(x)) - sinboc11_slope_abs(x) * x; }

// ------------------------------------------------------------------ lock detectors
float cn0_svn_estimator(const cf32* p, int length, float T)
{
    float psig = 0.0F, ptot = 0.0F;
    for (int i = 0; i < length; i++)
        {
            psig += std::abs(p[i].real());
            ptot += p[i].imag() * p[i].imag() + p[i].real() * p[i].real();
        }
    psig /= static_cast<float>(length);
    psig = psig * psig;
    ptot /= static_cast<float>(length);
    const float snr = psig / (ptot - psig);
    return 10.0F * std::log10(snr) - 10.0F * std::log10(T);
}

float cn0_m2m4_estimator(const cf32* p, int length, float T)
{
    float psig = 0.0F, m2 = 0.0F, m4 = 0.0F;
    const auto n = static_cast<float>(length);
    for (int i = 0; i < length; i++)
        {
            psig += std::abs(p[i].real());
            const float pw = p[i].imag() * p[i].imag() + p[i].real() * p[i].real();
            m2 += pw;
            m4 += pw * pw;
        }
    psig /= n;
    psig = psig * psig;
    m2 /= n;
    m4 /= n;
    const float root = std::sqrt(2.0F * m2 * m2 - m4);
    // negative radicand (short or noisy buffers): fall back to the signal-to-noise-variance form
    const float snr = std::isnan(root) ? psig / (m2 - psig) : root / (m2 - root);
    return 10.0F * std::log10(snr) - 10.0F * std::log10(T);
}

float carrier_lock_detector(const cf32* p, int length)
{
    float si = 0.0F, sq = 0.0F;
    for (int i = 0; i < length; i++)
        {
            si += p[i].real();
            sq += p[i].imag();
        }
    const float nbp = si * si + sq * sq;
    const float nbd = si * si - sq * sq;
    return nbd / nbp;
}

// ------------------------------------------------------------------ DLL loop filter
Tracking_loop_filter::Tracking_loop_filter() : bw_(15.0F), T_(0.001F), order_(2), last_integrator_(false) { update_coefficients(); }

Tracking_loop_filter::Tracking_loop_filter(float update_interval, float noise_bandwidth, int loop_order, bool include_last_integrator)
    : bw_(noise_bandwidth), T_(update_interval), order_(loop_order), last_integrator_(include_last_integrator)
{
    update_coefficients();
}

void Tracking_loop_filter::set_noise_bandwidth(float v)
{
    bw_ = v;
    update_coefficients();
}
void Tracking_loop_filter::set_update_interval(float v)
{
    T_ = v;
    update_coefficients();
}
void Tracking_loop_filter::set_include_last_integrator(bool v)
{
    last_integrator_ = v;
    update_coefficients();
}
void Tracking_loop_filter::set_order(int loop_order)
{
    if (loop_order < 1 || loop_order > 3) return;  // the reference logs a warning and keeps the current order
    order_ = loop_order;
    update_coefficients();
}

void Tracking_loop_filter::initialize(float initial_output)
{
    for (int i = 0; i < HISTORY; i++)
        {
            x_[i] = 0.0F;
            y_[i] = initial_output;
        }
    head_ = HISTORY - 1;
}

// Direct-form difference equation; history rings are indexed newest-first from head_.  Coefficients are stored as float
// after being formed in double where the reference's expressions mix float and double literals (T / 2.0, ...).
float Tracking_loop_filter::apply(float current_input)
{
    float result = 0.0F;
    for (size_t i = 0; i < out_coef_.size(); i++) result += out_coef_[i] * y_[(head_ + i) % HISTORY];
    head_ = (head_ + HISTORY - 1) % HISTORY;
    x_[head_] = current_input;
    for (size_t i = 0; i < in_coef_.size(); i++) result += in_coef_[i] * x_[(head_ + i) % HISTORY];
    y_[head_] = result;
    return result;
}

void Tracking_loop_filter::update_coefficients()
{
    const float T = T_;
    const float zeta = 1.0F / std::sqrt(2.0F);
    float g1, g2, g3, wn;
    switch (order_)
        {
        case 1:
            wn = bw_ * 4.0F;
            g1 = wn;
            if (last_integrator_)
                {
                    in_coef_ = {static_cast<float>(g1 * T / 2.0), static_cast<float>(g1 * T / 2.0)};
                    out_coef_ = {1.0F};
                }
            else
                {
                    in_coef_ = {g1};
                    out_coef_.clear();
                }
            break;
        case 2:
            wn = bw_ * (8.0F * zeta) / (4.0F * zeta * zeta + 1.0F);
            g1 = wn * wn;
            g2 = wn * 2.0F * zeta;
            if (last_integrator_)
                {
                    in_coef_ = {static_cast<float>(T / 2.0 * (g1 * T / 2.0 + g2)), static_cast<float>(T * T / 2.0 * g1), static_cast<float>(T / 2.0 * (g1 * T / 2.0 - g2))};
                    out_coef_ = {2.0F, -1.0F};
                }
            else
                {
                    in_coef_ = {static_cast<float>(g1 * T / 2.0 + g2), static_cast<float>(g1 * T / 2.0 - g2)};
                    out_coef_ = {1.0F};
                }
            break;
        case 3:
        default:
            {
                wn = bw_ / 0.7845F;  // Kaplan
                const float a3 = 1.1F, b3 = 2.4F;
                g1 = wn * wn * wn;
                g2 = a3 * wn * wn;
                g3 = b3 * wn;
                if (last_integrator_)
                    {
                        in_coef_ = {static_cast<float>(T / 2.0 * (g3 + T / 2.0 * (g2 + T / 2.0 * g1))), static_cast<float>(T / 2.0 * (-g3 + T / 2.0 * (g2 + 3.0 * T / 2.0 * g1))),
                            static_cast<float>(T / 2.0 * (-g3 - T / 2.0 * (g2 - 3.0 * T / 2.0 * g1))), static_cast<float>(T / 2.0 * (g3 - T / 2.0 * (g2 - T / 2.0 * g1)))};
                        out_coef_ = {3.0F, -3.0F, 1.0F};
                    }
                else
                    {
                        in_coef_ = {static_cast<float>(g3 + T / 2.0 * (g2 + T / 2.0 * g1)), static_cast<float>(g1 * T * T / 2.0 - 2.0 * g3), static_cast<float>(g3 + T / 2.0 * (-g2 + T / 2.0 * g1))};
                        out_coef_ = {2.0F, -1.0F};
                    }
                break;
            }
        }
}

// ------------------------------------------------------------------ carrier loop filter (Kaplan 2nd ed. fig. 5.x: FLL-assisted PLL)
void Tracking_FLL_PLL_filter::set_params(float fll_bw_hz, float pll_bw_hz, int order)
{
    order_ = order;
    a2_ = 1.414F;
    if (order_ == 3)
        {
            b3_ = 2.400F;
            a3_ = 1.100F;
            w0p_ = pll_bw_hz / 0.7845F;
            w0p2_ = w0p_ * w0p_;
            w0p3_ = w0p2_ * w0p_;
            w0f_ = fll_bw_hz / 0.53F;
            w0f2_ = w0f_ * w0f_;
        }
    else
        {
            w0p_ = pll_bw_hz / 0.53F;
            w0p2_ = w0p_ * w0p_;
            w0f_ = fll_bw_hz / 0.25F;
        }
}

void Tracking_FLL_PLL_filter::initialize(float doppler_hz)
{
    if (order_ == 3)
        {
            x_ = 2.0F * doppler_hz;
            w_ = 0;
        }
    else
        {
            w_ = doppler_hz;
            x_ = 0;
        }
}

float Tracking_FLL_PLL_filter::get_carrier_error(float fll, float pll, float Tc)
{
    if (order_ == 3)
        {
            w_ = w_ + Tc * (w0p3_ * pll + w0f2_ * fll);
            x_ = x_ + Tc * (0.5F * w_ + a2_ * w0f_ * fll + a3_ * w0p2_ * pll);
            return 0.5F * x_ + b3_ * w0p_ * pll;
        }
    const float w_new = w_ + pll * w0p2_ * Tc + fll * w0f_ * Tc;
    const float out = 0.5F * (w_new + w_) + a2_ * w0p_ * pll;
    w_ = w_new;
    return out;
}

// ------------------------------------------------------------------ smoother
void Exponential_Smoother::set_alpha(float alpha)
{
    alpha_ = alpha < 0 ? 0 : (alpha > 1 ? 1 : alpha);
    one_minus_alpha_ = 1.0F - alpha_;
}

void Exponential_Smoother::set_samples_for_initialization(int num_samples)
{
    samples_for_initialization_ = num_samples <= 0 ? 1 : num_samples;
    init_buffer_.reserve(samples_for_initialization_);
}

void Exponential_Smoother::reset()
{
    initializing_ = true;
    init_counter_ = 0;
    init_buffer_.clear();
}

float Exponential_Smoother::smooth(float raw)
{
    if (!initializing_)
        {
            old_value_ = alpha_ * raw + one_minus_alpha_ * old_value_;
            return old_value_;
        }
    init_buffer_.push_back(raw);
    if (++init_counter_ == samples_for_initialization_)
        {
            old_value_ = std::accumulate(init_buffer_.begin(), init_buffer_.end(), 0.0F) / static_cast<float>(init_buffer_.size());
            if (old_value_ < min_value_ + offset_)  // mean too low to be trusted: start the warm-up again
                {
                    init_counter_ = 0;
                    init_buffer_.clear();
                }
            else
                {
                    initializing_ = false;
                }
        }
    return raw;
}
}  // namespace gb200h
