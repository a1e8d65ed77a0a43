// Test-infrastructure stub (NOT product code): boost::math::gamma_p_inv(a, p) — the x with P(a, x) = p, P the regularised lower
// incomplete gamma function — by bisection + Newton on a series / continued-fraction evaluation of P (float64; Boost is not installed).
// Only pcps_acquisition::calculate_threshold uses it; the product's threshold is checked against scipy.special.gammaincinv separately.
#ifndef ORACLE_SHIM_ACQ_BOOST_GAMMA_H
#define ORACLE_SHIM_ACQ_BOOST_GAMMA_H
#include <cmath>
namespace boost
{
namespace math
{
inline double oracle_gamma_p(double a, double x)
{
    if (x <= 0.0) return 0.0;
    const double lg = std::lgamma(a);
    if (x < a + 1.0)
        {
            double ap = a, sum = 1.0 / a, del = sum;
            for (int n = 0; n < 2000; n++)
                {
                    ap += 1.0;
                    del *= x / ap;
                    sum += del;
                    if (std::fabs(del) < std::fabs(sum) * 1e-17) break;
                }
            return sum * std::exp(-x + a * std::log(x) - lg);
        }
    double b = x + 1.0 - a, c = 1.0 / 1e-300, d = 1.0 / b, h = d;
    for (int i = 1; i < 2000; i++)
        {
            const double an = -i * (i - a);
            b += 2.0;
            d = an * d + b;
            if (std::fabs(d) < 1e-300) d = 1e-300;
            c = b + an / c;
            if (std::fabs(c) < 1e-300) c = 1e-300;
            d = 1.0 / d;
            const double del = d * c;
            h *= del;
            if (std::fabs(del - 1.0) < 1e-17) break;
        }
    return 1.0 - std::exp(-x + a * std::log(x) - lg) * h;
}
inline double gamma_p_inv(double a, double p)
{
    if (p <= 0.0) return 0.0;
    double lo = 0.0, hi = a + 10.0;
    while (oracle_gamma_p(a, hi) < p && hi < 1e12) hi *= 2.0;
    for (int i = 0; i < 200; i++)
        {
            const double mid = 0.5 * (lo + hi);
            if (oracle_gamma_p(a, mid) < p)
                lo = mid;
            else
                hi = mid;
            if (hi - lo <= 1e-15 * hi) break;
        }
    return 0.5 * (lo + hi);
}
}  // namespace math
}  // namespace boost
#endif
