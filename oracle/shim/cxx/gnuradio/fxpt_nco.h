// Test-infrastructure stub (NOT product code): only complex_exp_gen*, which
// the hot path never calls, uses gr::fxpt_nco; a libm stand-in keeps
// gnss_signal_replica.cc linkable.
#ifndef ORACLE_SHIM_FXPT_NCO_H
#define ORACLE_SHIM_FXPT_NCO_H
#include <cmath>
#include <complex>
namespace gr {
class fxpt_nco {
public:
    void set_freq(float f) { d_inc = f; }
    void sincos(std::complex<float>* out, int n, double ampl = 1.0)
    {
        double ph = 0.0;
        for (int i = 0; i < n; i++) { out[i] = std::complex<float>(std::cos(ph) * ampl, std::sin(ph) * ampl); ph += d_inc; }
    }
private:
    double d_inc{0.0};
};
}
#endif
