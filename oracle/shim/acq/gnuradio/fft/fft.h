// Test-infrastructure stub (NOT product code): gr::fft::fft_complex_fwd / _rev (FFTW single precision in GNU Radio) replaced by the
// oracle's mixed-radix DFT (oracle/oracle_fft.c): unnormalised in both directions like FFTW.  oracle_shim_fft_f64 != 0 runs the
// transform in float64 and rounds the result to float32 (the index assertions of the tests must hold for both).
#ifndef ORACLE_SHIM_ACQ_GR_FFT_H
#define ORACLE_SHIM_ACQ_GR_FFT_H
#include <complex>
#include <vector>
extern "C" int oracle_shim_fft_f64;
extern "C" void oracle_shim_fft(const std::complex<float>* in, std::complex<float>* out, int n, int inverse);
namespace gr
{
namespace fft
{
template <bool forward>
class fft_complex_t
{
public:
    explicit fft_complex_t(int size, int nthreads = 1) : d_in(size), d_out(size) { (void)nthreads; }
    std::complex<float>* get_inbuf() { return d_in.data(); }
    std::complex<float>* get_outbuf() { return d_out.data(); }
    int inbuf_length() const { return static_cast<int>(d_in.size()); }
    int outbuf_length() const { return static_cast<int>(d_out.size()); }
    void execute() { oracle_shim_fft(d_in.data(), d_out.data(), static_cast<int>(d_in.size()), forward ? 0 : 1); }

private:
    std::vector<std::complex<float>> d_in, d_out;
};
typedef fft_complex_t<true> fft_complex_fwd;
typedef fft_complex_t<false> fft_complex_rev;
}  // namespace fft
}  // namespace gr
#endif
