/* TEST INFRASTRUCTURE ONLY (oracle/_ref build): gr::fft::fft_complex is FFTW3f inside GNU Radio; neither is installed, so the
 * reference's acquisition block is compiled against this stand-in, which evaluates the same un-normalised DFT (forward e^-j,
 * backward e^+j, as FFTW_FORWARD / FFTW_BACKWARD) with the oracle's mixed-radix transform go_fft_f32 (oracle/gnss_oracle.c).
 * Everything around the two transforms is then the reference's own code. */
#ifndef GB_ORACLE_SHIM_GR_FFT_H
#define GB_ORACLE_SHIM_GR_FFT_H
#include <complex>
#include <cstring>
#include <vector>
extern "C" void go_fft_f32(float* data_iq, int n, int sign);
namespace gr { namespace fft {
class fft_complex
{
    std::vector<std::complex<float>> in_, out_;
    int n_;
    bool fwd_;

public:
    fft_complex(int fft_size, bool forward = true, int = 1) : in_(fft_size), out_(fft_size), n_(fft_size), fwd_(forward) {}
    std::complex<float>* get_inbuf() { return in_.data(); }
    std::complex<float>* get_outbuf() { return out_.data(); }
    int inbuf_length() const { return n_; }
    int outbuf_length() const { return n_; }
    void execute()
    {
        std::memcpy(out_.data(), in_.data(), sizeof(std::complex<float>) * n_);
        go_fft_f32(reinterpret_cast<float*>(out_.data()), n_, fwd_ ? -1 : +1);
    }
};
} }
#endif
