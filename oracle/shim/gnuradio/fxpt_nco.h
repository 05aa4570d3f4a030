/* TEST INFRASTRUCTURE ONLY: minimal gr::fxpt_nco so gnss_signal_processing.cc compiles.
 * Only complex_exp_gen() uses it and nothing on the hot path calls that. */
#ifndef GB_ORACLE_SHIM_FXPT_NCO_H
#define GB_ORACLE_SHIM_FXPT_NCO_H
#include <cmath>
#include <complex>
namespace gr {
class fxpt_nco {
    double d_phase = 0.0, d_inc = 0.0;
public:
    void set_freq(float angle_rate) { d_inc = angle_rate; }
    void set_phase(float angle) { d_phase = angle; }
    void sincos(std::complex<float>* out, int n, double ampl = 1.0) {
        for (int i = 0; i < n; i++) { out[i] = std::complex<float>(float(std::cos(d_phase) * ampl), float(std::sin(d_phase) * ampl)); d_phase += d_inc; }
    }
};
}
#endif
