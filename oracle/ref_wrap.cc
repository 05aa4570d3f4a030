// TEST INFRASTRUCTURE ONLY -- never linked into the product library.
//
// extern "C" doorway into the UNMODIFIED reference sources, compiled where they
// lie under /root/reference by oracle/Makefile into oracle/_ref/libgnss_ref.so.
// Used (a) to validate the C restatement in oracle/gnss_oracle.c, (b) to
// generate tests/golden/*.npz (tests/golden/make_golden.py) and (c) as the
// timed CPU baseline ("kind": "reference") of bench.py for the tracking path.
//
// Nothing here restates an algorithm: every function forwards to a reference
// symbol (class Cpu_Multicorrelator_Real_Codes, volk_gnsssdr protokernels,
// PRN code generators, discriminators / loop filters / lock detectors).
#include "cpu_multicorrelator.h"
#include "cpu_multicorrelator_16sc.h"
#include "cpu_multicorrelator_real_codes.h"
#include "beidou_b1i_signal_processing.h"
#include "beidou_b3i_signal_processing.h"
#include "galileo_e1_signal_processing.h"
#include "galileo_e5_signal_processing.h"
#include "glonass_l1_signal_processing.h"
#include "gps_l2c_signal.h"
#include "gps_l5_signal.h"
#include "gps_sdr_signal_processing.h"
#include "lock_detectors.h"
#include "tracking_2nd_DLL_filter.h"
#include "tracking_2nd_PLL_filter.h"
#include "tracking_FLL_PLL_filter.h"
#include "tracking_discriminators.h"
#include "tracking_loop_filter.h"
#include <volk_gnsssdr/volk_gnsssdr.h>
#include <chrono>
#include <complex>
#include <cstring>
#include <thread>
#include <vector>

extern "C" {

/* gb_ref_volk_arch and the stand-alone protokernel wrappers live in ref_volk.c (C) */

size_t volk_gnsssdr_get_alignment(void) { return 64; }
void* volk_gnsssdr_malloc(size_t size, size_t alignment)
{
    size_t a = alignment < 64 ? 64 : alignment;
    // Zero-filled, like the fresh pages a large first allocation gets from the kernel: pcps_acquisition reads d_tmp_buffer beyond
    // the part it has written in first_vs_second_peak_statistic (memcpy of d_fft_size BYTES, :647), so with recycled heap memory
    // the reference's result there is undefined; zeros are the defined behaviour the restatement and the engine reproduce.
    void* p = aligned_alloc(a, (size + a - 1) / a * a);
    if (p) memset(p, 0, (size + a - 1) / a * a);
    return p;
}
void volk_gnsssdr_free(void* p) { free(p); }

void gbref_set_arch(int arch) { gb_ref_volk_arch = arch; }

// ---- Cpu_Multicorrelator_Real_Codes (cpu_multicorrelator_real_codes.cc:62-152) ----
int gbref_multicorrelator_real(const float* sig_in_iq, int n_samples, const float* code, int code_len,
    const float* shifts, int n_taps, float rem_carr_phase_rad, float phase_step_rad, float phase_rate_step_rad,
    float rem_code_phase_chips, float code_phase_step_chips, float code_phase_rate_step_chips,
    int high_dyn, float* corr_out_iq)
{
    Cpu_Multicorrelator_Real_Codes mc;
    std::vector<float> sh(shifts, shifts + n_taps);
    std::vector<std::complex<float>> out(n_taps);
    mc.init(2 * n_samples, n_taps);
    mc.set_high_dynamics_resampler(high_dyn != 0);
    mc.set_local_code_and_taps(code_len, code, sh.data());
    mc.set_input_output_vectors(out.data(), reinterpret_cast<const std::complex<float>*>(sig_in_iq));
    mc.Carrier_wipeoff_multicorrelator_resampler(rem_carr_phase_rad, phase_step_rad, phase_rate_step_rad,
        rem_code_phase_chips, code_phase_step_chips, code_phase_rate_step_chips, n_samples);
    std::memcpy(corr_out_iq, out.data(), sizeof(float) * 2 * n_taps);
    mc.free();
    return 0;
}

// ---- Cpu_Multicorrelator, complex codes (cpu_multicorrelator.cc:99-126) ----
int gbref_multicorrelator_cplx(const float* sig_in_iq, int n_samples, const float* code_iq, int code_len,
    const float* shifts, int n_taps, float rem_carr_phase_rad, float phase_step_rad,
    float rem_code_phase_chips, float code_phase_step_chips, float* corr_out_iq)
{
    Cpu_Multicorrelator mc;
    std::vector<float> sh(shifts, shifts + n_taps);
    std::vector<std::complex<float>> out(n_taps);
    mc.init(2 * n_samples, n_taps);
    mc.set_local_code_and_taps(code_len, reinterpret_cast<const std::complex<float>*>(code_iq), sh.data());
    mc.set_input_output_vectors(out.data(), reinterpret_cast<const std::complex<float>*>(sig_in_iq));
    mc.Carrier_wipeoff_multicorrelator_resampler(rem_carr_phase_rad, phase_step_rad, rem_code_phase_chips, code_phase_step_chips, n_samples);
    std::memcpy(corr_out_iq, out.data(), sizeof(float) * 2 * n_taps);
    mc.free();
    return 0;
}

// ---- Cpu_Multicorrelator_16sc (cpu_multicorrelator_16sc.cc:78-105) ----
int gbref_multicorrelator_16sc(const int16_t* sig_in_iq, int n_samples, const int16_t* code_iq, int code_len,
    const float* shifts, int n_taps, float rem_carr_phase_rad, float phase_step_rad,
    float rem_code_phase_chips, float code_phase_step_chips, int16_t* corr_out_iq)
{
    Cpu_Multicorrelator_16sc mc;
    std::vector<float> sh(shifts, shifts + n_taps);
    std::vector<lv_16sc_t> out(n_taps);
    mc.init(2 * n_samples, n_taps);
    mc.set_local_code_and_taps(code_len, reinterpret_cast<const lv_16sc_t*>(code_iq), sh.data());
    mc.set_input_output_vectors(out.data(), reinterpret_cast<const lv_16sc_t*>(sig_in_iq));
    mc.Carrier_wipeoff_multicorrelator_resampler(rem_carr_phase_rad, phase_step_rad, rem_code_phase_chips, code_phase_step_chips, n_samples);
    std::memcpy(corr_out_iq, out.data(), sizeof(int16_t) * 2 * n_taps);
    mc.free();
    return 0;
}

// ---- PRN code generators ----
void gbref_gps_l1_ca_code_float(float* dest, int prn, unsigned chip_shift) { gps_l1_ca_code_gen_float(dest, prn, chip_shift); }
void gbref_gps_l1_ca_code_sampled(float* dest_iq, unsigned prn, int fs, unsigned chip_shift)
{
    gps_l1_ca_code_gen_complex_sampled(reinterpret_cast<std::complex<float>*>(dest_iq), prn, fs, chip_shift);
}
void gbref_galileo_e1_code_sinboc11_float(float* dest, const char* sig, unsigned prn)
{
    char s[3] = {sig[0], sig[1], 0};
    galileo_e1_code_gen_sinboc11_float(dest, s, prn);
}
void gbref_galileo_e1_code_sampled(float* dest_iq, const char* sig, int cboc, unsigned prn, int fs, unsigned chip_shift, int secondary)
{
    char s[3] = {sig[0], sig[1], 0};
    galileo_e1_code_gen_complex_sampled(reinterpret_cast<std::complex<float>*>(dest_iq), s, cboc != 0, prn, fs, chip_shift, secondary != 0);
}
void gbref_gps_l5i_code_float(float* dest, unsigned prn) { gps_l5i_code_gen_float(dest, prn); }
void gbref_gps_l5q_code_float(float* dest, unsigned prn) { gps_l5q_code_gen_float(dest, prn); }
void gbref_gps_l5i_code_sampled(float* dest_iq, unsigned prn, int fs) { gps_l5i_code_gen_complex_sampled(reinterpret_cast<std::complex<float>*>(dest_iq), prn, fs); }
void gbref_gps_l2c_m_code_float(float* dest, unsigned prn) { gps_l2c_m_code_gen_float(dest, prn); }
void gbref_galileo_e5a_code_primary(float* dest_iq, int prn, const char* sig)
{
    char s[3] = {sig[0], sig[1], 0};
    galileo_e5_a_code_gen_complex_primary(reinterpret_cast<std::complex<float>*>(dest_iq), prn, s);
}
void gbref_galileo_e5a_code_sampled(float* dest_iq, const char* sig, unsigned prn, int fs, unsigned chip_shift)
{
    char s[3] = {sig[0], sig[1], 0};
    galileo_e5_a_code_gen_complex_sampled(reinterpret_cast<std::complex<float>*>(dest_iq), s, prn, fs, chip_shift);
}
void gbref_beidou_b1i_code_float(float* dest, int prn, unsigned chip_shift) { beidou_b1i_code_gen_float(dest, prn, chip_shift); }
void gbref_beidou_b3i_code_float(float* dest, int prn, unsigned chip_shift) { beidou_b3i_code_gen_float(dest, prn, chip_shift); }
void gbref_beidou_b1i_code_sampled(float* dest_iq, unsigned prn, int fs, unsigned chip_shift) { beidou_b1i_code_gen_complex_sampled(reinterpret_cast<std::complex<float>*>(dest_iq), prn, fs, chip_shift); }
void gbref_glonass_l1_ca_code(float* dest_iq, unsigned chip_shift) { glonass_l1_ca_code_gen_complex(reinterpret_cast<std::complex<float>*>(dest_iq), chip_shift); }
void gbref_glonass_l1_ca_code_sampled(float* dest_iq, int fs, unsigned chip_shift) { glonass_l1_ca_code_gen_complex_sampled(reinterpret_cast<std::complex<float>*>(dest_iq), fs, chip_shift); }

// ---- discriminators / lock detectors (tracking_discriminators.cc:49-128, lock_detectors.cc:68-111) ----
double gbref_fll_four_quadrant_atan(float i1, float q1, float i2, float q2, double t1, double t2) { return fll_four_quadrant_atan(gr_complex(i1, q1), gr_complex(i2, q2), t1, t2); }
double gbref_pll_four_quadrant_atan(float i, float q) { return pll_four_quadrant_atan(gr_complex(i, q)); }
double gbref_pll_cloop_two_quadrant_atan(float i, float q) { return pll_cloop_two_quadrant_atan(gr_complex(i, q)); }
double gbref_dll_nc_e_minus_l_normalized(float ei, float eq, float li, float lq) { return dll_nc_e_minus_l_normalized(gr_complex(ei, eq), gr_complex(li, lq)); }
double gbref_dll_nc_vemlp_normalized(const float* ve, const float* e, const float* l, const float* vl)
{
    return dll_nc_vemlp_normalized(gr_complex(ve[0], ve[1]), gr_complex(e[0], e[1]), gr_complex(l[0], l[1]), gr_complex(vl[0], vl[1]));
}
float gbref_cn0_svn_estimator(const float* prompt_iq, int length, double coh_time_s) { return cn0_svn_estimator(reinterpret_cast<const gr_complex*>(prompt_iq), length, coh_time_s); }
float gbref_carrier_lock_detector(float* prompt_iq, int length) { return carrier_lock_detector(reinterpret_cast<gr_complex*>(prompt_iq), length); }

// ---- loop filters: run a whole input sequence through a fresh filter ----
void gbref_tracking_loop_filter_run(float update_interval, float bw, int order, int include_last_integrator, float init_out, const float* in, float* out, int n)
{
    Tracking_loop_filter f(update_interval, bw, order, include_last_integrator != 0);
    f.initialize(init_out);
    for (int i = 0; i < n; i++) out[i] = f.apply(in[i]);
}
void gbref_fll_pll_filter_run(float fll_bw, float pll_bw, int order, float acq_doppler_hz, const float* fll_disc, const float* pll_disc, float corr_time_s, float* out, int n)
{
    Tracking_FLL_PLL_filter f;
    f.set_params(fll_bw, pll_bw, order);
    f.initialize(acq_doppler_hz);
    for (int i = 0; i < n; i++) out[i] = f.get_carrier_error(fll_disc[i], pll_disc[i], corr_time_s);
}

// the older second-order filters (GLONASS / C_Aid trackers): which = 0 Tracking_2nd_DLL_filter, 1 Tracking_2nd_PLL_filter
void gbref_2nd_filter_run(int which, float bw_hz, float pdi_s, const float* disc, float* out, int n)
{
    if (which == 0)
        {
            Tracking_2nd_DLL_filter f(pdi_s);
            f.set_DLL_BW(bw_hz);
            f.initialize();
            for (int i = 0; i < n; i++) out[i] = f.get_code_nco(disc[i]);
        }
    else
        {
            Tracking_2nd_PLL_filter f(pdi_s);
            f.set_PLL_BW(bw_hz);
            f.initialize();
            for (int i = 0; i < n; i++) out[i] = f.get_carrier_nco(disc[i]);
        }
}

// ---- timed CPU baseline: the shape of cpu_multicorrelator_real_codes_test.cc:133-163, but with the
//      block's default kernel pair (standard resampler + standard rotator, dll_pll_veml_tracking.cc:419) ----
double gbref_bench_multicorrelator_real(const float* sig_in_iq, int n_samples, const float* code, int code_len,
    const float* shifts, int n_taps, int n_threads, int iters_per_thread, int avx, const float* data_code)
{
    gb_ref_volk_arch = avx;
    std::vector<std::thread> th;
    std::vector<double> sink(n_threads, 0.0);
    auto t0 = std::chrono::steady_clock::now();
    for (int t = 0; t < n_threads; t++)
        th.emplace_back([&, t]() {
            Cpu_Multicorrelator_Real_Codes mc;
            std::vector<float> sh(shifts, shifts + n_taps);
            std::vector<std::complex<float>> out(n_taps);
            mc.init(2 * n_samples, n_taps);
            mc.set_high_dynamics_resampler(false);
            mc.set_local_code_and_taps(code_len, code, sh.data());
            mc.set_input_output_vectors(out.data(), reinterpret_cast<const std::complex<float>*>(sig_in_iq));
            // pilot tracking: second 1-tap instance on the data code with identical NCO parameters
            // (correlator_data_cpu, dll_pll_veml_tracking.cc:900-910)
            Cpu_Multicorrelator_Real_Codes md;
            float dshift[1] = {shifts[n_taps / 2]};
            std::complex<float> dout[1];
            if (data_code)
                {
                    md.init(2 * n_samples, 1);
                    md.set_high_dynamics_resampler(false);
                    md.set_local_code_and_taps(code_len, data_code, dshift);
                    md.set_input_output_vectors(dout, reinterpret_cast<const std::complex<float>*>(sig_in_iq));
                }
            float step = float(code_len) / float(n_samples);
            for (int i = 0; i < iters_per_thread; i++)
                {
                    mc.Carrier_wipeoff_multicorrelator_resampler(0.1f * float(i % 7), 0.0123f + 1e-4f * float(t), 0.0f, 0.25f * float(i % 3), step, 0.0f, n_samples);
                    sink[t] += out[n_taps / 2].real();
                    if (data_code)
                        {
                            md.Carrier_wipeoff_multicorrelator_resampler(0.1f * float(i % 7), 0.0123f + 1e-4f * float(t), 0.0f, 0.25f * float(i % 3), step, 0.0f, n_samples);
                            sink[t] += dout[0].real();
                        }
                }
            mc.free();
            if (data_code) md.free();
        });
    for (auto& x : th) x.join();
    auto t1 = std::chrono::steady_clock::now();
    volatile double s = 0;
    for (double v : sink) s += v;
    return std::chrono::duration<double>(t1 - t0).count();
}

}  // extern "C"
