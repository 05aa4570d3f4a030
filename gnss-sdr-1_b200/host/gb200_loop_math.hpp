// Host-side scalar loop maths of the tracking adapters: discriminators, loop filters, CN0 / lock detectors.
// Mirrors (semantics and number types, not code) the reference's scalar tracking libraries so that an adapter built on
// the B200 correlator closes the loop exactly like dll_pll_veml_tracking does:
//   src/algorithms/tracking/libs/tracking_discriminators.cc:49-128   (arguments are complex<float>; which operations run
//                                                                      in float and which in double is kept)
//   src/algorithms/tracking/libs/tracking_loop_filter.cc:74-225      (Kaplan & Hegarty table 5.6 loops, bilinear integrators)
//   src/algorithms/tracking/libs/tracking_FLL_PLL_filter.cc:54-133   (FLL-assisted PLL, 2nd / 3rd order)
//   src/algorithms/tracking/libs/lock_detectors.cc:68-111            (SNV C/N0 estimator, NBD/NBP carrier lock test)
// Pinned by tests/test_host_cpu.py against tests/golden/loop_golden.npz, which was generated from the reference's own
// sources (oracle/_ref).  Header-only, no dependencies beyond the C++ standard library.
#pragma once
#include <array>
#include <cmath>
#include <complex>

namespace gb200
{
using cfloat = std::complex<float>;
constexpr double kTwoPi = 6.283185307179586476925286766559;

// ---- discriminators ------------------------------------------------------------------------------------------------
// tracking_discriminators.cc:49-55: dot and cross are formed in float, the angle in double
inline double fll_four_quadrant_atan(cfloat p1, cfloat p2, double t1, double t2)
{
    const double dot = p1.real() * p2.real() + p1.imag() * p2.imag();
    const double cross = p1.real() * p2.imag() - p2.real() * p1.imag();
    return std::atan2(cross, dot) / (t2 - t1);
}
// :62-65
inline double pll_four_quadrant_atan(cfloat p) { return static_cast<double>(std::atan2(p.imag(), p.real())); }
// :72-79  Costas discriminator (float division and float atan)
inline double pll_cloop_two_quadrant_atan(cfloat p)
{
    if (p.real() != 0.0f) return static_cast<double>(std::atan(p.imag() / p.real()));
    return 0.0;
}
// :88-98
inline double dll_nc_e_minus_l_normalized(cfloat early, cfloat late)
{
    const double pe = std::abs(early), pl = std::abs(late);
    if (pe + pl == 0.0) return 0.0;
    return 0.5 * (pe - pl) / (pe + pl);
}
// :107-117
inline double dll_nc_vemlp_normalized(cfloat ve, cfloat e, cfloat l, cfloat vl)
{
    const double pe = std::sqrt(std::norm(ve) + std::norm(e));
    const double pl = std::sqrt(std::norm(vl) + std::norm(l));
    if (pe + pl == 0.0) return 0.0;
    return (pe - pl) / (pe + pl);
}

// ---- generic loop filter (DLL) -------------------------------------------------------------------------------------
// y[n] = sum_k a_k y[n-1-k] + sum_k b_k x[n-k]; coefficients from the analog prototype with every 1/s replaced by the
// bilinear integrator T/2 (1+z^-1)/(1-z^-1).  State and coefficients are float like the reference's.
class LoopFilter
{
public:
    LoopFilter(float update_interval = 0.001f, float noise_bandwidth = 15.0f, int order = 2, bool last_integrator = false)
        : T_(update_interval), bw_(noise_bandwidth), order_(order), last_(last_integrator)
    {
        design();
        initialize(0.0f);
    }
    void set_noise_bandwidth(float bw)
    {
        bw_ = bw;
        design();
    }
    void set_update_interval(float t)
    {
        T_ = t;
        design();
    }
    void set_include_last_integrator(bool v)
    {
        last_ = v;
        design();
    }
    void set_order(int order)
    {
        if (order < 1 || order > 3) return;  // the reference logs a warning and keeps the old order (:256-262)
        order_ = order;
        design();
    }
    int order() const { return order_; }
    // tracking_loop_filter.cc:220-225: inputs cleared, every stored output preset
    void initialize(float initial_output = 0.0f)
    {
        x_.fill(0.0f);
        y_.fill(initial_output);
        head_ = kHist - 1;
    }
    // tracking_loop_filter.cc:74-108: the feedback sum uses the outputs BEFORE the history pointer moves
    float apply(float in)
    {
        float acc = 0.0f;
        for (int k = 0; k < na_; k++) acc += a_[k] * y_[(head_ + k) % kHist];
        head_ = (head_ + kHist - 1) % kHist;
        x_[head_] = in;
        for (int k = 0; k < nb_; k++) acc += b_[k] * x_[(head_ + k) % kHist];
        y_[head_] = acc;
        return acc;
    }

private:
    static constexpr int kHist = 4;
    // :111-214.  Natural frequency from the noise bandwidth: order 1 wn = 4 Bn; order 2 wn = Bn 8 zeta / (4 zeta^2 + 1),
    // zeta = 1/sqrt(2); order 3 wn = Bn / 0.7845 with a3 = 1.1, b3 = 2.4.  The intermediate products are float.
    void design()
    {
        const float T = T_;
        const float zeta = static_cast<float>(1.0 / std::sqrt(2.0));
        float g1, g2, g3, wn;
        na_ = nb_ = 0;
        if (order_ == 1)
            {
                wn = static_cast<float>(bw_ * 4.0);
                g1 = wn;
                if (last_)
                    {
                        set_b({static_cast<float>(g1 * T / 2.0), static_cast<float>(g1 * T / 2.0)});
                        set_a({1.0f});
                    }
                else
                    set_b({g1});
            }
        else if (order_ == 2)
            {
                wn = static_cast<float>(bw_ * (8.0 * zeta) / (4.0 * zeta * zeta + 1.0));
                g1 = wn * wn;
                g2 = static_cast<float>(wn * 2.0 * zeta);
                if (last_)
                    {
                        set_b({static_cast<float>(T / 2.0 * (g1 * T / 2.0 + g2)), static_cast<float>(T * T / 2.0 * g1),
                            static_cast<float>(T / 2.0 * (g1 * T / 2.0 - g2))});
                        set_a({2.0f, -1.0f});
                    }
                else
                    {
                        set_b({static_cast<float>(g1 * T / 2.0 + g2), static_cast<float>(g1 * T / 2.0 - g2)});
                        set_a({1.0f});
                    }
            }
        else
            {
                wn = static_cast<float>(bw_ / 0.7845);
                const float a3 = 1.1f, b3 = 2.4f;
                g1 = wn * wn * wn;
                g2 = a3 * wn * wn;
                g3 = b3 * wn;
                if (last_)
                    {
                        set_b({static_cast<float>(T / 2.0 * (g3 + T / 2.0 * (g2 + T / 2.0 * g1))),
                            static_cast<float>(T / 2.0 * (-g3 + T / 2.0 * (g2 + 3.0 * T / 2.0 * g1))),
                            static_cast<float>(T / 2.0 * (-g3 - T / 2.0 * (g2 - 3.0 * T / 2.0 * g1))),
                            static_cast<float>(T / 2.0 * (g3 - T / 2.0 * (g2 - T / 2.0 * g1)))});
                        set_a({3.0f, -3.0f, 1.0f});
                    }
                else
                    {
                        set_b({static_cast<float>(g3 + T / 2.0 * (g2 + T / 2.0 * g1)), static_cast<float>(g1 * T * T / 2.0 - 2.0 * g3),
                            static_cast<float>(g3 + T / 2.0 * (-g2 + T / 2.0 * g1))});
                        set_a({2.0f, -1.0f});
                    }
            }
    }
    void set_a(std::initializer_list<float> v)
    {
        na_ = 0;
        for (float f : v) a_[na_++] = f;
    }
    void set_b(std::initializer_list<float> v)
    {
        nb_ = 0;
        for (float f : v) b_[nb_++] = f;
    }
    float T_, bw_;
    int order_;
    bool last_;
    std::array<float, kHist> a_{}, b_{}, x_{}, y_{};
    int na_ = 0, nb_ = 0, head_ = kHist - 1;
};

// ---- FLL-assisted PLL (carrier) ------------------------------------------------------------------------------------
class FllPllFilter
{
public:
    // tracking_FLL_PLL_filter.cc:54-86
    void set_params(float fll_bw_hz, float pll_bw_hz, int order)
    {
        order_ = order;
        if (order_ == 3)
            {
                b3_ = 2.400f;
                a3_ = 1.100f;
                a2_ = 1.414f;
                w0p_ = static_cast<float>(pll_bw_hz / 0.7845);
                w0p2_ = w0p_ * w0p_;
                w0p3_ = w0p2_ * w0p_;
                w0f_ = static_cast<float>(fll_bw_hz / 0.53);
                w0f2_ = w0f_ * w0f_;
            }
        else
            {
                a2_ = 1.414f;
                w0p_ = static_cast<float>(pll_bw_hz / 0.53);
                w0p2_ = w0p_ * w0p_;
                w0f_ = static_cast<float>(fll_bw_hz / 0.25);
            }
    }
    // :89-101
    void initialize(float acq_doppler_hz)
    {
        if (order_ == 3)
            {
                x_ = static_cast<float>(2.0 * acq_doppler_hz);
                w_ = 0.0f;
            }
        else
            {
                w_ = acq_doppler_hz;
                x_ = 0.0f;
            }
    }
    // :105-133
    float get_carrier_error(float fll, float pll, float t)
    {
        if (order_ == 3)
            {
                w_ = w_ + t * (w0p3_ * pll + w0f2_ * fll);
                x_ = static_cast<float>(x_ + t * (0.5 * w_ + a2_ * w0f_ * fll + a3_ * w0p2_ * pll));
                return static_cast<float>(0.5 * x_ + b3_ * w0p_ * pll);
            }
        const float w_new = w_ + pll * w0p2_ * t + fll * w0f_ * t;
        const float out = static_cast<float>(0.5 * (w_new + w_) + a2_ * w0p_ * pll);
        w_ = w_new;
        return out;
    }

private:
    int order_ = 0;
    float w_ = 0, x_ = 0, a2_ = 0, a3_ = 0, b3_ = 0, w0p_ = 0, w0p2_ = 0, w0p3_ = 0, w0f_ = 0, w0f2_ = 0;
};

// ---- lock detectors ------------------------------------------------------------------------------------------------
// lock_detectors.cc:68-86 : signal-to-noise-variance estimator, sums in double
// src/algorithms/tracking/libs/tracking_2nd_DLL_filter.cc:43-108 and tracking_2nd_PLL_filter.cc:42-105: the older second-order
// loop filters the GLONASS / C_Aid / KF-free trackers use.  All state and coefficients are float32 as there; the
// double literals in the coefficient expressions promote exactly where the reference's do.
class Tracking2ndDllFilter
{
public:
    explicit Tracking2ndDllFilter(float pdi_code = 0.001f) : pdi_(pdi_code) {}
    void set_DLL_BW(float dll_bw_hz)
    {
        bw_ = dll_bw_hz;
        const float Wn = bw_ * 8.0 * zeta_ / (4.0 * zeta_ * zeta_ + 1.0);
        tau1_ = 1.0f / (Wn * Wn);  // k / (Wn * Wn) with k = 1.0 passed as float
        tau2_ = 2.0 * zeta_ / Wn;
    }
    void set_pdi(float pdi_code) { pdi_ = pdi_code; }
    void initialize()
    {
        old_nco_ = 0.0f;
        old_error_ = 0.0f;
    }
    float get_code_nco(float disc)
    {
        const float nco = old_nco_ + (tau2_ / tau1_) * (disc - old_error_) + (disc + old_error_) * (pdi_ / (2.0 * tau1_));
        old_nco_ = nco;
        old_error_ = disc;
        return nco;
    }

private:
    float tau1_ = 0.0f, tau2_ = 0.0f, pdi_, bw_ = 0.0f, zeta_ = 0.7, old_error_ = 0.0f, old_nco_ = 0.0f;
};

class Tracking2ndPllFilter
{
public:
    explicit Tracking2ndPllFilter(float pdi_carr = 0.001f) : pdi_(pdi_carr) {}
    void set_PLL_BW(float pll_bw_hz)
    {
        bw_ = pll_bw_hz;
        const float Wn = bw_ * 8.0 * zeta_ / (4.0 * zeta_ * zeta_ + 1.0);
        tau1_ = 0.25f / (Wn * Wn);  // k = 0.25 passed as float
        tau2_ = 2.0 * zeta_ / Wn;
    }
    void set_pdi(float pdi_carr) { pdi_ = pdi_carr; }
    void initialize()
    {
        old_nco_ = 0.0f;
        old_error_ = 0.0f;
    }
    float get_carrier_nco(float disc)
    {
        const float nco = old_nco_ + (tau2_ / tau1_) * (disc - old_error_) + (disc + old_error_) * (pdi_ / (2.0 * tau1_));
        old_nco_ = nco;
        old_error_ = disc;
        return nco;
    }

private:
    float tau1_ = 0.0f, tau2_ = 0.0f, pdi_, bw_ = 0.0f, zeta_ = 0.7, old_error_ = 0.0f, old_nco_ = 0.0f;
};

inline float cn0_svn_estimator(const cfloat* prompt, int n, double coh_integration_time_s)
{
    double psig = 0.0, ptot = 0.0;
    for (int i = 0; i < n; i++)
        {
            psig += std::abs(static_cast<double>(prompt[i].real()));
            ptot += static_cast<double>(prompt[i].imag()) * static_cast<double>(prompt[i].imag()) +
                    static_cast<double>(prompt[i].real()) * static_cast<double>(prompt[i].real());
        }
    psig /= static_cast<double>(n);
    psig *= psig;
    ptot /= static_cast<double>(n);
    const double snr = psig / (ptot - psig);
    return static_cast<float>(10.0 * std::log10(snr) - 10.0 * std::log10(coh_integration_time_s));
}
// lock_detectors.cc:98-111 : cos(2 phi) estimate, sums in float
inline float carrier_lock_detector(const cfloat* prompt, int n)
{
    float si = 0.0f, sq = 0.0f;
    for (int i = 0; i < n; i++)
        {
            si += prompt[i].real();
            sq += prompt[i].imag();
        }
    const float nbp = si * si + sq * sq;
    const float nbd = si * si - sq * sq;
    return nbd / nbp;
}
}  // namespace gb200
