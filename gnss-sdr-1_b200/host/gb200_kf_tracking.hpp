// GpsL1CaKfTracking<Correlator>: host mirror of Gps_L1_Ca_Kf_Tracking_cc
// (src/algorithms/tracking/gnuradio_blocks/gps_l1_ca_kf_tracking_cc.cc): GPS L1 C/A tracker whose carrier loop is a 2- or 3-state
// Kalman filter (phase, Doppler[, Doppler rate]) driven by the two-quadrant atan of the prompt, with a second-order DLL
// (Tracking_2nd_DLL_filter) aided by the Kalman Doppler, the gflags-sized CN0 / lock test, and the optional Bayesian estimator of
// the measurement covariance (src/algorithms/tracking/libs/bayesian_estimation.cc).  It uses the SAME correlator call as
// dll_pll_veml_tracking (Cpu_Multicorrelator_Real_Codes, :676-683), so the B200 multicorrelator serves it unchanged.
// GNU Radio's scheduler is replaced by step(): one call = one general_work, samples named by absolute index.
//
// The reference writes the filter with Armadillo; the matrices are 1x1 .. 3x3, so this mirror keeps them in a fixed 3x3 array
// (KfMat) and evaluates every product as the reference's source reads: left to right, each element a k-ordered sum.
//
// Correlator concept (= Cpu_Multicorrelator_Real_Codes, cpu_multicorrelator_real_codes.h:49-77):
//   bool set_local_code_and_taps(int code_len, const float* code, const float* shifts, int n_taps);
//   bool correlate(uint64_t first_sample, float rem_carr_rad, float phase_step_rad, float rem_code_chips, float code_step_chips,
//                  float code_rate_step_chips, int n_samples, std::complex<float>* out);      // the six-argument overload (:155-170)
// The block never calls set_high_dynamics_resampler(), so the reference correlator stays at its constructor default
// (d_use_high_dynamics_resampler = true, cpu_multicorrelator_real_codes.cc:49): the code replica comes from the high-dynamics
// resampler (tap 0 by formula with a zero rate term, taps >= 1 circular SAMPLE shifts of tap 0) while the carrier uses the plain
// rotator.  The correlator handed to this class must be configured the same way: set_high_dynamics_resampler(true) before init().
#pragma once
#include <cmath>
#include <complex>
#include <cstdint>
#include <vector>

#include "gb200_loop_math.hpp"

namespace gb200
{
struct KfTrkConf
{
    uint32_t order = 2;  // 2: phase + Doppler; 3: + Doppler rate (gps_l1_ca_kf_tracking.cc:70 "order", block ctor :233-256)
    int64_t fs_in = 4000000;
    uint32_t vector_length = 4000;
    float dll_bw_hz = 2.0f, early_late_space_chips = 0.5f;
    // Bayesian covariance estimation (adapter keys bce_run / bce_ptrans / bce_strans / bce_nu / bce_kappa, gps_l1_ca_kf_tracking.cc:86-99)
    bool bce_run = false;
    uint32_t bce_ptrans = 0, bce_strans = 0;
    int32_t bce_nu = 0, bce_kappa = 0;
    // gnss_sdr_flags.cc:49-59 defaults (FLAGS_cn0_samples, FLAGS_cn0_min, FLAGS_max_lock_fail, FLAGS_carrier_lock_th)
    int cn0_samples = 20, cn0_min = 25, max_lock_fail = 50;
    double carrier_lock_th = 0.85;
};

struct KfEpoch
{
    uint64_t tracking_sample_counter = 0;
    double prompt_i = 0, prompt_q = 0, carrier_phase_rads = 0, carrier_doppler_hz = 0, code_phase_samples = 0, cn0_db_hz = 0;
    float abs_e = 0, abs_p = 0, abs_l = 0;
    bool valid_symbol = false, loss_of_lock = false;
    int32_t prn_length_samples = 0;
    double carrier_dopplerrate_hz2 = 0, carr_phase_sigma2 = 0, rem_carr_phase_rad = 0;
};

// dense r x c matrix, r, c <= 3
struct KfMat
{
    int r = 0, c = 0;
    double d[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
    KfMat() = default;
    KfMat(int rows, int cols) : r(rows), c(cols) {}
    static KfMat eye(int n)
    {
        KfMat m(n, n);
        for (int i = 0; i < n; i++) m.d[i][i] = 1.0;
        return m;
    }
    KfMat resized(int rows, int cols) const  // arma::resize: top-left block kept, zeros elsewhere
    {
        KfMat m(rows, cols);
        for (int i = 0; i < rows && i < r; i++)
            for (int j = 0; j < cols && j < c; j++) m.d[i][j] = d[i][j];
        return m;
    }
    KfMat t() const
    {
        KfMat m(c, r);
        for (int i = 0; i < r; i++)
            for (int j = 0; j < c; j++) m.d[j][i] = d[i][j];
        return m;
    }
};
inline KfMat operator*(const KfMat& a, const KfMat& b)
{
    KfMat m(a.r, b.c);
    for (int i = 0; i < a.r; i++)
        for (int j = 0; j < b.c; j++)
            {
                double s = 0.0;
                for (int k = 0; k < a.c; k++) s += a.d[i][k] * b.d[k][j];
                m.d[i][j] = s;
            }
    return m;
}
inline KfMat operator*(const KfMat& a, double k)
{
    KfMat m(a.r, a.c);
    for (int i = 0; i < a.r; i++)
        for (int j = 0; j < a.c; j++) m.d[i][j] = a.d[i][j] * k;
    return m;
}
inline KfMat operator+(const KfMat& a, const KfMat& b)
{
    KfMat m(a.r, a.c);
    for (int i = 0; i < a.r; i++)
        for (int j = 0; j < a.c; j++) m.d[i][j] = a.d[i][j] + b.d[i][j];
    return m;
}
inline KfMat operator-(const KfMat& a, const KfMat& b)
{
    KfMat m(a.r, a.c);
    for (int i = 0; i < a.r; i++)
        for (int j = 0; j < a.c; j++) m.d[i][j] = a.d[i][j] - b.d[i][j];
    return m;
}

// Bayesian_estimator for one scalar measurement per update (ny = 1, K = 1: the only way the tracker calls it, :711)
class BayesianEstimatorScalar
{
public:
    // = Bayesian_estimator::init (bayesian_estimation.cc:79-88)
    void init(double mu_prior_0, int kappa_prior_0, int nu_prior_0, double psi_prior_0)
    {
        mu_prior_ = mu_prior_0;
        kappa_prior_ = kappa_prior_0;
        nu_prior_ = nu_prior_0;
        psi_prior_ = psi_prior_0;
        mu_est_ = mu_prior_;
        psi_est_ = psi_prior_;
    }
    // = update_sequential(data) (:94-136) with a 1 x 1 data matrix
    void update_sequential(double data)
    {
        const int K = 1, ny = 1;
        const double y_mean = data / static_cast<double>(K);
        double psi_n = 0.0;
        psi_n = psi_n + (data - y_mean) * (data - y_mean);
        const double mu_posterior = (kappa_prior_ * mu_prior_ + K * y_mean) / (kappa_prior_ + K);
        const int kappa_posterior = kappa_prior_ + K;
        const int nu_posterior = nu_prior_ + K;
        // (kappa_prior * K) / (kappa_prior + K) is an INTEGER division in the reference (:120)
        const int w = (kappa_prior_ * K) / (kappa_prior_ + K);
        const double psi_posterior = psi_prior_ + psi_n + (w * (y_mean - mu_prior_)) * (y_mean - mu_prior_);
        mu_est_ = mu_posterior;
        if ((nu_posterior - ny - 1) > 0)
            psi_est_ = psi_posterior / (nu_posterior - ny - 1);
        else
            psi_est_ = psi_posterior / (nu_posterior + ny + 1);
        mu_prior_ = mu_posterior;
        kappa_prior_ = kappa_posterior;
        nu_prior_ = nu_posterior;
        psi_prior_ = psi_posterior;
    }
    double psi_est() const { return psi_est_; }
    double mu_est() const { return mu_est_; }

private:
    double mu_est_ = 0, psi_est_ = 0, mu_prior_ = 0, psi_prior_ = 0;
    int kappa_prior_ = 0, nu_prior_ = 0;
};

template <class Correlator>
class GpsL1CaKfTracking
{
public:
    GpsL1CaKfTracking(const KfTrkConf& c, Correlator* corr) : c_(c), corr_(corr)
    {
        // ctor (:90-272)
        const double kTwoPi = 6.283185307179586;   // GPS_TWO_PI
        const double kCodePeriod = 0.001;          // GPS_L1_CA_CODE_PERIOD
        current_prn_length_samples_ = static_cast<int32_t>(c_.vector_length);
        code_loop_filter_.set_DLL_BW(c_.dll_bw_hz);
        early_late_spc_chips_ = c_.early_late_space_chips;
        shifts_[0] = -early_late_spc_chips_;
        shifts_[1] = 0.0;
        shifts_[2] = early_late_spc_chips_;
        code_freq_chips_ = 1.023e6;
        prompt_buffer_.resize(c_.cn0_samples);
        carrier_lock_threshold_ = c_.carrier_lock_th;
        // Kalman filter initialization (:196-256)
        const double CN_dB_Hz = 30;
        const double CN_lin = pow(10, CN_dB_Hz / 10.0);
        const double sigma2_phase_detector_cycles2 = (1.0 / (2.0 * CN_lin * kCodePeriod)) * (1.0 + 1.0 / (2.0 * CN_lin * kCodePeriod));
        const double sigma2_carrier_phase = kTwoPi / 4;
        const double sigma2_doppler = 450;
        const double sigma2_doppler_rate = pow(4.0 * kTwoPi, 2) / 12.0;
        P_x_ini_ = KfMat(2, 2);
        P_x_ini_.d[0][0] = sigma2_carrier_phase;
        P_x_ini_.d[1][1] = sigma2_doppler;
        R_ = KfMat(1, 1);
        R_.d[0][0] = sigma2_phase_detector_cycles2;
        Q_ = KfMat(2, 2);
        Q_.d[0][0] = pow(kCodePeriod, 4);
        Q_.d[1][1] = kCodePeriod;
        F_ = KfMat(2, 2);
        F_.d[0][0] = 1.0;
        F_.d[0][1] = kTwoPi * kCodePeriod;
        F_.d[1][0] = 0.0;
        F_.d[1][1] = 1.0;
        H_ = KfMat(1, 2);
        H_.d[0][0] = 1.0;
        x_ = KfMat(2, 1);
        y_ = KfMat(1, 1);
        P_y_ = KfMat(1, 1);
        if (c_.order == 3)
            {
                P_x_ini_ = P_x_ini_.resized(3, 3);
                P_x_ini_.d[2][2] = sigma2_doppler_rate;
                Q_ = KfMat(3, 3);
                Q_.d[0][0] = pow(kCodePeriod, 4);
                Q_.d[1][1] = kCodePeriod;
                Q_.d[2][2] = kCodePeriod;
                F_ = F_.resized(3, 3);
                F_.d[0][2] = 0.5 * kTwoPi * pow(kCodePeriod, 2);
                F_.d[1][2] = kCodePeriod;
                F_.d[2][0] = 0.0;
                F_.d[2][1] = 0.0;
                F_.d[2][2] = 1.0;
                H_ = H_.resized(1, 3);
                H_.d[0][2] = 0.0;
                x_ = x_.resized(3, 1);
                x_.d[2][0] = 0.0;
            }
        kf_iter_ = 0;
        R_est_ = R_;
        init_bayes();
    }
    void set_sample_counter(uint64_t now) { sample_counter_ = now; }
    uint64_t sample_counter() const { return sample_counter_; }

    // = start_tracking() (:275-353); code = gps_l1_ca_code_gen_float(.., PRN, 0); acq_doppler_step = Gnss_Synchro::Acq_doppler_step
    bool start_tracking(const float* code, double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp, uint32_t acq_doppler_step)
    {
        const double kTwoPi = 6.283185307179586, kL1 = 1.57542e9, kRate = 1.023e6, kChips = 1023.0;
        acq_code_phase_samples_ = acq_delay_samples;
        acq_carrier_doppler_hz_ = acq_doppler_hz;
        acq_sample_stamp_ = acq_samplestamp;
        acq_carrier_doppler_step_hz_ = static_cast<double>(acq_doppler_step);
        // Kalman covariance from the acquisition's Doppler step (3 sigma) (:285-290)
        if (acq_doppler_step > 0)
            {
                P_x_ini_.d[1][1] = pow(acq_carrier_doppler_step_hz_ / 3.0, 2);
                init_bayes();
            }
        const int64_t acq_trk_diff_samples = static_cast<int64_t>(sample_counter_) - static_cast<int64_t>(acq_sample_stamp_);
        const double acq_trk_diff_seconds = static_cast<float>(acq_trk_diff_samples) / static_cast<float>(c_.fs_in);
        const double radial_velocity = (kL1 + acq_carrier_doppler_hz_) / kL1;
        code_freq_chips_ = radial_velocity * kRate;
        code_phase_step_chips_ = static_cast<double>(code_freq_chips_) / static_cast<double>(c_.fs_in);
        const double T_chip_mod_seconds = 1 / code_freq_chips_;
        const double T_prn_mod_seconds = T_chip_mod_seconds * kChips;
        const double T_prn_mod_samples = T_prn_mod_seconds * static_cast<double>(c_.fs_in);
        current_prn_length_samples_ = round(T_prn_mod_samples);
        const double T_prn_true_seconds = kChips / kRate;
        const double T_prn_true_samples = T_prn_true_seconds * static_cast<double>(c_.fs_in);
        const double T_prn_diff_seconds = T_prn_true_seconds - T_prn_mod_seconds;
        const double N_prn_diff = acq_trk_diff_seconds / T_prn_true_seconds;
        double corrected_acq_phase_samples = fmod((acq_code_phase_samples_ + T_prn_diff_seconds * N_prn_diff * static_cast<double>(c_.fs_in)), T_prn_true_samples);
        if (corrected_acq_phase_samples < 0) corrected_acq_phase_samples = T_prn_mod_samples + corrected_acq_phase_samples;
        acq_code_phase_samples_ = corrected_acq_phase_samples;
        carrier_doppler_hz_ = acq_carrier_doppler_hz_;
        carrier_dopplerrate_hz2_ = 0;
        carrier_phase_step_rad_ = kTwoPi * carrier_doppler_hz_ / static_cast<double>(c_.fs_in);
        code_loop_filter_.initialize();
        if (!corr_->set_local_code_and_taps(static_cast<int>(kChips), code, shifts_, 3)) return false;
        for (auto& o : outs_) o = cfloat(0, 0);
        carrier_lock_fail_counter_ = 0;
        rem_code_phase_samples_ = 0;
        rem_carr_phase_rad_ = 0.0;
        rem_code_phase_chips_ = 0.0;
        acc_carrier_phase_rad_ = 0.0;
        carr_phase_sigma2_ = 0.0;
        code_phase_samples_ = acq_code_phase_samples_;
        pull_in_ = true;
        enable_tracking_ = true;
        return true;
    }

    // = general_work (:778-966).  false: the correlator refused the request (samples not in the ring)
    bool step(KfEpoch* e)
    {
        const double kPi2 = 2.0 * 3.1415926535897932, kL1 = 1.57542e9, kRate = 1.023e6, kChips = 1023.0, kCodePeriod = 0.001;
        *e = KfEpoch();
        if (enable_tracking_)
            {
                if (pull_in_)
                    {
                        // (:796-834)
                        const uint64_t acq_to_trk_delay_samples = sample_counter_ - acq_sample_stamp_;
                        const double acq_trk_shif_correction_samples = static_cast<double>(current_prn_length_samples_) -
                                                                       std::fmod(static_cast<double>(acq_to_trk_delay_samples), static_cast<double>(current_prn_length_samples_));
                        int32_t samples_offset = std::round(acq_code_phase_samples_ + acq_trk_shif_correction_samples);
                        if (samples_offset < 0) samples_offset = 0;
                        acc_carrier_phase_rad_ -= carrier_phase_step_rad_ * acq_code_phase_samples_;
                        sample_counter_ += samples_offset;
                        pull_in_ = false;
                        // the item is the acquisition's Gnss_Synchro with these two fields: its Tracking_sample_counter is whatever the
                        // channel's object holds (0, nothing writes it)
                        e->carrier_phase_rads = acc_carrier_phase_rad_;
                        e->carrier_doppler_hz = carrier_doppler_hz_;
                        e->prn_length_samples = samples_offset;
                        P_x_ = P_x_ini_;
                        x_.d[0][0] = carrier_phase_step_rad_ * samples_offset;
                        x_.d[1][0] = carrier_doppler_hz_;
                        if (x_.r > 2) x_.d[2][0] = carrier_dopplerrate_hz2_;
                        kf_iter_ = 0;
                        init_bayes();
                        return true;  // consume_each(samples_offset)
                    }
                // the six-argument overload (cpu_multicorrelator_real_codes.cc:155-170)
                if (!corr_->correlate(sample_counter_, rem_carr_phase_rad_, carrier_phase_step_rad_, rem_code_phase_chips_, code_phase_step_chips_,
                        code_phase_rate_step_chips_, current_prn_length_samples_, outs_))
                    return false;
                // Kalman state prediction (:688-689)
                const KfMat x_pre = F_ * x_;
                const KfMat P_x_pre = F_ * P_x_ * F_.t() + Q_;
                const double carr_phase_error_rad = pll_cloop_two_quadrant_atan(outs_[1]);
                // measurement update (:694-730)
                const double CN_lin = pow(10, cn0_snv_db_hz_ / 10.0);
                const double sigma2_phase_detector_cycles2 = (1.0 / (2.0 * CN_lin * kCodePeriod)) * (1.0 + 1.0 / (2.0 * CN_lin * kCodePeriod));
                y_.d[0][0] = carr_phase_error_rad;
                R_.d[0][0] = sigma2_phase_detector_cycles2;
                if (c_.bce_run && (kf_iter_ >= c_.bce_ptrans)) bayes_.update_sequential(y_.d[0][0]);
                if (c_.bce_run && (kf_iter_ >= (c_.bce_ptrans + c_.bce_strans)))
                    {
                        P_y_.d[0][0] = bayes_.psi_est();
                        R_est_ = P_y_ - H_ * P_x_pre * H_.t();
                    }
                else
                    {
                        P_y_ = H_ * P_x_pre * H_.t() + R_;
                        R_est_ = R_;
                    }
                KfMat inv_P_y(1, 1);
                inv_P_y.d[0][0] = 1.0 / P_y_.d[0][0];
                const KfMat K = (P_x_pre * H_.t()) * inv_P_y;
                x_ = x_pre + K * y_;
                P_x_ = (KfMat::eye(P_x_pre.r) - K * H_) * P_x_pre;
                rem_carr_phase_rad_ = x_.d[0][0];
                carrier_doppler_hz_ = x_.d[1][0];
                carrier_dopplerrate_hz2_ = x_.r > 2 ? x_.d[2][0] : 0;
                carr_phase_sigma2_ = R_est_.d[0][0];
                // DLL (:745-754)
                code_freq_chips_ = kRate + ((carrier_doppler_hz_ * kRate) / kL1);
                const double code_error_chips = dll_nc_e_minus_l_normalized(outs_[0], outs_[2]);
                const double code_error_filt_chips = code_loop_filter_.get_code_nco(code_error_chips);
                const double T_chip_seconds = 1.0 / static_cast<double>(code_freq_chips_);
                const double T_prn_seconds = T_chip_seconds * kChips;
                const double code_error_filt_secs = (T_prn_seconds * code_error_filt_chips * T_chip_seconds);
                // NCO alignment (:759-775)
                const double T_prn_samples = T_prn_seconds * static_cast<double>(c_.fs_in);
                const double K_blk_samples = T_prn_samples + rem_code_phase_samples_ + code_error_filt_secs * static_cast<double>(c_.fs_in);
                current_prn_length_samples_ = static_cast<int>(round(K_blk_samples));
                carrier_phase_step_rad_ = kPi2 * carrier_doppler_hz_ / static_cast<double>(c_.fs_in);
                acc_carrier_phase_rad_ -= carrier_phase_step_rad_ * static_cast<double>(current_prn_length_samples_);
                code_phase_step_chips_ = code_freq_chips_ / static_cast<double>(c_.fs_in);
                rem_code_phase_samples_ = K_blk_samples - static_cast<double>(current_prn_length_samples_);
                rem_code_phase_chips_ = code_freq_chips_ * (rem_code_phase_samples_ / static_cast<double>(c_.fs_in));
                // CN0 and lock detector (:778-818)
                if (cn0_estimation_counter_ < c_.cn0_samples)
                    {
                        prompt_buffer_[cn0_estimation_counter_] = outs_[1];
                        cn0_estimation_counter_++;
                    }
                else
                    {
                        cn0_estimation_counter_ = 0;
                        cn0_snv_db_hz_ = cn0_svn_estimator(prompt_buffer_.data(), c_.cn0_samples, kCodePeriod);
                        carrier_lock_test_ = carrier_lock_detector(prompt_buffer_.data(), c_.cn0_samples);
                        if (carrier_lock_test_ < carrier_lock_threshold_ or cn0_snv_db_hz_ < c_.cn0_min)
                            carrier_lock_fail_counter_++;
                        else if (carrier_lock_fail_counter_ > 0)
                            carrier_lock_fail_counter_--;
                        if (carrier_lock_fail_counter_ > c_.max_lock_fail)
                            {
                                e->loss_of_lock = true;  // event 3
                                carrier_lock_fail_counter_ = 0;
                                enable_tracking_ = false;
                            }
                    }
                e->prompt_i = static_cast<double>(outs_[1].real());
                e->prompt_q = static_cast<double>(outs_[1].imag());
                e->tracking_sample_counter = sample_counter_ + static_cast<uint64_t>(current_prn_length_samples_);
                e->code_phase_samples = rem_code_phase_samples_;
                e->carrier_phase_rads = acc_carrier_phase_rad_;
                e->carrier_doppler_hz = carrier_doppler_hz_;
                e->cn0_db_hz = cn0_snv_db_hz_;
                e->valid_symbol = true;
                e->abs_e = std::abs<float>(outs_[0]);
                e->abs_p = std::abs<float>(outs_[1]);
                e->abs_l = std::abs<float>(outs_[2]);
                e->carrier_dopplerrate_hz2 = carrier_dopplerrate_hz2_;
                e->carr_phase_sigma2 = carr_phase_sigma2_;
                e->rem_carr_phase_rad = rem_carr_phase_rad_;
                kf_iter_++;
            }
        else
            {
                for (auto& o : outs_) o = cfloat(0, 0);
                e->tracking_sample_counter = sample_counter_ + static_cast<uint64_t>(current_prn_length_samples_);
            }
        // consume_each(d_current_prn_length_samples); d_sample_counter += d_current_prn_length_samples (:963-964)
        e->prn_length_samples = current_prn_length_samples_;
        sample_counter_ += current_prn_length_samples_;
        return true;
    }
    const cfloat* correlator_outs() const { return outs_; }
    bool tracking_enabled() const { return enable_tracking_; }

private:
    // bayes_estimator.init(zeros(1,1), kappa, nu, (H * P_x_ini * H' + R) * (nu + 2)) (:271, :289, :831)
    void init_bayes()
    {
        const KfMat psi = (H_ * P_x_ini_ * H_.t() + R_) * static_cast<double>(c_.bce_nu + 2);
        bayes_.init(0.0, c_.bce_kappa, c_.bce_nu, psi.d[0][0]);
    }
    KfTrkConf c_;
    Correlator* corr_;
    Tracking2ndDllFilter code_loop_filter_;
    BayesianEstimatorScalar bayes_;
    KfMat P_x_ini_, P_x_, P_y_, F_, H_, R_, Q_, x_, y_, R_est_;
    uint32_t kf_iter_ = 0;
    cfloat outs_[3];
    float shifts_[3];
    std::vector<cfloat> prompt_buffer_;
    double early_late_spc_chips_ = 0, rem_code_phase_samples_ = 0, rem_code_phase_chips_ = 0, rem_carr_phase_rad_ = 0;
    double acq_code_phase_samples_ = 0, acq_carrier_doppler_hz_ = 0, acq_carrier_doppler_step_hz_ = 0;
    double code_freq_chips_ = 0, code_phase_step_chips_ = 0, code_phase_rate_step_chips_ = 0, carrier_doppler_hz_ = 0, carrier_dopplerrate_hz2_ = 0;
    double carrier_phase_step_rad_ = 0, acc_carrier_phase_rad_ = 0, carr_phase_sigma2_ = 0, code_phase_samples_ = 0;
    int32_t current_prn_length_samples_ = 0;
    uint64_t sample_counter_ = 0, acq_sample_stamp_ = 0;
    int32_t cn0_estimation_counter_ = 0;
    double carrier_lock_test_ = 1, cn0_snv_db_hz_ = 0, carrier_lock_threshold_ = 0.85;
    int32_t carrier_lock_fail_counter_ = 0;
    bool enable_tracking_ = false, pull_in_ = false;
};
}  // namespace gb200

// flat images for the ctypes-facing test libraries (host_abi.cc and tests/native/host_oracle_loop.cc)
// conf = {order, fs_in, vector_length, dll_bw_hz, early_late_space_chips, bce_run, bce_ptrans, bce_strans, bce_nu, bce_kappa,
//         cn0_samples, cn0_min, max_lock_fail, carrier_lock_th}
inline void gbh_kf_conf(const double* conf, gb200::KfTrkConf* c)
{
    c->order = static_cast<uint32_t>(conf[0]);
    c->fs_in = static_cast<int64_t>(conf[1]);
    c->vector_length = static_cast<uint32_t>(conf[2]);
    c->dll_bw_hz = static_cast<float>(conf[3]);
    c->early_late_space_chips = static_cast<float>(conf[4]);
    c->bce_run = conf[5] != 0.0;
    c->bce_ptrans = static_cast<uint32_t>(conf[6]);
    c->bce_strans = static_cast<uint32_t>(conf[7]);
    c->bce_nu = static_cast<int32_t>(conf[8]);
    c->bce_kappa = static_cast<int32_t>(conf[9]);
    c->cn0_samples = static_cast<int>(conf[10]);
    c->cn0_min = static_cast<int>(conf[11]);
    c->max_lock_fail = static_cast<int>(conf[12]);
    c->carrier_lock_th = conf[13];
}
// epoch = {tracking_sample_counter, prompt_i, prompt_q, carrier_doppler_hz, code_phase_samples, carrier_phase_rads, cn0_db_hz, abs_e, abs_p,
//          abs_l, valid_symbol + 2 * loss_of_lock, prn_length_samples, carrier_dopplerrate_hz2, carr_phase_sigma2, rem_carr_phase_rad}
inline void gbh_kf_fill(const gb200::KfEpoch& e, double* o)
{
    o[0] = static_cast<double>(e.tracking_sample_counter);
    o[1] = e.prompt_i;
    o[2] = e.prompt_q;
    o[3] = e.carrier_doppler_hz;
    o[4] = e.code_phase_samples;
    o[5] = e.carrier_phase_rads;
    o[6] = e.cn0_db_hz;
    o[7] = e.abs_e;
    o[8] = e.abs_p;
    o[9] = e.abs_l;
    o[10] = (e.valid_symbol ? 1 : 0) + (e.loss_of_lock ? 2 : 0);
    o[11] = e.prn_length_samples;
    o[12] = e.carrier_dopplerrate_hz2;
    o[13] = e.carr_phase_sigma2;
    o[14] = e.rem_carr_phase_rad;
}
