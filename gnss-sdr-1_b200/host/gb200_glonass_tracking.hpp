// GlonassL1CaDllPllTracking<Correlator>: host mirror of Glonass_L1_Ca_Dll_Pll_Tracking_cc
// (src/algorithms/tracking/gnuradio_blocks/glonass_l1_ca_dll_pll_tracking_cc.cc), the tracker BASELINE.json's C4 needs next
// to dll_pll_veml_tracking: complex local code (Cpu_Multicorrelator), FDMA carrier, second-order DLL / PLL filters
// (Tracking_2nd_DLL_filter / Tracking_2nd_PLL_filter), 10-prompt CN0 / lock test.  The GLONASS L2 C/A block
// (glonass_l2_ca_dll_pll_tracking_cc.cc) is the same loop with the L2 band constants: pass them in GlonassTrkConf.
// GNU Radio's scheduler is replaced by step(): one call = one general_work, samples named by absolute index.
//
// Correlator concept (= Cpu_Multicorrelator, cpu_multicorrelator.h:47-70):
//   bool set_local_code_and_taps(int code_len, const std::complex<float>* code, const float* shifts, int n_taps);
//   bool correlate(uint64_t first_sample, float rem_carr_rad, float phase_step_rad, float rem_code_chips, float code_step_chips,
//                  int n_samples, std::complex<float>* out);
#pragma once
#include <cmath>
#include <complex>
#include <cstdint>
#include <vector>

#include "gb200_loop_math.hpp"

namespace gb200
{
struct GlonassTrkConf
{
    int64_t fs_in = 4000000;
    uint32_t vector_length = 4000;
    float pll_bw_hz = 50.0f, dll_bw_hz = 2.0f, early_late_space_chips = 0.5f;
    // band constants (GLONASS_L1_L2_CA.h:92-96; L2: FREQ2_GLO / DFRQ2_GLO)
    double band_freq_hz = 1.602e9, band_dfreq_hz = 0.5625e6;
    double code_rate_hz = 0.511e6, code_length_chips = 511.0, code_period_s = 0.001;
    // gnss_sdr_flags.cc:53-59 defaults
    int cn0_min = 25, max_lock_fail = 50;
    double carrier_lock_th = 0.85;
};

struct GlonassEpoch
{
    uint64_t tracking_sample_counter = 0;
    double prompt_i = 0, prompt_q = 0, carrier_phase_rads = 0, carrier_doppler_hz = 0, code_phase_samples = 0, cn0_db_hz = 0;
    float abs_e = 0, abs_p = 0, abs_l = 0;
    bool valid_symbol = false, loss_of_lock = false;
    int32_t prn_length_samples = 0;
};

template <class Correlator>
class GlonassL1CaDllPllTracking
{
public:
    static constexpr int kCn0EstimationSamples = 10;  // CN0_ESTIMATION_SAMPLES (:57)
    GlonassL1CaDllPllTracking(const GlonassTrkConf& c, Correlator* corr) : c_(c), corr_(corr)
    {
        // ctor (:99-175)
        current_prn_length_samples_ = static_cast<int32_t>(c_.vector_length);
        code_loop_filter_.set_DLL_BW(c_.dll_bw_hz);
        carrier_loop_filter_.set_PLL_BW(c_.pll_bw_hz);
        early_late_spc_chips_ = c_.early_late_space_chips;
        shifts_[0] = -early_late_spc_chips_;
        shifts_[1] = 0.0;
        shifts_[2] = early_late_spc_chips_;
        code_freq_chips_ = c_.code_rate_hz;
        prompt_buffer_.resize(kCn0EstimationSamples);
        carrier_lock_threshold_ = c_.carrier_lock_th;
    }
    void set_sample_counter(uint64_t now) { sample_counter_ = now; }
    uint64_t sample_counter() const { return sample_counter_; }

    // = start_tracking() (:205-293); freq_channel = GLONASS_PRN.at(PRN), code = glonass_l1_ca_code_gen_complex(.., 0)
    bool start_tracking(const std::complex<float>* code, int freq_channel, double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp)
    {
        const double kTwoPi = 6.283185307179586;  // GLONASS_TWO_PI
        acq_code_phase_samples_ = acq_delay_samples;
        acq_carrier_doppler_hz_ = acq_doppler_hz;
        acq_sample_stamp_ = acq_samplestamp;
        const int64_t acq_trk_diff_samples = static_cast<int64_t>(sample_counter_) - static_cast<int64_t>(acq_sample_stamp_);
        const double acq_trk_diff_seconds = static_cast<float>(acq_trk_diff_samples) / static_cast<float>(c_.fs_in);
        glonass_freq_ch_ = c_.band_freq_hz + (c_.band_dfreq_hz * freq_channel);  // int64_t member, truncates
        const double radial_velocity = (glonass_freq_ch_ + acq_carrier_doppler_hz_) / glonass_freq_ch_;
        code_freq_chips_ = radial_velocity * c_.code_rate_hz;
        code_phase_step_chips_ = static_cast<double>(code_freq_chips_) / static_cast<double>(c_.fs_in);
        const double T_chip_mod_seconds = 1 / code_freq_chips_;
        const double T_prn_mod_seconds = T_chip_mod_seconds * c_.code_length_chips;
        const double T_prn_mod_samples = T_prn_mod_seconds * static_cast<double>(c_.fs_in);
        current_prn_length_samples_ = round(T_prn_mod_samples);
        const double T_prn_true_seconds = c_.code_length_chips / c_.code_rate_hz;
        const double T_prn_true_samples = T_prn_true_seconds * static_cast<double>(c_.fs_in);
        const double T_prn_diff_seconds = T_prn_true_seconds - T_prn_mod_seconds;
        const double N_prn_diff = acq_trk_diff_seconds / T_prn_true_seconds;
        double corrected_acq_phase_samples = fmod((acq_code_phase_samples_ + T_prn_diff_seconds * N_prn_diff * static_cast<double>(c_.fs_in)), T_prn_true_samples);
        if (corrected_acq_phase_samples < 0) corrected_acq_phase_samples = T_prn_mod_samples + corrected_acq_phase_samples;
        acq_code_phase_samples_ = corrected_acq_phase_samples;
        carrier_frequency_hz_ = acq_carrier_doppler_hz_ + (c_.band_dfreq_hz * freq_channel);
        carrier_doppler_hz_ = acq_carrier_doppler_hz_;
        carrier_phase_step_rad_ = kTwoPi * carrier_frequency_hz_ / static_cast<double>(c_.fs_in);
        carrier_doppler_phase_step_rad_ = kTwoPi * carrier_doppler_hz_ / static_cast<double>(c_.fs_in);
        carrier_loop_filter_.initialize();
        code_loop_filter_.initialize();
        if (!corr_->set_local_code_and_taps(static_cast<int32_t>(c_.code_length_chips), code, shifts_, 3)) return false;
        for (auto& o : outs_) o = cfloat(0, 0);
        carrier_lock_fail_counter_ = 0;
        rem_code_phase_samples_ = 0;
        rem_carr_phase_rad_ = 0.0;
        rem_code_phase_chips_ = 0.0;
        acc_carrier_phase_rad_ = 0.0;
        pull_in_ = true;
        enable_tracking_ = true;
        return true;
    }

    // = general_work (:553-765).  false: the correlator refused the request (samples not in the ring)
    bool step(GlonassEpoch* e)
    {
        const double kTwoPi = 6.283185307179586;
        *e = GlonassEpoch();
        if (enable_tracking_)
            {
                if (pull_in_)
                    {
                        const int32_t acq_to_trk_delay_samples = sample_counter_ - acq_sample_stamp_;
                        const double acq_trk_shif_correction_samples =
                            current_prn_length_samples_ - std::fmod(static_cast<float>(acq_to_trk_delay_samples), static_cast<float>(current_prn_length_samples_));
                        const int32_t samples_offset = round(acq_code_phase_samples_ + acq_trk_shif_correction_samples);
                        e->tracking_sample_counter = sample_counter_ + static_cast<uint64_t>(samples_offset);
                        sample_counter_ = sample_counter_ + static_cast<uint64_t>(samples_offset);
                        pull_in_ = false;
                        acc_carrier_phase_rad_ -= carrier_doppler_phase_step_rad_ * samples_offset;
                        e->carrier_phase_rads = acc_carrier_phase_rad_;
                        e->carrier_doppler_hz = carrier_doppler_hz_;
                        e->prn_length_samples = samples_offset;
                        return true;  // consume_each(samples_offset)
                    }
                if (!corr_->correlate(sample_counter_, rem_carr_phase_rad_, carrier_phase_step_rad_, rem_code_phase_chips_, code_phase_step_chips_,
                        current_prn_length_samples_, outs_))
                    return false;
                const double carr_error_hz = pll_cloop_two_quadrant_atan(outs_[1]) / kTwoPi;
                const double carr_error_filt_hz = carrier_loop_filter_.get_carrier_nco(carr_error_hz);
                carrier_frequency_hz_ += carr_error_filt_hz;
                carrier_doppler_hz_ += carr_error_filt_hz;
                code_freq_chips_ = c_.code_rate_hz + ((carrier_doppler_hz_ * c_.code_rate_hz) / glonass_freq_ch_);
                const double code_error_chips = dll_nc_e_minus_l_normalized(outs_[0], outs_[2]);
                const double code_error_filt_chips = code_loop_filter_.get_code_nco(code_error_chips);
                const double T_chip_seconds = 1.0 / static_cast<double>(code_freq_chips_);
                const double T_prn_seconds = T_chip_seconds * c_.code_length_chips;
                const double code_error_filt_secs = (T_prn_seconds * code_error_filt_chips * T_chip_seconds);
                const double T_prn_samples = T_prn_seconds * static_cast<double>(c_.fs_in);
                const double K_blk_samples = T_prn_samples + rem_code_phase_samples_ + code_error_filt_secs * static_cast<double>(c_.fs_in);
                const int32_t this_block = current_prn_length_samples_;
                current_prn_length_samples_ = round(K_blk_samples);
                carrier_doppler_phase_step_rad_ = kTwoPi * carrier_doppler_hz_ / static_cast<double>(c_.fs_in);
                carrier_phase_step_rad_ = kTwoPi * carrier_frequency_hz_ / static_cast<double>(c_.fs_in);
                rem_carr_phase_rad_ = rem_carr_phase_rad_ + carrier_phase_step_rad_ * current_prn_length_samples_;
                rem_carr_phase_rad_ = fmod(rem_carr_phase_rad_, kTwoPi);
                acc_carrier_phase_rad_ -= carrier_doppler_phase_step_rad_ * current_prn_length_samples_;
                code_phase_step_chips_ = code_freq_chips_ / static_cast<double>(c_.fs_in);
                rem_code_phase_samples_ = K_blk_samples - current_prn_length_samples_;
                rem_code_phase_chips_ = code_freq_chips_ * (rem_code_phase_samples_ / static_cast<double>(c_.fs_in));
                if (cn0_estimation_counter_ < kCn0EstimationSamples)
                    {
                        prompt_buffer_[cn0_estimation_counter_] = outs_[1];
                        cn0_estimation_counter_++;
                    }
                else
                    {
                        cn0_estimation_counter_ = 0;
                        cn0_snv_db_hz_ = cn0_svn_estimator(prompt_buffer_.data(), kCn0EstimationSamples, c_.code_period_s);
                        carrier_lock_test_ = carrier_lock_detector(prompt_buffer_.data(), kCn0EstimationSamples);
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
                (void)this_block;
            }
        else
            {
                for (auto& o : outs_) o = cfloat(0, 0);
                e->tracking_sample_counter = sample_counter_ + static_cast<uint64_t>(current_prn_length_samples_);
            }
        // consume_each(d_current_prn_length_samples); d_sample_counter += d_current_prn_length_samples  (:762-763): the block
        // consumes the NEXT block's length, as the reference does
        e->prn_length_samples = current_prn_length_samples_;
        sample_counter_ += current_prn_length_samples_;
        return true;
    }
    const cfloat* correlator_outs() const { return outs_; }
    bool tracking_enabled() const { return enable_tracking_; }

private:
    GlonassTrkConf c_;
    Correlator* corr_;
    Tracking2ndDllFilter code_loop_filter_;
    Tracking2ndPllFilter carrier_loop_filter_;
    cfloat outs_[3];
    float shifts_[3];
    std::vector<cfloat> prompt_buffer_;
    int64_t glonass_freq_ch_ = 0;
    double early_late_spc_chips_ = 0, rem_code_phase_samples_ = 0, rem_code_phase_chips_ = 0, rem_carr_phase_rad_ = 0;
    double acq_code_phase_samples_ = 0, acq_carrier_doppler_hz_ = 0;
    double code_freq_chips_ = 0, code_phase_step_chips_ = 0, carrier_doppler_hz_ = 0, carrier_doppler_phase_step_rad_ = 0;
    double carrier_frequency_hz_ = 0, carrier_phase_step_rad_ = 0, acc_carrier_phase_rad_ = 0;
    int32_t current_prn_length_samples_ = 0;
    uint64_t sample_counter_ = 0, acq_sample_stamp_ = 0;
    int32_t cn0_estimation_counter_ = 0;
    double carrier_lock_test_ = 1, cn0_snv_db_hz_ = 0, carrier_lock_threshold_ = 0.85;
    int32_t carrier_lock_fail_counter_ = 0;
    bool enable_tracking_ = false, pull_in_ = false;
};
}  // namespace gb200

// flat images for the ctypes-facing test libraries (host_abi.cc and tests/native/host_oracle_loop.cc)
inline void gbh_glo_conf(const double* conf, gb200::GlonassTrkConf* c)
{
    c->fs_in = static_cast<int64_t>(conf[0]);
    c->vector_length = static_cast<uint32_t>(conf[1]);
    c->pll_bw_hz = static_cast<float>(conf[2]);
    c->dll_bw_hz = static_cast<float>(conf[3]);
    c->early_late_space_chips = static_cast<float>(conf[4]);
    c->band_freq_hz = conf[5];
    c->band_dfreq_hz = conf[6];
}
inline void gbh_glo_fill(const gb200::GlonassEpoch& e, double* o)
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
}

