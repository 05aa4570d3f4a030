// Host-side DLL/PLL(-VEML) epoch loop on top of a multicorrelator, the part of dll_pll_veml_tracking that stays on the
// CPU when the correlations move to the B200.
//
// Mirrors src/algorithms/tracking/gnuradio_blocks/dll_pll_veml_tracking.cc without GNU Radio: instead of being scheduled
// with a window of the input stream, the loop keeps the absolute sample counter and asks the correlator for the samples
// at that index (the device ring holds the stream).  Implemented: start_tracking (:549-747), the pull-in alignment
// (state 1, :1568-1600), wide tracking (state 2, :1601-1773) with do_correlation_step (:886-911), the lock test
// (:839-878), run_dll_pll (:914-973) and update_tracking_vars (:998-1070), pilot tracking with the data-prompt tap, and
// the Gnss_Synchro fields of one epoch (:1720-1725, 1897-1904).  Not implemented here (the reference keeps doing them):
// secondary-code / preamble synchronisation, extended integration (states 3/4), dump files, message ports.
//
// Correlator concept (same calls as Cpu_Multicorrelator_Real_Codes, cpu_multicorrelator_real_codes.h:46-58):
//   bool set_local_code_and_taps(int code_length_samples, const float* code, const float* shifts, int n_taps);
//   bool set_data_code(const float* code, int code_length_samples, int prompt_tap);      // optional second code
//   bool correlate(uint64_t first_sample, float rem_carr_phase_rad, float phase_step_rad, float phase_rate_step_rad,
//                  float rem_code_phase_chips, float code_phase_step_chips, float code_phase_rate_step_chips,
//                  int n_samples, std::complex<float>* out /* n_taps (+1) */);
#pragma once
#include <cstdint>
#include <vector>

#include "gb200_loop_math.hpp"

namespace gb200
{
// = Dll_Pll_Conf (src/algorithms/tracking/libs/dll_pll_conf.h:39-80) restricted to what the implemented states read,
// plus the per-signal constants start_tracking derives from system / signal (:113-336)
struct DllPllConf
{
    double fs_in = 4000000.0;
    int vector_length = 4000;
    float pll_bw_hz = 35.0f, dll_bw_hz = 2.0f, fll_bw_hz = 35.0f;
    int pll_filter_order = 3, dll_filter_order = 2;
    bool enable_fll_pull_in = false, enable_fll_steady_state = false;
    float pull_in_time_s = 2.0f;
    float early_late_space_chips = 0.5f, very_early_late_space_chips = 0.5f;
    int cn0_samples = 20;
    float cn0_min = 25.0f;
    int max_lock_fail = 50;
    float carrier_lock_th = 0.85f;
    bool track_pilot = false;
    bool veml = false;               // five taps VE,E,P,L,VL (Galileo E1) instead of E,P,L
    double signal_carrier_freq = 1575.42e6;
    double code_chip_rate = 1.023e6;
    int code_length_chips = 1023;
    int code_samples_per_chip = 1;   // 2 for the sinBOC(1,1) tables
    double code_period = 0.001;
};

// the Gnss_Synchro fields one epoch writes (gnss_synchro.h:54-75)
struct TrackingEpoch
{
    bool flag_valid_symbol_output = false;
    uint64_t tracking_sample_counter = 0;
    double prompt_i = 0, prompt_q = 0;
    double code_phase_samples = 0, carrier_phase_rads = 0, carrier_doppler_hz = 0, cn0_db_hz = 0;
    double code_error_chips = 0, carr_phase_error_hz = 0, code_freq_chips = 0;
    int current_prn_length_samples = 0;
    bool loss_of_lock = false;  // the "events" message 3 (:873)
    cfloat taps[6];
};

template <class Correlator>
class DllPllVemlTracking
{
public:
    DllPllVemlTracking(const DllPllConf& conf, Correlator* correlator) : c_(conf), corr_(correlator)
    {
        n_taps_ = c_.veml ? 5 : 3;
        prompt_ = c_.veml ? 2 : 1;
        shifts_.assign(n_taps_, 0.0f);
        prompt_buffer_.assign(c_.cn0_samples, cfloat(0, 0));
        // constructor part of the reference (:455-470): filters configured once
        code_filter_ = LoopFilter(static_cast<float>(c_.code_period), c_.dll_bw_hz, c_.dll_filter_order, false);
        carrier_filter_.set_params(c_.fll_bw_hz, c_.pll_bw_hz, c_.pll_filter_order);
    }

    // = start_tracking (:549-747).  `code` is the 1-or-2-samples-per-chip float table the reference generates there.
    bool start_tracking(const float* code, const float* data_code, double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp,
        uint64_t sample_counter_now)
    {
        acq_code_phase_samples_ = acq_delay_samples;
        acq_carrier_doppler_hz_ = acq_doppler_hz;
        acq_sample_stamp_ = acq_samplestamp;
        sample_counter_ = sample_counter_now;
        carrier_doppler_hz_ = acq_doppler_hz;
        carrier_phase_step_rad_ = kTwoPi * carrier_doppler_hz_ / c_.fs_in;
        carrier_phase_rate_step_rad_ = 0.0;
        carrier_filter_.initialize(static_cast<float>(acq_doppler_hz));
        code_filter_.initialize();
        const float spc = static_cast<float>(c_.code_samples_per_chip);
        if (c_.veml)
            {
                shifts_[0] = -c_.very_early_late_space_chips * spc;
                shifts_[1] = -c_.early_late_space_chips * spc;
                shifts_[2] = 0.0f;
                shifts_[3] = c_.early_late_space_chips * spc;
                shifts_[4] = c_.very_early_late_space_chips * spc;
            }
        else
            {
                shifts_[0] = -c_.early_late_space_chips * spc;
                shifts_[1] = 0.0f;
                shifts_[2] = c_.early_late_space_chips * spc;
            }
        if (!corr_->set_local_code_and_taps(c_.code_samples_per_chip * c_.code_length_chips, code, shifts_.data(), n_taps_)) return false;
        if (c_.track_pilot && data_code && !corr_->set_data_code(data_code, c_.code_samples_per_chip * c_.code_length_chips, prompt_)) return false;
        carrier_lock_fail_counter_ = 0;
        rem_code_phase_samples_ = 0.0;
        rem_carr_phase_rad_ = 0.0f;
        rem_code_phase_chips_ = 0.0;
        acc_carrier_phase_rad_ = 0.0;
        cn0_estimation_counter_ = 0;
        carrier_lock_test_ = 1.0f;
        cn0_db_hz_ = 0.0f;
        current_correlation_time_s_ = c_.code_period;
        code_filter_.set_noise_bandwidth(c_.dll_bw_hz);
        code_filter_.set_update_interval(static_cast<float>(c_.code_period));
        p_accu_old_ = cfloat(0, 0);
        code_phase_rate_step_chips_ = 0.0;
        state_ = 1;
        cloop_ = true;
        pull_in_transitory_ = true;
        return true;
    }

    void stop_tracking() { state_ = 0; }
    int state() const { return state_; }
    uint64_t sample_counter() const { return sample_counter_; }

    // One call of general_work (:1544-1907).  Returns false when the correlator failed (samples not in the ring).
    bool step(TrackingEpoch* out)
    {
        *out = TrackingEpoch();
        if (pull_in_transitory_)
            {
                // (:1552-1558) integer division of the elapsed samples by the integer sampling rate, as written there
                if (c_.pull_in_time_s < static_cast<float>((sample_counter_ - acq_sample_stamp_) / static_cast<uint64_t>(static_cast<int>(c_.fs_in))))
                    pull_in_transitory_ = false;
            }
        if (state_ == 0) return true;
        if (state_ == 1)
            {
                // signal alignment: skip samples until the incoming code starts with the local replica (:1568-1600)
                const int64_t acq_trk_diff_samples = static_cast<int64_t>(sample_counter_) - static_cast<int64_t>(acq_sample_stamp_);
                const double delta = static_cast<double>(acq_trk_diff_samples) - acq_code_phase_samples_;
                code_freq_chips_ = c_.code_chip_rate;
                code_phase_step_chips_ = code_freq_chips_ / c_.fs_in;
                code_phase_rate_step_chips_ = 0.0;
                const double t_prn_mod_samples = (1.0 / code_freq_chips_) * static_cast<double>(c_.code_length_chips) * c_.fs_in;
                acq_code_phase_samples_ = t_prn_mod_samples - std::fmod(delta, t_prn_mod_samples);
                current_prn_length_samples_ = static_cast<int32_t>(std::round(t_prn_mod_samples));
                const int32_t samples_offset = static_cast<int32_t>(std::round(acq_code_phase_samples_));
                acc_carrier_phase_rad_ -= carrier_phase_step_rad_ * static_cast<double>(samples_offset);
                state_ = 2;
                sample_counter_ += static_cast<uint64_t>(samples_offset);
                return true;
            }
        // ---- state 2: wide tracking ----
        cfloat taps[6] = {};
        const float spc = static_cast<float>(c_.code_samples_per_chip);
        // do_correlation_step (:886-911): float casts of the double NCO state at the call, vector_length samples
        if (!corr_->correlate(sample_counter_, rem_carr_phase_rad_, static_cast<float>(carrier_phase_step_rad_), static_cast<float>(carrier_phase_rate_step_rad_),
                static_cast<float>(rem_code_phase_chips_) * spc, static_cast<float>(code_phase_step_chips_) * spc,
                static_cast<float>(code_phase_rate_step_chips_) * spc, c_.vector_length, taps))
            return false;
        cfloat ve(0, 0), vl(0, 0), e, p, l;
        if (c_.veml)
            {
                ve = taps[0];
                e = taps[1];
                p = taps[2];
                l = taps[3];
                vl = taps[4];
            }
        else
            {
                e = taps[0];
                p = taps[1];
                l = taps[2];
            }
        for (int t = 0; t < 6; t++) out->taps[t] = taps[t];
        if (!lock_status(p, c_.code_period))
            {
                // loss of lock (:1616-1620)
                out->loss_of_lock = true;
                p_accu_old_ = cfloat(0, 0);
                state_ = 0;
            }
        else
            {
                run_dll_pll(ve, e, p, l, vl);
                update_tracking_vars();
                const cfloat prompt_out = c_.track_pilot ? taps[n_taps_] : p;  // data component when tracking the pilot (:1707-1718)
                out->prompt_i = static_cast<double>(prompt_out.real());
                out->prompt_q = static_cast<double>(prompt_out.imag());
                out->code_phase_samples = rem_code_phase_samples_;
                out->carrier_phase_rads = acc_carrier_phase_rad_;
                out->carrier_doppler_hz = carrier_doppler_hz_;
                out->cn0_db_hz = cn0_db_hz_;
                out->code_error_chips = code_error_chips_;
                out->carr_phase_error_hz = carr_phase_error_hz_;
                out->code_freq_chips = code_freq_chips_;
                out->flag_valid_symbol_output = true;
            }
        // common tail (:1897-1906)
        out->current_prn_length_samples = current_prn_length_samples_;
        sample_counter_ += static_cast<uint64_t>(current_prn_length_samples_);
        out->tracking_sample_counter = sample_counter_;
        return true;
    }

private:
    // = cn0_and_tracking_lock_status (:839-878)
    bool lock_status(cfloat p_accu, double coh_integration_time_s)
    {
        if (cn0_estimation_counter_ < c_.cn0_samples)
            {
                prompt_buffer_[cn0_estimation_counter_++] = p_accu;
                return true;
            }
        cn0_estimation_counter_ = 0;
        cn0_db_hz_ = cn0_svn_estimator(prompt_buffer_.data(), c_.cn0_samples, coh_integration_time_s);
        carrier_lock_test_ = carrier_lock_detector(prompt_buffer_.data(), c_.cn0_samples);
        if (!pull_in_transitory_)
            {
                if (carrier_lock_test_ < c_.carrier_lock_th || cn0_db_hz_ < c_.cn0_min)
                    carrier_lock_fail_counter_++;
                else if (carrier_lock_fail_counter_ > 0)
                    carrier_lock_fail_counter_--;
            }
        if (carrier_lock_fail_counter_ > c_.max_lock_fail)
            {
                carrier_lock_fail_counter_ = 0;
                return false;
            }
        return true;
    }

    // = run_dll_pll (:914-973)
    void run_dll_pll(cfloat ve, cfloat e, cfloat p, cfloat l, cfloat vl)
    {
        // d_cloop is set in start_tracking and only cleared by save_correlation_results (states 3/4), so wide tracking always runs the Costas form
        carr_phase_error_hz_ = cloop_ ? pll_cloop_two_quadrant_atan(p) / kTwoPi : pll_four_quadrant_atan(p) / kTwoPi;
        const float t = static_cast<float>(current_correlation_time_s_);
        if ((pull_in_transitory_ && c_.enable_fll_pull_in) || c_.enable_fll_steady_state)
            {
                const double carr_freq_error_hz = fll_four_quadrant_atan(p_accu_old_, p, 0, current_correlation_time_s_) / kTwoPi;
                p_accu_old_ = p;
                if (pull_in_transitory_ && c_.enable_fll_pull_in)
                    carr_error_filt_hz_ = carrier_filter_.get_carrier_error(static_cast<float>(carr_freq_error_hz), 0.0f, t);
                else
                    carr_error_filt_hz_ = carrier_filter_.get_carrier_error(static_cast<float>(carr_freq_error_hz), static_cast<float>(carr_phase_error_hz_), t);
            }
        else
            carr_error_filt_hz_ = carrier_filter_.get_carrier_error(0.0f, static_cast<float>(carr_phase_error_hz_), t);
        carrier_doppler_hz_ = carr_error_filt_hz_;
        code_error_chips_ = c_.veml ? dll_nc_vemlp_normalized(ve, e, l, vl) : dll_nc_e_minus_l_normalized(e, l);
        code_error_filt_chips_ = code_filter_.apply(static_cast<float>(code_error_chips_));
        code_freq_chips_ = (1.0 + (carrier_doppler_hz_ / c_.signal_carrier_freq)) * c_.code_chip_rate - code_error_filt_chips_;
    }

    // = update_tracking_vars (:998-1070), high_dyn smoothing excluded
    void update_tracking_vars()
    {
        const double t_chip_seconds = 1.0 / code_freq_chips_;
        const double t_prn_seconds = t_chip_seconds * static_cast<double>(c_.code_length_chips);
        const double t_prn_samples = t_prn_seconds * c_.fs_in;
        const double k_blk_samples = t_prn_samples + rem_code_phase_samples_;
        current_prn_length_samples_ = static_cast<int32_t>(std::floor(k_blk_samples));
        carrier_phase_step_rad_ = kTwoPi * carrier_doppler_hz_ / c_.fs_in;
        const double n = static_cast<double>(current_prn_length_samples_);
        // the remnant carrier phase is a FLOAT member in the reference (:1036-1037): float +=, then fmod in double, stored float
        rem_carr_phase_rad_ += static_cast<float>(carrier_phase_step_rad_ * n + 0.5 * carrier_phase_rate_step_rad_ * n * n);
        rem_carr_phase_rad_ = static_cast<float>(std::fmod(rem_carr_phase_rad_, kTwoPi));
        acc_carrier_phase_rad_ -= (carrier_phase_step_rad_ * n + 0.5 * carrier_phase_rate_step_rad_ * n * n);
        code_phase_step_chips_ = code_freq_chips_ / c_.fs_in;
        rem_code_phase_samples_ = k_blk_samples - n;
        rem_code_phase_chips_ = code_freq_chips_ * rem_code_phase_samples_ / c_.fs_in;
    }

    DllPllConf c_;
    Correlator* corr_;
    int n_taps_ = 3, prompt_ = 1;
    std::vector<float> shifts_;
    std::vector<cfloat> prompt_buffer_;
    LoopFilter code_filter_;
    FllPllFilter carrier_filter_;
    // state named after the reference's members
    int state_ = 0;
    bool cloop_ = true, pull_in_transitory_ = false;
    uint64_t sample_counter_ = 0, acq_sample_stamp_ = 0;
    double acq_code_phase_samples_ = 0, acq_carrier_doppler_hz_ = 0;
    double carrier_doppler_hz_ = 0, carrier_phase_step_rad_ = 0, carrier_phase_rate_step_rad_ = 0, acc_carrier_phase_rad_ = 0;
    float rem_carr_phase_rad_ = 0;
    double code_freq_chips_ = 0, code_phase_step_chips_ = 0, code_phase_rate_step_chips_ = 0;
    double rem_code_phase_samples_ = 0, rem_code_phase_chips_ = 0;
    double carr_phase_error_hz_ = 0, carr_error_filt_hz_ = 0, code_error_chips_ = 0, code_error_filt_chips_ = 0;
    double current_correlation_time_s_ = 0.001;
    int32_t current_prn_length_samples_ = 0;
    int cn0_estimation_counter_ = 0, carrier_lock_fail_counter_ = 0;
    float cn0_db_hz_ = 0, carrier_lock_test_ = 1;
    cfloat p_accu_old_{0, 0};
};
}  // namespace gb200
