// Host-side DLL/PLL(-VEML) epoch loop on top of a multicorrelator, the part of dll_pll_veml_tracking that stays on the
// CPU when the correlations move to the B200.
//
// Mirrors src/algorithms/tracking/gnuradio_blocks/dll_pll_veml_tracking.cc without GNU Radio: instead of being scheduled
// with a window of the input stream, the loop keeps the absolute sample counter and asks the correlator for the samples
// at that index (the device ring holds the stream).  Implemented: start_tracking (:549-747), the pull-in alignment
// (state 1, :1568-1600), wide tracking (state 2, :1601-1773) with do_correlation_step (:886-911), secondary-code lock
// (acquire_secondary, :798-836) and the preamble search for signals without one (:1643-1683), the extended coherent
// integration (state 3, :1774-1826; save_correlation_results, :1073-1125), narrow tracking with the narrow tap spacings and
// bandwidths (state 4, :1827-1895; switch at :1737-1764), the lock test (:839-878), run_dll_pll (:914-973),
// update_tracking_vars (:998-1070) including the high-dynamics rate smoothing, pilot tracking with the data-prompt tap,
// the Gnss_Synchro fields of one epoch (:1720-1725, 1897-1904) and the binary dump record (log_data, :1128-1250).
// Not here: GNU Radio message ports (loss of lock is a flag of the epoch), the .mat conversion of the dump.
//
// Correlator concept (same calls as Cpu_Multicorrelator_Real_Codes, cpu_multicorrelator_real_codes.h:46-58):
//   bool set_local_code_and_taps(int code_length_samples, const float* code, const float* shifts, int n_taps);
//   bool set_data_code(const float* code, int code_length_samples, int prompt_tap);      // optional second code
//   bool set_taps(const float* shifts, int n_taps);   // the reference edits d_local_code_shift_chips in place (:1753-1764)
//   bool correlate(uint64_t first_sample, float rem_carr_phase_rad, float phase_step_rad, float phase_rate_step_rad,
//                  float rem_code_phase_chips, float code_phase_step_chips, float code_phase_rate_step_chips,
//                  int n_samples, std::complex<float>* out /* n_taps (+1) */);
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <deque>
#include <string>
#include <utility>
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
    // narrow tracking / extended integration (dll_pll_conf.cc:56-63)
    float pll_bw_narrow_hz = 5.0f, dll_bw_narrow_hz = 0.75f;
    float early_late_space_narrow_chips = 0.1f, very_early_late_space_narrow_chips = 0.1f;
    int extend_correlation_symbols = 1;
    // high dynamics (:38-39): carrier / code rate from a two-window moving average of the NCO steps
    bool high_dyn = false;
    int smoother_length = 10;
    // what the constructor derives from system / signal (:113-336)
    bool secondary = false;            // the tracked component carries a secondary code
    std::string secondary_code;        // '0' / '1' string, e.g. GALILEO_E1_C_SECONDARY_CODE
    int symbols_per_bit = 1;           // > 1 without secondary code: wait for the preamble before leaving state 2
    std::vector<int32_t> preamble_symbols;  // +-1 per symbol (GPS L1 C/A: 8 bits x 20)
    bool interchange_iq = false;       // pilot in quadrature with the data component (L5, E5a)
    int correlation_length_ms = 1;
    uint32_t prn = 0;
    // log_data (:1128-1250): one 96-byte record per DLL/PLL cycle appended to this file when not empty
    std::string dump_filename;
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
    int correlation_length_ms = 0;
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
        // (:388-396)
        if (c_.extend_correlation_symbols > 1)
            enable_extended_integration_ = true;
        else
            {
                enable_extended_integration_ = false;
                c_.extend_correlation_symbols = 1;
            }
        if (c_.secondary_code.empty()) c_.secondary = false;
        if (c_.symbols_per_bit < 1) c_.symbols_per_bit = 1;
        secondary_code_length_ = c_.secondary ? static_cast<uint32_t>(c_.secondary_code.size()) : 0U;
        history_capacity_ = c_.smoother_length > 0 ? static_cast<size_t>(c_.smoother_length) * 2 : 1;  // (:440-449)
    }
    ~DllPllVemlTracking()
    {
        if (dump_) std::fclose(dump_);
    }
    DllPllVemlTracking(const DllPllVemlTracking&) = delete;
    DllPllVemlTracking& operator=(const DllPllVemlTracking&) = delete;

    // = set_channel (:1498-1525): the dump file is <dump_filename><channel>.dat
    void set_channel(uint32_t channel)
    {
        channel_ = channel;
        if (!c_.dump_filename.empty() && !dump_)
            {
                std::string name = c_.dump_filename;
                const size_t slash = name.find_last_of('/');
                const size_t dot = name.find_last_of('.');
                if (dot != std::string::npos && dot > 0 && (slash == std::string::npos || dot > slash + 1)) name = name.substr(0, dot);  // (:489-492)
                name += std::to_string(channel_) + ".dat";
                dump_ = std::fopen(name.c_str(), "wb");
            }
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
        carr_ph_history_.clear();
        code_ph_history_.clear();
        carrier_filter_.initialize(static_cast<float>(acq_doppler_hz));
        code_filter_.initialize();
        set_wide_shifts();
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
        // the reference re-arms the carrier filter's bandwidth only in its constructor; a restart after narrow tracking keeps the
        // narrow PLL bandwidth there (:1751) and therefore here
        p_accu_old_ = cfloat(0, 0);
        code_phase_rate_step_chips_ = 0.0;
        state_ = 1;
        cloop_ = true;
        pull_in_transitory_ = true;
        prompt_circular_buffer_.clear();
        return true;
    }

    void stop_tracking() { state_ = 0; }
    // the "telemetry_to_trk" message port (dll_pll_veml_tracking.cc:516-536): int 1 = telemetry fault in this channel, the
    // decoder asks for a re-acquisition by forcing the loss-of-lock condition at the next lock test
    void telemetry_to_trk(int tlm_event)
    {
        if (tlm_event == 1) carrier_lock_fail_counter_ = 10000;
    }
    int state() const { return state_; }
    uint64_t sample_counter() const { return sample_counter_; }
    double current_correlation_time_s() const { return current_correlation_time_s_; }

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
        switch (state_)
            {
            case 0:
                return true;
            case 1:
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
            case 2:  // wide tracking and symbol synchronisation (:1601-1773)
                {
                    if (!do_correlation_step(out)) return false;
                    if (c_.veml)
                        {
                            ve_accu_ = taps_[0];
                            vl_accu_ = taps_[4];
                        }
                    e_accu_ = taps_[prompt_ - 1];
                    p_accu_ = taps_[prompt_];
                    l_accu_ = taps_[prompt_ + 1];
                    if (!lock_status(c_.code_period))
                        {
                            clear_tracking_vars();
                            out->loss_of_lock = true;
                            state_ = 0;
                            break;
                        }
                    bool next_state = false;
                    run_dll_pll();
                    update_tracking_vars();
                    log_data(false);
                    if (c_.secondary)
                        {
                            // secondary code lock (:1628-1641): a sliding window of the last `length` prompts
                            prompt_circular_buffer_.push_back(taps_[prompt_]);
                            if (prompt_circular_buffer_.size() > secondary_code_length_) prompt_circular_buffer_.pop_front();
                            if (prompt_circular_buffer_.size() == secondary_code_length_) next_state = acquire_secondary();
                        }
                    else if (c_.symbols_per_bit > 1)
                        {
                            // no secondary code: look for the telemetry preamble in the prompt signs after 10 s (:1643-1683)
                            const float current_tracking_time_s = static_cast<float>(sample_counter_ - acq_sample_stamp_) / static_cast<float>(c_.fs_in);
                            if (current_tracking_time_s > 10)
                                {
                                    const size_t len = c_.preamble_symbols.size();
                                    symbol_history_.push_back(taps_[prompt_].real());
                                    if (symbol_history_.size() > len) symbol_history_.pop_front();
                                    int32_t corr_value = 0;
                                    if (symbol_history_.size() == len)
                                        for (size_t i = 0; i < len; i++) corr_value += symbol_history_[i] < 0.0f ? -c_.preamble_symbols[i] : c_.preamble_symbols[i];
                                    next_state = len > 0 && corr_value == static_cast<int32_t>(len);
                                }
                        }
                    else
                        next_state = true;
                    fill_epoch(out);
                    if (next_state)
                        {
                            reset_accumulators();
                            prompt_circular_buffer_.clear();
                            current_symbol_ = 0;
                            if (enable_extended_integration_)
                                {
                                    // (:1737-1764) longer integration, narrow loop bandwidths, narrow tap spacings
                                    extend_correlation_symbols_count_ = 0;
                                    current_correlation_time_s_ = static_cast<float>(c_.extend_correlation_symbols) * static_cast<float>(c_.code_period);
                                    state_ = 3;
                                    code_filter_.set_update_interval(static_cast<float>(current_correlation_time_s_));
                                    code_filter_.set_noise_bandwidth(c_.dll_bw_narrow_hz);
                                    carrier_filter_.set_params(c_.fll_bw_hz, c_.pll_bw_narrow_hz, c_.pll_filter_order);
                                    const float spc = static_cast<float>(c_.code_samples_per_chip);
                                    if (c_.veml)
                                        {
                                            shifts_[0] = -c_.very_early_late_space_narrow_chips * spc;
                                            shifts_[1] = -c_.early_late_space_narrow_chips * spc;
                                            shifts_[3] = c_.early_late_space_narrow_chips * spc;
                                            shifts_[4] = c_.very_early_late_space_narrow_chips * spc;
                                        }
                                    else
                                        {
                                            shifts_[0] = -c_.early_late_space_narrow_chips * spc;
                                            shifts_[2] = c_.early_late_space_narrow_chips * spc;
                                        }
                                    if (!corr_->set_taps(shifts_.data(), n_taps_)) return false;
                                }
                            else
                                state_ = 4;
                        }
                    break;
                }
            case 3:  // coherent integration, no loop update (:1774-1826)
                {
                    if (!do_correlation_step(out)) return false;
                    update_tracking_vars();
                    save_correlation_results();
                    fill_epoch(out);
                    extend_correlation_symbols_count_++;
                    if (extend_correlation_symbols_count_ == c_.extend_correlation_symbols - 1)
                        {
                            extend_correlation_symbols_count_ = 0;
                            state_ = 4;
                        }
                    log_data(true);
                    break;
                }
            case 4:  // narrow tracking: the loops run on the accumulated correlators (:1827-1895)
                {
                    if (!do_correlation_step(out)) return false;
                    save_correlation_results();
                    if (!lock_status(c_.code_period * static_cast<double>(c_.extend_correlation_symbols)))
                        {
                            clear_tracking_vars();
                            out->loss_of_lock = true;
                            state_ = 0;
                            break;
                        }
                    run_dll_pll();
                    update_tracking_vars();
                    fill_epoch(out);
                    log_data(false);
                    reset_accumulators();
                    if (enable_extended_integration_) state_ = 3;
                    break;
                }
            default:
                return true;
            }
        // common tail (:1897-1906)
        out->current_prn_length_samples = current_prn_length_samples_;
        sample_counter_ += static_cast<uint64_t>(current_prn_length_samples_);
        out->tracking_sample_counter = sample_counter_;
        return true;
    }

private:
    void set_wide_shifts()
    {
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
    }

    // = do_correlation_step (:886-911): float casts of the double NCO state at the call, vector_length samples
    bool do_correlation_step(TrackingEpoch* out)
    {
        const float spc = static_cast<float>(c_.code_samples_per_chip);
        for (auto& t : taps_) t = cfloat(0, 0);
        if (!corr_->correlate(sample_counter_, rem_carr_phase_rad_, static_cast<float>(carrier_phase_step_rad_), static_cast<float>(carrier_phase_rate_step_rad_),
                static_cast<float>(rem_code_phase_chips_) * spc, static_cast<float>(code_phase_step_chips_) * spc,
                static_cast<float>(code_phase_rate_step_chips_) * spc, c_.vector_length, taps_))
            return false;
        for (int t = 0; t < 6; t++) out->taps[t] = taps_[t];
        return true;
    }

    // the Gnss_Synchro fields every valid epoch carries (:1685-1725 and twins in states 3 / 4)
    void fill_epoch(TrackingEpoch* out) const
    {
        const cfloat prompt_out = c_.track_pilot ? taps_[n_taps_] : taps_[prompt_];  // data component when tracking the pilot
        out->prompt_i = static_cast<double>(c_.interchange_iq ? prompt_out.imag() : prompt_out.real());
        out->prompt_q = static_cast<double>(c_.interchange_iq ? prompt_out.real() : prompt_out.imag());
        out->code_phase_samples = rem_code_phase_samples_;
        out->carrier_phase_rads = acc_carrier_phase_rad_;
        out->carrier_doppler_hz = carrier_doppler_hz_;
        out->cn0_db_hz = cn0_db_hz_;
        out->code_error_chips = code_error_chips_;
        out->carr_phase_error_hz = carr_phase_error_hz_;
        out->code_freq_chips = code_freq_chips_;
        out->correlation_length_ms = c_.correlation_length_ms;
        out->flag_valid_symbol_output = true;
    }

    void reset_accumulators() { ve_accu_ = e_accu_ = p_accu_ = l_accu_ = vl_accu_ = cfloat(0, 0); }

    // = acquire_secondary (:798-836): every prompt sign agrees (or every one disagrees) with the secondary code
    bool acquire_secondary() const
    {
        int32_t corr_value = 0;
        for (uint32_t i = 0; i < secondary_code_length_; i++)
            {
                const bool zero = c_.secondary_code[i] == '0';
                if (prompt_circular_buffer_[i].real() < 0.0f)
                    corr_value += zero ? 1 : -1;
                else
                    corr_value += zero ? -1 : 1;
            }
        return std::abs(corr_value) == static_cast<int32_t>(secondary_code_length_);
    }

    // = save_correlation_results (:1073-1125)
    void save_correlation_results()
    {
        float sign = 1.0f;
        if (c_.secondary)
            {
                if (c_.secondary_code[current_symbol_] != '0') sign = -1.0f;
                current_symbol_ = (current_symbol_ + 1) % secondary_code_length_;
            }
        else
            current_symbol_ = (current_symbol_ + 1) % static_cast<uint32_t>(c_.symbols_per_bit);
        if (sign > 0)
            {
                if (c_.veml)
                    {
                        ve_accu_ += taps_[0];
                        vl_accu_ += taps_[4];
                    }
                e_accu_ += taps_[prompt_ - 1];
                p_accu_ += taps_[prompt_];
                l_accu_ += taps_[prompt_ + 1];
            }
        else
            {
                if (c_.veml)
                    {
                        ve_accu_ -= taps_[0];
                        vl_accu_ -= taps_[4];
                    }
                e_accu_ -= taps_[prompt_ - 1];
                p_accu_ -= taps_[prompt_];
                l_accu_ -= taps_[prompt_ + 1];
            }
        cloop_ = !c_.track_pilot;  // tracking the pilot: the Costas discriminator is no longer needed
    }

    // = clear_tracking_vars (:976-995)
    void clear_tracking_vars()
    {
        for (auto& t : taps_) t = cfloat(0, 0);
        p_accu_old_ = cfloat(0, 0);
        carr_phase_error_hz_ = 0.0;
        carr_freq_error_hz_ = 0.0;
        carr_error_filt_hz_ = 0.0;
        code_error_chips_ = 0.0;
        code_error_filt_chips_ = 0.0;
        current_symbol_ = 0;
        prompt_circular_buffer_.clear();
        carrier_phase_rate_step_rad_ = 0.0;
        code_phase_rate_step_chips_ = 0.0;
        carr_ph_history_.clear();
        code_ph_history_.clear();
    }

    // = cn0_and_tracking_lock_status (:839-878)
    bool lock_status(double coh_integration_time_s)
    {
        if (cn0_estimation_counter_ < c_.cn0_samples)
            {
                prompt_buffer_[cn0_estimation_counter_++] = p_accu_;
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

    // = run_dll_pll (:914-973), on the accumulated correlators
    void run_dll_pll()
    {
        // Costas form until the secondary code is wiped (save_correlation_results clears d_cloop when tracking a pilot)
        carr_phase_error_hz_ = cloop_ ? pll_cloop_two_quadrant_atan(p_accu_) / kTwoPi : pll_four_quadrant_atan(p_accu_) / kTwoPi;
        const float t = static_cast<float>(current_correlation_time_s_);
        if ((pull_in_transitory_ && c_.enable_fll_pull_in) || c_.enable_fll_steady_state)
            {
                carr_freq_error_hz_ = fll_four_quadrant_atan(p_accu_old_, p_accu_, 0, current_correlation_time_s_) / kTwoPi;
                p_accu_old_ = p_accu_;
                if (pull_in_transitory_ && c_.enable_fll_pull_in)
                    carr_error_filt_hz_ = carrier_filter_.get_carrier_error(static_cast<float>(carr_freq_error_hz_), 0.0f, t);
                else
                    carr_error_filt_hz_ = carrier_filter_.get_carrier_error(static_cast<float>(carr_freq_error_hz_), static_cast<float>(carr_phase_error_hz_), t);
            }
        else
            carr_error_filt_hz_ = carrier_filter_.get_carrier_error(0.0f, static_cast<float>(carr_phase_error_hz_), t);
        carrier_doppler_hz_ = carr_error_filt_hz_;
        code_error_chips_ = c_.veml ? dll_nc_vemlp_normalized(ve_accu_, e_accu_, l_accu_, vl_accu_) : dll_nc_e_minus_l_normalized(e_accu_, l_accu_);
        code_error_filt_chips_ = code_filter_.apply(static_cast<float>(code_error_chips_));
        code_freq_chips_ = (1.0 + (carrier_doppler_hz_ / c_.signal_carrier_freq)) * c_.code_chip_rate - code_error_filt_chips_;
    }

    // the high-dynamics rate estimate (:1018-1036, 1049-1066): mean step of the newer half of the history minus the mean of
    // the older half, over the samples the newer half spans
    using History = std::deque<std::pair<double, double>>;
    bool push_history(History& h, double step, double samples, double* rate) const
    {
        h.emplace_back(step, samples);
        if (h.size() > history_capacity_) h.pop_front();
        if (h.size() < history_capacity_) return false;
        const unsigned len = static_cast<unsigned>(c_.smoother_length);
        double cp1 = 0.0, cp2 = 0.0, tmp_samples = 0.0;
        for (unsigned k = 0; k < len; k++)
            {
                cp1 += h[k].first;
                cp2 += h[len * 2 - k - 1].first;
                tmp_samples += h[len * 2 - k - 1].second;
            }
        cp1 /= static_cast<double>(len);
        cp2 /= static_cast<double>(len);
        *rate = (cp2 - cp1) / tmp_samples;
        return true;
    }

    // = update_tracking_vars (:998-1070)
    void update_tracking_vars()
    {
        const double t_chip_seconds = 1.0 / code_freq_chips_;
        const double t_prn_seconds = t_chip_seconds * static_cast<double>(c_.code_length_chips);
        const double t_prn_samples = t_prn_seconds * c_.fs_in;
        const double k_blk_samples = t_prn_samples + rem_code_phase_samples_;
        current_prn_length_samples_ = static_cast<int32_t>(std::floor(k_blk_samples));
        carrier_phase_step_rad_ = kTwoPi * carrier_doppler_hz_ / c_.fs_in;
        const double n = static_cast<double>(current_prn_length_samples_);
        if (c_.high_dyn && c_.smoother_length > 0) push_history(carr_ph_history_, carrier_phase_step_rad_, n, &carrier_phase_rate_step_rad_);
        // the remnant carrier phase is a FLOAT member in the reference (:1036-1037): float +=, then fmod in double, stored float
        rem_carr_phase_rad_ += static_cast<float>(carrier_phase_step_rad_ * n + 0.5 * carrier_phase_rate_step_rad_ * n * n);
        rem_carr_phase_rad_ = static_cast<float>(std::fmod(rem_carr_phase_rad_, kTwoPi));
        acc_carrier_phase_rad_ -= (carrier_phase_step_rad_ * n + 0.5 * carrier_phase_rate_step_rad_ * n * n);
        code_phase_step_chips_ = code_freq_chips_ / c_.fs_in;
        if (c_.high_dyn && c_.smoother_length > 0) push_history(code_ph_history_, code_phase_step_chips_, n, &code_phase_rate_step_chips_);
        rem_code_phase_samples_ = k_blk_samples - n;
        rem_code_phase_chips_ = code_freq_chips_ * rem_code_phase_samples_ / c_.fs_in;
    }

    // = log_data (:1128-1250): the 96-byte record tracking_dump_reader.cc and dll_pll_veml_read_tracking_dump.m parse
    void log_data(bool integrating)
    {
        if (!dump_) return;
        const cfloat prompt_out = c_.track_pilot ? taps_[n_taps_] : taps_[prompt_];
        const float prompt_i = c_.interchange_iq ? prompt_out.imag() : prompt_out.real();
        const float prompt_q = c_.interchange_iq ? prompt_out.real() : prompt_out.imag();
        float m[5] = {c_.veml ? std::abs(ve_accu_) : 0.0f, std::abs(e_accu_), std::abs(p_accu_), std::abs(l_accu_), c_.veml ? std::abs(vl_accu_) : 0.0f};
        if (integrating && extend_correlation_symbols_count_ > 0)
            {
                const float scale = static_cast<float>(c_.extend_correlation_symbols) / static_cast<float>(extend_correlation_symbols_count_);
                for (float& v : m) v *= scale;
            }
        unsigned char rec[96];
        size_t o = 0;
        auto put = [&](const void* p, size_t nbytes) {
            std::memcpy(rec + o, p, nbytes);
            o += nbytes;
        };
        auto putf = [&](double v) {
            const float f = static_cast<float>(v);
            put(&f, sizeof f);
        };
        for (float v : m) put(&v, sizeof v);
        put(&prompt_i, sizeof prompt_i);
        put(&prompt_q, sizeof prompt_q);
        const uint64_t stamp = sample_counter_ + static_cast<uint64_t>(current_prn_length_samples_);
        put(&stamp, sizeof stamp);
        putf(acc_carrier_phase_rad_);
        putf(carrier_doppler_hz_);
        putf(carrier_phase_rate_step_rad_ * c_.fs_in * c_.fs_in / kTwoPi);
        putf(code_freq_chips_);
        putf(code_phase_rate_step_chips_ * c_.fs_in * c_.fs_in);
        putf(carr_phase_error_hz_);
        putf(carr_error_filt_hz_);
        putf(code_error_chips_);
        putf(code_error_filt_chips_);
        putf(cn0_db_hz_);
        putf(carrier_lock_test_);
        putf(rem_code_phase_samples_);
        const double stamp_d = static_cast<double>(sample_counter_ + static_cast<uint64_t>(current_prn_length_samples_));
        put(&stamp_d, sizeof stamp_d);
        put(&c_.prn, sizeof c_.prn);
        std::fwrite(rec, 1, o, dump_);
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
    uint32_t channel_ = 0;
    bool cloop_ = true, pull_in_transitory_ = false, enable_extended_integration_ = false;
    uint64_t sample_counter_ = 0, acq_sample_stamp_ = 0;
    double acq_code_phase_samples_ = 0, acq_carrier_doppler_hz_ = 0;
    double carrier_doppler_hz_ = 0, carrier_phase_step_rad_ = 0, carrier_phase_rate_step_rad_ = 0, acc_carrier_phase_rad_ = 0;
    float rem_carr_phase_rad_ = 0;
    double code_freq_chips_ = 0, code_phase_step_chips_ = 0, code_phase_rate_step_chips_ = 0;
    double rem_code_phase_samples_ = 0, rem_code_phase_chips_ = 0;
    double carr_phase_error_hz_ = 0, carr_freq_error_hz_ = 0, carr_error_filt_hz_ = 0, code_error_chips_ = 0, code_error_filt_chips_ = 0;
    double current_correlation_time_s_ = 0.001;
    int32_t current_prn_length_samples_ = 0;
    int cn0_estimation_counter_ = 0, carrier_lock_fail_counter_ = 0;
    float cn0_db_hz_ = 0, carrier_lock_test_ = 1;
    cfloat p_accu_old_{0, 0};
    cfloat taps_[6] = {};
    cfloat ve_accu_{0, 0}, e_accu_{0, 0}, p_accu_{0, 0}, l_accu_{0, 0}, vl_accu_{0, 0};
    uint32_t secondary_code_length_ = 0, current_symbol_ = 0;
    int extend_correlation_symbols_count_ = 0;
    std::deque<cfloat> prompt_circular_buffer_;
    std::deque<float> symbol_history_;
    History carr_ph_history_, code_ph_history_;
    size_t history_capacity_ = 1;
    std::FILE* dump_ = nullptr;
};
}  // namespace gb200
