// GpsL1CaDllPllCAidTracking<Correlator>: host mirror of the carrier-aided GPS L1 C/A trackers
//   gps_l1_ca_dll_pll_c_aid_tracking_cc  (src/algorithms/tracking/gnuradio_blocks/gps_l1_ca_dll_pll_c_aid_tracking_cc.cc,
//                                          Cpu_Multicorrelator, gr_complex outputs)
//   gps_l1_ca_dll_pll_c_aid_tracking_sc  (.../gps_l1_ca_dll_pll_c_aid_tracking_sc.cc, Cpu_Multicorrelator_16sc, lv_16sc_t
//                                          samples, codes and outputs)
// (SURVEY.md 8f N3).  The two blocks run the same loop; they differ in the correlator and in the type of its outputs, which is
// the template parameter here (Correlator::out_t).  Loop: FLL-assisted PLL filter in its 2nd-order form with the FLL input
// held at 0 (Tracking_FLL_PLL_filter), 2nd-order DLL filter (Tracking_2nd_DLL_filter), the block length recomputed from the
// code NCO every epoch, optional extended integration over `extend_correlation_ms` prompts once a preamble time is known.
// GNU Radio's scheduler is replaced by step(): one call = one general_work, samples named by absolute index.
//
// Correlator concept:
//   using out_t = std::complex<float> | std::complex<int16_t>;
//   bool set_local_code_and_taps(int code_len, const std::complex<float>* code, const float* shifts, int n_taps);
//        (the _sc block converts the float code with volk_gnsssdr_32fc_convert_16ic first, :252-255: that is the correlator's job)
//   bool correlate(uint64_t first_sample, float rem_carr_rad, float phase_step_rad, float rem_code_chips, float code_step_chips,
//                  int n_samples, out_t* out);
#pragma once
#include <type_traits>
#include <cmath>
#include <complex>
#include <cstdint>
#include <deque>
#include <vector>

#include "gb200_loop_math.hpp"

namespace gb200
{
struct CAidTrkConf
{
    int64_t fs_in = 4000000;
    uint32_t vector_length = 4000;
    float pll_bw_hz = 50.0f, dll_bw_hz = 2.0f, pll_bw_narrow_hz = 20.0f, dll_bw_narrow_hz = 2.0f, early_late_space_chips = 0.5f;
    int32_t extend_correlation_ms = 1;
    // gnss_sdr_flags.cc:53-59 defaults
    int cn0_samples = 20, cn0_min = 25, max_lock_fail = 50;
    double carrier_lock_th = 0.85;
};

struct CAidEpoch
{
    uint64_t tracking_sample_counter = 0;
    double prompt_i = 0, prompt_q = 0, carrier_phase_rads = 0, carrier_doppler_hz = 0, code_phase_samples = 0, cn0_db_hz = 0;
    float abs_e = 0, abs_p = 0, abs_l = 0;
    bool valid_symbol = false, loss_of_lock = false;
    int32_t correlation_length_samples = 0, correlation_length_ms = 0;
};

template <class Correlator>
class GpsL1CaDllPllCAidTracking
{
public:
    using out_t = typename Correlator::out_t;
    // complex-float correlator outputs = the _cc block, int16 outputs = the _sc block
    static constexpr bool kRetunesDllPdi = std::is_same<out_t, std::complex<float>>::value;
    GpsL1CaDllPllCAidTracking(const CAidTrkConf& c, Correlator* corr) : c_(c), corr_(corr)
    {
        // ctor (gps_l1_ca_dll_pll_c_aid_tracking_sc.cc:110-197)
        correlation_length_samples_ = static_cast<int32_t>(c_.vector_length);
        code_loop_filter_.set_DLL_BW(c_.dll_bw_hz);
        carrier_loop_filter_.set_params(10.0f, c_.pll_bw_hz, 2);
        shifts_[0] = -c_.early_late_space_chips;
        shifts_[1] = 0.0;
        shifts_[2] = c_.early_late_space_chips;
        code_freq_chips_ = kCodeRateHz;
        prompt_buffer_.resize(c_.cn0_samples);
        carrier_lock_threshold_ = c_.carrier_lock_th;
    }
    void set_sample_counter(uint64_t now) { sample_counter_ = now; }
    uint64_t sample_counter() const { return sample_counter_; }
    // = msg_handler_preamble_index (:90-107): the telemetry decoder reports a preamble time; extended integration starts
    void set_preamble_timestamp(double preamble_timestamp_s)
    {
        if (enable_extended_integration_ == false)
            {
                preamble_timestamp_s_ = preamble_timestamp_s;
                enable_extended_integration_ = true;
                preamble_synchronized_ = false;
            }
    }

    // = start_tracking() (:200-285); code = gps_l1_ca_code_gen_complex(.., PRN, 0), 1023 chips
    bool start_tracking(const std::complex<float>* code, double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp)
    {
        acq_code_phase_samples_ = acq_delay_samples;
        acq_carrier_doppler_hz_ = acq_doppler_hz;
        acq_sample_stamp_ = acq_samplestamp;
        const int64_t acq_trk_diff_samples = static_cast<int64_t>(sample_counter_) - static_cast<int64_t>(acq_sample_stamp_);
        const double acq_trk_diff_seconds = static_cast<double>(acq_trk_diff_samples) / static_cast<double>(c_.fs_in);
        const double radial_velocity = (kL1FreqHz + acq_carrier_doppler_hz_) / kL1FreqHz;
        code_freq_chips_ = radial_velocity * kCodeRateHz;
        code_phase_step_chips_ = static_cast<double>(code_freq_chips_) / static_cast<double>(c_.fs_in);
        const double T_chip_mod_seconds = 1.0 / code_freq_chips_;
        const double T_prn_mod_seconds = T_chip_mod_seconds * kCodeLengthChips;
        const double T_prn_mod_samples = T_prn_mod_seconds * static_cast<double>(c_.fs_in);
        correlation_length_samples_ = round(T_prn_mod_samples);
        const double T_prn_true_seconds = kCodeLengthChips / kCodeRateHz;
        const double T_prn_true_samples = T_prn_true_seconds * static_cast<double>(c_.fs_in);
        const double T_prn_diff_seconds = T_prn_true_seconds - T_prn_mod_seconds;
        const double N_prn_diff = acq_trk_diff_seconds / T_prn_true_seconds;
        double corrected_acq_phase_samples = fmod((acq_code_phase_samples_ + T_prn_diff_seconds * N_prn_diff * static_cast<double>(c_.fs_in)), T_prn_true_samples);
        if (corrected_acq_phase_samples < 0) corrected_acq_phase_samples = T_prn_mod_samples + corrected_acq_phase_samples;
        acq_code_phase_samples_ = corrected_acq_phase_samples;
        carrier_doppler_hz_ = acq_carrier_doppler_hz_;
        carrier_phase_step_rad_ = kTwoPi * carrier_doppler_hz_ / static_cast<double>(c_.fs_in);
        carrier_loop_filter_.initialize(acq_carrier_doppler_hz_);  // the carrier loop filter implements the Doppler accumulator
        code_loop_filter_.initialize();
        if (!corr_->set_local_code_and_taps(static_cast<int32_t>(kCodeLengthChips), code, shifts_, 3)) return false;
        for (auto& o : outs_) o = out_t(0, 0);
        carrier_lock_fail_counter_ = 0;
        rem_code_phase_samples_ = 0.0;
        rem_carrier_phase_rad_ = 0.0;
        rem_code_phase_chips_ = 0.0;
        acc_carrier_phase_cycles_ = 0.0;
        pll_to_dll_assist_secs_Ti_ = 0.0;
        pull_in_ = true;
        enable_tracking_ = true;
        enable_extended_integration_ = false;
        preamble_synchronized_ = false;
        E_history_.clear();
        P_history_.clear();
        L_history_.clear();
        return true;
    }

    // = general_work (:578-915).  false: the correlator refused the request (samples not in the ring)
    bool step(CAidEpoch* e)
    {
        *e = CAidEpoch();
        double code_error_filt_secs_Ti = 0.0;
        double CURRENT_INTEGRATION_TIME_S = 0.0;
        if (enable_tracking_)
            {
                if (pull_in_)
                    {
                        const int32_t acq_to_trk_delay_samples = sample_counter_ - acq_sample_stamp_;
                        const double acq_trk_shif_correction_samples =
                            correlation_length_samples_ - fmod(static_cast<double>(acq_to_trk_delay_samples), static_cast<double>(correlation_length_samples_));
                        const int32_t samples_offset = round(acq_code_phase_samples_ + acq_trk_shif_correction_samples);
                        e->tracking_sample_counter = sample_counter_ + static_cast<uint64_t>(samples_offset);
                        sample_counter_ += static_cast<uint64_t>(samples_offset);
                        pull_in_ = false;
                        acc_carrier_phase_cycles_ -= carrier_phase_step_rad_ * samples_offset / kTwoPi;
                        e->carrier_phase_rads = acc_carrier_phase_cycles_ * kTwoPi;
                        e->carrier_doppler_hz = carrier_doppler_hz_;
                        e->correlation_length_samples = samples_offset;
                        return true;  // consume_each(samples_offset); return 1
                    }
                if (!corr_->correlate(sample_counter_, rem_carrier_phase_rad_, carrier_phase_step_rad_, rem_code_phase_chips_, code_phase_step_chips_,
                        correlation_length_samples_, outs_))
                    return false;
                const int32_t this_block = correlation_length_samples_;
                // history for the extended correlator (:619-628)
                E_history_.push_back(outs_[0]);
                P_history_.push_back(outs_[1]);
                L_history_.push_back(outs_[2]);
                if (static_cast<int32_t>(P_history_.size()) > c_.extend_correlation_ms)
                    {
                        E_history_.pop_front();
                        P_history_.pop_front();
                        L_history_.pop_front();
                    }
                bool enable_dll_pll;
                if (enable_extended_integration_ == true)
                    {
                        const int64_t symbol_diff =
                            round(1000.0 * ((static_cast<double>(sample_counter_) + rem_code_phase_samples_) / static_cast<double>(c_.fs_in) - preamble_timestamp_s_));
                        if (symbol_diff > 0 and symbol_diff % c_.extend_correlation_ms == 0)
                            {
                                // coherent sum of the last prompts, in the correlator's own output type (:639-647)
                                outs_[0] = out_t(0, 0);
                                outs_[1] = out_t(0, 0);
                                outs_[2] = out_t(0, 0);
                                for (int32_t n = 0; n < c_.extend_correlation_ms; n++)
                                    {
                                        outs_[0] += E_history_.at(n);
                                        outs_[1] += P_history_.at(n);
                                        outs_[2] += L_history_.at(n);
                                    }
                                if (preamble_synchronized_ == false)
                                    {
                                        code_loop_filter_.set_DLL_BW(c_.dll_bw_narrow_hz);
                                        carrier_loop_filter_.set_params(10.0f, c_.pll_bw_narrow_hz, 2);
                                        preamble_synchronized_ = true;
                                    }
                                CURRENT_INTEGRATION_TIME_S = static_cast<double>(c_.extend_correlation_ms) * kCodePeriodS;
                                // the _cc block also retunes the DLL filter's integration time here (gps_l1_ca_dll_pll_c_aid_tracking_cc.cc:669);
                                // the _sc block does not (gps_l1_ca_dll_pll_c_aid_tracking_sc.cc:666-668) -- found by running both reference blocks
                                if (kRetunesDllPdi) code_loop_filter_.set_pdi(static_cast<float>(CURRENT_INTEGRATION_TIME_S));
                                enable_dll_pll = true;
                            }
                        else if (preamble_synchronized_ == true)
                            {
                                // between updates: keep the NCOs running with the last estimates (:666-697)
                                const double T_chip_seconds = 1.0 / code_freq_chips_;
                                const double T_prn_seconds = T_chip_seconds * kCodeLengthChips;
                                const double T_prn_samples = T_prn_seconds * static_cast<double>(c_.fs_in);
                                const int32_t K_prn_samples = round(T_prn_samples);
                                const double K_T_prn_error_samples = K_prn_samples - T_prn_samples;
                                rem_code_phase_samples_ = rem_code_phase_samples_ - K_T_prn_error_samples;
                                rem_code_phase_integer_samples_ = round(rem_code_phase_samples_);
                                correlation_length_samples_ = K_prn_samples + rem_code_phase_integer_samples_;
                                rem_code_phase_samples_ = rem_code_phase_samples_ - rem_code_phase_integer_samples_;
                                code_phase_step_chips_ = code_freq_chips_ / static_cast<double>(c_.fs_in);
                                rem_code_phase_chips_ = rem_code_phase_samples_ * (code_freq_chips_ / static_cast<double>(c_.fs_in));
                                rem_carrier_phase_rad_ = fmod(rem_carrier_phase_rad_ + carrier_phase_step_rad_ * static_cast<double>(correlation_length_samples_), kTwoPi);
                                acc_carrier_phase_cycles_ -= carrier_phase_step_rad_ * correlation_length_samples_ / kTwoPi;
                                enable_dll_pll = false;
                            }
                        else
                            {
                                CURRENT_INTEGRATION_TIME_S = static_cast<double>(correlation_length_samples_) / static_cast<double>(c_.fs_in);
                                if (kRetunesDllPdi) code_loop_filter_.set_pdi(static_cast<float>(CURRENT_INTEGRATION_TIME_S));  // (_cc.cc:706)
                                enable_dll_pll = true;
                            }
                    }
                else
                    {
                        CURRENT_INTEGRATION_TIME_S = static_cast<double>(correlation_length_samples_) / static_cast<double>(c_.fs_in);
                        enable_dll_pll = true;
                    }
                const cfloat E(outs_[0].real(), outs_[0].imag()), P(outs_[1].real(), outs_[1].imag()), L(outs_[2].real(), outs_[2].imag());
                if (enable_dll_pll == true)
                    {
                        // PLL: two-quadrant arctan, FLL input 0 (:710-715)
                        carr_phase_error_secs_Ti_ = pll_cloop_two_quadrant_atan(P) / kTwoPi;
                        carrier_doppler_hz_ = carrier_loop_filter_.get_carrier_error(0.0, carr_phase_error_secs_Ti_, CURRENT_INTEGRATION_TIME_S);
                        pll_to_dll_assist_secs_Ti_ = (carrier_doppler_hz_ * CURRENT_INTEGRATION_TIME_S) / kL1FreqHz;
                        code_freq_chips_ = kCodeRateHz + ((carrier_doppler_hz_ * kCodeRateHz) / kL1FreqHz);
                        // DLL (:718-723)
                        code_error_chips_Ti_ = dll_nc_e_minus_l_normalized(E, L);
                        code_error_filt_chips_s_ = code_loop_filter_.get_code_nco(code_error_chips_Ti_);
                        code_error_filt_chips_Ti_ = code_error_filt_chips_s_ * CURRENT_INTEGRATION_TIME_S;
                        code_error_filt_secs_Ti = code_error_filt_chips_Ti_ / code_freq_chips_;
                        // next block length and NCO remainders (:727-752)
                        const double T_chip_seconds = 1.0 / code_freq_chips_;
                        const double T_prn_seconds = T_chip_seconds * kCodeLengthChips;
                        const double T_prn_samples = T_prn_seconds * static_cast<double>(c_.fs_in);
                        const double K_prn_samples = round(T_prn_samples);
                        const double K_T_prn_error_samples = K_prn_samples - T_prn_samples;
                        rem_code_phase_samples_ = rem_code_phase_samples_ - K_T_prn_error_samples + code_error_filt_secs_Ti * static_cast<double>(c_.fs_in);
                        rem_code_phase_integer_samples_ = round(rem_code_phase_samples_);
                        correlation_length_samples_ = K_prn_samples + rem_code_phase_integer_samples_;
                        rem_code_phase_samples_ = rem_code_phase_samples_ - rem_code_phase_integer_samples_;
                        carrier_phase_step_rad_ = kTwoPi * carrier_doppler_hz_ / static_cast<double>(c_.fs_in);
                        acc_carrier_phase_cycles_ -= carrier_phase_step_rad_ * correlation_length_samples_ / kTwoPi;
                        const double CORRECTED_INTEGRATION_TIME_S = (static_cast<double>(correlation_length_samples_) / static_cast<double>(c_.fs_in));
                        rem_carrier_phase_rad_ = fmod(rem_carrier_phase_rad_ + kTwoPi * carrier_doppler_hz_ * CORRECTED_INTEGRATION_TIME_S, kTwoPi);
                        code_phase_step_chips_ = code_freq_chips_ / static_cast<double>(c_.fs_in);
                        rem_code_phase_chips_ = rem_code_phase_samples_ * (code_freq_chips_ / static_cast<double>(c_.fs_in));
                        // CN0 and lock test over FLAGS_cn0_samples prompts (:756-787)
                        if (cn0_estimation_counter_ < c_.cn0_samples)
                            {
                                prompt_buffer_[cn0_estimation_counter_] = P;
                                cn0_estimation_counter_++;
                            }
                        else
                            {
                                cn0_estimation_counter_ = 0;
                                cn0_snv_db_hz_ = cn0_svn_estimator(prompt_buffer_.data(), c_.cn0_samples, kCodePeriodS);
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
                        e->valid_symbol = true;
                        e->correlation_length_ms = preamble_synchronized_ ? c_.extend_correlation_ms : 1;
                    }
                e->prompt_i = static_cast<double>(outs_[1].real());
                e->prompt_q = static_cast<double>(outs_[1].imag());
                e->tracking_sample_counter = sample_counter_ + static_cast<uint64_t>(correlation_length_samples_);
                e->code_phase_samples = rem_code_phase_samples_;
                e->carrier_phase_rads = kTwoPi * acc_carrier_phase_cycles_;
                e->carrier_doppler_hz = carrier_doppler_hz_;
                e->cn0_db_hz = cn0_snv_db_hz_;
                e->abs_e = std::abs<float>(E);
                e->abs_p = std::abs<float>(P);
                e->abs_l = std::abs<float>(L);
                (void)this_block;
            }
        else
            {
                for (auto& o : outs_) o = out_t(0, 0);
                e->tracking_sample_counter = sample_counter_ + static_cast<uint64_t>(correlation_length_samples_);
            }
        // consume_each(d_correlation_length_samples); d_sample_counter += d_correlation_length_samples (:912-913): like the
        // reference, what is consumed is the length just computed for the NEXT block
        e->correlation_length_samples = correlation_length_samples_;
        sample_counter_ += correlation_length_samples_;
        return true;
    }
    const out_t* correlator_outs() const { return outs_; }
    bool tracking_enabled() const { return enable_tracking_; }

private:
    static constexpr double kTwoPi = 6.283185307179586;      // GPS_TWO_PI
    static constexpr double kL1FreqHz = 1.57542e9;           // GPS_L1_FREQ_HZ
    static constexpr double kCodeRateHz = 1.023e6;           // GPS_L1_CA_CODE_RATE_HZ
    static constexpr double kCodeLengthChips = 1023.0;       // GPS_L1_CA_CODE_LENGTH_CHIPS
    static constexpr double kCodePeriodS = 0.001;            // GPS_L1_CA_CODE_PERIOD
    CAidTrkConf c_;
    Correlator* corr_;
    Tracking2ndDllFilter code_loop_filter_;
    FllPllFilter carrier_loop_filter_;
    out_t outs_[3];
    float shifts_[3];
    std::deque<out_t> E_history_, P_history_, L_history_;
    std::vector<cfloat> prompt_buffer_;
    double rem_code_phase_samples_ = 0, rem_code_phase_chips_ = 0, rem_carrier_phase_rad_ = 0;
    int32_t rem_code_phase_integer_samples_ = 0;
    double acq_code_phase_samples_ = 0, acq_carrier_doppler_hz_ = 0;
    double code_freq_chips_ = 0, code_phase_step_chips_ = 0, carrier_doppler_hz_ = 0, carrier_phase_step_rad_ = 0, acc_carrier_phase_cycles_ = 0;
    double carr_phase_error_secs_Ti_ = 0, pll_to_dll_assist_secs_Ti_ = 0;
    double code_error_chips_Ti_ = 0, code_error_filt_chips_s_ = 0, code_error_filt_chips_Ti_ = 0;
    double preamble_timestamp_s_ = 0;
    int32_t correlation_length_samples_ = 0;
    uint64_t sample_counter_ = 0, acq_sample_stamp_ = 0;
    int32_t cn0_estimation_counter_ = 0;
    double carrier_lock_test_ = 1, cn0_snv_db_hz_ = 0, carrier_lock_threshold_ = 0.85;
    int32_t carrier_lock_fail_counter_ = 0;
    bool enable_tracking_ = false, pull_in_ = false, enable_extended_integration_ = false, preamble_synchronized_ = false;
};
}  // namespace gb200

// flat images for the ctypes-facing test libraries (host_abi.cc and tests/native/host_oracle_loop.cc)
// conf = {fs_in, vector_length, pll_bw_hz, dll_bw_hz, pll_bw_narrow_hz, dll_bw_narrow_hz, early_late_space_chips, extend_correlation_ms}
inline void gbh_caid_conf(const double* conf, gb200::CAidTrkConf* c)
{
    c->fs_in = static_cast<int64_t>(conf[0]);
    c->vector_length = static_cast<uint32_t>(conf[1]);
    c->pll_bw_hz = static_cast<float>(conf[2]);
    c->dll_bw_hz = static_cast<float>(conf[3]);
    c->pll_bw_narrow_hz = static_cast<float>(conf[4]);
    c->dll_bw_narrow_hz = static_cast<float>(conf[5]);
    c->early_late_space_chips = static_cast<float>(conf[6]);
    c->extend_correlation_ms = static_cast<int32_t>(conf[7]);
}
// epoch = {tracking_sample_counter, prompt_i, prompt_q, carrier_doppler_hz, code_phase_samples, carrier_phase_rads, cn0_db_hz,
//          abs_e, abs_p, abs_l, valid_symbol + 2 * loss_of_lock, correlation_length_samples, correlation_length_ms}
inline void gbh_caid_fill(const gb200::CAidEpoch& e, double* o)
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
    o[11] = e.correlation_length_samples;
    o[12] = e.correlation_length_ms;
}
