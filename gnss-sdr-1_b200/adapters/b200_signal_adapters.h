// Table-driven *_B200 adapters: every signal whose reference adapter wraps pcps_acquisition or dll_pll_veml_tracking.
//
// The reference has one adapter file per signal (src/algorithms/acquisition/adapters/*_pcps_acquisition.cc,
// src/algorithms/tracking/adapters/*_dll_pll_tracking.cc), all wrapping the same two GNU Radio blocks (pcps_acquisition,
// dll_pll_veml_tracking) and differing in a handful of constants, defaults and rounding rules.  Here those differences are
// data: one B200PcpsAcquisitionAdapter and one B200DllPllTrackingAdapter class, and a table row per signal that records, with
// the reference line it comes from, every place where that signal's adapter departs from the others.  Same .conf keys, same
// defaults, same derived Acq_Conf / Dll_Pll_Conf; the blocks behind them are b200_pcps_acquisition / b200_dll_pll_veml_tracking.
//
//   implementation name                      replaces                        reference adapter
//   GPS_L1_CA_PCPS_Acquisition_B200          GPS_L1_CA_PCPS_Acquisition      gps_l1_ca_pcps_acquisition.cc
//   Galileo_E1_PCPS_Ambiguous_Acquisition_B200  Galileo_E1_PCPS_Ambiguous_Acquisition  galileo_e1_pcps_ambiguous_acquisition.cc
//   GPS_L5i_PCPS_Acquisition_B200            GPS_L5i_PCPS_Acquisition        gps_l5i_pcps_acquisition.cc
//   Galileo_E5a_Pcps_Acquisition_B200        Galileo_E5a_Pcps_Acquisition    galileo_e5a_pcps_acquisition.cc
//   BEIDOU_B1I_PCPS_Acquisition_B200         BEIDOU_B1I_PCPS_Acquisition     beidou_b1i_pcps_acquisition.cc
//   GLONASS_L1_CA_PCPS_Acquisition_B200      GLONASS_L1_CA_PCPS_Acquisition  glonass_l1_ca_pcps_acquisition.cc
//   GPS_L1_CA_DLL_PLL_Tracking_B200          GPS_L1_CA_DLL_PLL_Tracking      gps_l1_ca_dll_pll_tracking.cc
//   Galileo_E1_DLL_PLL_VEML_Tracking_B200    Galileo_E1_DLL_PLL_VEML_Tracking  galileo_e1_dll_pll_veml_tracking.cc
//   GPS_L5_DLL_PLL_Tracking_B200             GPS_L5_DLL_PLL_Tracking         gps_l5_dll_pll_tracking.cc
//   Galileo_E5a_DLL_PLL_Tracking_B200        Galileo_E5a_DLL_PLL_Tracking    galileo_e5a_dll_pll_tracking.cc
//   BEIDOU_B1I_DLL_PLL_Tracking_B200         BEIDOU_B1I_DLL_PLL_Tracking     beidou_b1i_dll_pll_tracking.cc
//   GPS_L2_M_PCPS_Acquisition_B200, BEIDOU_B3I_PCPS_Acquisition_B200, GLONASS_L2_CA_PCPS_Acquisition_B200, GPS_L2_M_DLL_PLL_Tracking_B200,
//   BEIDOU_B3I_DLL_PLL_Tracking_B200         (gps_l2_m_*, beidou_b3i_*, glonass_l2_ca_pcps_acquisition.cc)
#ifndef GNSS_SDR_B200_SIGNAL_ADAPTERS_H
#define GNSS_SDR_B200_SIGNAL_ADAPTERS_H

#include "acq_conf.h"
#include "acquisition_interface.h"
#include "b200_dll_pll_veml_tracking.h"
#include "b200_pcps_acquisition.h"
#include "channel_fsm.h"
#include "gnss_synchro.h"
#include "tracking_interface.h"
#include <complex>
#include <memory>
#include <string>
#include <vector>

class ConfigurationInterface;

struct B200AcqFlags
{
    bool acquire_pilot = false, acquire_iq = false;  // Galileo E1 (pilot) and E5a (pilot / both)
    bool cboc = false;                               // Galileo E1: "Acquisition" + channel + ".cboc", read when the code is made
};

// what distinguishes one *_pcps_acquisition.cc adapter from another
struct B200AcqTraits
{
    const char* implementation;
    const char* replaces;
    double code_rate_hz, code_length_chips;
    double code_period_ms;            // Acq_Conf::samples_per_code = samples_per_ms * code_period_ms (when samples_per_code_from_period)
    double opt_acq_fs_hz;             // the resampler branch's target rate; 0: this adapter has no resampler branch
    int64_t default_fs;               // default of GNSS-SDR.internal_fs_hz
    bool default_cfar;                // default of .use_CFAR_algorithm
    bool reads_sampled_ms;            // .coherent_integration_time_ms is a key (E5a fixes sampled_ms = 1)
    bool modern;                      // floor() code length + float samples_per_ms + dump_channel / blocking_on_standby keys; else the older
                                      // layout: round() code length, samples_per_ms = code_length, no resampler keys
    bool final_round_code_length;     // E5a recomputes code_length_ from fs_in with round() after the resampler branch (:123)
    int ms_per_code;                  // Acq_Conf::ms_per_code; 0: never assigned (BeiDou B1I / B3I: the struct's default stays).  > 1 (GPS L2C M: 20): it is also
                                      // the default of coherent_integration_time_ms, which must be a multiple of it (gps_l2_m_pcps_acquisition.cc:81-87)
    bool samples_per_code_is_length;  // BeiDou B1I: samples_per_code = code_length_ (:106)
    bool has_blocking_on_standby, has_dump_channel;
    double modern_chip_period_s;      // modern layout: samples_per_chip = ceil(chip period * float(fs)); 0: the adapter writes (1.0 / code rate) instead of a constant
    bool vector_from_requested_ms;    // Galileo E1: vector_length_ and the number of code copies use the coherent_integration_time_ms AS READ, even when
                                      // Acq_Conf::sampled_ms was reset to 4 because it is not a multiple of 4 (galileo_e1_pcps_ambiguous_acquisition.cc:68-75, :123)
    bool leaves_resampled_fs_zero;    // Galileo E1: without the resampler Acq_Conf::resampled_fs is never assigned (the other modern adapters set it to fs_in)
    double old_layout_chip_period_s;  // older layout only: samples_per_chip = ceil(chip period * float(fs)) with the adapter's own chip-period constant
                                      // (glonass_l1_ca_pcps_acquisition.cc:64, GLONASS_L1_CA_CHIP_PERIOD); 0: never assigned (BeiDou B1I)
    bool per_channel_pfa;             // role + channel + ".pfa" first (L5i, E5a)
    bool inclusive_loop_bins;         // calculate_threshold counts the bins with the for loop; else (2 max + step) / step
    const char* default_dump_filename;
    // the adapter's set_local_code: one code period at `fs` into dst (code_length_ items)
    void (*gen)(std::complex<float>* dst, const Gnss_Synchro& syn, int64_t fs, const B200AcqFlags& flags);
};
const B200AcqTraits* b200_acq_traits(const std::string& implementation);

class B200PcpsAcquisitionAdapter : public AcquisitionInterface
{
public:
    B200PcpsAcquisitionAdapter(ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams, const B200AcqTraits& traits);
    ~B200PcpsAcquisitionAdapter() = default;
    inline std::string role() override { return role_; }
    inline std::string implementation() override { return t_.implementation; }
    inline size_t item_size() override { return item_size_; }
    void connect(gr::top_block_sptr top_block) override;
    void disconnect(gr::top_block_sptr top_block) override;
    gr::basic_block_sptr get_left_block() override { return acquisition_; }
    gr::basic_block_sptr get_right_block() override { return acquisition_; }
    void set_gnss_synchro(Gnss_Synchro* p_gnss_synchro) override;
    void set_channel(unsigned int channel) override;
    inline void set_channel_fsm(std::shared_ptr<ChannelFsm> channel_fsm) override
    {
        channel_fsm_ = channel_fsm;
        acquisition_->set_channel_fsm(channel_fsm);
    }
    void set_threshold(float threshold) override;
    void set_doppler_max(unsigned int doppler_max) override;
    void set_doppler_step(unsigned int doppler_step) override;
    void init() override { acquisition_->init(); }
    void set_local_code() override;
    signed int mag() override { return acquisition_->mag(); }
    void reset() override { acquisition_->set_active(true); }
    void set_state(int state) override { acquisition_->set_state(state); }
    void stop_acquisition() override {}
    void set_resampler_latency(uint32_t latency_samples) override { acquisition_->set_resampler_latency(latency_samples); }
    b200_pcps_acquisition_sptr block() const { return acquisition_; }
    const Acq_Conf& conf() const { return acq_parameters_; }

private:
    float calculate_threshold(float pfa);
    const B200AcqTraits& t_;
    ConfigurationInterface* configuration_;
    b200_pcps_acquisition_sptr acquisition_;
    Acq_Conf acq_parameters_;
    B200AcqFlags flags_;
    size_t item_size_ = 0;
    std::string item_type_;
    unsigned int vector_length_ = 0, code_length_ = 0, num_codes_ = 1, requested_ms_ = 1;
    unsigned int channel_ = 0;
    std::shared_ptr<ChannelFsm> channel_fsm_;
    float threshold_ = 0.0f;
    unsigned int doppler_max_ = 0, doppler_step_ = 0;
    int64_t fs_in_ = 0;
    std::vector<std::complex<float>> code_;
    Gnss_Synchro* gnss_synchro_ = nullptr;
    std::string role_;
    unsigned int in_streams_, out_streams_;
};

// what distinguishes one *_dll_pll_tracking.cc adapter from another
struct B200TrkTraits
{
    const char* implementation;
    const char* replaces;
    const char* what;  // name in the adapter's warnings
    char system;
    const char* signal;
    double code_rate_hz, code_length_chips;
    // defaults of pll_bw_hz, dll_bw_hz, pll_bw_narrow_hz, dll_bw_narrow_hz, early_late_space_chips, early_late_space_narrow_chips, cn0_min,
    // carrier_lock_th, GNSS-SDR.internal_fs_hz
    double pll_bw_hz, dll_bw_hz, pll_bw_narrow_hz, dll_bw_narrow_hz, early_late_space_chips, early_late_space_narrow_chips;
    int cn0_min;
    double carrier_lock_th;
    int default_fs;
    bool reads_dump_mat;       // .dump_mat is a key (the BeiDou adapters do not read it)
    bool reads_high_dyn;       // .high_dyn and .smoother_length are keys (GPS L5, Galileo E5a; not GPS L2C, BeiDou)
    bool reads_narrow;         // .pll_bw_narrow_hz, .dll_bw_narrow_hz, .early_late_space_narrow_chips are keys (not GPS L2C)
    double very_early_late_space_chips, very_early_late_space_narrow_chips;  // defaults of the two VEML keys; < 0: the adapter assigns 0 and reads no key
    int extend_mode;           // extend_correlation_symbols: 0 clamp to max_extend_data only when the data component is tracked (L5, E5a); 1 clamp to
                               // [1, max_extend_data] always (BeiDou); 2 forced to 1 with a warning (GPS L2C)
    bool has_pilot;            // false: track_pilot is forced off with a warning and extend_correlation_symbols is clamped to [1, max_extend]
    int max_extend_data;       // clamp of extend_correlation_symbols when the data component is tracked
};
const B200TrkTraits* b200_trk_traits(const std::string& implementation);

class B200DllPllTrackingAdapter : public TrackingInterface
{
public:
    B200DllPllTrackingAdapter(ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams, const B200TrkTraits& traits);
    ~B200DllPllTrackingAdapter() = default;
    inline std::string role() override { return role_; }
    inline std::string implementation() override { return t_.implementation; }
    inline size_t item_size() override { return item_size_; }
    void connect(gr::top_block_sptr top_block) override;
    void disconnect(gr::top_block_sptr top_block) override;
    gr::basic_block_sptr get_left_block() override { return tracking_; }
    gr::basic_block_sptr get_right_block() override { return tracking_; }
    void set_channel(unsigned int channel) override
    {
        channel_ = channel;
        tracking_->set_channel(channel);
    }
    void set_gnss_synchro(Gnss_Synchro* p_gnss_synchro) override { tracking_->set_gnss_synchro(p_gnss_synchro); }
    void start_tracking() override { tracking_->start_tracking(); }
    void stop_tracking() override { tracking_->stop_tracking(); }
    b200_dll_pll_veml_tracking_sptr block() const { return tracking_; }
    const Dll_Pll_Conf& conf() const { return trk_param_; }

private:
    const B200TrkTraits& t_;
    b200_dll_pll_veml_tracking_sptr tracking_;
    Dll_Pll_Conf trk_param_;
    size_t item_size_ = 0;
    unsigned int channel_ = 0;
    std::string role_;
    unsigned int in_streams_, out_streams_;
};

#endif
