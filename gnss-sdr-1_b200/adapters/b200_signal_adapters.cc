// See b200_signal_adapters.h.  The two constructors walk the configuration the way the reference adapters do (same keys, same
// defaults, same clamps); every branch on a B200AcqTraits / B200TrkTraits field names the reference adapter line it reproduces.
// The gflags overrides (FLAGS_doppler_max, FLAGS_pll_bw_hz, ...) are the receiver's command line and stay with the reference's main().
#include "b200_signal_adapters.h"
#include "Beidou_B1I.h"
#include "Beidou_B3I.h"
#include "GPS_L1_CA.h"
#include "GPS_L2C.h"
#include "Galileo_E1.h"
#include "galileo_e1_signal_processing.h"
#include "gps_sdr_signal_processing.h"
#include "GLONASS_L1_L2_CA.h"
#include "GPS_L5.h"
#include "Galileo_E5a.h"
#include "b200_service_singleton.h"
#include "beidou_b1i_signal_processing.h"
#include "beidou_b3i_signal_processing.h"
#include "glonass_l2_signal_processing.h"
#include "gps_l2c_signal.h"
#include "configuration_interface.h"
#include "galileo_e5_signal_processing.h"
#include "glonass_l1_signal_processing.h"
#include "gps_l5_signal.h"
#include <glog/logging.h>
#include <cmath>
#include <cstring>
#include <iostream>

// ------------------------------------------------------------------------------------------------ the signal tables
namespace
{
void gen_l1ca(std::complex<float>* dst, const Gnss_Synchro& syn, int64_t fs, const B200AcqFlags&)
{
    gps_l1_ca_code_gen_complex_sampled(dst, syn.PRN, static_cast<int32_t>(fs), 0);  // gps_l1_ca_pcps_acquisition.cc:227-247
}
void gen_e1(std::complex<float>* dst, const Gnss_Synchro& syn, int64_t fs, const B200AcqFlags& f)
{
    // galileo_e1_pcps_ambiguous_acquisition.cc:240-282: the pilot replica on request, else the channel's own signal ("1B")
    char signal_[3];
    std::memcpy(signal_, f.acquire_pilot ? "1C" : syn.Signal, 3);
    signal_[2] = 0;
    galileo_e1_code_gen_complex_sampled(dst, signal_, f.cboc, syn.PRN, static_cast<int32_t>(fs), 0, false);
}
void gen_l5i(std::complex<float>* dst, const Gnss_Synchro& syn, int64_t fs, const B200AcqFlags&)
{
    gps_l5i_code_gen_complex_sampled(dst, syn.PRN, static_cast<int32_t>(fs));  // gps_l5i_pcps_acquisition.cc:237-252
}
void gen_e5a(std::complex<float>* dst, const Gnss_Synchro& syn, int64_t fs, const B200AcqFlags& f)
{
    char signal_[3];
    std::strcpy(signal_, f.acquire_iq ? "5X" : (f.acquire_pilot ? "5Q" : "5I"));  // galileo_e5a_pcps_acquisition.cc:239-252
    galileo_e5_a_code_gen_complex_sampled(dst, signal_, syn.PRN, static_cast<int32_t>(fs), 0);
}
void gen_b1i(std::complex<float>* dst, const Gnss_Synchro& syn, int64_t fs, const B200AcqFlags&)
{
    beidou_b1i_code_gen_complex_sampled(dst, syn.PRN, static_cast<int32_t>(fs), 0);  // beidou_b1i_pcps_acquisition.cc:209
}
void gen_glo_l1(std::complex<float>* dst, const Gnss_Synchro&, int64_t fs, const B200AcqFlags&)
{
    glonass_l1_ca_code_gen_complex_sampled(dst, static_cast<int32_t>(fs), 0);  // glonass_l1_ca_pcps_acquisition.cc:211 (one code for all satellites)
}

void gen_l2m(std::complex<float>* dst, const Gnss_Synchro& syn, int64_t fs, const B200AcqFlags&)
{
    gps_l2c_m_code_gen_complex_sampled(dst, syn.PRN, static_cast<int32_t>(fs));  // gps_l2_m_pcps_acquisition.cc:240-252
}
void gen_b3i(std::complex<float>* dst, const Gnss_Synchro& syn, int64_t fs, const B200AcqFlags&)
{
    beidou_b3i_code_gen_complex_sampled(dst, syn.PRN, static_cast<int32_t>(fs), 0);  // beidou_b3i_pcps_acquisition.cc:206
}
void gen_glo_l2(std::complex<float>* dst, const Gnss_Synchro&, int64_t fs, const B200AcqFlags&)
{
    glonass_l2_ca_code_gen_complex_sampled(dst, static_cast<int32_t>(fs), 0);  // glonass_l2_ca_pcps_acquisition.cc:210
}

// columns: implementation, replaces, code rate, code length, period ms, resampler target, default fs, default CFAR, reads sampled_ms,
//          modern layout, final round(), ms_per_code, samples_per_code = length, blocking_on_standby key, dump_channel key,
//          chip period of the modern samples_per_chip (0: 1 / rate), vector length from the requested ms, resampled_fs left 0, samples_per_chip in the older layout,
//          per-channel pfa, loop-counted bins, default dump file name, generator
const B200AcqTraits kAcq[] = {
    {"GPS_L1_CA_PCPS_Acquisition_B200", "GPS_L1_CA_PCPS_Acquisition", GPS_L1_CA_CODE_RATE_HZ, GPS_L1_CA_CODE_LENGTH_CHIPS, GPS_L1_CA_CODE_PERIOD * 1000.0, GPS_L1_CA_OPT_ACQ_FS_HZ,
        2048000, true, true, true, false, 1, false, true, true, GPS_L1_CA_CHIP_PERIOD, false, false, 0.0, false, true, "./acquisition.mat", gen_l1ca},
    {"Galileo_E1_PCPS_Ambiguous_Acquisition_B200", "Galileo_E1_PCPS_Ambiguous_Acquisition", GALILEO_E1_CODE_CHIP_RATE_HZ, GALILEO_E1_B_CODE_LENGTH_CHIPS,
        GALILEO_E1_CODE_PERIOD_MS, GALILEO_E1_OPT_ACQ_FS_HZ, 4000000, true, true, true, false, 4, false, true, true, 0.0, true, true, 0.0, true, true, "./acquisition.mat", gen_e1},
    {"GPS_L5i_PCPS_Acquisition_B200", "GPS_L5i_PCPS_Acquisition", GPS_L5I_CODE_RATE_HZ, GPS_L5I_CODE_LENGTH_CHIPS, GPS_L5I_PERIOD * 1000.0, GPS_L5_OPT_ACQ_FS_HZ,
        2048000, true, true, true, false, 1, false, true, true, 0.0, false, false, 0.0, true, true, "./acquisition.mat", gen_l5i},
    {"GPS_L2_M_PCPS_Acquisition_B200", "GPS_L2_M_PCPS_Acquisition", GPS_L2_M_CODE_RATE_HZ, GPS_L2_M_CODE_LENGTH_CHIPS, GPS_L2_M_PERIOD * 1000.0, GPS_L2C_OPT_ACQ_FS_HZ,
        2048000, true, true, true, false, 20, false, true, true, 0.0, false, false, 0.0, true, true, "./acquisition.mat", gen_l2m},
    {"Galileo_E5a_Pcps_Acquisition_B200", "Galileo_E5a_Pcps_Acquisition", GALILEO_E5A_CODE_CHIP_RATE_HZ, GALILEO_E5A_CODE_LENGTH_CHIPS, GALILEO_E5A_CODE_PERIOD_MS,
        GALILEO_E5A_OPT_ACQ_FS_HZ, 32000000, false, false, true, true, 1, false, true, true, 0.0, false, false, 0.0, true, true, "./acquisition.mat", gen_e5a},
    {"BEIDOU_B1I_PCPS_Acquisition_B200", "BEIDOU_B1I_PCPS_Acquisition", BEIDOU_B1I_CODE_RATE_HZ, BEIDOU_B1I_CODE_LENGTH_CHIPS, 1.0, 0.0, 2048000, true, true, false,
        false, 0, true, false, false, 0.0, false, false, 0.0, false, false, "./data/acquisition.dat", gen_b1i},
    {"BEIDOU_B3I_PCPS_Acquisition_B200", "BEIDOU_B3I_PCPS_Acquisition", BEIDOU_B3I_CODE_RATE_HZ, BEIDOU_B3I_CODE_LENGTH_CHIPS, 1.0, 0.0, 2048000, true, true, false,
        false, 0, true, false, false, 0.0, false, false, 0.0, false, false, "./data/acquisition.dat", gen_b3i},
    {"GLONASS_L1_CA_PCPS_Acquisition_B200", "GLONASS_L1_CA_PCPS_Acquisition", GLONASS_L1_CA_CODE_RATE_HZ, GLONASS_L1_CA_CODE_LENGTH_CHIPS,
        GLONASS_L1_CA_CODE_PERIOD * 1000.0, 0.0, 2048000, true, true, false, false, 1, false, true, true, 0.0, false, false, GLONASS_L1_CA_CHIP_PERIOD, false, false, "./data/acquisition.dat", gen_glo_l1},
    {"GLONASS_L2_CA_PCPS_Acquisition_B200", "GLONASS_L2_CA_PCPS_Acquisition", GLONASS_L2_CA_CODE_RATE_HZ, GLONASS_L2_CA_CODE_LENGTH_CHIPS,
        GLONASS_L2_CA_CODE_PERIOD * 1000.0, 0.0, 2048000, true, true, false, false, 1, false, true, true, 0.0, false, false, GLONASS_L2_CA_CHIP_PERIOD, false, false, "./data/acquisition.dat", gen_glo_l2},
};

// columns: implementation, replaces, name in warnings, system, signal, code rate, code length, pll_bw, dll_bw, pll_bw_narrow, dll_bw_narrow,
//          early_late_space, early_late_space_narrow, cn0_min, carrier_lock_th, default fs, reads dump_mat, reads high_dyn + smoother_length, reads the narrow keys, very_early_late_space defaults (wide, narrow; -1: assigned 0), extend mode, has pilot,
//          data-component clamp
const B200TrkTraits kTrk[] = {
    {"GPS_L1_CA_DLL_PLL_Tracking_B200", "GPS_L1_CA_DLL_PLL_Tracking", "GPS L1 C/A", 'G', "1C", GPS_L1_CA_CODE_RATE_HZ, GPS_L1_CA_CODE_LENGTH_CHIPS, 50.0, 2.0, 20.0, 2.0, 0.5,
        0.5, 30, 0.80, 2048000, true, true, true, -1.0, -1.0, 1, false, 20},  // gps_l1_ca_dll_pll_tracking.cc:55-191
    {"Galileo_E1_DLL_PLL_VEML_Tracking_B200", "Galileo_E1_DLL_PLL_VEML_Tracking", "Galileo E1", 'E', "1B", GALILEO_E1_CODE_CHIP_RATE_HZ, GALILEO_E1_B_CODE_LENGTH_CHIPS, 5.0,
        0.5, 2.0, 0.25, 0.15, 0.15, 25, 0.85, 2048000, true, true, true, 0.6, 0.6, 0, true, 1},  // galileo_e1_dll_pll_veml_tracking.cc:55-190
    {"GPS_L5_DLL_PLL_Tracking_B200", "GPS_L5_DLL_PLL_Tracking", "GPS L5", 'G', "L5", GPS_L5I_CODE_RATE_HZ, GPS_L5I_CODE_LENGTH_CHIPS, 50.0, 2.0, 2.0, 0.25, 0.5, 0.15,
        25, 0.75, 2048000, true, true, true, -1.0, -1.0, 0, true, GPS_L5I_NH_CODE_LENGTH},  // gps_l5_dll_pll_tracking.cc:78-178
    {"GPS_L2_M_DLL_PLL_Tracking_B200", "GPS_L2_M_DLL_PLL_Tracking", "GPS L2", 'G', "2S", GPS_L2_M_CODE_RATE_HZ, GPS_L2_M_CODE_LENGTH_CHIPS, 2.0, 0.75, 0.0, 0.0, 0.5, 0.0,
        25, 0.85, 2048000, true, false, false, -1.0, -1.0, 2, false, 1},  // gps_l2_m_dll_pll_tracking.cc:56-150 (narrow values assigned 0.0, not read)
    {"Galileo_E5a_DLL_PLL_Tracking_B200", "Galileo_E5a_DLL_PLL_Tracking", "Galileo E5a", 'E', "5X", GALILEO_E5A_CODE_CHIP_RATE_HZ, GALILEO_E5A_CODE_LENGTH_CHIPS, 20.0,
        20.0, 5.0, 2.0, 0.5, 0.15, 25, 0.85, 12000000, true, true, true, -1.0, -1.0, 0, true, GALILEO_E5A_I_SECONDARY_CODE_LENGTH},  // galileo_e5a_dll_pll_tracking.cc:72-172
    {"BEIDOU_B1I_DLL_PLL_Tracking_B200", "BEIDOU_B1I_DLL_PLL_Tracking", "BEIDOU B1I", 'C', "B1", BEIDOU_B1I_CODE_RATE_HZ, BEIDOU_B1I_CODE_LENGTH_CHIPS, 50.0, 2.0, 20.0,
        2.0, 0.5, 0.5, 25, 0.85, 2048000, false, false, true, -1.0, -1.0, 1, false, 20},  // beidou_b1i_dll_pll_tracking.cc:60-160
    {"BEIDOU_B3I_DLL_PLL_Tracking_B200", "BEIDOU_B3I_DLL_PLL_Tracking", "BEIDOU B3I", 'C', "B3", BEIDOU_B3I_CODE_RATE_HZ, BEIDOU_B3I_CODE_LENGTH_CHIPS, 50.0, 2.0, 20.0,
        2.0, 0.5, 0.5, 25, 0.85, 2048000, false, false, true, -1.0, -1.0, 1, false, 20},  // beidou_b3i_dll_pll_tracking.cc:58-158
};
}  // namespace

const B200AcqTraits* b200_acq_traits(const std::string& implementation)
{
    for (const auto& t : kAcq)
        if (implementation == t.implementation) return &t;
    return nullptr;
}

const B200TrkTraits* b200_trk_traits(const std::string& implementation)
{
    for (const auto& t : kTrk)
        if (implementation == t.implementation) return &t;
    return nullptr;
}

// ------------------------------------------------------------------------------------------------ acquisition
B200PcpsAcquisitionAdapter::B200PcpsAcquisitionAdapter(ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams,
    unsigned int out_streams, const B200AcqTraits& traits) : t_(traits), configuration_(configuration), role_(role), in_streams_(in_streams), out_streams_(out_streams)
{
    acq_parameters_ = Acq_Conf();
    std::string default_item_type = "gr_complex";
    std::string default_dump_filename = t_.default_dump_filename;
    DLOG(INFO) << "role " << role;
    item_type_ = configuration_->property(role + ".item_type", default_item_type);
    int64_t fs_in_deprecated = configuration_->property("GNSS-SDR.internal_fs_hz", t_.default_fs);
    fs_in_ = configuration_->property("GNSS-SDR.internal_fs_sps", fs_in_deprecated);
    acq_parameters_.fs_in = fs_in_;
    flags_.acquire_pilot = configuration_->property(role + ".acquire_pilot", false);  // only the E5a generator looks at these (galileo_e5a_pcps_acquisition.cc:60-65)
    flags_.acquire_iq = configuration_->property(role + ".acquire_iq", false);
    if (flags_.acquire_iq) flags_.acquire_pilot = false;
    acq_parameters_.dump = configuration_->property(role + ".dump", false);
    if (t_.has_dump_channel) acq_parameters_.dump_channel = configuration_->property(role + ".dump_channel", 0);
    acq_parameters_.blocking = configuration_->property(role + ".blocking", true);
    doppler_max_ = configuration_->property(role + ".doppler_max", 5000);
    acq_parameters_.doppler_max = doppler_max_;
    acq_parameters_.sampled_ms = t_.reads_sampled_ms ? configuration_->property(role + ".coherent_integration_time_ms", std::max(1, t_.ms_per_code)) : 1;
    requested_ms_ = acq_parameters_.sampled_ms;
    if (t_.ms_per_code > 1 && (acq_parameters_.sampled_ms % t_.ms_per_code) != 0)
        {
            LOG(WARNING) << "Parameter coherent_integration_time_ms should be a multiple of " << t_.ms_per_code << ". Setting it to " << t_.ms_per_code;
            acq_parameters_.sampled_ms = t_.ms_per_code;
        }
    acq_parameters_.bit_transition_flag = configuration_->property(role + ".bit_transition_flag", false);
    acq_parameters_.use_CFAR_algorithm_flag = configuration_->property(role + ".use_CFAR_algorithm", t_.default_cfar);
    acq_parameters_.max_dwells = configuration_->property(role + ".max_dwells", 1);
    acq_parameters_.dump_filename = configuration_->property(role + ".dump_filename", default_dump_filename);
    if (t_.ms_per_code > 0) acq_parameters_.ms_per_code = t_.ms_per_code;
    num_codes_ = (t_.vector_from_requested_ms ? requested_ms_ : acq_parameters_.sampled_ms) / std::max(1, t_.ms_per_code);
    acq_parameters_.num_doppler_bins_step2 = configuration_->property(role + ".second_nbins", 4);
    acq_parameters_.doppler_step2 = configuration_->property(role + ".second_doppler_step", 125.0);
    acq_parameters_.make_2_steps = configuration_->property(role + ".make_two_steps", false);
    if (t_.has_blocking_on_standby) acq_parameters_.blocking_on_standby = configuration_->property(role + ".blocking_on_standby", false);
    const double code_period_hz = t_.code_rate_hz / t_.code_length_chips;
    if (t_.modern)
        {
            // gps_l5i_pcps_acquisition.cc:100-135 / galileo_e5a_pcps_acquisition.cc:88-124 / gps_l1_ca_pcps_acquisition.cc:88-123
            const double chip_period = t_.modern_chip_period_s > 0.0 ? t_.modern_chip_period_s : (1.0 / t_.code_rate_hz);  // GPS L1: the constant GPS_L1_CA_CHIP_PERIOD
            acq_parameters_.use_automatic_resampler = t_.opt_acq_fs_hz > 0.0 && configuration_->property("GNSS-SDR.use_acquisition_resampler", false);
            if (acq_parameters_.use_automatic_resampler == true and item_type_ != "gr_complex")
                {
                    LOG(WARNING) << t_.replaces << " disabled the automatic resampler feature because its item_type is not set to gr_complex";
                    acq_parameters_.use_automatic_resampler = false;
                }
            if (acq_parameters_.use_automatic_resampler)
                {
                    if (acq_parameters_.fs_in > t_.opt_acq_fs_hz)
                        {
                            acq_parameters_.resampler_ratio = floor(static_cast<float>(acq_parameters_.fs_in) / t_.opt_acq_fs_hz);
                            uint32_t decimation = acq_parameters_.fs_in / t_.opt_acq_fs_hz;
                            while (acq_parameters_.fs_in % decimation > 0) decimation--;
                            acq_parameters_.resampler_ratio = decimation;
                            acq_parameters_.resampled_fs = acq_parameters_.fs_in / static_cast<int>(acq_parameters_.resampler_ratio);
                        }
                    code_length_ = static_cast<unsigned int>(std::floor(static_cast<double>(acq_parameters_.resampled_fs) / code_period_hz));
                    acq_parameters_.samples_per_ms = static_cast<float>(acq_parameters_.resampled_fs) * 0.001;
                    acq_parameters_.samples_per_chip = static_cast<unsigned int>(ceil(chip_period * static_cast<float>(acq_parameters_.resampled_fs)));
                }
            else
                {
                    if (!t_.leaves_resampled_fs_zero) acq_parameters_.resampled_fs = fs_in_;
                    code_length_ = static_cast<unsigned int>(std::floor(static_cast<double>(fs_in_) / code_period_hz));
                    acq_parameters_.samples_per_ms = static_cast<float>(fs_in_) * 0.001;
                    acq_parameters_.samples_per_chip = static_cast<unsigned int>(ceil(chip_period * static_cast<float>(acq_parameters_.fs_in)));
                }
            acq_parameters_.samples_per_code = acq_parameters_.samples_per_ms * static_cast<float>(t_.code_period_ms);
            if (t_.final_round_code_length)
                {
                    // galileo_e5a_pcps_acquisition.cc:123-124: recomputed from fs_in with round(), and no bit-transition factor
                    code_length_ = static_cast<unsigned int>(std::round(static_cast<double>(fs_in_) / t_.code_rate_hz * static_cast<double>(t_.code_length_chips)));
                    vector_length_ = code_length_ * acq_parameters_.sampled_ms;
                }
            else
                vector_length_ = std::floor((t_.vector_from_requested_ms ? requested_ms_ : acq_parameters_.sampled_ms) * acq_parameters_.samples_per_ms) *
                                 (acq_parameters_.bit_transition_flag ? 2 : 1);
        }
    else
        {
            // beidou_b1i_pcps_acquisition.cc:84-106 / glonass_l1_ca_pcps_acquisition.cc:87-110
            code_length_ = static_cast<unsigned int>(std::round(static_cast<double>(fs_in_) / code_period_hz));
            vector_length_ = code_length_ * acq_parameters_.sampled_ms;
            if (acq_parameters_.bit_transition_flag) vector_length_ *= 2;
            if (t_.old_layout_chip_period_s > 0.0)
                acq_parameters_.samples_per_chip = static_cast<unsigned int>(ceil(t_.old_layout_chip_period_s * static_cast<float>(acq_parameters_.fs_in)));
            if (t_.samples_per_code_is_length)
                {
                    acq_parameters_.samples_per_ms = code_length_;
                    acq_parameters_.samples_per_code = code_length_;
                }
            else
                {
                    acq_parameters_.samples_per_ms = static_cast<float>(fs_in_) * 0.001;
                    acq_parameters_.samples_per_code = acq_parameters_.samples_per_ms * static_cast<float>(t_.code_period_ms);
                }
        }
    code_.assign(std::max<size_t>(vector_length_, static_cast<size_t>(code_length_) * num_codes_), std::complex<float>(0.0f, 0.0f));
    // item_type: "gr_complex" and "cshort" as in the reference; "cbyte" items stay packed on the device as well (the reference
    // converts them to gr_complex with two more blocks in front of the acquisition)
    const int wire = B200Service::item_type_of(item_type_);
    item_size_ = B200Service::item_size_of(wire);
    acq_parameters_.it_size = item_size_;
    const int stream_id = configuration_->property(role + ".RF_channel_ID", 0);
    acquisition_ = b200_pcps_acquisition::make(B200Service::get(configuration_), stream_id, acq_parameters_, wire);
    DLOG(INFO) << "acquisition(" << acquisition_->unique_id() << ")";
    if (in_streams_ > 1) LOG(ERROR) << "This implementation only supports one input stream";
    if (out_streams_ > 0) LOG(ERROR) << "This implementation does not provide an output stream";
}

void B200PcpsAcquisitionAdapter::set_channel(unsigned int channel)
{
    channel_ = channel;
    acquisition_->set_channel(channel_);
}

void B200PcpsAcquisitionAdapter::set_threshold(float threshold)
{
    float pfa = t_.per_channel_pfa ? configuration_->property(role_ + std::to_string(channel_) + ".pfa", 0.0) : 0.0;
    if (pfa == 0.0) pfa = configuration_->property(role_ + ".pfa", 0.0);
    threshold_ = pfa == 0.0 ? threshold : calculate_threshold(pfa);
    DLOG(INFO) << "Channel " << channel_ << " Threshold = " << threshold_;
    acquisition_->set_threshold(threshold_);
}

void B200PcpsAcquisitionAdapter::set_doppler_max(unsigned int doppler_max)
{
    doppler_max_ = doppler_max;
    acquisition_->set_doppler_max(doppler_max_);
}

void B200PcpsAcquisitionAdapter::set_doppler_step(unsigned int doppler_step)
{
    doppler_step_ = doppler_step;
    acquisition_->set_doppler_step(doppler_step_);
}

void B200PcpsAcquisitionAdapter::set_gnss_synchro(Gnss_Synchro* gnss_synchro)
{
    gnss_synchro_ = gnss_synchro;
    acquisition_->set_gnss_synchro(gnss_synchro_);
}

void B200PcpsAcquisitionAdapter::set_local_code()
{
    // one code period at the acquisition's sample rate, repeated sampled_ms times (e.g. gps_l5i_pcps_acquisition.cc:237-252)
    std::vector<std::complex<float>> code(code_length_);
    flags_.cboc = configuration_->property("Acquisition" + std::to_string(channel_) + ".cboc", false);  // galileo_e1_pcps_ambiguous_acquisition.cc:242 (sic: no underscore, no signal)
    t_.gen(code.data(), *gnss_synchro_, acq_parameters_.use_automatic_resampler ? acq_parameters_.resampled_fs : fs_in_, flags_);
    for (unsigned int i = 0; i < num_codes_; i++) std::memcpy(&code_[static_cast<size_t>(i) * code_length_], code.data(), sizeof(std::complex<float>) * code_length_);
    acquisition_->set_local_code(code_.data());
}

float B200PcpsAcquisitionAdapter::calculate_threshold(float pfa)
{
    unsigned int frequency_bins = 0;
    if (t_.inclusive_loop_bins)
        for (int doppler = static_cast<int>(-doppler_max_); doppler <= static_cast<int>(doppler_max_); doppler += doppler_step_) frequency_bins++;
    else
        frequency_bins = (2 * doppler_max_ + doppler_step_) / doppler_step_;  // beidou_b1i_pcps_acquisition.cc:245, glonass_l1_ca_pcps_acquisition.cc:247
    DLOG(INFO) << "Channel " << channel_ << "  Pfa = " << pfa;
    unsigned int ncells = vector_length_ * frequency_bins;
    double exponent = 1 / static_cast<double>(ncells);
    double val = pow(1.0 - pfa, exponent);
    auto lambda = static_cast<double>(vector_length_);
    return static_cast<float>(-std::log1p(-val) / lambda);  // quantile(boost::math::exponential_distribution(lambda), val) in closed form
}

void B200PcpsAcquisitionAdapter::connect(gr::top_block_sptr top_block)
{
    if (top_block) { /* nothing to connect internally: the block takes the conditioned stream in its wire type */ }
}

void B200PcpsAcquisitionAdapter::disconnect(gr::top_block_sptr top_block)
{
    if (top_block) { /* nothing to disconnect */ }
}

// ------------------------------------------------------------------------------------------------ tracking
B200DllPllTrackingAdapter::B200DllPllTrackingAdapter(ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams,
    unsigned int out_streams, const B200TrkTraits& traits) : t_(traits), role_(role), in_streams_(in_streams), out_streams_(out_streams)
{
    DLOG(INFO) << "role " << role;
    Dll_Pll_Conf trk_param = Dll_Pll_Conf();
    std::string default_item_type = "gr_complex";
    std::string item_type = configuration->property(role + ".item_type", default_item_type);
    int fs_in_deprecated = configuration->property("GNSS-SDR.internal_fs_hz", t_.default_fs);
    int fs_in = configuration->property("GNSS-SDR.internal_fs_sps", fs_in_deprecated);
    trk_param.fs_in = fs_in;
    trk_param.dump = configuration->property(role + ".dump", false);
    std::string default_dump_filename = "./track_ch";
    trk_param.dump_filename = configuration->property(role + ".dump_filename", default_dump_filename);
    if (t_.reads_dump_mat) trk_param.dump_mat = configuration->property(role + ".dump_mat", true);
    if (t_.reads_high_dyn)
        {
            trk_param.high_dyn = configuration->property(role + ".high_dyn", false);
            if (configuration->property(role + ".smoother_length", 10) < 1)
                {
                    trk_param.smoother_length = 1;
                    std::cout << "WARNING: " << t_.what << ". smoother_length must be bigger than 0. It has been set to 1" << std::endl;
                }
            else
                trk_param.smoother_length = configuration->property(role + ".smoother_length", 10);
        }
    trk_param.pll_bw_hz = configuration->property(role + ".pll_bw_hz", t_.pll_bw_hz);
    trk_param.dll_bw_hz = configuration->property(role + ".dll_bw_hz", t_.dll_bw_hz);
    if (t_.reads_narrow)
        {
            trk_param.pll_bw_narrow_hz = configuration->property(role + ".pll_bw_narrow_hz", t_.pll_bw_narrow_hz);
            trk_param.dll_bw_narrow_hz = configuration->property(role + ".dll_bw_narrow_hz", t_.dll_bw_narrow_hz);
        }
    else
        {
            trk_param.pll_bw_narrow_hz = t_.pll_bw_narrow_hz;  // assigned, not read (gps_l2_m_dll_pll_tracking.cc:139-140: 0.0)
            trk_param.dll_bw_narrow_hz = t_.dll_bw_narrow_hz;
        }
    trk_param.early_late_space_chips = configuration->property(role + ".early_late_space_chips", t_.early_late_space_chips);
    int dll_filter_order = configuration->property(role + ".dll_filter_order", 2);
    if (dll_filter_order < 1)
        {
            LOG(WARNING) << "dll_filter_order parameter must be 1, 2 or 3. Set to 1.";
            dll_filter_order = 1;
        }
    if (dll_filter_order > 3)
        {
            LOG(WARNING) << "dll_filter_order parameter must be 1, 2 or 3. Set to 3.";
            dll_filter_order = 3;
        }
    trk_param.dll_filter_order = dll_filter_order;
    int pll_filter_order = configuration->property(role + ".pll_filter_order", 3);
    if (pll_filter_order < 2)
        {
            LOG(WARNING) << "pll_filter_order parameter must be 2 or 3. Set to 2.";
            pll_filter_order = 2;
        }
    if (pll_filter_order > 3)
        {
            LOG(WARNING) << "pll_filter_order parameter must be 2 or 3. Set to 3.";
            pll_filter_order = 3;
        }
    trk_param.pll_filter_order = pll_filter_order;
    if (pll_filter_order == 2) trk_param.fll_filter_order = 1;
    if (pll_filter_order == 3) trk_param.fll_filter_order = 2;
    trk_param.enable_fll_pull_in = configuration->property(role + ".enable_fll_pull_in", false);
    trk_param.fll_bw_hz = configuration->property(role + ".fll_bw_hz", 35.0);
    trk_param.pull_in_time_s = configuration->property(role + ".pull_in_time_s", 2.0);
    trk_param.vector_length = std::round(static_cast<double>(fs_in) / (t_.code_rate_hz / t_.code_length_chips));
    int extend_correlation_symbols = configuration->property(role + ".extend_correlation_symbols", 1);
    trk_param.early_late_space_narrow_chips =
        t_.reads_narrow ? configuration->property(role + ".early_late_space_narrow_chips", t_.early_late_space_narrow_chips) : t_.early_late_space_narrow_chips;  // (:79: 0.0)
    bool track_pilot = configuration->property(role + ".track_pilot", false);
    if (t_.extend_mode == 2)
        {
            // gps_l2_m_dll_pll_tracking.cc:113-118
            if (extend_correlation_symbols != 1)
                std::cout << "WARNING: Extended coherent integration is not allowed in " << t_.what << ". Coherent integration has been set to one code period (1 symbol)" << std::endl;
            extend_correlation_symbols = 1;
        }
    else if (extend_correlation_symbols < 1)
        {
            extend_correlation_symbols = 1;
            std::cout << "WARNING: " << t_.what << ". extend_correlation_symbols must be bigger than 0. Coherent integration has been set to 1 symbol (1 ms)" << std::endl;
        }
    else if ((t_.extend_mode == 1 or !track_pilot) and extend_correlation_symbols > t_.max_extend_data)
        {
            // gps_l5_dll_pll_tracking.cc:143-147 / galileo_e5a_dll_pll_tracking.cc:137-141 (data component only); beidou_b1i_dll_pll_tracking.cc:114-118 (always)
            extend_correlation_symbols = t_.max_extend_data;
            std::cout << "WARNING: " << t_.what << ". extend_correlation_symbols must be lower than " << t_.max_extend_data + 1
                      << " when tracking the data component. Coherent integration has been set to " << t_.max_extend_data << " symbols" << std::endl;
        }
    if (!t_.has_pilot and track_pilot)
        {
            std::cout << "WARNING: " << t_.what << " does not have pilot signal. Data tracking has been enabled" << std::endl;
            track_pilot = false;
        }
    if ((extend_correlation_symbols > 1) and (trk_param.pll_bw_narrow_hz > trk_param.pll_bw_hz or trk_param.dll_bw_narrow_hz > trk_param.dll_bw_hz))
        std::cout << "WARNING: " << t_.what << ". PLL or DLL narrow tracking bandwidth is higher than wide tracking one" << std::endl;
    trk_param.extend_correlation_symbols = extend_correlation_symbols;
    trk_param.track_pilot = track_pilot;
    // VEML spacings: keys of the Galileo E1 adapter only (:137-141); every other adapter assigns 0
    trk_param.very_early_late_space_chips = t_.very_early_late_space_chips < 0.0 ? 0.0 : configuration->property(role + ".very_early_late_space_chips", t_.very_early_late_space_chips);
    trk_param.very_early_late_space_narrow_chips =
        t_.very_early_late_space_narrow_chips < 0.0 ? 0.0 : configuration->property(role + ".very_early_late_space_narrow_chips", t_.very_early_late_space_narrow_chips);
    trk_param.system = t_.system;
    std::memcpy(trk_param.signal, t_.signal, 3);
    trk_param.cn0_samples = configuration->property(role + ".cn0_samples", 20);
    trk_param.cn0_min = configuration->property(role + ".cn0_min", t_.cn0_min);
    trk_param.max_lock_fail = configuration->property(role + ".max_lock_fail", 50);
    trk_param.carrier_lock_th = configuration->property(role + ".carrier_lock_th", t_.carrier_lock_th);
    trk_param_ = trk_param;
    const int wire = B200Service::item_type_of(item_type);
    item_size_ = B200Service::item_size_of(wire);
    const int stream_id = configuration->property(role + ".RF_channel_ID", 0);
    tracking_ = b200_dll_pll_veml_tracking::make(B200Service::get(configuration), stream_id, trk_param, wire);
    DLOG(INFO) << "tracking(" << tracking_->unique_id() << ")";
    if (in_streams_ > 1) LOG(ERROR) << "This implementation only supports one input stream";
    if (out_streams_ > 1) LOG(ERROR) << "This implementation only supports one output stream";
}

void B200DllPllTrackingAdapter::connect(gr::top_block_sptr top_block)
{
    if (top_block) { /* nothing to connect */ }
}

void B200DllPllTrackingAdapter::disconnect(gr::top_block_sptr top_block)
{
    if (top_block) { /* nothing to disconnect */ }
}
