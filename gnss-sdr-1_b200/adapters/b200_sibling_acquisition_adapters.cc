// See b200_sibling_acquisition_adapters.h.
#include "b200_sibling_acquisition_adapters.h"
#include "GPS_L1_CA.h"
#include "Galileo_E1.h"
#include "Galileo_E5a.h"
#include "galileo_e5_signal_processing.h"
#include "galileo_e1_signal_processing.h"
#include "b200_service_singleton.h"
#include "configuration_interface.h"
#include "gps_sdr_signal_processing.h"
#include <glog/logging.h>
#include <cmath>
#include <cstring>

template <class Policy>
B200SiblingAcquisitionAdapter<Policy>::B200SiblingAcquisitionAdapter(ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams,
    unsigned int out_streams, const std::string& implementation, const typename Policy::Conf& conf, unsigned int code_length, unsigned int num_codes, int pfa_mode,
    unsigned int pfa_cells_per_bin, int code_kind, int zero_padding)
    : configuration_(configuration), fs_in_(conf.fs_in), code_length_(code_length), num_codes_(num_codes), doppler_max_(conf.doppler_max), pfa_cells_per_bin_(pfa_cells_per_bin),
      pfa_mode_(pfa_mode), code_kind_(code_kind), zero_padding_(zero_padding), role_(role), implementation_(implementation)
{
    const int stream_id = configuration->property(role + ".RF_channel_ID", 0);
    acquisition_ = b200_sibling_acquisition<Policy>::make(B200Service::get(configuration), stream_id, conf);
    code_.assign(static_cast<size_t>(code_length_) * std::max(1u, num_codes_), std::complex<float>(0.0f, 0.0f));
    DLOG(INFO) << "acquisition(" << acquisition_->unique_id() << ")";
    if (in_streams > 1) LOG(ERROR) << "This implementation only supports one input stream";
    if (out_streams > 0) LOG(ERROR) << "This implementation does not provide an output stream";
}

template <class Policy>
void B200SiblingAcquisitionAdapter<Policy>::set_threshold(float threshold)
{
    float pfa = 0.0;
    if (pfa_mode_ == 1)
        {
            pfa = configuration_->property(role_ + std::to_string(channel_) + ".pfa", 0.0);
            if (pfa == 0.0) pfa = configuration_->property(role_ + ".pfa", 0.0);
        }
    float t = threshold;
    if (pfa != 0.0)
        {
            // calculate_threshold (gps_l1_ca_pcps_tong_acquisition.cc:235-254, gps_l1_ca_pcps_quicksync_acquisition.cc:275-292); the
            // exponential quantile in closed form
            unsigned int frequency_bins = 0;
            for (int doppler = static_cast<int>(-doppler_max_); doppler <= static_cast<int>(doppler_max_); doppler += doppler_step_) frequency_bins++;
            const unsigned int ncells = pfa_cells_per_bin_ * frequency_bins;
            const double val = pow(1.0 - pfa, 1.0 / static_cast<double>(ncells));
            t = static_cast<float>(-std::log1p(-val) / static_cast<double>(pfa_cells_per_bin_));
        }
    DLOG(INFO) << "Channel " << channel_ << "  Threshold = " << t;
    acquisition_->set_threshold(t);
}

template <class Policy>
void B200SiblingAcquisitionAdapter<Policy>::set_local_code()
{
    // one code period at fs_in, repeated (e.g. gps_l1_ca_pcps_tong_acquisition.cc:196-214)
    std::vector<std::complex<float>> code(code_length_);
    const unsigned int copies = std::max(1u, num_codes_);
    if (code_kind_ == 2)
        {
            // galileo_e5a_noncoherent_iq_acquisition_caf.cc:222-280: "5X" -> the I and the Q replica; Zero_padding -> one period + zeros
            std::vector<std::complex<float>> code_q(code_length_);
            const bool both = gnss_synchro_->Signal[0] == '5' && gnss_synchro_->Signal[1] == 'X';
            char a[3];
            if (both)
                {
                    std::strcpy(a, "5I");
                    galileo_e5_a_code_gen_complex_sampled(code.data(), a, gnss_synchro_->PRN, fs_in_, 0);
                    std::strcpy(a, "5Q");
                    galileo_e5_a_code_gen_complex_sampled(code_q.data(), a, gnss_synchro_->PRN, fs_in_, 0);
                }
            else
                galileo_e5_a_code_gen_complex_sampled(code.data(), gnss_synchro_->Signal, gnss_synchro_->PRN, fs_in_, 0);
            std::fill(code_.begin(), code_.end(), std::complex<float>(0.0f, 0.0f));
            code_q_.assign(code_.size(), std::complex<float>(0.0f, 0.0f));
            for (unsigned int i = 0; i < (zero_padding_ == 0 ? copies : 1u); i++)
                {
                    std::memcpy(&code_[static_cast<size_t>(i) * code_length_], code.data(), sizeof(std::complex<float>) * code_length_);
                    if (both) std::memcpy(&code_q_[static_cast<size_t>(i) * code_length_], code_q.data(), sizeof(std::complex<float>) * code_length_);
                }
            acquisition_->set_local_code(code_.data(), code_q_.data());
            return;
        }
    if (code_kind_ == 1)
        {
            // galileo_e1_pcps_8ms_ambiguous_acquisition.cc:216-222 (the key really is "Acquisition" + channel + ".cboc" there)
            const bool cboc = configuration_->property("Acquisition" + std::to_string(channel_) + ".cboc", false);
            galileo_e1_code_gen_complex_sampled(code.data(), gnss_synchro_->Signal, cboc, gnss_synchro_->PRN, fs_in_, 0, false);
        }
    else
        gps_l1_ca_code_gen_complex_sampled(code.data(), gnss_synchro_->PRN, fs_in_, 0);
    for (unsigned int i = 0; i < copies; i++) std::memcpy(&code_[static_cast<size_t>(i) * code_length_], code.data(), sizeof(std::complex<float>) * code_length_);
    acquisition_->set_local_code(code_.data());
}

std::unique_ptr<AcquisitionInterface> GetB200SiblingAcqBlock(ConfigurationInterface* configuration, const std::string& role, const std::string& implementation,
    unsigned int in_streams, unsigned int out_streams)
{
    const bool tong = implementation == "GPS_L1_CA_PCPS_Tong_Acquisition_B200", qs = implementation == "GPS_L1_CA_PCPS_QuickSync_Acquisition_B200",
               fd = implementation == "GPS_L1_CA_PCPS_Acquisition_Fine_Doppler_B200", g8 = implementation == "Galileo_E1_PCPS_8ms_Ambiguous_Acquisition_B200",
               caf = implementation == "Galileo_E5a_Noncoherent_IQ_Acquisition_CAF_B200";
    if (!tong && !qs && !fd && !g8 && !caf) return nullptr;
    std::string default_item_type = "gr_complex";
    const std::string item_type = configuration->property(role + ".item_type", default_item_type);
    if (item_type != "gr_complex")
        {
            LOG(WARNING) << item_type << " unknown acquisition item type";  // as the reference adapters
            return nullptr;
        }
    int64_t fs_in_deprecated = configuration->property("GNSS-SDR.internal_fs_hz", g8 ? 4000000 : (caf ? 32000000 : 2048000));
    const int64_t fs_in = configuration->property("GNSS-SDR.internal_fs_sps", fs_in_deprecated);
    const unsigned int doppler_max = configuration->property(role + ".doppler_max", 5000);
    if (g8)
        {
            // galileo_e1_pcps_8ms_ambiguous_acquisition.cc:54-88: sampled_ms a multiple of 4 (default 4), code_length = one 4 ms code period,
            // the block gets samples_per_ms = code_length / 4 and samples_per_code = code_length
            gb200::AcqVariantConf c;
            c.sampled_ms = configuration->property(role + ".coherent_integration_time_ms", 4);
            if (c.sampled_ms % 4 != 0)
                {
                    c.sampled_ms = static_cast<int>(c.sampled_ms / 4) * 4;
                    LOG(WARNING) << "coherent_integration_time should be multiple of Galileo code length (4 ms). coherent_integration_time = " << c.sampled_ms << " ms will be used.";
                }
            c.max_dwells = configuration->property(role + ".max_dwells", 1);
            c.doppler_max = doppler_max;
            c.fs_in = fs_in;
            const unsigned int e1_code_length = round(fs_in / (GALILEO_E1_CODE_CHIP_RATE_HZ / GALILEO_E1_B_CODE_LENGTH_CHIPS));
            c.samples_per_ms = static_cast<int>(e1_code_length / 4);
            c.samples_per_code = static_cast<int>(e1_code_length);
            const unsigned int vector_length = e1_code_length * (c.sampled_ms / 4);
            return std::unique_ptr<AcquisitionInterface>(new B200SiblingAcquisitionAdapter<B200Galileo8msPolicy>(configuration, role, in_streams, out_streams, implementation, c,
                e1_code_length, c.sampled_ms / 4, 1, vector_length, 1));
        }
    if (caf)
        {
            // galileo_e5a_noncoherent_iq_acquisition_caf.cc:58-105
            gb200::AcqVariantConf c;
            c.caf_window_hz = configuration->property(role + ".CAF_window_hz", 0);
            c.zero_padding = configuration->property(role + ".Zero_padding", 0);
            c.sampled_ms = configuration->property(role + ".coherent_integration_time_ms", 1);
            if (c.sampled_ms > 3)
                {
                    c.sampled_ms = 3;
                    LOG(WARNING) << "coherent_integration_time should be 3 ms or less. coherent_integration_time = " << c.sampled_ms << " ms will be used.";
                }
            if (c.zero_padding > 0)
                {
                    c.sampled_ms = 2;
                    LOG(WARNING) << "Zero padding activated. Changing to 1ms code + 1ms zero padding";
                }
            c.max_dwells = configuration->property(role + ".max_dwells", 1);
            c.bit_transition_flag = configuration->property(role + ".bit_transition_flag", false);
            c.doppler_max = doppler_max;
            c.fs_in = fs_in;
            const unsigned int e5_code_length = round(static_cast<double>(fs_in) / GALILEO_E5A_CODE_CHIP_RATE_HZ * static_cast<double>(GALILEO_E5A_CODE_LENGTH_CHIPS));
            c.samples_per_ms = static_cast<int>(e5_code_length);
            c.samples_per_code = static_cast<int>(e5_code_length);
            const std::string sig_ = configuration->property("Channel.signal", std::string("5X"));
            c.both_signal_components = sig_.size() >= 2 && sig_.at(0) == '5' && sig_.at(1) == 'X';
            return std::unique_ptr<AcquisitionInterface>(new B200SiblingAcquisitionAdapter<B200E5aCafPolicy>(configuration, role, in_streams, out_streams, implementation, c,
                e5_code_length, c.sampled_ms, 1, e5_code_length * c.sampled_ms, 2, c.zero_padding));
        }
    const unsigned int code_length = round(fs_in / (GPS_L1_CA_CODE_RATE_HZ / GPS_L1_CA_CODE_LENGTH_CHIPS));
    if (tong)
        {
            // gps_l1_ca_pcps_tong_acquisition.cc:52-80
            gb200::AcqVariantConf c;
            c.sampled_ms = configuration->property(role + ".coherent_integration_time_ms", 1);
            c.doppler_max = doppler_max;
            c.fs_in = fs_in;
            c.samples_per_ms = static_cast<int>(code_length);
            c.samples_per_code = static_cast<int>(code_length);
            c.tong_init_val = configuration->property(role + ".tong_init_val", 1);
            c.tong_max_val = configuration->property(role + ".tong_max_val", 2);
            c.tong_max_dwells = configuration->property(role + ".tong_max_dwells", static_cast<int>(c.tong_max_val) + 1);
            return std::unique_ptr<AcquisitionInterface>(new B200SiblingAcquisitionAdapter<B200TongPolicy>(configuration, role, in_streams, out_streams, implementation, c,
                code_length, c.sampled_ms, 1, code_length * c.sampled_ms));
        }
    if (fd)
        {
            // gps_l1_ca_pcps_acquisition_fine_doppler.cc:52-80
            gb200::AcqVariantConf c;
            c.fs_in = fs_in;
            c.doppler_max = doppler_max;
            c.sampled_ms = configuration->property(role + ".coherent_integration_time_ms", 1);
            c.max_dwells = configuration->property(role + ".max_dwells", 1);
            c.samples_per_ms = static_cast<int>(code_length);
            c.samples_per_code = static_cast<int>(code_length);
            return std::unique_ptr<AcquisitionInterface>(new B200SiblingAcquisitionAdapter<B200FineDopplerPolicy>(configuration, role, in_streams, out_streams, implementation, c,
                code_length, 1, 0, code_length));
        }
    // gps_l1_ca_pcps_quicksync_acquisition.cc:56-125
    gb200::QuickSyncConf c;
    c.fs_in = fs_in;
    c.doppler_max = doppler_max;
    c.sampled_ms = configuration->property(role + ".coherent_integration_time_ms", 4);
    const auto temp = static_cast<unsigned int>(ceil(sqrt(log2(code_length))));
    c.folding_factor = configuration->property(role + ".folding_factor", temp);
    if (c.sampled_ms % c.folding_factor != 0)
        {
            LOG(WARNING) << "QuickSync Algorithm requires a coherent_integration_time multiple of " << c.folding_factor << "ms, Value entered " << c.sampled_ms << " ms";
            c.sampled_ms = c.sampled_ms < c.folding_factor ? c.folding_factor : (c.sampled_ms / c.folding_factor) * c.folding_factor;
            LOG(WARNING) << " Coherent_integration_time of " << c.sampled_ms << " ms will be used instead.";
        }
    c.bit_transition_flag = configuration->property(role + ".bit_transition_flag", false);
    c.max_dwells = c.bit_transition_flag ? 2 : configuration->property(role + ".max_dwells", 1);
    c.samples_per_ms = static_cast<int32_t>(code_length);
    c.samples_per_code = static_cast<int32_t>(code_length);
    // the adapter's code_ holds sampled_ms / folding_factor copies (:243-247) though it is allocated for one (:107); the block copies one period
    return std::unique_ptr<AcquisitionInterface>(new B200SiblingAcquisitionAdapter<B200QuickSyncPolicy>(configuration, role, in_streams, out_streams, implementation, c,
        code_length, 1, 1, code_length / c.folding_factor));
}
