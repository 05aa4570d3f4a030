// See gps_l1_ca_pcps_acquisition_b200.h.  Key by key the constructor follows gps_l1_ca_pcps_acquisition.cc:46-136 (the gflags
// overrides FLAGS_doppler_max etc. are the receiver's command line and stay with the reference's main()).
#include "gps_l1_ca_pcps_acquisition_b200.h"
#include "GPS_L1_CA.h"
#include "b200_service_singleton.h"
#include "configuration_interface.h"
#include "gps_sdr_signal_processing.h"
#include <glog/logging.h>
#include <cmath>
#include <cstring>

GpsL1CaPcpsAcquisitionB200::GpsL1CaPcpsAcquisitionB200(ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams,
    unsigned int out_streams) : role_(role), in_streams_(in_streams), out_streams_(out_streams)
{
    configuration_ = configuration;
    std::string default_item_type = "gr_complex";
    std::string default_dump_filename = "./acquisition.mat";
    DLOG(INFO) << "role " << role;
    item_type_ = configuration_->property(role + ".item_type", default_item_type);
    int64_t fs_in_deprecated = configuration_->property("GNSS-SDR.internal_fs_hz", 2048000);
    fs_in_ = configuration_->property("GNSS-SDR.internal_fs_sps", fs_in_deprecated);
    acq_parameters_.fs_in = fs_in_;
    acq_parameters_.dump = configuration_->property(role + ".dump", false);
    acq_parameters_.dump_channel = configuration_->property(role + ".dump_channel", 0);
    acq_parameters_.blocking = configuration_->property(role + ".blocking", true);
    doppler_max_ = configuration_->property(role + ".doppler_max", 5000);
    acq_parameters_.doppler_max = doppler_max_;
    sampled_ms_ = configuration_->property(role + ".coherent_integration_time_ms", 1);
    acq_parameters_.sampled_ms = sampled_ms_;
    acq_parameters_.ms_per_code = 1;
    bit_transition_flag_ = configuration_->property(role + ".bit_transition_flag", false);
    acq_parameters_.bit_transition_flag = bit_transition_flag_;
    use_CFAR_algorithm_flag_ = configuration_->property(role + ".use_CFAR_algorithm", true);
    acq_parameters_.use_CFAR_algorithm_flag = use_CFAR_algorithm_flag_;
    max_dwells_ = configuration_->property(role + ".max_dwells", 1);
    acq_parameters_.max_dwells = max_dwells_;
    acq_parameters_.dump_filename = configuration_->property(role + ".dump_filename", default_dump_filename);
    acq_parameters_.num_doppler_bins_step2 = configuration_->property(role + ".second_nbins", 4);
    acq_parameters_.doppler_step2 = configuration_->property(role + ".second_doppler_step", 125.0);
    acq_parameters_.make_2_steps = configuration_->property(role + ".make_two_steps", false);
    acq_parameters_.use_automatic_resampler = configuration_->property("GNSS-SDR.use_acquisition_resampler", false);
    if (acq_parameters_.use_automatic_resampler == true and item_type_ != "gr_complex")
        {
            LOG(WARNING) << "GPS L1 CA acquisition disabled the automatic resampler feature because its item_type is not set to gr_complex";
            acq_parameters_.use_automatic_resampler = false;
        }
    if (acq_parameters_.use_automatic_resampler)
        {
            if (acq_parameters_.fs_in > GPS_L1_CA_OPT_ACQ_FS_HZ)
                {
                    acq_parameters_.resampler_ratio = floor(static_cast<float>(acq_parameters_.fs_in) / GPS_L1_CA_OPT_ACQ_FS_HZ);
                    uint32_t decimation = acq_parameters_.fs_in / GPS_L1_CA_OPT_ACQ_FS_HZ;
                    while (acq_parameters_.fs_in % decimation > 0)
                        {
                            decimation--;
                        };
                    acq_parameters_.resampler_ratio = decimation;
                    acq_parameters_.resampled_fs = acq_parameters_.fs_in / static_cast<int>(acq_parameters_.resampler_ratio);
                }
            // -- Find number of samples per spreading code -------------------------
            code_length_ = static_cast<unsigned int>(std::floor(static_cast<double>(acq_parameters_.resampled_fs) / (GPS_L1_CA_CODE_RATE_HZ / GPS_L1_CA_CODE_LENGTH_CHIPS)));
            acq_parameters_.samples_per_ms = static_cast<float>(acq_parameters_.resampled_fs) * 0.001;
            acq_parameters_.samples_per_chip = static_cast<unsigned int>(ceil(GPS_L1_CA_CHIP_PERIOD * static_cast<float>(acq_parameters_.resampled_fs)));
        }
    else
        {
            acq_parameters_.resampled_fs = fs_in_;
            // -- Find number of samples per spreading code -------------------------
            code_length_ = static_cast<unsigned int>(std::floor(static_cast<double>(fs_in_) / (GPS_L1_CA_CODE_RATE_HZ / GPS_L1_CA_CODE_LENGTH_CHIPS)));
            acq_parameters_.samples_per_ms = static_cast<float>(fs_in_) * 0.001;
            acq_parameters_.samples_per_chip = static_cast<unsigned int>(ceil(GPS_L1_CA_CHIP_PERIOD * static_cast<float>(acq_parameters_.fs_in)));
        }
    acq_parameters_.samples_per_code = acq_parameters_.samples_per_ms * static_cast<float>(GPS_L1_CA_CODE_PERIOD * 1000.0);
    vector_length_ = std::floor(acq_parameters_.sampled_ms * acq_parameters_.samples_per_ms) * (acq_parameters_.bit_transition_flag ? 2 : 1);
    code_ = new std::complex<float>[vector_length_]();
    // item_type: "gr_complex" and "cshort" as in the reference; "cbyte" items stay packed on the device as well (the reference
    // converts them to gr_complex with two more blocks in front of the acquisition, :292-299)
    const int wire = B200Service::item_type_of(item_type_);
    item_size_ = B200Service::item_size_of(wire);
    acq_parameters_.it_size = item_size_;
    acq_parameters_.blocking_on_standby = configuration_->property(role + ".blocking_on_standby", false);
    const int stream_id = configuration_->property(role + ".RF_channel_ID", 0);
    acquisition_ = b200_pcps_acquisition::make(B200Service::get(configuration_), stream_id, acq_parameters_, wire);
    DLOG(INFO) << "acquisition(" << acquisition_->unique_id() << ")";
    channel_ = 0;
    threshold_ = 0.0;
    doppler_step_ = 0;
    gnss_synchro_ = nullptr;
    if (in_streams_ > 1) LOG(ERROR) << "This implementation only supports one input stream";
    if (out_streams_ > 0) LOG(ERROR) << "This implementation does not provide an output stream";
}

GpsL1CaPcpsAcquisitionB200::~GpsL1CaPcpsAcquisitionB200() { delete[] code_; }

void GpsL1CaPcpsAcquisitionB200::stop_acquisition() {}

void GpsL1CaPcpsAcquisitionB200::set_channel(unsigned int channel)
{
    channel_ = channel;
    acquisition_->set_channel(channel_);
}

void GpsL1CaPcpsAcquisitionB200::set_threshold(float threshold)
{
    float pfa = configuration_->property(role_ + ".pfa", 0.0);
    if (pfa == 0.0)
        threshold_ = threshold;
    else
        threshold_ = calculate_threshold(pfa);
    DLOG(INFO) << "Channel " << channel_ << " Threshold = " << threshold_;
    acquisition_->set_threshold(threshold_);
}

void GpsL1CaPcpsAcquisitionB200::set_doppler_max(unsigned int doppler_max)
{
    doppler_max_ = doppler_max;
    acquisition_->set_doppler_max(doppler_max_);
}

void GpsL1CaPcpsAcquisitionB200::set_doppler_step(unsigned int doppler_step)
{
    doppler_step_ = doppler_step;
    acquisition_->set_doppler_step(doppler_step_);
}

void GpsL1CaPcpsAcquisitionB200::set_gnss_synchro(Gnss_Synchro* gnss_synchro)
{
    gnss_synchro_ = gnss_synchro;
    acquisition_->set_gnss_synchro(gnss_synchro_);
}

signed int GpsL1CaPcpsAcquisitionB200::mag() { return acquisition_->mag(); }

void GpsL1CaPcpsAcquisitionB200::init() { acquisition_->init(); }

void GpsL1CaPcpsAcquisitionB200::set_local_code()  // (:227-247)
{
    auto* code = new std::complex<float>[code_length_];
    if (acq_parameters_.use_automatic_resampler)
        gps_l1_ca_code_gen_complex_sampled(code, gnss_synchro_->PRN, acq_parameters_.resampled_fs, 0);
    else
        gps_l1_ca_code_gen_complex_sampled(code, gnss_synchro_->PRN, fs_in_, 0);
    for (unsigned int i = 0; i < sampled_ms_; i++) memcpy(&(code_[i * code_length_]), code, sizeof(std::complex<float>) * code_length_);
    acquisition_->set_local_code(code_);
    delete[] code;
}

void GpsL1CaPcpsAcquisitionB200::reset() { acquisition_->set_active(true); }

void GpsL1CaPcpsAcquisitionB200::set_state(int state) { acquisition_->set_state(state); }

float GpsL1CaPcpsAcquisitionB200::calculate_threshold(float pfa)  // (:262-279); the exponential quantile in closed form
{
    unsigned int frequency_bins = 0;
    for (int doppler = static_cast<int>(-doppler_max_); doppler <= static_cast<int>(doppler_max_); doppler += doppler_step_) frequency_bins++;
    DLOG(INFO) << "Channel " << channel_ << "  Pfa = " << pfa;
    unsigned int ncells = vector_length_ * frequency_bins;
    double exponent = 1 / static_cast<double>(ncells);
    double val = pow(1.0 - pfa, exponent);
    auto lambda = static_cast<double>(vector_length_);
    return static_cast<float>(-std::log1p(-val) / lambda);  // quantile(exponential_distribution(lambda), val)
}

void GpsL1CaPcpsAcquisitionB200::connect(gr::top_block_sptr top_block)
{
    if (top_block) { /* nothing to connect internally: the block takes the conditioned stream in its wire type */ }
}

void GpsL1CaPcpsAcquisitionB200::disconnect(gr::top_block_sptr top_block)
{
    if (top_block) { /* nothing to disconnect */ }
}

gr::basic_block_sptr GpsL1CaPcpsAcquisitionB200::get_left_block() { return acquisition_; }

gr::basic_block_sptr GpsL1CaPcpsAcquisitionB200::get_right_block() { return acquisition_; }

void GpsL1CaPcpsAcquisitionB200::set_resampler_latency(uint32_t latency_samples) { acquisition_->set_resampler_latency(latency_samples); }
