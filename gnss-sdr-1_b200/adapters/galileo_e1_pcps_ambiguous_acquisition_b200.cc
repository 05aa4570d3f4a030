// See galileo_e1_pcps_ambiguous_acquisition_b200.h.  Key by key the constructor follows galileo_e1_pcps_ambiguous_acquisition.cc:44-160.
#include "galileo_e1_pcps_ambiguous_acquisition_b200.h"
#include "Galileo_E1.h"
#include "b200_service_singleton.h"
#include "configuration_interface.h"
#include "galileo_e1_signal_processing.h"
#include <glog/logging.h>
#include <cmath>
#include <cstring>

GalileoE1PcpsAmbiguousAcquisitionB200::GalileoE1PcpsAmbiguousAcquisitionB200(ConfigurationInterface* configuration, const std::string& role,
    unsigned int in_streams, unsigned int out_streams) : role_(role), in_streams_(in_streams), out_streams_(out_streams)
{
    configuration_ = configuration;
    std::string default_item_type = "gr_complex";
    std::string default_dump_filename = "./acquisition.mat";
    DLOG(INFO) << "role " << role;
    item_type_ = configuration_->property(role + ".item_type", default_item_type);
    int64_t fs_in_deprecated = configuration_->property("GNSS-SDR.internal_fs_hz", 4000000);
    fs_in_ = configuration_->property("GNSS-SDR.internal_fs_sps", fs_in_deprecated);
    acq_parameters_.fs_in = fs_in_;
    doppler_max_ = configuration_->property(role + ".doppler_max", 5000);
    acq_parameters_.doppler_max = doppler_max_;
    acq_parameters_.ms_per_code = 4;
    sampled_ms_ = configuration_->property(role + ".coherent_integration_time_ms", acq_parameters_.ms_per_code);
    acq_parameters_.sampled_ms = sampled_ms_;
    if ((acq_parameters_.sampled_ms % acq_parameters_.ms_per_code) != 0)
        {
            LOG(WARNING) << "Parameter coherent_integration_time_ms should be a multiple of 4. Setting it to 4";
            acq_parameters_.sampled_ms = acq_parameters_.ms_per_code;
        }
    bit_transition_flag_ = configuration_->property(role + ".bit_transition_flag", false);
    acq_parameters_.bit_transition_flag = bit_transition_flag_;
    use_CFAR_algorithm_flag_ = configuration_->property(role + ".use_CFAR_algorithm", true);
    acq_parameters_.use_CFAR_algorithm_flag = use_CFAR_algorithm_flag_;
    acquire_pilot_ = configuration_->property(role + ".acquire_pilot", false);
    max_dwells_ = configuration_->property(role + ".max_dwells", 1);
    acq_parameters_.max_dwells = max_dwells_;
    acq_parameters_.dump = configuration_->property(role + ".dump", false);
    acq_parameters_.dump_channel = configuration_->property(role + ".dump_channel", 0);
    acq_parameters_.blocking = configuration_->property(role + ".blocking", true);
    acq_parameters_.dump_filename = configuration_->property(role + ".dump_filename", default_dump_filename);
    acq_parameters_.use_automatic_resampler = configuration_->property("GNSS-SDR.use_acquisition_resampler", false);
    if (acq_parameters_.use_automatic_resampler == true and item_type_ != "gr_complex")
        {
            LOG(WARNING) << "Galileo E1 acqisition disabled the automatic resampler feature because its item_type is not set to gr_complex";
            acq_parameters_.use_automatic_resampler = false;
        }
    if (acq_parameters_.use_automatic_resampler)
        {
            if (acq_parameters_.fs_in > GALILEO_E1_OPT_ACQ_FS_HZ)
                {
                    acq_parameters_.resampler_ratio = floor(static_cast<float>(acq_parameters_.fs_in) / GALILEO_E1_OPT_ACQ_FS_HZ);
                    uint32_t decimation = acq_parameters_.fs_in / GALILEO_E1_OPT_ACQ_FS_HZ;
                    while (acq_parameters_.fs_in % decimation > 0)
                        {
                            decimation--;
                        };
                    acq_parameters_.resampler_ratio = decimation;
                    acq_parameters_.resampled_fs = acq_parameters_.fs_in / static_cast<int>(acq_parameters_.resampler_ratio);
                }
            // -- Find number of samples per spreading code (4 ms)  -----------------
            code_length_ = static_cast<unsigned int>(std::floor(static_cast<double>(acq_parameters_.resampled_fs) / (GALILEO_E1_CODE_CHIP_RATE_HZ / GALILEO_E1_B_CODE_LENGTH_CHIPS)));
            acq_parameters_.samples_per_ms = static_cast<float>(acq_parameters_.resampled_fs) * 0.001;
            acq_parameters_.samples_per_chip = static_cast<unsigned int>(ceil((1.0 / GALILEO_E1_CODE_CHIP_RATE_HZ) * static_cast<float>(acq_parameters_.resampled_fs)));
        }
    else
        {
            // -- Find number of samples per spreading code (4 ms)  -----------------
            code_length_ = static_cast<unsigned int>(std::floor(static_cast<double>(fs_in_) / (GALILEO_E1_CODE_CHIP_RATE_HZ / GALILEO_E1_B_CODE_LENGTH_CHIPS)));
            acq_parameters_.samples_per_ms = static_cast<float>(fs_in_) * 0.001;
            acq_parameters_.samples_per_chip = static_cast<unsigned int>(ceil((1.0 / GALILEO_E1_CODE_CHIP_RATE_HZ) * static_cast<float>(acq_parameters_.fs_in)));
        }
    acq_parameters_.samples_per_code = acq_parameters_.samples_per_ms * static_cast<float>(GALILEO_E1_CODE_PERIOD_MS);
    vector_length_ = sampled_ms_ * acq_parameters_.samples_per_ms;
    if (bit_transition_flag_) vector_length_ *= 2;
    code_ = new std::complex<float>[vector_length_]();
    const int wire = B200Service::item_type_of(item_type_);
    item_size_ = B200Service::item_size_of(wire);
    acq_parameters_.it_size = item_size_;
    acq_parameters_.num_doppler_bins_step2 = configuration_->property(role + ".second_nbins", 4);
    acq_parameters_.doppler_step2 = configuration_->property(role + ".second_doppler_step", 125.0);
    acq_parameters_.make_2_steps = configuration_->property(role + ".make_two_steps", false);
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

GalileoE1PcpsAmbiguousAcquisitionB200::~GalileoE1PcpsAmbiguousAcquisitionB200() { delete[] code_; }

void GalileoE1PcpsAmbiguousAcquisitionB200::stop_acquisition() {}

void GalileoE1PcpsAmbiguousAcquisitionB200::set_channel(unsigned int channel)
{
    channel_ = channel;
    acquisition_->set_channel(channel_);
}

void GalileoE1PcpsAmbiguousAcquisitionB200::set_threshold(float threshold)  // (:177-199): per-channel pfa key first
{
    float pfa = configuration_->property(role_ + std::to_string(channel_) + ".pfa", 0.0);
    if (pfa == 0.0) pfa = configuration_->property(role_ + ".pfa", 0.0);
    if (pfa == 0.0)
        threshold_ = threshold;
    else
        threshold_ = calculate_threshold(pfa);
    DLOG(INFO) << "Channel " << channel_ << " Threshold = " << threshold_;
    acquisition_->set_threshold(threshold_);
}

void GalileoE1PcpsAmbiguousAcquisitionB200::set_doppler_max(unsigned int doppler_max)
{
    doppler_max_ = doppler_max;
    acquisition_->set_doppler_max(doppler_max_);
}

void GalileoE1PcpsAmbiguousAcquisitionB200::set_doppler_step(unsigned int doppler_step)
{
    doppler_step_ = doppler_step;
    acquisition_->set_doppler_step(doppler_step_);
}

void GalileoE1PcpsAmbiguousAcquisitionB200::set_gnss_synchro(Gnss_Synchro* gnss_synchro)
{
    gnss_synchro_ = gnss_synchro;
    acquisition_->set_gnss_synchro(gnss_synchro_);
}

signed int GalileoE1PcpsAmbiguousAcquisitionB200::mag() { return acquisition_->mag(); }

void GalileoE1PcpsAmbiguousAcquisitionB200::init() { acquisition_->init(); }

void GalileoE1PcpsAmbiguousAcquisitionB200::set_local_code()  // (:240-282)
{
    bool cboc = configuration_->property("Acquisition" + std::to_string(channel_) + ".cboc", false);
    auto* code = new std::complex<float>[code_length_];
    const int32_t fs = static_cast<int32_t>(acq_parameters_.use_automatic_resampler ? acq_parameters_.resampled_fs : fs_in_);
    if (acquire_pilot_ == true)
        {
            char pilot_signal[3] = "1C";
            galileo_e1_code_gen_complex_sampled(code, pilot_signal, cboc, gnss_synchro_->PRN, fs, 0, false);
        }
    else
        galileo_e1_code_gen_complex_sampled(code, gnss_synchro_->Signal, cboc, gnss_synchro_->PRN, fs, 0, false);
    for (unsigned int i = 0; i < sampled_ms_ / 4; i++) memcpy(&(code_[i * code_length_]), code, sizeof(std::complex<float>) * code_length_);
    acquisition_->set_local_code(code_);
    delete[] code;
}

void GalileoE1PcpsAmbiguousAcquisitionB200::reset() { acquisition_->set_active(true); }

void GalileoE1PcpsAmbiguousAcquisitionB200::set_state(int state) { acquisition_->set_state(state); }

float GalileoE1PcpsAmbiguousAcquisitionB200::calculate_threshold(float pfa)  // (:299-317)
{
    unsigned int frequency_bins = 0;
    for (int doppler = static_cast<int>(-doppler_max_); doppler <= static_cast<int>(doppler_max_); doppler += doppler_step_) frequency_bins++;
    DLOG(INFO) << "Channel " << channel_ << "  Pfa = " << pfa;
    unsigned int ncells = vector_length_ * frequency_bins;
    double exponent = 1 / static_cast<double>(ncells);
    double val = pow(1.0 - pfa, exponent);
    auto lambda = static_cast<double>(vector_length_);
    return static_cast<float>(-std::log1p(-val) / lambda);
}

void GalileoE1PcpsAmbiguousAcquisitionB200::connect(gr::top_block_sptr top_block)
{
    if (top_block) { /* nothing to connect internally */ }
}

void GalileoE1PcpsAmbiguousAcquisitionB200::disconnect(gr::top_block_sptr top_block)
{
    if (top_block) { /* nothing to disconnect */ }
}

gr::basic_block_sptr GalileoE1PcpsAmbiguousAcquisitionB200::get_left_block() { return acquisition_; }

gr::basic_block_sptr GalileoE1PcpsAmbiguousAcquisitionB200::get_right_block() { return acquisition_; }

void GalileoE1PcpsAmbiguousAcquisitionB200::set_resampler_latency(uint32_t latency_samples) { acquisition_->set_resampler_latency(latency_samples); }
