// GpsL1CaPcpsAcquisitionB200: implementation name "GPS_L1_CA_PCPS_Acquisition_B200".
// Drop-in for GpsL1CaPcpsAcquisition (src/algorithms/acquisition/adapters/gps_l1_ca_pcps_acquisition.h): the same
// AcquisitionInterface (src/core/interfaces/acquisition_interface.h:56-72), the same configuration keys with the same defaults
// (gps_l1_ca_pcps_acquisition.cc:46-136), the same local-code generator; the search runs on the shared B200 correlator service.
#ifndef GNSS_SDR_GPS_L1_CA_PCPS_ACQUISITION_B200_H
#define GNSS_SDR_GPS_L1_CA_PCPS_ACQUISITION_B200_H

#include "acq_conf.h"
#include "acquisition_interface.h"
#include "b200_pcps_acquisition.h"
#include "channel_fsm.h"
#include "gnss_synchro.h"
#include <memory>
#include <string>

class ConfigurationInterface;

class GpsL1CaPcpsAcquisitionB200 : public AcquisitionInterface
{
public:
    GpsL1CaPcpsAcquisitionB200(ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams);
    ~GpsL1CaPcpsAcquisitionB200();

    inline std::string role() override { return role_; }
    inline std::string implementation() override { return "GPS_L1_CA_PCPS_Acquisition_B200"; }
    inline size_t item_size() override { return item_size_; }
    void connect(gr::top_block_sptr top_block) override;
    void disconnect(gr::top_block_sptr top_block) override;
    gr::basic_block_sptr get_left_block() override;
    gr::basic_block_sptr get_right_block() override;

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
    void init() override;
    void set_local_code() override;
    signed int mag() override;
    void reset() override;
    void set_state(int state) override;
    void stop_acquisition() override;
    void set_resampler_latency(uint32_t latency_samples) override;

    b200_pcps_acquisition_sptr block() const { return acquisition_; }

private:
    float calculate_threshold(float pfa);
    ConfigurationInterface* configuration_;
    b200_pcps_acquisition_sptr acquisition_;
    Acq_Conf acq_parameters_;
    size_t item_size_;
    std::string item_type_;
    unsigned int vector_length_, code_length_;
    bool bit_transition_flag_, use_CFAR_algorithm_flag_;
    unsigned int channel_;
    std::shared_ptr<ChannelFsm> channel_fsm_;
    float threshold_;
    unsigned int doppler_max_, doppler_step_, sampled_ms_, max_dwells_;
    int64_t fs_in_;
    std::complex<float>* code_;
    Gnss_Synchro* gnss_synchro_;
    std::string role_;
    unsigned int in_streams_, out_streams_;
};

#endif
