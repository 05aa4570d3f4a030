// *_B200 adapters of the sibling acquisition blocks, AcquisitionInterface implementations made by name:
//   GPS_L1_CA_PCPS_Tong_Acquisition_B200           replaces GPS_L1_CA_PCPS_Tong_Acquisition           (gps_l1_ca_pcps_tong_acquisition.cc)
//   GPS_L1_CA_PCPS_QuickSync_Acquisition_B200      replaces GPS_L1_CA_PCPS_QuickSync_Acquisition      (gps_l1_ca_pcps_quicksync_acquisition.cc)
//   GPS_L1_CA_PCPS_Acquisition_Fine_Doppler_B200   replaces GPS_L1_CA_PCPS_Acquisition_Fine_Doppler   (gps_l1_ca_pcps_acquisition_fine_doppler.cc)
//   Galileo_E1_PCPS_8ms_Ambiguous_Acquisition_B200 replaces Galileo_E1_PCPS_8ms_Ambiguous_Acquisition (galileo_e1_pcps_8ms_ambiguous_acquisition.cc)
//   Galileo_E5a_Noncoherent_IQ_Acquisition_CAF_B200 replaces Galileo_E5a_Noncoherent_IQ_Acquisition_CAF (galileo_e5a_noncoherent_iq_acquisition_caf.cc)
// Same keys and defaults; gr_complex streams only, as in the reference adapters.
#ifndef GNSS_SDR_B200_SIBLING_ACQUISITION_ADAPTERS_H
#define GNSS_SDR_B200_SIBLING_ACQUISITION_ADAPTERS_H

#include "acquisition_interface.h"
#include "b200_sibling_acquisition.h"
#include <complex>
#include <memory>
#include <string>
#include <vector>

class ConfigurationInterface;

template <class Policy>
class B200SiblingAcquisitionAdapter : public AcquisitionInterface
{
public:
    // pfa_mode 0: no pfa key (fine Doppler); 1: role + channel + ".pfa" then role + ".pfa", cells = vector_length * inclusive bins (Tong);
    // 2: role + ".pfa", cells = (code_length / folding_factor) * inclusive bins (QuickSync)
    B200SiblingAcquisitionAdapter(ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams,
        const std::string& implementation, const typename Policy::Conf& conf, unsigned int code_length, unsigned int num_codes, int pfa_mode, unsigned int pfa_cells_per_bin,
        int code_kind = 0, int zero_padding = 0);
    inline std::string role() override { return role_; }
    inline std::string implementation() override { return implementation_; }
    inline size_t item_size() override { return sizeof(std::complex<float>); }
    void connect(gr::top_block_sptr top_block) override
    {
        if (top_block) { /* the block takes the plain stream: no stream_to_vector */ }
    }
    void disconnect(gr::top_block_sptr top_block) override
    {
        if (top_block) { /* nothing to disconnect */ }
    }
    gr::basic_block_sptr get_left_block() override { return acquisition_; }
    gr::basic_block_sptr get_right_block() override { return acquisition_; }
    void set_gnss_synchro(Gnss_Synchro* p_gnss_synchro) override
    {
        gnss_synchro_ = p_gnss_synchro;
        acquisition_->set_gnss_synchro(p_gnss_synchro);
    }
    void set_channel(unsigned int channel) override
    {
        channel_ = channel;
        acquisition_->set_channel(channel);
    }
    void set_channel_fsm(std::shared_ptr<ChannelFsm> channel_fsm) override { acquisition_->set_channel_fsm(channel_fsm); }
    void set_threshold(float threshold) override;
    void set_doppler_max(unsigned int doppler_max) override
    {
        doppler_max_ = doppler_max;
        acquisition_->set_doppler_max(doppler_max);
    }
    void set_doppler_step(unsigned int doppler_step) override
    {
        doppler_step_ = doppler_step;
        acquisition_->set_doppler_step(doppler_step);
    }
    void init() override { acquisition_->init(); }
    void set_local_code() override;
    signed int mag() override { return acquisition_->mag(); }
    void reset() override { acquisition_->set_active(true); }
    void set_state(int state) override { acquisition_->set_state(state); }
    void stop_acquisition() override {}
    void set_resampler_latency(uint32_t latency_samples __attribute__((unused))) override {}
    bool ok() const { return acquisition_ && acquisition_->ok(); }

private:
    ConfigurationInterface* configuration_;
    typename b200_sibling_acquisition<Policy>::sptr acquisition_;
    Gnss_Synchro* gnss_synchro_ = nullptr;
    int64_t fs_in_;
    unsigned int code_length_, num_codes_, channel_ = 0, doppler_max_, doppler_step_ = 0, pfa_cells_per_bin_;
    int pfa_mode_;
    int code_kind_ = 0;  // 0: GPS L1 C/A generator, 1: galileo_e1_code_gen_complex_sampled, 2: Galileo E5a I (+ Q) replicas
    int zero_padding_ = 0;
    std::vector<std::complex<float>> code_q_;
    std::vector<std::complex<float>> code_;
    std::string role_, implementation_;
};

// nullptr: not one of the three names (or an item_type they do not take)
std::unique_ptr<AcquisitionInterface> GetB200SiblingAcqBlock(ConfigurationInterface* configuration, const std::string& role, const std::string& implementation,
    unsigned int in_streams, unsigned int out_streams);

#endif
