// GalileoE1DllPllVemlTrackingB200: implementation name "Galileo_E1_DLL_PLL_VEML_Tracking_B200".
// Drop-in for GalileoE1DllPllVemlTracking (src/algorithms/tracking/adapters/galileo_e1_dll_pll_veml_tracking.h): the same
// TrackingInterface, the same keys and E1 defaults (galileo_e1_dll_pll_veml_tracking.cc:55-190: PLL 5 Hz / narrow 2 Hz, DLL 0.5 Hz /
// narrow 0.25 Hz, E-L 0.15 chip, VE-VL 0.6 chip, cn0_min 25, lock threshold 0.85, track_pilot, extend_correlation_symbols).
#ifndef GNSS_SDR_GALILEO_E1_DLL_PLL_VEML_TRACKING_B200_H
#define GNSS_SDR_GALILEO_E1_DLL_PLL_VEML_TRACKING_B200_H

#include "b200_dll_pll_veml_tracking.h"
#include "tracking_interface.h"
#include <string>

class ConfigurationInterface;

class GalileoE1DllPllVemlTrackingB200 : public TrackingInterface
{
public:
    GalileoE1DllPllVemlTrackingB200(ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams);
    virtual ~GalileoE1DllPllVemlTrackingB200() = default;
    inline std::string role() override { return role_; }
    inline std::string implementation() override { return "Galileo_E1_DLL_PLL_VEML_Tracking_B200"; }
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

private:
    b200_dll_pll_veml_tracking_sptr tracking_;
    size_t item_size_;
    unsigned int channel_;
    std::string role_;
    unsigned int in_streams_, out_streams_;
};

#endif
