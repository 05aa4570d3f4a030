// GpsL1CaDllPllTrackingB200: implementation name "GPS_L1_CA_DLL_PLL_Tracking_B200".
// Drop-in for GpsL1CaDllPllTracking (src/algorithms/tracking/adapters/gps_l1_ca_dll_pll_tracking.h): the same TrackingInterface
// (src/core/interfaces/tracking_interface.h:55-62), the same configuration keys with the same defaults
// (gps_l1_ca_dll_pll_tracking.cc:55-191); item_type additionally accepts cshort / cbyte (samples stay packed on the device).
#ifndef GNSS_SDR_GPS_L1_CA_DLL_PLL_TRACKING_B200_H
#define GNSS_SDR_GPS_L1_CA_DLL_PLL_TRACKING_B200_H

#include "b200_dll_pll_veml_tracking.h"
#include "tracking_interface.h"
#include <string>

class ConfigurationInterface;

// what the two adapters share: Dll_Pll_Conf from the configuration (the per-signal defaults are the arguments)
struct B200TrackingDefaults
{
    double pll_bw_hz, dll_bw_hz, pll_bw_narrow_hz, dll_bw_narrow_hz;
    double early_late_space_chips, very_early_late_space_chips, early_late_space_narrow_chips, very_early_late_space_narrow_chips;
    int cn0_min;
    double carrier_lock_th;
    int fs_in_default;
};
Dll_Pll_Conf b200_tracking_conf_from(ConfigurationInterface* configuration, const std::string& role, const B200TrackingDefaults& d, std::string* item_type);

class GpsL1CaDllPllTrackingB200 : public TrackingInterface
{
public:
    GpsL1CaDllPllTrackingB200(ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams);
    virtual ~GpsL1CaDllPllTrackingB200() = default;
    inline std::string role() override { return role_; }
    inline std::string implementation() override { return "GPS_L1_CA_DLL_PLL_Tracking_B200"; }
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
