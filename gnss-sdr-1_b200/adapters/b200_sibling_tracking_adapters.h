// *_B200 adapters of the older per-signal trackers, TrackingInterface implementations made by name:
//   GLONASS_L1_CA_DLL_PLL_Tracking_B200         replaces GLONASS_L1_CA_DLL_PLL_Tracking        (glonass_l1_ca_dll_pll_tracking.cc)
//   GLONASS_L2_CA_DLL_PLL_Tracking_B200         replaces GLONASS_L2_CA_DLL_PLL_Tracking        (glonass_l2_ca_dll_pll_tracking.cc)
//   GPS_L1_CA_DLL_PLL_C_Aid_Tracking_B200       replaces GPS_L1_CA_DLL_PLL_C_Aid_Tracking      (gps_l1_ca_dll_pll_c_aid_tracking.cc; item_type
//                                                gr_complex -> the _cc loop, cshort -> the _sc loop on the int16 correlator, as there)
//   GPS_L1_CA_KF_Tracking_B200                  replaces GPS_L1_CA_KF_Tracking                 (gps_l1_ca_kf_tracking.cc)
// Same keys, same defaults; the gflags values (cn0_samples, cn0_min, max_lock_fail, carrier_lock_th) are gnss_sdr_flags.cc's defaults.
#ifndef GNSS_SDR_B200_SIBLING_TRACKING_ADAPTERS_H
#define GNSS_SDR_B200_SIBLING_TRACKING_ADAPTERS_H

#include "b200_sibling_tracking.h"
#include "tracking_interface.h"
#include <memory>
#include <string>

class ConfigurationInterface;

template <class Policy>
class B200SiblingTrackingAdapter : public TrackingInterface
{
public:
    B200SiblingTrackingAdapter(ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams,
        const std::string& implementation, const typename Policy::Conf& conf, int item_type);
    inline std::string role() override { return role_; }
    inline std::string implementation() override { return implementation_; }
    inline size_t item_size() override { return item_size_; }
    void connect(gr::top_block_sptr top_block) override
    {
        if (top_block) { /* nothing to connect */ }
    }
    void disconnect(gr::top_block_sptr top_block) override
    {
        if (top_block) { /* nothing to disconnect */ }
    }
    gr::basic_block_sptr get_left_block() override { return tracking_; }
    gr::basic_block_sptr get_right_block() override { return tracking_; }
    void set_channel(unsigned int channel) override { tracking_->set_channel(channel); }
    void set_gnss_synchro(Gnss_Synchro* p_gnss_synchro) override { tracking_->set_gnss_synchro(p_gnss_synchro); }
    void start_tracking() override { tracking_->start_tracking(); }
    void stop_tracking() override { tracking_->stop_tracking(); }
    bool ok() const { return tracking_ && tracking_->ok(); }

private:
    typename b200_sibling_tracking<Policy>::sptr tracking_;
    size_t item_size_;
    std::string role_, implementation_;
};

// nullptr: not one of the three names (or an item_type the named tracker does not take)
std::unique_ptr<TrackingInterface> GetB200SiblingTrkBlock(ConfigurationInterface* configuration, const std::string& role, const std::string& implementation,
    unsigned int in_streams, unsigned int out_streams);

#endif
