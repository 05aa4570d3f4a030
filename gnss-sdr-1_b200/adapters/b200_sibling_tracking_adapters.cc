// See b200_sibling_tracking_adapters.h.
#include "b200_sibling_tracking_adapters.h"
#include "b200_service_singleton.h"
#include "configuration_interface.h"
#include <glog/logging.h>
#include <cmath>

template <class Policy>
B200SiblingTrackingAdapter<Policy>::B200SiblingTrackingAdapter(ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams,
    unsigned int out_streams, const std::string& implementation, const typename Policy::Conf& conf, int item_type) : role_(role), implementation_(implementation)
{
    item_size_ = B200Service::item_size_of(item_type);
    const int stream_id = configuration->property(role + ".RF_channel_ID", 0);
    tracking_ = b200_sibling_tracking<Policy>::make(B200Service::get(configuration), stream_id, conf, item_type);
    DLOG(INFO) << "tracking(" << tracking_->unique_id() << ")";
    if (in_streams > 1) LOG(ERROR) << "This implementation only supports one input stream";
    if (out_streams > 1) LOG(ERROR) << "This implementation only supports one output stream";
}

std::unique_ptr<TrackingInterface> GetB200SiblingTrkBlock(ConfigurationInterface* configuration, const std::string& role, const std::string& implementation,
    unsigned int in_streams, unsigned int out_streams)
{
    const bool glo2 = implementation == "GLONASS_L2_CA_DLL_PLL_Tracking_B200";
    const bool glo = implementation == "GLONASS_L1_CA_DLL_PLL_Tracking_B200" || glo2, caid = implementation == "GPS_L1_CA_DLL_PLL_C_Aid_Tracking_B200",
               kf = implementation == "GPS_L1_CA_KF_Tracking_B200";
    if (!glo && !caid && !kf) return nullptr;
    std::string default_item_type = "gr_complex";
    const std::string item_type = configuration->property(role + ".item_type", default_item_type);
    const int wire = B200Service::item_type_of(item_type);
    int fs_in_deprecated = configuration->property("GNSS-SDR.internal_fs_hz", 2048000);
    const int fs_in = configuration->property("GNSS-SDR.internal_fs_sps", fs_in_deprecated);
    if (glo)
        {
            // glonass_l1_ca_dll_pll_tracking.cc:61-78
            gb200::GlonassTrkConf c;
            c.fs_in = fs_in;
            c.pll_bw_hz = configuration->property(role + ".pll_bw_hz", 50.0);
            c.dll_bw_hz = configuration->property(role + ".dll_bw_hz", 2.0);
            c.early_late_space_chips = configuration->property(role + ".early_late_space_chips", 0.5);
            c.vector_length = std::round(fs_in / (GLONASS_L1_CA_CODE_RATE_HZ / GLONASS_L1_CA_CODE_LENGTH_CHIPS));
            if (glo2)
                {
                    // glonass_l2_ca_dll_pll_tracking.cc: the same keys and defaults; the block uses the L2 band (GLONASS_L1_L2_CA.h:84-85)
                    c.band_freq_hz = GLONASS_L2_CA_FREQ_HZ;
                    c.band_dfreq_hz = GLONASS_L2_CA_DFREQ_HZ;
                    return std::unique_ptr<TrackingInterface>(new B200SiblingTrackingAdapter<B200GlonassL2Policy>(configuration, role, in_streams, out_streams, implementation, c, wire));
                }
            auto* a = new B200SiblingTrackingAdapter<B200GlonassL1Policy>(configuration, role, in_streams, out_streams, implementation, c, wire);
            return std::unique_ptr<TrackingInterface>(a);
        }
    if (caid)
        {
            // gps_l1_ca_dll_pll_c_aid_tracking.cc:62-85; item_type selects the block (:88-125)
            gb200::CAidTrkConf c;
            c.fs_in = fs_in;
            c.pll_bw_hz = configuration->property(role + ".pll_bw_hz", 50.0);
            c.dll_bw_hz = configuration->property(role + ".dll_bw_hz", 2.0);
            c.pll_bw_narrow_hz = configuration->property(role + ".pll_bw_narrow_hz", 20.0);
            c.dll_bw_narrow_hz = configuration->property(role + ".dll_bw_narrow_hz", 2.0);
            c.extend_correlation_ms = configuration->property(role + ".extend_correlation_ms", 1);
            c.early_late_space_chips = configuration->property(role + ".early_late_space_chips", 0.5);
            c.vector_length = std::round(fs_in / (GPS_L1_CA_CODE_RATE_HZ / GPS_L1_CA_CODE_LENGTH_CHIPS));
            if (item_type == "gr_complex")
                return std::unique_ptr<TrackingInterface>(new B200SiblingTrackingAdapter<B200CAidCcPolicy>(configuration, role, in_streams, out_streams, implementation, c, wire));
            if (item_type == "cshort")
                return std::unique_ptr<TrackingInterface>(new B200SiblingTrackingAdapter<B200CAidScPolicy>(configuration, role, in_streams, out_streams, implementation, c, wire));
            LOG(WARNING) << item_type << " unknown tracking item type.";
            return nullptr;
        }
    // gps_l1_ca_kf_tracking.cc:70-90
    gb200::KfTrkConf c;
    c.order = configuration->property(role + ".order", 2);
    c.fs_in = fs_in;
    c.dll_bw_hz = configuration->property(role + ".dll_bw_hz", 2.0);
    c.early_late_space_chips = configuration->property(role + ".early_late_space_chips", 0.5);
    c.vector_length = std::round(fs_in / (GPS_L1_CA_CODE_RATE_HZ / GPS_L1_CA_CODE_LENGTH_CHIPS));
    c.bce_run = configuration->property(role + ".bce_run", false);
    c.bce_ptrans = configuration->property(role + ".p_transient", 0);
    c.bce_strans = configuration->property(role + ".s_transient", 0);
    c.bce_nu = configuration->property(role + ".bce_nu", 0);
    c.bce_kappa = configuration->property(role + ".bce_kappa", 0);
    return std::unique_ptr<TrackingInterface>(new B200SiblingTrackingAdapter<B200KfPolicy>(configuration, role, in_streams, out_streams, implementation, c, wire));
}
