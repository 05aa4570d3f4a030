// See gps_l1_ca_dll_pll_tracking_b200.h.  b200_tracking_conf_from reads the keys in the order and with the clamps of
// gps_l1_ca_dll_pll_tracking.cc:55-191 / galileo_e1_dll_pll_veml_tracking.cc:55-190 (gflags overrides stay with main()).
#include "gps_l1_ca_dll_pll_tracking_b200.h"
#include "GPS_L1_CA.h"
#include "b200_service_singleton.h"
#include "configuration_interface.h"
#include <glog/logging.h>
#include <cmath>
#include <cstring>
#include <iostream>

Dll_Pll_Conf b200_tracking_conf_from(ConfigurationInterface* configuration, const std::string& role, const B200TrackingDefaults& d, std::string* item_type)
{
    Dll_Pll_Conf trk_param = Dll_Pll_Conf();
    std::string default_item_type = "gr_complex";
    *item_type = configuration->property(role + ".item_type", default_item_type);
    int fs_in_deprecated = configuration->property("GNSS-SDR.internal_fs_hz", d.fs_in_default);
    int fs_in = configuration->property("GNSS-SDR.internal_fs_sps", fs_in_deprecated);
    trk_param.fs_in = fs_in;
    trk_param.high_dyn = configuration->property(role + ".high_dyn", false);
    if (configuration->property(role + ".smoother_length", 10) < 1)
        {
            trk_param.smoother_length = 1;
            std::cout << "WARNING: smoother_length must be bigger than 0. It has been set to 1" << std::endl;
        }
    else
        trk_param.smoother_length = configuration->property(role + ".smoother_length", 10);
    trk_param.dump = configuration->property(role + ".dump", false);
    std::string default_dump_filename = "./track_ch";
    trk_param.dump_filename = configuration->property(role + ".dump_filename", default_dump_filename);
    trk_param.dump_mat = configuration->property(role + ".dump_mat", true);
    trk_param.pll_bw_hz = configuration->property(role + ".pll_bw_hz", d.pll_bw_hz);
    trk_param.pll_bw_narrow_hz = configuration->property(role + ".pll_bw_narrow_hz", d.pll_bw_narrow_hz);
    trk_param.dll_bw_narrow_hz = configuration->property(role + ".dll_bw_narrow_hz", d.dll_bw_narrow_hz);
    trk_param.dll_bw_hz = configuration->property(role + ".dll_bw_hz", d.dll_bw_hz);
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
    trk_param.early_late_space_chips = configuration->property(role + ".early_late_space_chips", d.early_late_space_chips);
    trk_param.early_late_space_narrow_chips = configuration->property(role + ".early_late_space_narrow_chips", d.early_late_space_narrow_chips);
    trk_param.very_early_late_space_chips = configuration->property(role + ".very_early_late_space_chips", d.very_early_late_space_chips);
    trk_param.very_early_late_space_narrow_chips = configuration->property(role + ".very_early_late_space_narrow_chips", d.very_early_late_space_narrow_chips);
    trk_param.extend_correlation_symbols = configuration->property(role + ".extend_correlation_symbols", 1);
    trk_param.track_pilot = configuration->property(role + ".track_pilot", false);
    trk_param.cn0_samples = configuration->property(role + ".cn0_samples", 20);
    trk_param.cn0_min = configuration->property(role + ".cn0_min", d.cn0_min);
    trk_param.max_lock_fail = configuration->property(role + ".max_lock_fail", 50);
    trk_param.carrier_lock_th = configuration->property(role + ".carrier_lock_th", d.carrier_lock_th);
    return trk_param;
}

GpsL1CaDllPllTrackingB200::GpsL1CaDllPllTrackingB200(ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams,
    unsigned int out_streams) : role_(role), in_streams_(in_streams), out_streams_(out_streams)
{
    DLOG(INFO) << "role " << role;
    std::string item_type;
    // (:76, :81, :83, :85, :135, :137, :174, :186): PLL 50 Hz, narrow 20 Hz, DLL 2 Hz, narrow 2 Hz, E-L 0.5 chip, cn0_min 30, lock threshold 0.80
    Dll_Pll_Conf trk_param = b200_tracking_conf_from(configuration, role, B200TrackingDefaults{50.0, 2.0, 20.0, 2.0, 0.5, 0.0, 0.5, 0.0, 30, 0.80, 2048000}, &item_type);
    trk_param.vector_length = std::round(trk_param.fs_in / (GPS_L1_CA_CODE_RATE_HZ / GPS_L1_CA_CODE_LENGTH_CHIPS));
    if (trk_param.extend_correlation_symbols < 1)
        {
            trk_param.extend_correlation_symbols = 1;
            std::cout << "WARNING: GPS L1 C/A. extend_correlation_symbols must be bigger than 1. Coherent integration has been set to 1 symbol (1 ms)" << std::endl;
        }
    else if (trk_param.extend_correlation_symbols > 20)
        {
            trk_param.extend_correlation_symbols = 20;
            std::cout << "WARNING: GPS L1 C/A. extend_correlation_symbols must be lower than 21. Coherent integration has been set to 20 symbols (20 ms)" << std::endl;
        }
    if (trk_param.track_pilot) std::cout << "WARNING: GPS L1 C/A does not have pilot signal. Data tracking has been enabled" << std::endl;
    trk_param.very_early_late_space_chips = 0.0;
    trk_param.very_early_late_space_narrow_chips = 0.0;
    trk_param.track_pilot = false;
    trk_param.system = 'G';
    char sig_[3] = "1C";
    std::memcpy(trk_param.signal, sig_, 3);
    const int wire = B200Service::item_type_of(item_type);
    item_size_ = B200Service::item_size_of(wire);
    const int stream_id = configuration->property(role + ".RF_channel_ID", 0);
    tracking_ = b200_dll_pll_veml_tracking::make(B200Service::get(configuration), stream_id, trk_param, wire);
    channel_ = 0;
    DLOG(INFO) << "tracking(" << tracking_->unique_id() << ")";
    if (in_streams_ > 1) LOG(ERROR) << "This implementation only supports one input stream";
    if (out_streams_ > 1) LOG(ERROR) << "This implementation only supports one output stream";
}

void GpsL1CaDllPllTrackingB200::connect(gr::top_block_sptr top_block)
{
    if (top_block) { /* nothing to connect */ }
}

void GpsL1CaDllPllTrackingB200::disconnect(gr::top_block_sptr top_block)
{
    if (top_block) { /* nothing to disconnect */ }
}
