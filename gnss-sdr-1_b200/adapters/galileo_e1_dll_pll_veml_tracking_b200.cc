// See galileo_e1_dll_pll_veml_tracking_b200.h.
#include "galileo_e1_dll_pll_veml_tracking_b200.h"
#include "Galileo_E1.h"
#include "b200_service_singleton.h"
#include "configuration_interface.h"
#include "gps_l1_ca_dll_pll_tracking_b200.h"  // b200_tracking_conf_from
#include <glog/logging.h>
#include <cmath>
#include <cstring>
#include <iostream>

GalileoE1DllPllVemlTrackingB200::GalileoE1DllPllVemlTrackingB200(ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams,
    unsigned int out_streams) : role_(role), in_streams_(in_streams), out_streams_(out_streams)
{
    DLOG(INFO) << "role " << role;
    std::string item_type;
    Dll_Pll_Conf trk_param = b200_tracking_conf_from(configuration, role, B200TrackingDefaults{5.0, 0.5, 2.0, 0.25, 0.15, 0.6, 0.15, 0.6, 25, 0.85, 2048000}, &item_type);
    if (trk_param.extend_correlation_symbols < 1)
        {
            trk_param.extend_correlation_symbols = 1;
            std::cout << "WARNING: Galileo E1. extend_correlation_symbols must be bigger than 0. Coherent integration has been set to 1 symbol (4 ms)" << std::endl;
        }
    else if (!trk_param.track_pilot and trk_param.extend_correlation_symbols > 1)
        {
            trk_param.extend_correlation_symbols = 1;
            std::cout << "WARNING: Galileo E1. Extended coherent integration is not allowed when tracking the data component. Coherent integration has been set to 4 ms (1 symbol)" << std::endl;
        }
    if ((trk_param.extend_correlation_symbols > 1) and (trk_param.pll_bw_narrow_hz > trk_param.pll_bw_hz or trk_param.dll_bw_narrow_hz > trk_param.dll_bw_hz))
        std::cout << "WARNING: Galileo E1. PLL or DLL narrow tracking bandwidth is higher than wide tracking one" << std::endl;
    trk_param.vector_length = std::round(trk_param.fs_in / (GALILEO_E1_CODE_CHIP_RATE_HZ / GALILEO_E1_B_CODE_LENGTH_CHIPS));
    trk_param.system = 'E';
    char sig_[3] = "1B";
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

void GalileoE1DllPllVemlTrackingB200::connect(gr::top_block_sptr top_block)
{
    if (top_block) { /* nothing to connect */ }
}

void GalileoE1DllPllVemlTrackingB200::disconnect(gr::top_block_sptr top_block)
{
    if (top_block) { /* nothing to disconnect */ }
}
