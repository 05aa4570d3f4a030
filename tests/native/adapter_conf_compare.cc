// TEST INFRASTRUCTURE: the reference's OWN adapter classes next to the *_B200 adapters, on the same configuration.
//
// "Same keys, same defaults" is a claim about src/algorithms/{acquisition,tracking}/adapters/*.cc, so it is checked against those
// files themselves: they are compiled unmodified from /root/reference (with the stand-in GNU Radio headers of oracle/shim; the two
// GNU Radio blocks they construct, pcps_acquisition and dll_pll_veml_tracking, come from oracle/_ref/libgnss_ref.so), constructed on
// the same FileConfiguration as the B200 adapter, and the Acq_Conf / Dll_Pll_Conf each pair ends up with -- read from the BLOCKS the
// adapters made, after the blocks' own constructors touched them -- is flattened into two arrays the Python test compares.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <complex>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <utility>
#include <vector>
#include <gnuradio/block.h>
#include <gnuradio/fft/fft.h>
#include <pmt/pmt.h>

#define private public
#define protected public
#include "b200_block_factory.h"
#include "b200_service_singleton.h"
#include "b200_signal_adapters.h"
#include "beidou_b1i_dll_pll_tracking.h"
#include "beidou_b1i_pcps_acquisition.h"
#include "beidou_b3i_dll_pll_tracking.h"
#include "beidou_b3i_pcps_acquisition.h"
#include "file_configuration.h"
#include "galileo_e1_dll_pll_veml_tracking.h"
#include "galileo_e1_pcps_ambiguous_acquisition.h"
#include "galileo_e5a_dll_pll_tracking.h"
#include "galileo_e5a_pcps_acquisition.h"
#include "glonass_l1_ca_pcps_acquisition.h"
#include "glonass_l2_ca_pcps_acquisition.h"
#include "gps_l1_ca_dll_pll_tracking.h"
#include "gps_l1_ca_pcps_acquisition.h"
#include "gps_l2_m_dll_pll_tracking.h"
#include "gps_l2_m_pcps_acquisition.h"
#include "gps_l5_dll_pll_tracking.h"
#include "gps_l5i_pcps_acquisition.h"
#undef private
#undef protected

namespace
{
int flatten(const Acq_Conf& c, double threshold, size_t item_size, double* o)
{
    int k = 0;
    o[k++] = c.sampled_ms;
    o[k++] = c.ms_per_code;
    o[k++] = c.samples_per_chip;
    o[k++] = c.max_dwells;
    o[k++] = c.doppler_max;
    o[k++] = c.num_doppler_bins_step2;
    o[k++] = c.doppler_step2;
    o[k++] = static_cast<double>(c.fs_in);
    o[k++] = c.samples_per_ms;
    o[k++] = c.samples_per_code;
    o[k++] = c.bit_transition_flag;
    o[k++] = c.use_CFAR_algorithm_flag;
    o[k++] = c.dump;
    o[k++] = c.blocking;
    o[k++] = c.blocking_on_standby;
    o[k++] = c.make_2_steps;
    o[k++] = c.dump_channel;
    o[k++] = c.use_automatic_resampler;
    o[k++] = c.resampler_ratio;
    o[k++] = static_cast<double>(c.resampled_fs);
    o[k++] = threshold;
    o[k++] = static_cast<double>(item_size);
    return k;
}
int flatten(const Dll_Pll_Conf& c, size_t item_size, double* o)
{
    int k = 0;
    o[k++] = c.fs_in;
    o[k++] = c.vector_length;
    o[k++] = c.dump;
    o[k++] = c.dump_mat;
    o[k++] = c.pll_bw_hz;
    o[k++] = c.dll_bw_hz;
    o[k++] = c.fll_bw_hz;
    o[k++] = c.pll_bw_narrow_hz;
    o[k++] = c.dll_bw_narrow_hz;
    o[k++] = c.early_late_space_chips;
    o[k++] = c.very_early_late_space_chips;
    o[k++] = c.early_late_space_narrow_chips;
    o[k++] = c.very_early_late_space_narrow_chips;
    o[k++] = c.extend_correlation_symbols;
    o[k++] = c.high_dyn;
    o[k++] = c.cn0_samples;
    o[k++] = c.cn0_min;
    o[k++] = c.max_lock_fail;
    o[k++] = c.carrier_lock_th;
    o[k++] = c.track_pilot;
    o[k++] = c.system;
    o[k++] = c.signal[0];
    o[k++] = c.signal[1];
    o[k++] = c.pll_filter_order;
    o[k++] = c.dll_filter_order;
    o[k++] = c.fll_filter_order;
    o[k++] = c.enable_fll_pull_in;
    o[k++] = c.pull_in_time_s;
    o[k++] = c.smoother_length;
    o[k++] = static_cast<double>(item_size);
    return k;
}

template <class RefAdapter>
int ref_acq(ConfigurationInterface* cfg, const std::string& role, Gnss_Synchro* syn, unsigned doppler_step, float threshold, double* o)
{
    RefAdapter a(cfg, role, 1, 0);
    a.set_channel(0);
    a.set_gnss_synchro(syn);
    a.set_doppler_step(doppler_step);
    a.set_threshold(threshold);
    a.init();  // as Channel does; the block's destructor frees the step-two tables init() allocates without checking that it ran
    return flatten(a.acquisition_->acq_parameters, a.threshold_, a.item_size(), o);
}
template <class RefAdapter>
int ref_trk(ConfigurationInterface* cfg, const std::string& role, double* o)
{
    RefAdapter t(cfg, role, 1, 1);
    if (!t.tracking_) return -5;  // an item_type the reference adapter does not know: it makes no block
    return flatten(t.tracking_->trk_parameters, t.item_size(), o);
}
}  // namespace

extern "C" {

// kind 0: acquisition, 1: tracking.  Returns the number of values written to each array, or < 0 (unknown name / no device).
int b200_adapter_conf_compare(const char* conf_path, int kind, const char* role_c, const char* ref_impl_c, const char* b200_impl_c, unsigned doppler_step,
    float threshold, double* out_ref, double* out_b200)
{
    std::shared_ptr<ConfigurationInterface> config = std::make_shared<FileConfiguration>(conf_path);
    const std::string role(role_c), ref_impl(ref_impl_c), b200_impl(b200_impl_c);
    Gnss_Synchro syn = Gnss_Synchro();
    syn.PRN = 1;
    int n_ref = -1, n_b = -1;
    try
        {
            if (kind == 0)
                {
                    if (ref_impl == "GPS_L1_CA_PCPS_Acquisition") n_ref = ref_acq<GpsL1CaPcpsAcquisition>(config.get(), role, &syn, doppler_step, threshold, out_ref);
                    if (ref_impl == "Galileo_E1_PCPS_Ambiguous_Acquisition") n_ref = ref_acq<GalileoE1PcpsAmbiguousAcquisition>(config.get(), role, &syn, doppler_step, threshold, out_ref);
                    if (ref_impl == "GPS_L5i_PCPS_Acquisition") n_ref = ref_acq<GpsL5iPcpsAcquisition>(config.get(), role, &syn, doppler_step, threshold, out_ref);
                    if (ref_impl == "Galileo_E5a_Pcps_Acquisition") n_ref = ref_acq<GalileoE5aPcpsAcquisition>(config.get(), role, &syn, doppler_step, threshold, out_ref);
                    if (ref_impl == "BEIDOU_B1I_PCPS_Acquisition") n_ref = ref_acq<BeidouB1iPcpsAcquisition>(config.get(), role, &syn, doppler_step, threshold, out_ref);
                    if (ref_impl == "GLONASS_L1_CA_PCPS_Acquisition") n_ref = ref_acq<GlonassL1CaPcpsAcquisition>(config.get(), role, &syn, doppler_step, threshold, out_ref);
                    if (ref_impl == "GLONASS_L2_CA_PCPS_Acquisition") n_ref = ref_acq<GlonassL2CaPcpsAcquisition>(config.get(), role, &syn, doppler_step, threshold, out_ref);
                    if (ref_impl == "BEIDOU_B3I_PCPS_Acquisition") n_ref = ref_acq<BeidouB3iPcpsAcquisition>(config.get(), role, &syn, doppler_step, threshold, out_ref);
                    if (ref_impl == "GPS_L2_M_PCPS_Acquisition") n_ref = ref_acq<GpsL2MPcpsAcquisition>(config.get(), role, &syn, doppler_step, threshold, out_ref);
                    if (getenv("GB_CC_TRACE")) fprintf(stderr, "cc: reference adapter done (%d values)\n", n_ref);
                    std::unique_ptr<AcquisitionInterface> b = GetB200AcqBlock(config, role, b200_impl, 1, 0);
                    if (getenv("GB_CC_TRACE")) fprintf(stderr, "cc: b200 adapter made %p\n", (void*)b.get());
                    if (!b) return -2;
                    b->set_channel(0);
                    b->set_gnss_synchro(&syn);
                    b->set_doppler_step(doppler_step);
                    b->set_threshold(threshold);
                    b->init();
                    if (getenv("GB_CC_TRACE")) fprintf(stderr, "cc: b200 adapter set up\n");
                    if (auto* g = dynamic_cast<B200PcpsAcquisitionAdapter*>(b.get())) n_b = flatten(g->acquisition_->acq_parameters, g->threshold_, g->item_size(), out_b200);
                }
            else
                {
                    if (ref_impl == "GPS_L1_CA_DLL_PLL_Tracking") n_ref = ref_trk<GpsL1CaDllPllTracking>(config.get(), role, out_ref);
                    if (ref_impl == "Galileo_E1_DLL_PLL_VEML_Tracking") n_ref = ref_trk<GalileoE1DllPllVemlTracking>(config.get(), role, out_ref);
                    if (ref_impl == "GPS_L5_DLL_PLL_Tracking") n_ref = ref_trk<GpsL5DllPllTracking>(config.get(), role, out_ref);
                    if (ref_impl == "Galileo_E5a_DLL_PLL_Tracking") n_ref = ref_trk<GalileoE5aDllPllTracking>(config.get(), role, out_ref);
                    if (ref_impl == "BEIDOU_B1I_DLL_PLL_Tracking") n_ref = ref_trk<BeidouB1iDllPllTracking>(config.get(), role, out_ref);
                    if (ref_impl == "BEIDOU_B3I_DLL_PLL_Tracking") n_ref = ref_trk<BeidouB3iDllPllTracking>(config.get(), role, out_ref);
                    if (ref_impl == "GPS_L2_M_DLL_PLL_Tracking") n_ref = ref_trk<GpsL2MDllPllTracking>(config.get(), role, out_ref);
                    std::unique_ptr<TrackingInterface> b = GetB200TrkBlock(config, role, b200_impl, 1, 1);
                    if (!b) return -2;
                    if (auto* g = dynamic_cast<B200DllPllTrackingAdapter*>(b.get())) n_b = flatten(g->tracking_->trk_parameters, g->item_size(), out_b200);
                }
        }
    catch (const std::exception& e)
        {
            std::cerr << "b200_adapter_conf_compare: " << e.what() << std::endl;
            return -3;
        }
    if (getenv("GB_CC_TRACE")) fprintf(stderr, "cc: adapters destroyed\n");
    B200Service::shutdown();
    if (getenv("GB_CC_TRACE")) fprintf(stderr, "cc: service shut down\n");
    if (n_ref < 0) return -1;
    if (n_b != n_ref) return -4;
    return n_ref;
}
}
