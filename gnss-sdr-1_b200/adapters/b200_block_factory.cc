#include "b200_block_factory.h"
#include "b200_sibling_acquisition_adapters.h"
#include "b200_sibling_tracking_adapters.h"
#include "b200_signal_adapters.h"
#include "galileo_e1_dll_pll_veml_tracking_b200.h"
#include "galileo_e1_pcps_ambiguous_acquisition_b200.h"
#include "gps_l1_ca_dll_pll_tracking_b200.h"
#include "gps_l1_ca_pcps_acquisition_b200.h"

std::unique_ptr<AcquisitionInterface> GetB200AcqBlock(const std::shared_ptr<ConfigurationInterface>& configuration, const std::string& role,
    const std::string& implementation, unsigned int in_streams, unsigned int out_streams)
{
    if (implementation == "GPS_L1_CA_PCPS_Acquisition_B200")
        return std::unique_ptr<AcquisitionInterface>(new GpsL1CaPcpsAcquisitionB200(configuration.get(), role, in_streams, out_streams));
    if (implementation == "Galileo_E1_PCPS_Ambiguous_Acquisition_B200")
        return std::unique_ptr<AcquisitionInterface>(new GalileoE1PcpsAmbiguousAcquisitionB200(configuration.get(), role, in_streams, out_streams));
    if (const B200AcqTraits* t = b200_acq_traits(implementation))  // GPS L5i, Galileo E5a, BeiDou B1I, GLONASS L1 C/A: one class, a table row each
        return std::unique_ptr<AcquisitionInterface>(new B200PcpsAcquisitionAdapter(configuration.get(), role, in_streams, out_streams, *t));
    return GetB200SiblingAcqBlock(configuration.get(), role, implementation, in_streams, out_streams);  // GPS L1 C/A Tong, QuickSync, fine-Doppler acquisitions
}

std::unique_ptr<TrackingInterface> GetB200TrkBlock(const std::shared_ptr<ConfigurationInterface>& configuration, const std::string& role,
    const std::string& implementation, unsigned int in_streams, unsigned int out_streams)
{
    if (implementation == "GPS_L1_CA_DLL_PLL_Tracking_B200")
        return std::unique_ptr<TrackingInterface>(new GpsL1CaDllPllTrackingB200(configuration.get(), role, in_streams, out_streams));
    if (implementation == "Galileo_E1_DLL_PLL_VEML_Tracking_B200")
        return std::unique_ptr<TrackingInterface>(new GalileoE1DllPllVemlTrackingB200(configuration.get(), role, in_streams, out_streams));
    if (const B200TrkTraits* t = b200_trk_traits(implementation))  // GPS L5, Galileo E5a, BeiDou B1I
        return std::unique_ptr<TrackingInterface>(new B200DllPllTrackingAdapter(configuration.get(), role, in_streams, out_streams, *t));
    return GetB200SiblingTrkBlock(configuration.get(), role, implementation, in_streams, out_streams);  // GLONASS L1 C/A, carrier-aided and Kalman GPS L1 C/A trackers
}

std::unique_ptr<GNSSBlockInterface> GetB200Block(const std::shared_ptr<ConfigurationInterface>& configuration, const std::string& role,
    const std::string& implementation, unsigned int in_streams, unsigned int out_streams)
{
    if (auto a = GetB200AcqBlock(configuration, role, implementation, in_streams, out_streams)) return a;
    if (auto t = GetB200TrkBlock(configuration, role, implementation, in_streams, out_streams)) return t;
    return nullptr;
}
