#include "b200_block_factory.h"
#include "b200_sibling_acquisition_adapters.h"
#include "b200_sibling_tracking_adapters.h"
#include "b200_signal_adapters.h"

std::unique_ptr<AcquisitionInterface> GetB200AcqBlock(const std::shared_ptr<ConfigurationInterface>& configuration, const std::string& role,
    const std::string& implementation, unsigned int in_streams, unsigned int out_streams)
{
    if (const B200AcqTraits* t = b200_acq_traits(implementation))  // every signal's pcps_acquisition adapter: one class, a table row each
        return std::unique_ptr<AcquisitionInterface>(new B200PcpsAcquisitionAdapter(configuration.get(), role, in_streams, out_streams, *t));
    return GetB200SiblingAcqBlock(configuration.get(), role, implementation, in_streams, out_streams);  // GPS L1 C/A Tong, QuickSync, fine-Doppler acquisitions
}

std::unique_ptr<TrackingInterface> GetB200TrkBlock(const std::shared_ptr<ConfigurationInterface>& configuration, const std::string& role,
    const std::string& implementation, unsigned int in_streams, unsigned int out_streams)
{
    if (const B200TrkTraits* t = b200_trk_traits(implementation))  // every signal's dll_pll_veml_tracking adapter: one class, a table row each
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
