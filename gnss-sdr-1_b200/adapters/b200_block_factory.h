// The factory entries a maintainer adds to GNSSBlockFactory (src/core/receiver/gnss_block_factory.cc): in GetBlock (:1552 / :1713,
// the acquisition and tracking chains) and in GetAcqBlock / GetTrkBlock (:1954 / :2131), next to the existing
// `#if CUDA_GPU_ACCEL ... "GPS_L1_CA_DLL_PLL_Tracking_GPU"` entries (:1773-1780, :2225-2232):
//
//     else if (std::unique_ptr<GNSSBlockInterface> b = GetB200Block(configuration, role, implementation, in_streams, out_streams))
//         {
//             block = std::move(b);
//         }
//
// Returns nullptr for every implementation name that is not one of the *_B200 names, so the reference's chain continues.
#ifndef GNSS_SDR_B200_BLOCK_FACTORY_H
#define GNSS_SDR_B200_BLOCK_FACTORY_H

#include "acquisition_interface.h"
#include "configuration_interface.h"
#include "gnss_block_interface.h"
#include "tracking_interface.h"
#include <memory>
#include <string>

std::unique_ptr<GNSSBlockInterface> GetB200Block(const std::shared_ptr<ConfigurationInterface>& configuration, const std::string& role,
    const std::string& implementation, unsigned int in_streams, unsigned int out_streams);
std::unique_ptr<AcquisitionInterface> GetB200AcqBlock(const std::shared_ptr<ConfigurationInterface>& configuration, const std::string& role,
    const std::string& implementation, unsigned int in_streams, unsigned int out_streams);
std::unique_ptr<TrackingInterface> GetB200TrkBlock(const std::shared_ptr<ConfigurationInterface>& configuration, const std::string& role,
    const std::string& implementation, unsigned int in_streams, unsigned int out_streams);

#endif
