// One gb200_service per process and GPU, shared by every *_B200 adapter of the receiver: "one shared GPU correlator service
// batches every channel's epoch work into single launches".  Created on first use from the GNSS-SDR section of the configuration:
//   GNSS-SDR.b200_device            CUDA ordinal (default 0)
//   GNSS-SDR.b200_batch_window_us   how long the first caller of an epoch waits for the other channels (default 200; 0 = no coalescing)
//   GNSS-SDR.b200_ring_items        items kept per stream on the device (default 2^24)
// Replaces the per-channel cudaSetDevice / stream / buffer set-up of cuda_multicorrelator::init_cuda_integrated_resampler
// (src/algorithms/tracking/libs/cuda_multicorrelator.cu:137-241).  Stream ids: one per RF channel of the signal conditioner.
#ifndef GNSS_SDR_B200_SERVICE_SINGLETON_H
#define GNSS_SDR_B200_SERVICE_SINGLETON_H

#include "configuration_interface.h"
#include <gnss_b200.h>
#include <mutex>
#include <stdexcept>
#include <string>

class B200Service
{
public:
    static gb200_service* get(ConfigurationInterface* configuration)
    {
        static std::mutex mu;
        std::lock_guard<std::mutex> lk(mu);
        gb200_service*& s = instance();
        if (s == nullptr)
            {
                gb200_service_cfg cfg{};
                cfg.device = configuration ? configuration->property("GNSS-SDR.b200_device", 0) : 0;
                cfg.max_streams = 8;
                cfg.ring_capacity = configuration ? static_cast<uint64_t>(configuration->property("GNSS-SDR.b200_ring_items", static_cast<int64_t>(1) << 24)) : (1ULL << 24);
                cfg.ring_slack = 1ULL << 21;
                cfg.batch_window_us = configuration ? static_cast<uint32_t>(configuration->property("GNSS-SDR.b200_batch_window_us", 200)) : 200U;
                if (gb200_service_create(&cfg, &s) != GB200_OK)
                    {
                        // no CPU fallback: a receiver configured with *_B200 blocks cannot start without the device
                        throw std::runtime_error(std::string("B200 correlator service: ") + gb200_last_error(nullptr));
                    }
            }
        return s;
    }
    // tests / orderly shutdown of the receiver
    static void shutdown()
    {
        gb200_service*& s = instance();
        if (s != nullptr) gb200_service_destroy(s);
        s = nullptr;
    }
    static int item_type_of(const std::string& item_type)  // the adapters' item_type key -> wire type kept on the device
    {
        if (item_type == "cshort" || item_type == "ishort") return GB200_ITEM_16IC;
        if (item_type == "cbyte" || item_type == "ibyte") return GB200_ITEM_8IC;
        return GB200_ITEM_32FC;
    }
    static size_t item_size_of(int item_type) { return item_type == GB200_ITEM_32FC ? 8 : (item_type == GB200_ITEM_16IC ? 4 : 2); }

private:
    static gb200_service*& instance()
    {
        static gb200_service* s = nullptr;
        return s;
    }
};

#endif
