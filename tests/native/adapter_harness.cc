// TEST INFRASTRUCTURE: a miniature receiver around the *_B200 adapters (gnss-sdr-1_b200/adapters).
// GNU Radio is not installed, so the adapters are compiled against the stand-in headers under oracle/shim (gr::block / top_block /
// pmt ...) and against the REFERENCE'S OWN interface headers, configuration classes and PRN generators from /root/reference
// (AcquisitionInterface, TrackingInterface, ConfigurationInterface, FileConfiguration + INIReader, Gnss_Synchro, Acq_Conf,
// Dll_Pll_Conf, gps_l1_ca_code_gen_*, galileo_e1_code_gen_*).  This file does what Channel, ChannelFsm and the scheduler do:
//   * FileConfiguration on a real .conf; blocks made through GetB200AcqBlock / GetB200TrkBlock from the configured names;
//   * Channel's constructor order (channel.cc:60-104): set_channel, set_channel_fsm, set_gnss_synchro, set_doppler_step,
//     set_threshold, init; Channel::set_signal -> set_local_code (:208-226); ChannelFsm::start_acquisition -> reset();
//     Event_valid_acquisition -> start_tracking (channel_fsm.cc:197-218); acquisition event 2 -> reset() again (repeat);
//   * the scheduler: every block's general_work is offered the same stream from its own read position.
#include "b200_block_factory.h"
#include "b200_service_singleton.h"
#include "channel_fsm.h"
#include "file_configuration.h"
#include "gnss_synchro.h"
#include <algorithm>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

namespace
{
struct Chan
{
    std::shared_ptr<AcquisitionInterface> acq;
    std::shared_ptr<TrackingInterface> trk;
    std::shared_ptr<ChannelFsm> fsm;
    Gnss_Synchro syn;
    gr::block* acq_blk = nullptr;
    gr::block* trk_blk = nullptr;
    long long acq_pos = 0, trk_pos = 0;
    int acq_events_pos = 0, acq_events_neg = 0, loss_of_lock = 0;
    bool tracking = false;
    std::vector<Gnss_Synchro> out;
    double first_acq_delay = -1, first_acq_doppler = 0;
    unsigned long long first_acq_stamp = 0;
};
}  // namespace

extern "C" {

// per channel out[16]: 0 positive acquisitions, 1 negative acquisitions, 2 Acq_delay_samples, 3 Acq_doppler_hz, 4 Acq_samplestamp,
// 5 tracking epochs out, 6 mean Doppler of the last quarter, 7 last CN0, 8 losses of lock, 9 last Tracking_sample_counter,
// 10 last Code_phase_samples, 11 item size of the acquisition adapter, 12 item size of the tracking adapter, 13 tracking state at the end
// returns 0, or a negative number when a block could not be made (unknown implementation name, no device)
int b200_adapter_receiver(const char* conf_path, const void* samples, long long n_items, int n_channels, const uint32_t* prns, const char* system_signal,
    int chunk_items, double* out, char* impl_names, int impl_cap)
{
    std::shared_ptr<ConfigurationInterface> config = std::make_shared<FileConfiguration>(conf_path);
    const std::string sig(system_signal + 1);          // "1C" / "1B"
    const std::string acq_role = "Acquisition_" + sig, trk_role = "Tracking_" + sig;
    const std::string acq_impl = config->property(acq_role + ".implementation", std::string("")), trk_impl = config->property(trk_role + ".implementation", std::string(""));
    if (impl_names && impl_cap > 0) snprintf(impl_names, impl_cap, "%s %s", acq_impl.c_str(), trk_impl.c_str());
    std::vector<std::unique_ptr<Chan>> ch;
    try
        {
            for (int k = 0; k < n_channels; k++)
                {
                    std::unique_ptr<Chan> c(new Chan());
                    std::unique_ptr<AcquisitionInterface> a = GetB200AcqBlock(config, acq_role, acq_impl, 1, 0);
                    std::unique_ptr<TrackingInterface> t = GetB200TrkBlock(config, trk_role, trk_impl, 1, 1);
                    if (!a || (!t && !trk_impl.empty())) return -1;  // no Tracking_XX.implementation in the conf: acquisition only
                    c->acq = std::move(a);  // Channel keeps them as shared_ptr (channel.cc:45-52)
                    c->trk = std::move(t);
                    c->fsm = std::make_shared<ChannelFsm>();
                    // ---- Channel::Channel (channel.cc:60-104)
                    c->acq->set_channel(k);
                    c->acq->set_channel_fsm(c->fsm);
                    if (c->trk) c->trk->set_channel(k);
                    c->syn = Gnss_Synchro();
                    c->syn.Channel_ID = k;
                    c->acq->set_gnss_synchro(&c->syn);
                    if (c->trk) c->trk->set_gnss_synchro(&c->syn);
                    uint32_t doppler_step = config->property("Acquisition_" + sig + std::to_string(k) + ".doppler_step", 0);
                    if (doppler_step == 0) doppler_step = config->property("Acquisition_" + sig + ".doppler_step", 500);
                    c->acq->set_doppler_step(doppler_step);
                    float threshold = config->property("Acquisition_" + sig + std::to_string(k) + ".threshold", 0.0);
                    if (threshold == 0.0) threshold = config->property("Acquisition_" + sig + ".threshold", 0.0);
                    c->acq->set_threshold(threshold);
                    c->acq->init();
                    // ---- GNSSFlowgraph::connect: the blocks the channel exposes
                    c->acq_blk = dynamic_cast<gr::block*>(c->acq->get_left_block().get());
                    c->trk_blk = c->trk ? dynamic_cast<gr::block*>(c->trk->get_left_block().get()) : nullptr;
                    if (!c->acq_blk || (c->trk && !c->trk_blk)) return -2;
                    gr::top_block_sptr top = gr::make_top_block("harness");
                    c->acq->connect(top);
                    if (c->trk) c->trk->connect(top);
                    // ---- Channel::set_signal (:208-226)
                    c->syn.System = system_signal[0];
                    std::memcpy(c->syn.Signal, sig.c_str(), 3);
                    c->syn.PRN = prns[k];
                    c->acq->set_local_code();
                    // ---- ChannelFsm: valid acquisition -> start tracking (channel_fsm.cc:197-218)
                    Chan* cp = c.get();
                    c->fsm->on_valid_acquisition = [cp]() {
                        cp->acq_events_pos++;
                        if (cp->first_acq_delay < 0)
                            {
                                cp->first_acq_delay = cp->syn.Acq_delay_samples;
                                cp->first_acq_doppler = cp->syn.Acq_doppler_hz;
                                cp->first_acq_stamp = cp->syn.Acq_samplestamp_samples;
                            }
                        if (cp->trk) cp->trk->start_tracking();
                        cp->tracking = cp->trk != nullptr;
                    };
                    // ---- ChannelFsm::start_acquisition -> acq_->reset() (:203-206)
                    c->acq->reset();
                    ch.push_back(std::move(c));
                }
        }
    catch (const std::exception& e)
        {
            if (impl_names && impl_cap > 0) snprintf(impl_names, impl_cap, "%s", e.what());
            return -3;
        }
    const size_t isz = ch[0]->acq->item_size();
    const char* base = static_cast<const char*>(samples);
    // ---- scheduler: round robin; a block is offered at most chunk_items items per call from its own read position
    // (a block may need a call that consumes nothing to change state -- pcps_acquisition moves from buffering to the search that way --
    // so the loop ends only after a few rounds in which nobody consumed anything)
    bool progress = true;
    int idle_rounds = 0;
    while (idle_rounds < 4)
        {
            idle_rounds = progress ? 0 : idle_rounds + 1;
            progress = false;
            for (auto& c : ch)
                {
                    // acquisition block
                    if (c->acq_pos < n_items)
                        {
                            gr_vector_int ninput(1, (int)std::min<long long>(chunk_items, n_items - c->acq_pos));
                            gr_vector_const_void_star in(1, base + isz * c->acq_pos);
                            gr_vector_void_star outv;
                            c->acq_blk->gb_consumed = 0;
                            c->acq_blk->gb_events.clear();
                            c->acq_blk->general_work(0, ninput, in, outv);
                            c->acq_pos += c->acq_blk->gb_consumed;
                            if (c->acq_blk->gb_consumed > 0) progress = true;
                            for (long ev : c->acq_blk->gb_events)
                                if (ev == 2)
                                    {
                                        c->acq_events_neg++;
                                        c->acq->reset();  // Event_failed_acquisition_repeat -> start_acquisition again
                                        progress = true;
                                    }
                        }
                    if (!c->trk_blk) continue;
                    // tracking block: forecast() asks for 2 * vector_length items
                    gr_vector_int need(1, 0);
                    c->trk_blk->forecast(1, need);
                    // A standby tracking block runs at full throttle in its own GNU Radio thread and stays at the head of the stream; here it
                    // is called until it has caught up with its channel's acquisition block, so that start_tracking() never finds the
                    // tracker's sample counter behind the acquisition stamp (gps_l1_ca_kf_tracking_cc subtracts the two as uint64_t).
                    while (n_items - c->trk_pos >= need[0])
                        {
                            gr_vector_int ninput(1, (int)std::min<long long>(std::max(chunk_items, need[0]), n_items - c->trk_pos));
                            gr_vector_const_void_star in(1, base + isz * c->trk_pos);
                            Gnss_Synchro item = Gnss_Synchro();
                            gr_vector_void_star outv(1, &item);
                            c->trk_blk->gb_consumed = 0;
                            c->trk_blk->gb_events.clear();
                            const int produced = c->trk_blk->general_work(1, ninput, in, outv);
                            c->trk_pos += c->trk_blk->gb_consumed;
                            if (c->trk_blk->gb_consumed > 0) progress = true;
                            if (produced == 1) c->out.push_back(item);
                            for (long ev : c->trk_blk->gb_events)
                                if (ev == 3)
                                    {
                                        c->loss_of_lock++;
                                        c->tracking = false;
                                        c->acq->reset();  // Event_failed_tracking_standby -> the flowgraph restarts the acquisition
                                    }
                            if (c->tracking || c->trk_blk->gb_consumed == 0 || c->trk_pos >= c->acq_pos) break;
                        }
                }
        }
    for (int k = 0; k < n_channels; k++)
        {
            Chan& c = *ch[k];
            double* o = out + 16 * k;
            std::fill(o, o + 16, 0.0);
            o[0] = c.acq_events_pos;
            o[1] = c.acq_events_neg;
            o[2] = c.first_acq_delay;
            o[3] = c.first_acq_doppler;
            o[4] = (double)c.first_acq_stamp;
            o[5] = (double)c.out.size();
            if (!c.out.empty())
                {
                    const size_t q0 = c.out.size() - std::max<size_t>(1, c.out.size() / 4);
                    double s = 0;
                    for (size_t i = q0; i < c.out.size(); i++) s += c.out[i].Carrier_Doppler_hz;
                    o[6] = s / (double)(c.out.size() - q0);
                    o[7] = c.out.back().CN0_dB_hz;
                    o[9] = (double)c.out.back().Tracking_sample_counter;
                    o[10] = c.out.back().Code_phase_samples;
                }
            o[8] = c.loss_of_lock;
            o[11] = (double)c.acq->item_size();
            o[12] = c.trk ? (double)c.trk->item_size() : 0.0;
        }
    ch.clear();
    B200Service::shutdown();
    return 0;
}

}  // extern "C"
