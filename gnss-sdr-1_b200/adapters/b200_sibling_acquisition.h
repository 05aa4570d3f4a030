// b200_sibling_acquisition<Policy>: ONE GNU Radio block template for the sibling PCPS acquisition blocks of the reference (each its
// own gr::block there): pcps_tong_acquisition_cc, pcps_quicksync_acquisition_cc, pcps_acquisition_fine_doppler_cc,
// galileo_pcps_8ms_acquisition_cc, galileo_e5a_noncoherentIQ_acquisition_caf_cc.  The searches are
// the host mirrors of gnss-sdr-1_b200/host/ (each pinned to its block) on the shared B200 service; a Policy names the
// mirror, its configuration and how many items one pass through the block's search state reads and consumes.  What the blocks
// share is here: standby consumes whatever arrives, an active block waits for one whole search window, runs it, copies the result
// into the channel's Gnss_Synchro and reports 1 (positive) / 2 (negative) -- to the ChannelFsm when one is set, as pcps_acquisition
// does (:432-440), else on the "events" port as these older blocks do.
// The reference blocks take vectors of sampled_ms * samples_per_ms items from a stream_to_vector in their adapters; this block
// takes the plain gr_complex stream (forecast() asks for one window), so no stream_to_vector is needed in front of it.
#ifndef GNSS_SDR_B200_SIBLING_ACQUISITION_H
#define GNSS_SDR_B200_SIBLING_ACQUISITION_H

#include "channel_fsm.h"
#include "gnss_synchro.h"
#include "host/gb200_acquisition_variants.hpp"
#include "host/gb200_quicksync.hpp"
#include <gnss_b200.h>
#include <gnuradio/block.h>
#include <gnuradio/io_signature.h>
#include <pmt/pmt.h>
#include <boost/shared_ptr.hpp>
#include <algorithm>
#include <complex>
#include <memory>

struct B200TongPolicy
{
    using Conf = gb200::AcqVariantConf;
    using Mirror = gb200::B200_Pcps_Tong_Acquisition;
    static const char* name() { return "b200_pcps_tong_acquisition"; }
    static uint32_t window_items(const Conf& c) { return c.sampled_ms * static_cast<uint32_t>(c.samples_per_ms); }  // one dwell = one vector
    static uint32_t code_items(const Conf& c) { return window_items(c); }
    static bool set_code(Mirror& m, const std::complex<float>* code, const std::complex<float>*, const Conf& c) { return m.set_local_code(code, code_items(c)); }
};
struct B200FineDopplerPolicy
{
    using Conf = gb200::AcqVariantConf;
    using Mirror = gb200::B200_Pcps_Acquisition_Fine_Doppler;
    static const char* name() { return "b200_pcps_acquisition_fine_doppler"; }
    // one run() is the whole attempt: max_dwells grid dwells, then the 10 ms the fine estimate reads (pcps_acquisition_fine_doppler_cc.cc:543-578)
    static uint32_t window_items(const Conf& c) { return std::max<uint32_t>(c.max_dwells, 10u) * static_cast<uint32_t>(c.samples_per_ms); }
    static uint32_t code_items(const Conf& c) { return static_cast<uint32_t>(c.samples_per_ms); }
    static bool set_code(Mirror& m, const std::complex<float>* code, const std::complex<float>*, const Conf& c) { return m.set_local_code(code, code_items(c)); }
};
struct B200Galileo8msPolicy
{
    using Conf = gb200::AcqVariantConf;
    using Mirror = gb200::B200_Galileo_Pcps_8ms_Acquisition;
    static const char* name() { return "b200_galileo_pcps_8ms_acquisition"; }
    static uint32_t window_items(const Conf& c) { return c.sampled_ms * static_cast<uint32_t>(c.samples_per_ms); }
    static uint32_t code_items(const Conf& c) { return window_items(c); }
    static bool set_code(Mirror& m, const std::complex<float>* code, const std::complex<float>*, const Conf& c) { return m.set_local_code(code, code_items(c)); }
};
struct B200E5aCafPolicy
{
    using Conf = gb200::AcqVariantConf;
    using Mirror = gb200::B200_Galileo_E5a_NoncoherentIQ_Acquisition_Caf;
    static const char* name() { return "b200_galileo_e5a_noncoherent_iq_acquisition_caf"; }
    static uint32_t window_items(const Conf& c) { return c.sampled_ms * static_cast<uint32_t>(c.samples_per_ms); }  // d_fft_size (:76)
    static uint32_t code_items(const Conf& c) { return window_items(c); }
    static bool set_code(Mirror& m, const std::complex<float>* codeI, const std::complex<float>* codeQ, const Conf& c) { return m.set_local_code(codeI, codeQ ? codeQ : codeI, code_items(c)); }
};
struct B200QuickSyncPolicy
{
    using Conf = gb200::QuickSyncConf;
    using Mirror = gb200::B200_Pcps_QuickSync_Acquisition;
    static const char* name() { return "b200_pcps_quicksync_acquisition"; }
    static uint32_t window_items(const Conf& c) { return c.sampled_ms * static_cast<uint32_t>(c.samples_per_ms); }
    static uint32_t code_items(const Conf& c) { return static_cast<uint32_t>(c.samples_per_code); }
    static bool set_code(Mirror& m, const std::complex<float>* code, const std::complex<float>*, const Conf& c) { return m.set_local_code(code, code_items(c)); }
};

template <class Policy>
class b200_sibling_acquisition : public gr::block
{
public:
    using sptr = boost::shared_ptr<b200_sibling_acquisition<Policy>>;
    static sptr make(gb200_service* service, int stream_id, const typename Policy::Conf& conf) { return sptr(new b200_sibling_acquisition<Policy>(service, stream_id, conf)); }
    ~b200_sibling_acquisition() = default;

    bool ok() const { return mirror_ && mirror_->ok(); }
    void set_gnss_synchro(Gnss_Synchro* p) { d_gnss_synchro = p; }
    void set_channel(uint32_t channel) { d_channel = channel; }
    void set_channel_fsm(std::shared_ptr<ChannelFsm> channel_fsm) { d_channel_fsm = std::move(channel_fsm); }
    void set_threshold(float threshold)
    {
        gr::thread::scoped_lock lock(d_setlock);
        mirror_->set_threshold(threshold);
    }
    void set_doppler_max(uint32_t doppler_max)
    {
        gr::thread::scoped_lock lock(d_setlock);
        mirror_->set_doppler_max(doppler_max);
    }
    void set_doppler_step(uint32_t doppler_step)
    {
        gr::thread::scoped_lock lock(d_setlock);
        mirror_->set_doppler_step(doppler_step);
    }
    void init()
    {
        gr::thread::scoped_lock lock(d_setlock);
        if (d_gnss_synchro)
            {
                d_gnss_synchro->Flag_valid_acquisition = false;
                d_gnss_synchro->Flag_valid_symbol_output = false;
                d_gnss_synchro->Flag_valid_pseudorange = false;
                d_gnss_synchro->Flag_valid_word = false;
                d_gnss_synchro->Acq_delay_samples = 0.0;
                d_gnss_synchro->Acq_doppler_hz = 0.0;
                d_gnss_synchro->Acq_samplestamp_samples = 0ULL;
                d_gnss_synchro->Acq_doppler_step = 0U;
            }
        mirror_->init();
    }
    void set_local_code(std::complex<float>* code, std::complex<float>* code_q = nullptr)  // code_q: the E5a block's second replica
    {
        gr::thread::scoped_lock lock(d_setlock);
        Policy::set_code(*mirror_, code, code_q, conf_);
    }
    void set_active(bool active)
    {
        gr::thread::scoped_lock lock(d_setlock);
        d_active = active;
        if (active) mirror_->set_state(1);
    }
    void set_state(int32_t state)
    {
        gr::thread::scoped_lock lock(d_setlock);
        mirror_->set_state(state);
        d_active = state == 1;
    }
    uint32_t mag() const { return 0; }
    float test_statistics() const { return mirror_->test_statistics(); }

    void forecast(int noutput_items __attribute__((unused)), gr_vector_int& ninput_items_required)
    {
        ninput_items_required[0] = d_active ? static_cast<int>(Policy::window_items(conf_)) : 1;
    }

    int general_work(int noutput_items __attribute__((unused)), gr_vector_int& ninput_items, gr_vector_const_void_star& input_items,
        gr_vector_void_star& output_items __attribute__((unused)))
    {
        gr::thread::scoped_lock lk(d_setlock);
        const char* in = reinterpret_cast<const char*>(input_items[0]);
        if (ninput_items[0] > 0) gb200_stream_push_shared(s_, stream_, in, static_cast<size_t>(ninput_items[0]), GB200_ITEM_32FC, d_sample_counter);
        const uint32_t window = Policy::window_items(conf_);
        if (!d_active)
            {
                d_sample_counter += static_cast<uint64_t>(ninput_items[0]);
                consume_each(ninput_items[0]);
                return 0;
            }
        if (static_cast<uint32_t>(ninput_items[0]) < window)
            {
                consume_each(0);  // the scheduler comes back with a whole window (forecast)
                return 0;
            }
        gb200::AcqSynchro syn;
        mirror_->set_gnss_synchro(&syn);
        const int ev = mirror_->run(d_sample_counter);
        mirror_->set_gnss_synchro(nullptr);
        d_sample_counter += window;
        consume_each(static_cast<int>(window));
        if (ev < 0 || ev == 0) return 0;  // error (window no longer in the ring) or more dwells needed: stay active
        if (d_gnss_synchro)
            {
                d_gnss_synchro->Acq_delay_samples = syn.Acq_delay_samples;
                d_gnss_synchro->Acq_doppler_hz = syn.Acq_doppler_hz;
                d_gnss_synchro->Acq_samplestamp_samples = syn.Acq_samplestamp_samples;
                d_gnss_synchro->Acq_doppler_step = syn.Acq_doppler_step;
                d_gnss_synchro->Flag_valid_acquisition = ev == 1;
            }
        d_active = false;
        if (ev == 1 && d_channel_fsm)
            d_channel_fsm->Event_valid_acquisition();
        else
            this->message_port_pub(pmt::mp("events"), pmt::from_long(ev));
        return 0;
    }

private:
    b200_sibling_acquisition(gb200_service* service, int stream_id, const typename Policy::Conf& conf)
        : gr::block(Policy::name(), gr::io_signature::make(1, 1, sizeof(gr_complex)), gr::io_signature::make(0, 0, sizeof(gr_complex))), s_(service), stream_(stream_id), conf_(conf)
    {
        this->message_port_register_out(pmt::mp("events"));
        mirror_.reset(new typename Policy::Mirror(service, conf, stream_id));
    }
    gb200_service* s_;
    int stream_;
    typename Policy::Conf conf_;
    std::unique_ptr<typename Policy::Mirror> mirror_;
    Gnss_Synchro* d_gnss_synchro = nullptr;
    std::shared_ptr<ChannelFsm> d_channel_fsm;
    uint32_t d_channel = 0;
    uint64_t d_sample_counter = 0;
    bool d_active = false;
};

#endif
