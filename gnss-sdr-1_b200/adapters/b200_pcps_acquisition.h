// b200_pcps_acquisition: the GNU Radio block behind the *_PCPS_Acquisition_B200 adapters -- the public surface and the
// general_work state machine of pcps_acquisition (src/algorithms/acquisition/gnuradio_blocks/pcps_acquisition.h:93-250,
// pcps_acquisition.cc:941-1054), with acquisition_core replaced by one gb200_acq_run on the shared correlator service.
// "Buffering" d_consumed_samples items means handing them to the service's device ring (gb200_stream_push_shared: the first
// block of the flowgraph to see an item copies it) and remembering the index of the first one.
#ifndef GNSS_SDR_B200_PCPS_ACQUISITION_H
#define GNSS_SDR_B200_PCPS_ACQUISITION_H

#include "GLONASS_L1_L2_CA.h"  // DFRQ1_GLO, DFRQ2_GLO, GLONASS_PRN
#include "acq_conf.h"
#include "channel_fsm.h"
#include "gnss_synchro.h"
#include <gnss_b200.h>
#include <gnuradio/block.h>
#include <gnuradio/io_signature.h>
#include <pmt/pmt.h>
#include <cmath>
#include <complex>
#include <cstdint>
#include <cstring>
#include <memory>

class b200_pcps_acquisition;
using b200_pcps_acquisition_sptr = boost::shared_ptr<b200_pcps_acquisition>;

class b200_pcps_acquisition : public gr::block
{
public:
    // conf_: the Acq_Conf the adapter fills exactly as for pcps_make_acquisition; item_type: GB200_ITEM_* of the input stream
    static b200_pcps_acquisition_sptr make(gb200_service* service, int stream_id, const Acq_Conf& conf_, int item_type, int64_t freq_offset_hz = 0)
    {
        return b200_pcps_acquisition_sptr(new b200_pcps_acquisition(service, stream_id, conf_, item_type, freq_offset_hz));
    }
    ~b200_pcps_acquisition()
    {
        if (h_) gb200_acq_destroy(h_);
    }

    void set_gnss_synchro(Gnss_Synchro* p_gnss_synchro)
    {
        gr::thread::scoped_lock lock(d_setlock);
        d_gnss_synchro = p_gnss_synchro;
    }
    uint32_t mag() const { return static_cast<uint32_t>(d_mag); }
    void init()  // (:313-370): the Doppler grid is derived from doppler_max / doppler_step; nothing is stored per bin here
    {
        gr::thread::scoped_lock lock(d_setlock);
        d_gnss_synchro->Flag_valid_acquisition = false;
        d_gnss_synchro->Flag_valid_symbol_output = false;
        d_gnss_synchro->Flag_valid_pseudorange = false;
        d_gnss_synchro->Flag_valid_word = false;
        d_gnss_synchro->Acq_doppler_step = 0U;
        d_gnss_synchro->Acq_delay_samples = 0.0;
        d_gnss_synchro->Acq_doppler_hz = 0.0;
        d_gnss_synchro->Acq_samplestamp_samples = 0ULL;
        d_mag = 0.0;
        d_input_power = 0.0;
        gb200_acq_set_doppler(h_, acq_parameters.doppler_max, d_doppler_step, freq_offset_hz_ + fdma_offset_hz_);
    }
    void set_local_code(std::complex<float>* code)  // (:239-274); `code` as the adapter built it (d_consumed_samples items, or N/2 with bit transition)
    {
        gr::thread::scoped_lock lock(d_setlock);
        // FDMA (:242-248, is_fdma :277-293): the satellite's channel offset is added to every Doppler bin of the wipe-off grid
        int64_t fdma = 0;
        if (d_gnss_synchro && std::strcmp(d_gnss_synchro->Signal, "1G") == 0 && GLONASS_PRN.count(d_gnss_synchro->PRN)) fdma = DFRQ1_GLO * GLONASS_PRN.at(d_gnss_synchro->PRN);
        if (d_gnss_synchro && std::strcmp(d_gnss_synchro->Signal, "2G") == 0 && GLONASS_PRN.count(d_gnss_synchro->PRN)) fdma = DFRQ2_GLO * GLONASS_PRN.at(d_gnss_synchro->PRN);
        if (fdma != fdma_offset_hz_)
            {
                fdma_offset_hz_ = fdma;
                gb200_acq_set_doppler(h_, acq_parameters.doppler_max, d_doppler_step, freq_offset_hz_ + fdma_offset_hz_);
            }
        const size_t n = acq_parameters.bit_transition_flag ? gb200_acq_fft_size(h_) / 2 : gb200_acq_consumed_samples(h_);
        gb200_acq_set_local_code(h_, reinterpret_cast<const float*>(code), n);
    }
    void set_active(bool active)
    {
        gr::thread::scoped_lock lock(d_setlock);
        d_active = active;
    }
    void set_state(int32_t state)  // (:393-415)
    {
        gr::thread::scoped_lock lock(d_setlock);
        d_state = state;
        if (d_state == 1)
            {
                d_gnss_synchro->Acq_delay_samples = 0.0;
                d_gnss_synchro->Acq_doppler_hz = 0.0;
                d_gnss_synchro->Acq_samplestamp_samples = 0ULL;
                d_gnss_synchro->Acq_doppler_step = 0U;
                d_mag = 0.0;
                d_input_power = 0.0;
                d_test_statistics = 0.0;
                d_active = true;
            }
    }
    void set_channel(uint32_t channel) { d_channel = channel; }
    void set_channel_fsm(std::shared_ptr<ChannelFsm> channel_fsm) { d_channel_fsm = channel_fsm; }
    void set_threshold(float threshold)
    {
        gr::thread::scoped_lock lock(d_setlock);
        d_threshold = threshold;
        gb200_acq_set_threshold(h_, threshold);
    }
    void set_doppler_max(uint32_t doppler_max)
    {
        gr::thread::scoped_lock lock(d_setlock);
        acq_parameters.doppler_max = doppler_max;
    }
    void set_doppler_step(uint32_t doppler_step)
    {
        gr::thread::scoped_lock lock(d_setlock);
        d_doppler_step = doppler_step;
    }
    void set_resampler_latency(uint32_t latency_samples)
    {
        gr::thread::scoped_lock lock(d_setlock);
        acq_parameters.resampler_latency_samples = latency_samples;
    }
    float test_statistics() const { return d_test_statistics; }
    uint64_t sample_counter() const { return d_sample_counter; }

    int general_work(int noutput_items __attribute__((unused)), gr_vector_int& ninput_items, gr_vector_const_void_star& input_items,
        gr_vector_void_star& output_items __attribute__((unused)))
    {
        gr::thread::scoped_lock lk(d_setlock);
        const char* in = reinterpret_cast<const char*>(input_items[0]);
        if (!d_active)
            {
                // standby (:961-975): the stream flows past; whatever this block sees is still offered to the shared ring
                if (!acq_parameters.blocking_on_standby)
                    {
                        hand_over(in, ninput_items[0]);
                        d_sample_counter += static_cast<uint64_t>(ninput_items[0]);
                        consume_each(ninput_items[0]);
                    }
                if (d_step_two)
                    {
                        // the service re-centres the fine grid on the coarse estimate by itself (gb200_acq_result::decision == 3)
                        d_state = 0;
                        d_active = true;
                    }
                return 0;
            }
        switch (d_state)
            {
            case 0:
                {
                    d_gnss_synchro->Acq_delay_samples = 0.0;
                    d_gnss_synchro->Acq_doppler_hz = 0.0;
                    d_gnss_synchro->Acq_samplestamp_samples = 0ULL;
                    d_gnss_synchro->Acq_doppler_step = 0U;
                    d_mag = 0.0;
                    d_input_power = 0.0;
                    d_test_statistics = 0.0;
                    d_state = 1;
                    d_buffer_count = 0U;
                    if (!acq_parameters.blocking_on_standby)
                        {
                            hand_over(in, ninput_items[0]);
                            d_sample_counter += static_cast<uint64_t>(ninput_items[0]);
                            consume_each(ninput_items[0]);
                        }
                    break;
                }
            case 1:
                {
                    // (:995-1034) the next d_consumed_samples items of the stream are the block to search
                    const uint32_t consumed = gb200_acq_consumed_samples(h_);
                    uint32_t buff_increment;
                    if ((static_cast<uint32_t>(ninput_items[0]) + d_buffer_count) <= consumed)
                        buff_increment = static_cast<uint32_t>(ninput_items[0]);
                    else
                        buff_increment = consumed - d_buffer_count;
                    if (d_buffer_count == 0U) d_block_first = d_sample_counter;
                    hand_over(in, static_cast<int>(buff_increment));
                    if (d_buffer_count >= consumed) d_state = 2;
                    d_buffer_count += buff_increment;
                    d_sample_counter += static_cast<uint64_t>(buff_increment);
                    consume_each(static_cast<int>(buff_increment));
                    break;
                }
            case 2:
                {
                    acquisition_core();
                    consume_each(0);
                    d_buffer_count = 0U;
                    break;
                }
            }
        return 0;
    }

private:
    b200_pcps_acquisition(gb200_service* service, int stream_id, const Acq_Conf& conf_, int item_type, int64_t freq_offset_hz)
        : gr::block("b200_pcps_acquisition", gr::io_signature::make(1, 1, conf_.it_size), gr::io_signature::make(0, 0, conf_.it_size)),
          s_(service), stream_(stream_id), item_type_(item_type), freq_offset_hz_(freq_offset_hz), acq_parameters(conf_)
    {
        if (acq_parameters.bit_transition_flag) acq_parameters.max_dwells = 1;  // (:113-117): the block's own copy, as pcps_acquisition's constructor leaves it
        this->message_port_register_out(pmt::mp("events"));
        gb200_acq_cfg c{};
        c.sampled_ms = conf_.sampled_ms;
        c.ms_per_code = conf_.ms_per_code;
        c.samples_per_chip = conf_.samples_per_chip;
        c.max_dwells = conf_.max_dwells ? conf_.max_dwells : 1;
        c.doppler_max = conf_.doppler_max;
        c.doppler_step = 250;  // set_doppler_step() + init() follow (channel.cc:84-104)
        c.num_doppler_bins_step2 = conf_.num_doppler_bins_step2;
        c.doppler_step2 = conf_.doppler_step2;
        c.fs_in = conf_.use_automatic_resampler ? conf_.resampled_fs : conf_.fs_in;
        c.samples_per_ms = conf_.samples_per_ms;
        c.samples_per_code = conf_.samples_per_code;
        c.bit_transition_flag = conf_.bit_transition_flag;
        c.use_CFAR_algorithm_flag = conf_.use_CFAR_algorithm_flag;
        c.make_2_steps = conf_.make_2_steps;
        c.item_type = item_type;
        c.threshold = 0.0f;
        c.freq_offset_hz = freq_offset_hz;
        if (gb200_acq_create(s_, &c, &h_) != GB200_OK) h_ = nullptr;
        if (conf_.dump) gb200_acq_keep_grid(h_, 1);
    }

    void hand_over(const char* in, int n_items)
    {
        if (n_items > 0) gb200_stream_push_shared(s_, stream_, in, static_cast<size_t>(n_items), item_type_, d_sample_counter);
    }

    // = acquisition_core (:668-929) + send_positive / send_negative_acquisition (:418-459): the search itself, the statistic,
    // the dwell / two-step / bit-transition decision logic live in the service (gb200_acq_run); here its outcome is written
    // into the shared Gnss_Synchro and turned into the block's events
    void acquisition_core()
    {
        gb200_acq_result r;
        if (gb200_acq_run(h_, stream_, d_block_first, &r) != GB200_OK)
            {
                // the block of samples is gone from the ring (this block lagged more than the ring holds): search the next one
                d_state = 1;
                d_buffer_count = 0U;
                return;
            }
        d_mag = r.peak_mag;
        d_input_power = r.input_power;
        d_test_statistics = r.test_statistic;
        d_gnss_synchro->Acq_delay_samples = r.acq_delay_samples;
        d_gnss_synchro->Acq_doppler_hz = r.acq_doppler_hz;
        d_gnss_synchro->Acq_samplestamp_samples = r.acq_samplestamp;
        d_gnss_synchro->Acq_doppler_step = d_step_two ? static_cast<uint32_t>(acq_parameters.doppler_step2) : d_doppler_step;
        if (acq_parameters.use_automatic_resampler)
            {
                // (:756-762, :814-820) back to the rate the tracking blocks run at
                d_gnss_synchro->Acq_delay_samples = static_cast<double>(std::fmod(static_cast<float>(r.code_phase_index), acq_parameters.samples_per_code)) * acq_parameters.resampler_ratio;
                d_gnss_synchro->Acq_delay_samples -= static_cast<double>(acq_parameters.resampler_latency_samples);
                d_gnss_synchro->Acq_samplestamp_samples = rint(static_cast<double>(r.acq_samplestamp) * acq_parameters.resampler_ratio);
            }
        switch (r.decision)
            {
            case 1:  // positive
                d_step_two = false;
                d_active = false;
                d_state = 0;
                if (d_channel_fsm)
                    d_channel_fsm->Event_valid_acquisition();
                else
                    this->message_port_pub(pmt::mp("events"), pmt::from_long(1));
                break;
            case 0:  // negative
                d_step_two = false;
                d_active = false;
                d_state = 0;
                this->message_port_pub(pmt::mp("events"), pmt::from_long(2));
                break;
            case 3:  // step two armed: standby until general_work re-activates the block (:846-851, :978-985)
                d_step_two = true;
                d_active = false;
                d_state = 0;
                break;
            default:  // more dwells wanted
                d_state = 1;
                break;
            }
    }

    gb200_service* s_;
    gb200_acq* h_ = nullptr;
    int stream_, item_type_;
    int64_t freq_offset_hz_;
    int64_t fdma_offset_hz_ = 0;  // GLONASS channel offset of the satellite being searched (set_local_code)
    Acq_Conf acq_parameters;
    bool d_active = false, d_step_two = false;
    int32_t d_state = 0;
    float d_threshold = 0.0f, d_mag = 0.0f, d_input_power = 0.0f, d_test_statistics = 0.0f;
    uint32_t d_channel = 0U, d_doppler_step = 250U, d_buffer_count = 0U;
    uint64_t d_sample_counter = 0ULL, d_block_first = 0ULL;
    std::shared_ptr<ChannelFsm> d_channel_fsm;
    Gnss_Synchro* d_gnss_synchro = nullptr;
};

#endif
