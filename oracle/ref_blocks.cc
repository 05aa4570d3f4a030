// TEST INFRASTRUCTURE ONLY -- never linked into the product library.
//
// Drives the reference's GNU Radio blocks themselves -- pcps_acquisition.cc and dll_pll_veml_tracking.cc, compiled UNMODIFIED from
// /root/reference into oracle/_ref/libgnss_ref.so -- without GNU Radio: oracle/shim/gnuradio/block.h stands in for gr::block (it
// records consume_each() and published events, nothing else), oracle/shim/{boost,pmt,matio.h,armadillo,volk} for the other
// libraries that are not installed.  This file plays the scheduler: it hands general_work() a pointer into the caller's sample
// array and the number of items available, advances by what the block consumed and copies out the Gnss_Synchro it produced.
// No loop, discriminator, state-machine or statistics code is written here -- that is all the reference's.
//
// `private` is opened up for these two headers only, to read the blocks' internal state (d_state, remnant phases, statistics)
// next to their outputs; every standard header is included first so the macro never touches the library.
#include <algorithm>
#include <any>
#include <cmath>
#include <complex>
#include <cstdint>
#include <cstring>
#include <deque>
#include <fstream>
#include <functional>
#include <iostream>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <string>
#include <thread>
#include <utility>
#include <vector>
#include <gnuradio/block.h>
#include <gnuradio/fft/fft.h>
#include <volk/volk.h>
#include <volk_gnsssdr/volk_gnsssdr.h>
#include "gnss_synchro.h"
#include "channel_fsm.h"
#include "tracking_dump_reader.h"

#define private public
#include "dll_pll_veml_tracking.h"
#include "pcps_acquisition.h"
#undef private

extern "C" {

// ------------------------------------------------------------------------------------------------ tracking block
struct GbRefTrkConf
{
    double fs_in;
    uint32_t vector_length;
    char system;
    char signal[3];
    int pll_filter_order, dll_filter_order, fll_filter_order;
    int enable_fll_pull_in, enable_fll_steady_state;
    uint32_t pull_in_time_s;
    float fll_bw_hz, pll_bw_hz, dll_bw_hz, pll_bw_narrow_hz, dll_bw_narrow_hz;
    float early_late_space_chips, very_early_late_space_chips, early_late_space_narrow_chips, very_early_late_space_narrow_chips;
    int32_t extend_correlation_symbols;
    int high_dyn;
    uint32_t smoother_length;
    int32_t cn0_samples, cn0_min, max_lock_fail;
    double carrier_lock_th;
    int track_pilot;
    int dump;
    char dump_filename[256];
};

// one general_work() call that produced an output item
struct GbRefTrkEpoch
{
    uint64_t tracking_sample_counter;
    uint64_t first_sample;  // index (in the caller's array) of the first sample the correlators read in this call
    int32_t n_samples;      // items the call consumed (d_current_prn_length_samples)
    int32_t state_after;    // d_state when the call returned
    double prompt_i, prompt_q, carrier_phase_rads, carrier_doppler_hz, code_phase_samples, cn0_db_hz;
    double code_freq_chips, rem_code_phase_chips, acc_carrier_phase_rad, carrier_lock_test;
    float rem_carr_phase_rad;
    int32_t correlation_length_ms;
    int32_t cloop, pull_in_transitory, current_symbol, lock_fail_counter;
    float corr[10];  // d_correlator_outs (VE,E,P,L,VL or E,P,L) as I,Q pairs
    float p_accu[2];
};

struct GbRefTrk
{
    dll_pll_veml_tracking_sptr blk;
    Gnss_Synchro sync;
    long long pos = 0;  // next unread item of the caller's array
    std::vector<long> events;
};

void* gbref_trk_block_create(const GbRefTrkConf* c)
{
    Dll_Pll_Conf p;
    p.fs_in = c->fs_in;
    p.vector_length = c->vector_length;
    p.system = c->system;
    std::memcpy(p.signal, c->signal, 3);
    p.pll_filter_order = c->pll_filter_order;
    p.dll_filter_order = c->dll_filter_order;
    p.fll_filter_order = c->fll_filter_order;
    p.enable_fll_pull_in = c->enable_fll_pull_in != 0;
    p.enable_fll_steady_state = c->enable_fll_steady_state != 0;
    p.pull_in_time_s = c->pull_in_time_s;
    p.fll_bw_hz = c->fll_bw_hz;
    p.pll_bw_hz = c->pll_bw_hz;
    p.dll_bw_hz = c->dll_bw_hz;
    p.pll_bw_narrow_hz = c->pll_bw_narrow_hz;
    p.dll_bw_narrow_hz = c->dll_bw_narrow_hz;
    p.early_late_space_chips = c->early_late_space_chips;
    p.very_early_late_space_chips = c->very_early_late_space_chips;
    p.early_late_space_narrow_chips = c->early_late_space_narrow_chips;
    p.very_early_late_space_narrow_chips = c->very_early_late_space_narrow_chips;
    p.extend_correlation_symbols = c->extend_correlation_symbols;
    p.high_dyn = c->high_dyn != 0;
    p.smoother_length = c->smoother_length;
    p.cn0_samples = c->cn0_samples;
    p.cn0_min = c->cn0_min;
    p.max_lock_fail = c->max_lock_fail;
    p.carrier_lock_th = c->carrier_lock_th;
    p.track_pilot = c->track_pilot != 0;
    p.dump = c->dump != 0;
    p.dump_mat = false;
    p.dump_filename = c->dump_filename;
    GbRefTrk* h = new GbRefTrk();
    h->blk = dll_pll_veml_make_tracking(p);
    h->sync = Gnss_Synchro();
    return h;
}

void gbref_trk_block_destroy(void* hv) { delete (GbRefTrk*)hv; }

// = what Channel / ChannelFsm do around a positive acquisition: set_gnss_synchro, then start_tracking()
void gbref_trk_block_start(void* hv, uint32_t prn, double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp, uint64_t first_item_index)
{
    GbRefTrk* h = (GbRefTrk*)hv;
    h->sync.System = h->blk->trk_parameters.system;
    std::memcpy(h->sync.Signal, h->blk->trk_parameters.signal, 3);
    h->sync.PRN = prn;
    h->sync.Acq_delay_samples = acq_delay_samples;
    h->sync.Acq_doppler_hz = acq_doppler_hz;
    h->sync.Acq_samplestamp_samples = acq_samplestamp;
    h->blk->set_channel(0);
    h->blk->set_gnss_synchro(&h->sync);
    h->blk->d_sample_counter = first_item_index;  // the block has been consuming the stream in standby up to here
    h->pos = 0;
    h->blk->start_tracking();
}

void gbref_trk_block_stop(void* hv) { ((GbRefTrk*)hv)->blk->stop_tracking(); }
void gbref_trk_block_telemetry_fault(void* hv) { ((GbRefTrk*)hv)->blk->gb_post("telemetry_to_trk", pmt::make_any(1)); }
int gbref_trk_block_state(void* hv) { return ((GbRefTrk*)hv)->blk->d_state; }
long long gbref_trk_block_position(void* hv) { return ((GbRefTrk*)hv)->pos; }

// Scheduler stand-in: calls general_work() while at least 2 * vector_length items are available (forecast(), :507-514) and space
// for more epochs remains; returns the number of epochs written.  `samples_iq`: the whole stream, item 0 = the item
// `first_item_index` given to gbref_trk_block_start.  events (3 = loss of lock) are appended to `events` (up to max_events).
int gbref_trk_block_run(void* hv, const float* samples_iq, long long n_items, GbRefTrkEpoch* out, int max_epochs, int32_t* events, int max_events, int* n_events)
{
    GbRefTrk* h = (GbRefTrk*)hv;
    dll_pll_veml_tracking* b = h->blk.get();
    int n_out = 0;
    if (n_events) *n_events = 0;
    const long long need = 2LL * (long long)b->trk_parameters.vector_length;
    while (n_out < max_epochs && n_items - h->pos >= need)
        {
            if (b->d_state == 0) break;  // standby: the block would only swallow the stream
            gr_vector_int ninput(1, (int)std::min<long long>(n_items - h->pos, 1LL << 30));
            gr_vector_const_void_star in(1, samples_iq + 2 * h->pos);
            Gnss_Synchro item = Gnss_Synchro();
            gr_vector_void_star outv(1, &item);
            b->gb_consumed = 0;
            const long long first = h->pos;
            const int produced = b->general_work(1, ninput, in, outv);
            h->pos += b->gb_consumed;
            if (produced == 1)
                {
                    GbRefTrkEpoch& e = out[n_out++];
                    std::memset(&e, 0, sizeof e);
                    e.tracking_sample_counter = item.Tracking_sample_counter;
                    e.first_sample = (uint64_t)first;
                    e.n_samples = (int32_t)b->gb_consumed;
                    e.state_after = b->d_state;
                    e.prompt_i = item.Prompt_I;
                    e.prompt_q = item.Prompt_Q;
                    e.carrier_phase_rads = item.Carrier_phase_rads;
                    e.carrier_doppler_hz = item.Carrier_Doppler_hz;
                    e.code_phase_samples = item.Code_phase_samples;
                    e.cn0_db_hz = item.CN0_dB_hz;
                    e.code_freq_chips = b->d_code_freq_chips;
                    e.rem_code_phase_chips = b->d_rem_code_phase_chips;
                    e.acc_carrier_phase_rad = b->d_acc_carrier_phase_rad;
                    e.carrier_lock_test = b->d_carrier_lock_test;
                    e.rem_carr_phase_rad = b->d_rem_carr_phase_rad;
                    e.correlation_length_ms = item.correlation_length_ms;
                    e.cloop = b->d_cloop;
                    e.pull_in_transitory = b->d_pull_in_transitory;
                    e.current_symbol = b->d_current_symbol;
                    e.lock_fail_counter = b->d_carrier_lock_fail_counter;
                    for (int t = 0; t < b->d_n_correlator_taps && t < 5; t++)
                        {
                            e.corr[2 * t] = b->d_correlator_outs[t].real();
                            e.corr[2 * t + 1] = b->d_correlator_outs[t].imag();
                        }
                    e.p_accu[0] = b->d_P_accu.real();
                    e.p_accu[1] = b->d_P_accu.imag();
                }
            for (long ev : b->gb_events)
                if (events && n_events && *n_events < max_events) events[(*n_events)++] = (int32_t)ev;
            b->gb_events.clear();
        }
    return n_out;
}

// ------------------------------------------------------------------------------------------------ acquisition block
struct GbRefAcqConf
{
    uint32_t sampled_ms, ms_per_code, samples_per_chip, max_dwells, doppler_max, doppler_step, num_doppler_bins_step2;
    float doppler_step2;
    int64_t fs_in;
    float samples_per_ms, samples_per_code;
    int bit_transition_flag, use_CFAR_algorithm_flag, make_2_steps, cshort;
    float threshold;
    char system;
    char signal[3];
    uint32_t prn;
    int use_fsm;  // give the block a ChannelFsm: positive acquisitions are then notified by a direct call instead of event 1
};

struct GbRefAcqState
{
    int32_t state, active, step_two, positive_acq;
    uint32_t dwell_counter, fft_size, consumed_samples, num_doppler_bins;
    float threshold, mag, input_power, test_statistics, doppler_center_step_two;
    uint64_t sample_counter;
    double acq_delay_samples, acq_doppler_hz;
    uint64_t acq_samplestamp_samples;
    uint32_t acq_doppler_step;
    int32_t fsm_valid_acquisitions;
    int64_t old_freq;
};

struct GbRefAcq
{
    pcps_acquisition_sptr blk;
    Gnss_Synchro sync;
    std::shared_ptr<ChannelFsm> fsm;
    std::vector<long> events;
};

void* gbref_acq_block_create(const GbRefAcqConf* c)
{
    Acq_Conf p;
    p.sampled_ms = c->sampled_ms;
    p.ms_per_code = c->ms_per_code;
    p.samples_per_chip = c->samples_per_chip;
    p.max_dwells = c->max_dwells;
    p.doppler_max = c->doppler_max;
    p.num_doppler_bins_step2 = c->num_doppler_bins_step2;
    p.doppler_step2 = c->doppler_step2;
    p.fs_in = c->fs_in;
    p.samples_per_ms = c->samples_per_ms;
    p.samples_per_code = c->samples_per_code;
    p.bit_transition_flag = c->bit_transition_flag != 0;
    p.use_CFAR_algorithm_flag = c->use_CFAR_algorithm_flag != 0;
    p.make_2_steps = c->make_2_steps != 0;
    p.blocking = true;              // acquisition_core runs inside general_work (no worker thread)
    p.blocking_on_standby = true;   // "enable it only for unit testing to avoid sample consume on idle status" (acq_conf.h)
    p.dump = false;
    p.it_size = c->cshort ? sizeof(lv_16sc_t) : sizeof(gr_complex);
    GbRefAcq* h = new GbRefAcq();
    h->blk = pcps_make_acquisition(p);
    h->sync = Gnss_Synchro();
    h->sync.System = c->system;
    std::memcpy(h->sync.Signal, c->signal, 3);
    h->sync.PRN = c->prn;
    // the order Channel's constructor uses (channel.cc:60-104)
    h->blk->set_channel(0);
    if (c->use_fsm)
        {
            h->fsm = std::make_shared<ChannelFsm>();
            h->blk->set_channel_fsm(h->fsm);
        }
    h->blk->set_gnss_synchro(&h->sync);
    h->blk->set_doppler_max(c->doppler_max);
    h->blk->set_doppler_step(c->doppler_step);
    h->blk->set_threshold(c->threshold);
    h->blk->init();
    return h;
}

void gbref_acq_block_destroy(void* hv) { delete (GbRefAcq*)hv; }
void gbref_acq_block_set_local_code(void* hv, float* code_iq) { ((GbRefAcq*)hv)->blk->set_local_code(reinterpret_cast<std::complex<float>*>(code_iq)); }
void gbref_acq_block_set_state(void* hv, int state) { ((GbRefAcq*)hv)->blk->set_state(state); }
void gbref_acq_block_set_active(void* hv, int active) { ((GbRefAcq*)hv)->blk->set_active(active != 0); }  // = reset() of the adapters
void gbref_acq_block_set_threshold(void* hv, float t) { ((GbRefAcq*)hv)->blk->set_threshold(t); }
void gbref_acq_block_set_sample_counter(void* hv, uint64_t v) { ((GbRefAcq*)hv)->blk->d_sample_counter = v; }
void gbref_set_accumulator_lanes(int lanes) { gb_volk_accumulator_lanes = lanes; }

// one general_work() call on `n_items` available items at `items` (gr_complex or lv_16sc_t); returns the items consumed;
// *events_out gets the last event published in the call (1 positive, 2 negative) or 0
int gbref_acq_block_work(void* hv, const void* items, int n_items, int* event_out)
{
    GbRefAcq* h = (GbRefAcq*)hv;
    gr_vector_int ninput(1, n_items);
    gr_vector_const_void_star in(1, items);
    gr_vector_void_star out;
    h->blk->gb_consumed = 0;
    h->blk->gb_events.clear();
    h->blk->general_work(0, ninput, in, out);
    if (event_out) *event_out = h->blk->gb_events.empty() ? 0 : (int)h->blk->gb_events.back();
    return (int)h->blk->gb_consumed;
}

void gbref_acq_block_get_state(void* hv, GbRefAcqState* s)
{
    GbRefAcq* h = (GbRefAcq*)hv;
    pcps_acquisition* b = h->blk.get();
    s->state = b->d_state;
    s->active = b->d_active;
    s->step_two = b->d_step_two;
    s->positive_acq = b->d_positive_acq;
    s->dwell_counter = b->d_num_noncoherent_integrations_counter;
    s->fft_size = b->d_fft_size;
    s->consumed_samples = b->d_consumed_samples;
    s->num_doppler_bins = b->d_num_doppler_bins;
    s->threshold = b->d_threshold;
    s->mag = b->d_mag;
    s->input_power = b->d_input_power;
    s->test_statistics = b->d_test_statistics;
    s->doppler_center_step_two = b->d_doppler_center_step_two;
    s->sample_counter = b->d_sample_counter;
    s->acq_delay_samples = h->sync.Acq_delay_samples;
    s->acq_doppler_hz = h->sync.Acq_doppler_hz;
    s->acq_samplestamp_samples = h->sync.Acq_samplestamp_samples;
    s->acq_doppler_step = h->sync.Acq_doppler_step;
    s->fsm_valid_acquisitions = h->fsm ? h->fsm->valid_acquisitions : -1;
    s->old_freq = b->d_old_freq;
}

// row `bin` of d_magnitude_grid (d_fft_size floats; the block zeroes it at the end of an attempt, :919-928)
const float* gbref_acq_block_grid_row(void* hv, uint32_t bin)
{
    pcps_acquisition* b = ((GbRefAcq*)hv)->blk.get();
    return bin < b->d_num_doppler_bins ? b->d_magnitude_grid[bin] : nullptr;
}

// ------------------------------------------------------------------------------------------------ dump reader
// The reference's own reader of the tracking blocks' .dat dumps (src/tests/unit-tests/signal-processing-blocks/libs/
// tracking_dump_reader.cc, pure std): every epoch's 21 fields as doubles, `max_epochs` rows at most; returns the epoch count
// the reader reports (num_epochs()), or -1 when it cannot open the file.
long long gbref_read_tracking_dump(const char* path, double* out, long long max_epochs)
{
    Tracking_Dump_Reader rd;
    if (!rd.open_obs_file(path)) return -1;
    const long long n = rd.num_epochs();
    long long k = 0;
    while (k < max_epochs && rd.read_binary_obs())
        {
            double* r = out + 21 * k++;
            r[0] = rd.abs_VE; r[1] = rd.abs_E; r[2] = rd.abs_P; r[3] = rd.abs_L; r[4] = rd.abs_VL;
            r[5] = rd.prompt_I; r[6] = rd.prompt_Q; r[7] = (double)rd.PRN_start_sample_count; r[8] = rd.acc_carrier_phase_rad;
            r[9] = rd.carrier_doppler_hz; r[10] = rd.carrier_doppler_rate_hz_s; r[11] = rd.code_freq_chips; r[12] = rd.code_freq_rate_chips;
            r[13] = rd.carr_error_hz; r[14] = rd.carr_error_filt_hz; r[15] = rd.code_error_chips; r[16] = rd.code_error_filt_chips;
            r[17] = rd.CN0_SNV_dB_Hz; r[18] = rd.carrier_lock_test; r[19] = rd.aux1; r[20] = rd.aux2;
        }
    return n;
}

}  // extern "C"
