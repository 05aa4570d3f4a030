// TEST INFRASTRUCTURE ONLY -- never linked into the product library.
//
// The reference's SIBLING blocks, compiled unmodified from /root/reference into oracle/_ref/libgnss_ref.so like
// the two main blocks (see ref_blocks.cc for the stand-in headers): pcps_tong_acquisition_cc, galileo_pcps_8ms_acquisition_cc,
// galileo_e5a_noncoherentIQ_acquisition_caf_cc, pcps_quicksync_acquisition_cc, pcps_acquisition_fine_doppler_cc,
// glonass_l1_ca_dll_pll_tracking_cc, gps_l1_ca_dll_pll_c_aid_tracking_cc / _sc.  This file only plays the scheduler and copies
// the blocks' state out after every general_work() call; it contains no signal processing.
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
#include <sstream>
#include <string>
#include <thread>
#include <utility>
#include <vector>
#include <gnuradio/block.h>
#include <gnuradio/fft/fft.h>
#include <volk/volk.h>
#include <volk_gnsssdr/volk_gnsssdr.h>
#include "acq_conf.h"
#include "channel_fsm.h"
#include "gnss_sdr_flags.h"
#include "gnss_synchro.h"

#define private public
#include "galileo_e5a_noncoherent_iq_acquisition_caf_cc.h"
#include "galileo_pcps_8ms_acquisition_cc.h"
#include "glonass_l1_ca_dll_pll_tracking_cc.h"
#include "gps_l1_ca_dll_pll_c_aid_tracking_cc.h"
#include "gps_l1_ca_dll_pll_c_aid_tracking_sc.h"
#include "gps_l1_ca_kf_tracking_cc.h"
#include "pcps_acquisition_fine_doppler_cc.h"
#include "pcps_quicksync_acquisition_cc.h"
#include "pcps_tong_acquisition_cc.h"
#undef private

// gflags stand-ins (shim/gnss_sdr_flags.h), gnss_sdr_flags.cc:49-65 defaults
int32_t FLAGS_cn0_samples = 20;
int32_t FLAGS_cn0_min = 25;
int32_t FLAGS_max_lock_fail = 50;
double FLAGS_carrier_lock_th = 0.85;
int32_t FLAGS_doppler_max = 0;
int32_t FLAGS_doppler_step = 0;
double FLAGS_dll_bw_hz = 0.0;
double FLAGS_pll_bw_hz = 0.0;

extern "C" {

void gbref_set_flags(int cn0_samples, int cn0_min, int max_lock_fail, double carrier_lock_th)
{
    FLAGS_cn0_samples = cn0_samples;
    FLAGS_cn0_min = cn0_min;
    FLAGS_max_lock_fail = max_lock_fail;
    FLAGS_carrier_lock_th = carrier_lock_th;
}

// ------------------------------------------------------------------------------------------------ acquisition variants
// one general_work() call
struct GbRefAcqvCall
{
    int32_t state_before, state_after, consumed, event;  // event: last value published on "events" in the call (0 none)
    float test_statistics, input_power, mag;
    uint32_t count, tong_count;  // dwell / well counter, Tong counter
    double acq_delay_samples, acq_doppler_hz;
    uint64_t acq_samplestamp_samples, sample_counter;
    uint32_t acq_doppler_step;
    int32_t active;
};

struct GbRefAcqv
{
    int variant = 0;  // 0 Tong, 1 Galileo 8 ms, 2 E5a non-coherent I/Q CAF, 3 QuickSync, 4 fine Doppler
    pcps_acquisition_fine_doppler_cc_sptr fd;
    pcps_tong_acquisition_cc_sptr tong;
    galileo_pcps_8ms_acquisition_cc_sptr g8;
    galileo_e5a_noncoherentIQ_acquisition_caf_cc_sptr e5a;
    pcps_quicksync_acquisition_cc_sptr qs;
    Gnss_Synchro sync;
    gr::block* blk = nullptr;
    uint32_t vector_items = 0;  // gr_complex items per input vector
};

// cfg: sampled_ms, max_dwells, doppler_max, doppler_step, fs_in, samples_per_ms, samples_per_code, threshold, tong_init_val, tong_max_val,
//      tong_max_dwells, bit_transition_flag, both_signal_components, caf_window_hz, zero_padding, folding_factor, prn
void* gbref_acqv_create(int variant, const double* cfg, char system, const char* signal)
{
    GbRefAcqv* h = new GbRefAcqv();
    h->variant = variant;
    const uint32_t sampled_ms = (uint32_t)cfg[0], max_dwells = (uint32_t)cfg[1], doppler_max = (uint32_t)cfg[2], doppler_step = (uint32_t)cfg[3];
    const int64_t fs_in = (int64_t)cfg[4];
    const int32_t samples_per_ms = (int32_t)cfg[5], samples_per_code = (int32_t)cfg[6];
    h->sync = Gnss_Synchro();
    h->sync.System = system;
    std::memcpy(h->sync.Signal, signal, 3);
    h->sync.PRN = (uint32_t)cfg[16];
#define GB_SETUP(B)                              \
    B->set_channel(0);                           \
    B->set_gnss_synchro(&h->sync);               \
    B->set_threshold((float)cfg[7]);             \
    B->set_doppler_max(doppler_max);             \
    B->set_doppler_step(doppler_step);           \
    h->blk = B.get();
    switch (variant)
        {
        case 0:
            h->tong = pcps_tong_make_acquisition_cc(sampled_ms, doppler_max, fs_in, samples_per_ms, samples_per_code, (uint32_t)cfg[8], (uint32_t)cfg[9], (uint32_t)cfg[10], false, "");
            GB_SETUP(h->tong)
            h->vector_items = sampled_ms * samples_per_ms;
            break;
        case 1:
            h->g8 = galileo_pcps_8ms_make_acquisition_cc(sampled_ms, max_dwells, doppler_max, fs_in, samples_per_ms, samples_per_code, false, "");
            GB_SETUP(h->g8)
            h->vector_items = sampled_ms * samples_per_ms;
            break;
        case 2:
            h->e5a = galileo_e5a_noncoherentIQ_make_acquisition_caf_cc(sampled_ms, max_dwells, doppler_max, fs_in, samples_per_ms, samples_per_code, cfg[11] != 0, false, "",
                cfg[12] != 0, (int)cfg[13], (int)cfg[14]);
            GB_SETUP(h->e5a)
            h->vector_items = 1;  // this block takes a plain gr_complex stream and buffers it itself
            break;
        case 4:
            {
                // the adapter's Acq_Conf (gps_l1_ca_pcps_acquisition_fine_doppler.cc:52-80): doppler_max, sampled_ms, max_dwells, samples_per_ms
                Acq_Conf ac = Acq_Conf();
                ac.fs_in = fs_in;
                ac.samples_per_chip = 1;
                ac.doppler_max = doppler_max;
                ac.sampled_ms = sampled_ms;
                ac.max_dwells = max_dwells;
                ac.samples_per_ms = (float)samples_per_ms;
                ac.dump = false;
                ac.blocking_on_standby = false;
                h->fd = pcps_make_acquisition_fine_doppler_cc(ac);
                h->fd->set_channel(0);
                h->fd->set_gnss_synchro(&h->sync);
                h->fd->set_threshold((float)cfg[7]);
                h->fd->set_doppler_max(doppler_max);
                h->fd->set_doppler_step(doppler_step);
                h->blk = h->fd.get();
                h->vector_items = 1;  // plain gr_complex stream
                break;
            }
        default:
            h->qs = pcps_quicksync_make_acquisition_cc((uint32_t)cfg[15], sampled_ms, max_dwells, doppler_max, fs_in, samples_per_ms, samples_per_code, cfg[11] != 0, false, "");
            GB_SETUP(h->qs)
            h->vector_items = sampled_ms * samples_per_ms;
            break;
        }
#undef GB_SETUP
    return h;
}

void gbref_acqv_destroy(void* hv) { delete (GbRefAcqv*)hv; }
uint32_t gbref_acqv_vector_items(void* hv) { return ((GbRefAcqv*)hv)->vector_items; }

void gbref_acqv_init(void* hv)
{
    GbRefAcqv* h = (GbRefAcqv*)hv;
    if (h->tong) h->tong->init();
    if (h->g8) h->g8->init();
    if (h->e5a) h->e5a->init();
    if (h->qs) h->qs->init();
    if (h->fd) h->fd->init();
}

void gbref_acqv_set_local_code(void* hv, float* code_iq, float* code_q_iq)
{
    GbRefAcqv* h = (GbRefAcqv*)hv;
    auto* c = reinterpret_cast<std::complex<float>*>(code_iq);
    if (h->tong) h->tong->set_local_code(c);
    if (h->g8) h->g8->set_local_code(c);
    if (h->e5a) h->e5a->set_local_code(c, reinterpret_cast<std::complex<float>*>(code_q_iq));
    if (h->qs) h->qs->set_local_code(c);
    if (h->fd) h->fd->set_local_code(c);
}

void gbref_acqv_set_state(void* hv, int state)
{
    GbRefAcqv* h = (GbRefAcqv*)hv;
    if (h->tong) h->tong->set_state(state);
    if (h->g8) h->g8->set_state(state);
    if (h->e5a) h->e5a->set_state(state);
    if (h->qs) h->qs->set_state(state);
    if (h->fd) h->fd->set_state(state);
}

void gbref_acqv_set_active(void* hv, int active)
{
    GbRefAcqv* h = (GbRefAcqv*)hv;
    if (h->tong) h->tong->set_active(active != 0);
    if (h->g8) h->g8->set_active(active != 0);
    if (h->e5a) h->e5a->set_active(active != 0);
    if (h->qs) h->qs->set_active(active != 0);
    if (h->fd) h->fd->set_active(active != 0);
}

// one general_work() on `n_in` input items (vectors for Tong / 8 ms / QuickSync, gr_complex items for the E5a block)
void gbref_acqv_work(void* hv, const float* items_iq, int n_in, GbRefAcqvCall* out)
{
    GbRefAcqv* h = (GbRefAcqv*)hv;
    std::memset(out, 0, sizeof *out);
#define GB_STATE(B, CNT, TONG) \
    out->state_before = B->d_state
    if (h->tong) out->state_before = h->tong->d_state;
    if (h->g8) out->state_before = h->g8->d_state;
    if (h->e5a) out->state_before = h->e5a->d_state;
    if (h->qs) out->state_before = h->qs->d_state;
    if (h->fd) out->state_before = h->fd->d_state;
#undef GB_STATE
    gr_vector_int ninput(1, n_in);
    gr_vector_const_void_star in(1, items_iq);
    gr_vector_void_star outv;
    h->blk->gb_consumed = 0;
    h->blk->gb_events.clear();
    h->blk->general_work(h->fd ? n_in : 0, ninput, in, outv);  // the fine-Doppler block reads its input count from noutput_items
    out->consumed = (int32_t)h->blk->gb_consumed;
    out->event = h->blk->gb_events.empty() ? 0 : (int32_t)h->blk->gb_events.back();
#define GB_COPY(B, CNT)                                \
    out->state_after = B->d_state;                     \
    out->test_statistics = B->d_test_statistics;       \
    out->input_power = B->d_input_power;               \
    out->mag = B->d_mag;                               \
    out->count = B->CNT;                               \
    out->sample_counter = B->d_sample_counter;         \
    out->active = B->d_active;
    if (h->tong)
        {
            GB_COPY(h->tong, d_dwell_count)
            out->tong_count = h->tong->d_tong_count;
        }
    if (h->g8) { GB_COPY(h->g8, d_well_count) }
    if (h->e5a) { GB_COPY(h->e5a, d_well_count) }
    if (h->qs) { GB_COPY(h->qs, d_well_count) }
    if (h->fd)
        {
            out->state_after = h->fd->d_state;
            out->test_statistics = h->fd->d_test_statistics;
            out->count = (uint32_t)h->fd->d_well_count;
            out->sample_counter = h->fd->d_sample_counter;
            out->active = h->fd->d_active;
        }
#undef GB_COPY
    out->acq_delay_samples = h->sync.Acq_delay_samples;
    out->acq_doppler_hz = h->sync.Acq_doppler_hz;
    out->acq_samplestamp_samples = h->sync.Acq_samplestamp_samples;
    out->acq_doppler_step = h->sync.Acq_doppler_step;
}

// the E5a block's CAF vector (d_CAF_vector, one float per Doppler bin) after the last dwell
int gbref_acqv_caf_vector(void* hv, float* dst, int cap)
{
    GbRefAcqv* h = (GbRefAcqv*)hv;
    if (!h->e5a || !h->e5a->d_CAF_vector) return 0;
    const int n = std::min<int>(cap, (int)h->e5a->d_num_doppler_bins);
    for (int i = 0; i < n; i++) dst[i] = h->e5a->d_CAF_vector[i];
    return n;
}

// ------------------------------------------------------------------------------------------------ sibling trackers
// kind 0: glonass_l1_ca_dll_pll_tracking_cc (gr_complex), 1: gps_l1_ca_dll_pll_c_aid_tracking_cc (gr_complex),
// 2: gps_l1_ca_dll_pll_c_aid_tracking_sc (lv_16sc_t), 3: gps_l1_ca_kf_tracking_cc (gr_complex; Armadillo stand-in, see shim/armadillo)
struct GbRefTrk2
{
    int kind = 0;
    glonass_l1_ca_dll_pll_tracking_cc_sptr glo;
    gps_l1_ca_dll_pll_c_aid_tracking_cc_sptr cc;
    gps_l1_ca_dll_pll_c_aid_tracking_sc_sptr sc;
    gps_l1_ca_kf_tracking_cc_sptr kf;
    gr::block* blk = nullptr;
    Gnss_Synchro sync;
    long long pos = 0;
    uint32_t vector_length = 0;
};

// conf: fs_in, vector_length, pll_bw_hz, dll_bw_hz, pll_bw_narrow_hz, dll_bw_narrow_hz, early_late_space_chips, extend_correlation_ms,
//       (kind 3 only) order, bce_run, bce_ptrans, bce_strans, bce_nu, bce_kappa
void* gbref_trk2_create(int kind, const double* c)
{
    GbRefTrk2* h = new GbRefTrk2();
    h->kind = kind;
    h->vector_length = (uint32_t)c[1];
    if (kind == 0)
        {
            h->glo = glonass_l1_ca_dll_pll_make_tracking_cc((int64_t)c[0], (uint32_t)c[1], false, "", (float)c[2], (float)c[3], (float)c[6]);
            h->blk = h->glo.get();
        }
    else if (kind == 1)
        {
            h->cc = gps_l1_ca_dll_pll_c_aid_make_tracking_cc((int64_t)c[0], (uint32_t)c[1], false, "", (float)c[2], (float)c[3], (float)c[4], (float)c[5], (int32_t)c[7], (float)c[6]);
            h->blk = h->cc.get();
        }
    else if (kind == 3)
        {
            h->kf = gps_l1_ca_kf_make_tracking_cc((uint32_t)c[8], 0, (int64_t)c[0], (uint32_t)c[1], false, "", (float)c[3], (float)c[6], c[9] != 0.0, (uint32_t)c[10],
                (uint32_t)c[11], (int32_t)c[12], (int32_t)c[13]);
            h->blk = h->kf.get();
        }
    else
        {
            h->sc = gps_l1_ca_dll_pll_c_aid_make_tracking_sc((int64_t)c[0], (uint32_t)c[1], false, "", (float)c[2], (float)c[3], (float)c[4], (float)c[5], (int32_t)c[7], (float)c[6]);
            h->blk = h->sc.get();
        }
    h->sync = Gnss_Synchro();
    return h;
}
void gbref_trk2_destroy(void* hv) { delete (GbRefTrk2*)hv; }

void gbref_trk2_start(void* hv, uint32_t prn, double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp, uint64_t first_item_index)
{
    GbRefTrk2* h = (GbRefTrk2*)hv;
    h->sync.System = h->kind == 0 ? 'R' : 'G';
    std::memcpy(h->sync.Signal, h->kind == 0 ? "1G" : "1C", 3);
    h->sync.PRN = prn;
    h->sync.Acq_delay_samples = acq_delay_samples;
    h->sync.Acq_doppler_hz = acq_doppler_hz;
    h->sync.Acq_samplestamp_samples = acq_samplestamp;
    h->pos = 0;
#define GB_START(B)                          \
    B->set_channel(0);                       \
    B->set_gnss_synchro(&h->sync);           \
    B->d_sample_counter = first_item_index;  \
    B->start_tracking();
    if (h->glo) { GB_START(h->glo) }
    if (h->cc) { GB_START(h->cc) }
    if (h->sc) { GB_START(h->sc) }
    if (h->kf) { GB_START(h->kf) }
#undef GB_START
}

// Acq_doppler_step of the acquisition result (the Kalman tracker sizes its initial Doppler variance from it); before gbref_trk2_start
void gbref_trk2_set_doppler_step(void* hv, uint32_t step) { ((GbRefTrk2*)hv)->sync.Acq_doppler_step = step; }

// the "preamble_timestamp_s" message of the telemetry decoder (c_aid trackers: starts the extended integration)
void gbref_trk2_preamble(void* hv, double preamble_timestamp_s)
{
    GbRefTrk2* h = (GbRefTrk2*)hv;
    h->blk->gb_post("preamble_timestamp_s", pmt::from_double(preamble_timestamp_s));
}

// n_calls general_work() calls (or until fewer than 2 * vector_length items remain); one row of 13 doubles per call:
// tracking_sample_counter, prompt_i, prompt_q, carrier_doppler_hz, code_phase_samples, carrier_phase_rads, cn0_db_hz, abs_e, abs_p, abs_l,
// flags (1 valid symbol, 2 loss of lock event), items consumed, correlation_length_ms
int gbref_trk2_run(void* hv, const void* samples, long long n_items, int n_calls, double* rows)
{
    GbRefTrk2* h = (GbRefTrk2*)hv;
    const size_t isz = h->kind == 2 ? sizeof(lv_16sc_t) : sizeof(gr_complex);
    int k = 0;
    for (; k < n_calls && n_items - h->pos >= 2LL * h->vector_length; k++)
        {
            gr_vector_int ninput(1, (int)std::min<long long>(n_items - h->pos, 1LL << 30));
            gr_vector_const_void_star in(1, static_cast<const char*>(samples) + isz * h->pos);
            Gnss_Synchro item = Gnss_Synchro();
            gr_vector_void_star outv(1, &item);
            h->blk->gb_consumed = 0;
            h->blk->gb_events.clear();
            const int produced = h->blk->general_work(1, ninput, in, outv);
            h->pos += h->blk->gb_consumed;
            double* r = rows + 13 * k;
            std::fill(r, r + 13, 0.0);
            if (produced == 1)
                {
                    r[0] = (double)item.Tracking_sample_counter;
                    r[1] = item.Prompt_I;
                    r[2] = item.Prompt_Q;
                    r[3] = item.Carrier_Doppler_hz;
                    r[4] = item.Code_phase_samples;
                    r[5] = item.Carrier_phase_rads;
                    r[6] = item.CN0_dB_hz;
                    r[10] = item.Flag_valid_symbol_output ? 1.0 : 0.0;
                    r[12] = item.correlation_length_ms;
                }
            if (h->glo)
                for (int t = 0; t < 3; t++) r[7 + t] = std::abs(h->glo->d_correlator_outs[t]);
            if (h->cc)
                for (int t = 0; t < 3; t++) r[7 + t] = std::abs(h->cc->d_correlator_outs[t]);
            if (h->kf)
                for (int t = 0; t < 3; t++) r[7 + t] = std::abs(h->kf->d_correlator_outs[t]);
            if (h->sc)
                for (int t = 0; t < 3; t++) r[7 + t] = std::abs(std::complex<float>(h->sc->d_correlator_outs_16sc[t].real(), h->sc->d_correlator_outs_16sc[t].imag()));
            for (long ev : h->blk->gb_events)
                if (ev == 3) r[10] += 2.0;
            r[11] = (double)h->blk->gb_consumed;
        }
    return k;
}

}  // extern "C"
