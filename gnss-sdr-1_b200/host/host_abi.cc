// extern "C" doors of the host-side mirror (libgnss_b200_host.so, plain g++): lets the tests drive the C++ adapter
// classes -- loop maths, the DLL/PLL epoch loop on the B200 multicorrelator, the acquisition block mirror -- from Python.
#include <cstring>
#include <memory>

#include "gb200_acquisition.hpp"
#include "gb200_acquisition_variants.hpp"
#include "gb200_c_aid_tracking.hpp"
#include "gb200_glonass_tracking.hpp"
#include "gb200_kf_tracking.hpp"
#include "gb200_quicksync.hpp"
#include "gb200_host_flat.h"
#include "gb200_multicorrelator.hpp"
#include "gb200_tracking.hpp"

using namespace gb200;

extern "C" {
// ---- loop maths (arguments: complex values as re, im floats) ----
double gbh_pll_cloop_two_quadrant_atan(float re, float im) { return pll_cloop_two_quadrant_atan(cfloat(re, im)); }
double gbh_pll_four_quadrant_atan(float re, float im) { return pll_four_quadrant_atan(cfloat(re, im)); }
double gbh_fll_four_quadrant_atan(float re1, float im1, float re2, float im2, double t1, double t2)
{
    return fll_four_quadrant_atan(cfloat(re1, im1), cfloat(re2, im2), t1, t2);
}
double gbh_dll_nc_e_minus_l_normalized(float ere, float eim, float lre, float lim) { return dll_nc_e_minus_l_normalized(cfloat(ere, eim), cfloat(lre, lim)); }
double gbh_dll_nc_vemlp_normalized(const float* ve, const float* e, const float* l, const float* vl)
{
    return dll_nc_vemlp_normalized(cfloat(ve[0], ve[1]), cfloat(e[0], e[1]), cfloat(l[0], l[1]), cfloat(vl[0], vl[1]));
}
float gbh_cn0_svn_estimator(const float* prompt_iq, int n, double t) { return cn0_svn_estimator(reinterpret_cast<const cfloat*>(prompt_iq), n, t); }
float gbh_carrier_lock_detector(const float* prompt_iq, int n) { return carrier_lock_detector(reinterpret_cast<const cfloat*>(prompt_iq), n); }
void gbh_loop_filter_run(float update_interval, float bw, int order, int last_integrator, float init_out, const float* in, float* out, int n)
{
    LoopFilter f(update_interval, bw, order, last_integrator != 0);
    f.initialize(init_out);
    for (int i = 0; i < n; i++) out[i] = f.apply(in[i]);
}
void gbh_fll_pll_filter_run(float fll_bw, float pll_bw, int order, float acq_doppler_hz, const float* fll, const float* pll, float t, float* out, int n)
{
    FllPllFilter f;
    f.set_params(fll_bw, pll_bw, order);
    f.initialize(acq_doppler_hz);
    for (int i = 0; i < n; i++) out[i] = f.get_carrier_error(fll[i], pll[i], t);
}

void gbh_2nd_filter_run(int which, float bw_hz, float pdi_s, const float* disc, float* out, int n)
{
    if (which == 0)
        {
            Tracking2ndDllFilter f(pdi_s);
            f.set_DLL_BW(bw_hz);
            f.initialize();
            for (int i = 0; i < n; i++) out[i] = f.get_code_nco(disc[i]);
        }
    else
        {
            Tracking2ndPllFilter f(pdi_s);
            f.set_PLL_BW(bw_hz);
            f.initialize();
            for (int i = 0; i < n; i++) out[i] = f.get_carrier_nco(disc[i]);
        }
}

// ---- DLL/PLL tracking channel on the B200 multicorrelator ----
struct gbh_trk
{
    B200_Multicorrelator_Real_Codes corr;
    std::unique_ptr<DllPllVemlTracking<B200_Multicorrelator_Real_Codes>> loop;
};

gbh_trk* gbh_trk_open(gb200_service* service, int stream_id, int item_type, const gbh_trk_conf* conf)
{
    if (!service || !conf) return nullptr;
    auto* h = new gbh_trk();
    const DllPllConf d = gbh_to_conf(conf);
    h->corr.set_high_dynamics_resampler(d.high_dyn);  // dll_pll_veml_tracking.cc:413
    if (!h->corr.init(service, stream_id, item_type, 2 * d.vector_length, d.veml ? 5 : 3))
        {
            delete h;
            return nullptr;
        }
    h->loop.reset(new DllPllVemlTracking<B200_Multicorrelator_Real_Codes>(d, &h->corr));
    return h;
}
void gbh_trk_set_channel(gbh_trk* h, unsigned channel)
{
    if (h) h->loop->set_channel(channel);
}
void gbh_trk_telemetry(gbh_trk* h, int tlm_event)
{
    if (h) h->loop->telemetry_to_trk(tlm_event);
}
int gbh_trk_start(gbh_trk* h, const float* code, const float* data_code, double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp,
    uint64_t sample_counter_now)
{
    return (h && h->loop->start_tracking(code, data_code, acq_delay_samples, acq_doppler_hz, acq_samplestamp, sample_counter_now)) ? 0 : -1;
}
int gbh_trk_step(gbh_trk* h, gbh_epoch* out)
{
    if (!h || !out) return -1;
    TrackingEpoch e;
    const bool ok = h->loop->step(&e);
    gbh_fill_epoch(e, h->loop->state(), out);
    return ok ? 0 : -2;
}
void gbh_trk_close(gbh_trk* h) { delete h; }

// the MAT-file writer on its own (the CPU tests read the file back with scipy.io.loadmat)
int gbh_matfile_write_grid(const char* filename, const float* grid, size_t rows, size_t cols, uint32_t prn, uint64_t sample_counter, int32_t positive)
{
    MatFileWriter m(filename);
    if (!m.ok()) return -1;
    bool ok = m.write("acq_grid", MatFileWriter::kSingle, grid, rows, cols);
    ok = ok && m.write_scalar<uint32_t>("PRN", MatFileWriter::kUint32, prn);
    ok = ok && m.write_scalar<uint64_t>("sample_counter", MatFileWriter::kUint64, sample_counter);
    ok = ok && m.write_scalar<int32_t>("d_positive_acq", MatFileWriter::kInt32, positive);
    return ok ? 0 : -2;
}

float gbh_pcps_threshold_from_pfa(float pfa, unsigned doppler_max, unsigned doppler_step, unsigned vector_length)
{
    return pcps_threshold_from_pfa(pfa, doppler_max, doppler_step, vector_length);
}

// ---- acquisition block mirror: set_state(1) + run(samp_count); returns the event (1 positive, 2 negative, 0 none, <0 error) ----
int gbh_acq_search_dump(gb200_service* service, const gb200_acq_cfg* conf, int stream_id, const float* code_iq, size_t code_items, uint64_t samp_count,
    int max_runs, double* delay_samples, double* doppler_hz, uint64_t* samplestamp, float* test_statistic, int* mag, const char* dump_filename,
    unsigned channel, char system, const char* signal, unsigned prn, char* dump_file_out, size_t dump_file_cap);

int gbh_acq_search(gb200_service* service, const gb200_acq_cfg* conf, int stream_id, const float* code_iq, size_t code_items, uint64_t samp_count,
    int max_runs, double* delay_samples, double* doppler_hz, uint64_t* samplestamp, float* test_statistic, int* mag)
{
    return gbh_acq_search_dump(service, conf, stream_id, code_iq, code_items, samp_count, max_runs, delay_samples, doppler_hz, samplestamp, test_statistic,
        mag, nullptr, 0, 'G', "1C", 0, nullptr, 0);
}

// the same search with dump=true on `channel` (pcps_acquisition.cc:160-194, 462-562); the written file's name comes back in dump_file_out
int gbh_acq_search_dump(gb200_service* service, const gb200_acq_cfg* conf, int stream_id, const float* code_iq, size_t code_items, uint64_t samp_count,
    int max_runs, double* delay_samples, double* doppler_hz, uint64_t* samplestamp, float* test_statistic, int* mag, const char* dump_filename,
    unsigned channel, char system, const char* signal, unsigned prn, char* dump_file_out, size_t dump_file_cap)
{
    B200_Pcps_Acquisition acq(service, *conf, stream_id);
    if (!acq.ok()) return -1;
    AcqSynchro syn;
    syn.System = system;
    syn.Signal[0] = signal ? signal[0] : '1';
    syn.Signal[1] = signal ? signal[1] : 'C';
    syn.PRN = prn;
    acq.set_channel(channel);
    if (dump_filename && dump_filename[0]) acq.set_dump(dump_filename, channel);
    acq.set_gnss_synchro(&syn);
    acq.set_doppler_step(conf->doppler_step);
    acq.set_threshold(conf->threshold);
    acq.init();
    if (!acq.set_local_code(reinterpret_cast<const std::complex<float>*>(code_iq), code_items)) return -1;
    acq.reset();
    int ev = 0;
    uint64_t at = samp_count;
    for (int k = 0; k < max_runs && ev == 0; k++)
        {
            gb200_acq_result r{};
            ev = acq.run(at, &r);
            if (ev == 0 && r.decision == 2) at += acq.consumed_samples();  // next dwell consumes the next block; step two re-runs the same one
        }
    *delay_samples = syn.Acq_delay_samples;
    *doppler_hz = syn.Acq_doppler_hz;
    *samplestamp = syn.Acq_samplestamp_samples;
    *test_statistic = acq.test_statistics();
    *mag = acq.mag();
    if (dump_file_out && dump_file_cap > 0)
        {
            std::strncpy(dump_file_out, acq.last_dump_file().c_str(), dump_file_cap - 1);
            dump_file_out[dump_file_cap - 1] = 0;
        }
    return ev;
}

// ---- acquisition resampler (GNSSFlowgraph + the adapters' Acq_Conf fields) ----
unsigned gbh_acq_resampler_decimation(long long fs_in, unsigned opt_acq_fs) { return acq_resampler_decimation(fs_in, opt_acq_fs); }
int gbh_attach_acq_resampler(gb200_service* service, int src_stream, int dst_stream, long long fs_in, unsigned decimation)
{
    return attach_acq_resampler(service, src_stream, dst_stream, fs_in, decimation);
}
// the search of gbh_acq_search on the resampled stream `stream_id` (conf describes that stream), results scaled back as the
// block does with use_automatic_resampler (pcps_acquisition.cc:756-762)
int gbh_acq_search_resampled(gb200_service* service, const gb200_acq_cfg* conf, int stream_id, const float* code_iq, size_t code_items, uint64_t samp_count,
    int max_runs, float resampler_ratio, unsigned latency_samples, double* delay_samples, double* doppler_hz, uint64_t* samplestamp, float* test_statistic)
{
    B200_Pcps_Acquisition acq(service, *conf, stream_id);
    if (!acq.ok()) return -1;
    AcqSynchro syn;
    acq.set_gnss_synchro(&syn);
    acq.set_automatic_resampler(resampler_ratio);
    acq.set_resampler_latency(latency_samples);
    acq.set_doppler_step(conf->doppler_step);
    acq.set_threshold(conf->threshold);
    acq.init();
    if (!acq.set_local_code(reinterpret_cast<const std::complex<float>*>(code_iq), code_items)) return -1;
    acq.reset();
    int ev = 0;
    uint64_t at = samp_count;
    for (int k = 0; k < max_runs && ev == 0; k++)
        {
            gb200_acq_result r{};
            ev = acq.run(at, &r);
            if (ev == 0 && r.decision == 2) at += acq.consumed_samples();
        }
    *delay_samples = syn.Acq_delay_samples;
    *doppler_hz = syn.Acq_doppler_hz;
    *samplestamp = syn.Acq_samplestamp_samples;
    *test_statistic = acq.test_statistics();
    return ev;
}

// ---- sibling PCPS blocks (gb200_acquisition_variants.hpp): a whole attempt, one record of 8 doubles per dwell ----
// cfg = {sampled_ms, doppler_max, doppler_step, fs_in, samples_per_ms, samples_per_code, threshold, max_dwells, tong_init_val,
//        tong_max_val, tong_max_dwells, item_type}; variant 0 = Tong, 1 = Galileo E1 8 ms.
// record = {event, code_phase_index, Acq_doppler_hz, test_statistic, input_power, tong_count, dwell_count, Acq_delay_samples}
int gbh_acqv_attempt(gb200_service* service, int variant, const double* cfg, int stream_id, const float* code_iq, size_t code_items, uint64_t samp_count,
    int max_runs, double* records, int* n_records)
{
    AcqVariantConf c;
    c.sampled_ms = (uint32_t)cfg[0];
    c.doppler_max = (uint32_t)cfg[1];
    c.doppler_step = (uint32_t)cfg[2];
    c.fs_in = (int64_t)cfg[3];
    c.samples_per_ms = (int)cfg[4];
    c.samples_per_code = (int)cfg[5];
    c.threshold = (float)cfg[6];
    c.max_dwells = (uint32_t)cfg[7];
    c.tong_init_val = (uint32_t)cfg[8];
    c.tong_max_val = (uint32_t)cfg[9];
    c.tong_max_dwells = (uint32_t)cfg[10];
    c.item_type = (int)cfg[11];
    AcqSynchro syn;
    int ev = 0, n = 0;
    uint64_t at = samp_count;
    auto drive = [&](auto& acq) -> int {
        if (!acq.ok()) return -1;
        acq.set_gnss_synchro(&syn);
        if (!acq.set_local_code(reinterpret_cast<const std::complex<float>*>(code_iq), code_items)) return -1;
        acq.reset();
        for (int k = 0; k < max_runs && ev == 0; k++)
            {
                ev = acq.run(at);
                if (ev < 0) return ev;
                double* r = records + 8 * n++;
                r[0] = ev;
                r[1] = acq.code_phase_index();
                r[2] = syn.Acq_doppler_hz;
                r[3] = acq.test_statistics();
                r[4] = acq.input_power();
                r[7] = syn.Acq_delay_samples;
                at += acq.fft_size();
            }
        return ev;
    };
    if (variant == 0)
        {
            B200_Pcps_Tong_Acquisition acq(service, c, stream_id);
            ev = drive(acq);
            // the counters of the last dwell (the per-dwell values are implied by the event sequence)
            if (n > 0)
                {
                    records[8 * (n - 1) + 5] = acq.tong_count();
                    records[8 * (n - 1) + 6] = acq.dwell_count();
                }
        }
    else
        {
            B200_Galileo_Pcps_8ms_Acquisition acq(service, c, stream_id);
            ev = drive(acq);
        }
    *n_records = n;
    return ev;
}

// ---- pcps_quicksync_acquisition_cc mirror (gb200_quicksync.hpp): a whole attempt, one record of 10 doubles per dwell ----
// cfg = {folding_factor, sampled_ms, max_dwells, doppler_max, doppler_step, fs_in, samples_per_ms, samples_per_code, bit_transition_flag, threshold}
// record = {event, folded_index, Acq_doppler_hz, test_statistic, input_power, mag, well_count, Acq_delay_samples, Acq_samplestamp, best |corr|^2}
int gbh_qs_attempt(gb200_service* service, const double* cfg, int stream_id, const float* code_iq, size_t code_items, uint64_t samp_count, int max_runs,
    double* records, int* n_records)
{
    QuickSyncConf c;
    c.folding_factor = (uint32_t)cfg[0];
    c.sampled_ms = (uint32_t)cfg[1];
    c.max_dwells = (uint32_t)cfg[2];
    c.doppler_max = (uint32_t)cfg[3];
    c.doppler_step = (uint32_t)cfg[4];
    c.fs_in = (int64_t)cfg[5];
    c.samples_per_ms = (int32_t)cfg[6];
    c.samples_per_code = (int32_t)cfg[7];
    c.bit_transition_flag = cfg[8] != 0.0;
    c.threshold = (float)cfg[9];
    B200_Pcps_QuickSync_Acquisition acq(service, c, stream_id);
    *n_records = 0;
    if (!acq.ok()) return -1;
    AcqSynchro syn;
    acq.set_gnss_synchro(&syn);
    if (!acq.set_local_code(reinterpret_cast<const std::complex<float>*>(code_iq), code_items)) return -1;
    acq.reset();
    int ev = 0, n = 0;
    uint64_t at = samp_count;
    for (int k = 0; k < max_runs && ev == 0; k++)
        {
            ev = acq.run(at);
            if (ev < 0) return ev;
            double* r = records + 10 * n++;
            r[0] = ev;
            r[1] = acq.folded_index();
            r[2] = syn.Acq_doppler_hz;
            r[3] = acq.test_statistics();
            r[4] = acq.input_power();
            r[5] = acq.mag();
            r[6] = acq.well_count();
            r[7] = syn.Acq_delay_samples;
            r[8] = (double)syn.Acq_samplestamp_samples;
            float best = 0.0f;
            for (float v : acq.corr_output()) best = v > best ? v : best;
            r[9] = best;
            at += acq.vector_items();
        }
    *n_records = n;
    return ev;
}

// ---- pcps_acquisition_fine_doppler_cc mirror: one attempt ----
// cfg = {samples_per_ms, max_dwells, doppler_max, doppler_step, fs_in, threshold, item_type}
// out = {event, index_time, index_doppler, first_peak, second_peak, test_statistic, Acq_delay_samples, Acq_doppler_hz, fine_index, fine_accepted}
int gbh_fd_attempt(gb200_service* service, const double* cfg, int stream_id, const float* code_iq, size_t code_items, uint64_t samp_count, double* out)
{
    AcqVariantConf c;
    c.samples_per_ms = (int)cfg[0];
    c.samples_per_code = (int)cfg[0];
    c.max_dwells = (uint32_t)cfg[1];
    c.doppler_max = (uint32_t)cfg[2];
    c.doppler_step = (uint32_t)cfg[3];
    c.fs_in = (int64_t)cfg[4];
    c.threshold = (float)cfg[5];
    c.item_type = (int)cfg[6];
    B200_Pcps_Acquisition_Fine_Doppler acq(service, c, stream_id);
    if (!acq.ok()) return -1;
    AcqSynchro syn;
    acq.set_gnss_synchro(&syn);
    if (!acq.set_local_code(reinterpret_cast<const std::complex<float>*>(code_iq), code_items)) return -1;
    acq.reset();
    const int ev = acq.run(samp_count);
    if (ev < 0) return ev;
    out[0] = ev;
    out[1] = acq.index_time();
    out[2] = acq.index_doppler();
    out[3] = acq.first_peak();
    out[4] = acq.second_peak();
    out[5] = acq.test_statistics();
    out[6] = syn.Acq_delay_samples;
    out[7] = syn.Acq_doppler_hz;
    out[8] = acq.fine_index();
    out[9] = acq.fine_accepted() ? 1.0 : 0.0;
    return ev;
}

// ---- Galileo E5a non-coherent I/Q acquisition with CAF filter (gb200_acquisition_variants.hpp) ----
// cfg = {sampled_ms, doppler_max, doppler_step, fs_in, samples_per_ms, samples_per_code, threshold, max_dwells, bit_transition_flag,
//        both_signal_components, caf_window_hz, zero_padding, item_type}
// record (10 doubles per dwell) = {event, code_phase_index, Acq_doppler_hz, doppler_bin, test_statistic, input_power, mag, well_count,
//        Acq_delay_samples, updated}; caf_out (may be null) receives the CAF vector of the last dwell.
int gbh_e5a_attempt(gb200_service* service, const double* cfg, int stream_id, const float* code_i_iq, const float* code_q_iq, size_t code_items,
    uint64_t samp_count, int max_runs, double* records, int* n_records, float* caf_out)
{
    AcqVariantConf c;
    c.sampled_ms = (uint32_t)cfg[0];
    c.doppler_max = (uint32_t)cfg[1];
    c.doppler_step = (uint32_t)cfg[2];
    c.fs_in = (int64_t)cfg[3];
    c.samples_per_ms = (int)cfg[4];
    c.samples_per_code = (int)cfg[5];
    c.threshold = (float)cfg[6];
    c.max_dwells = (uint32_t)cfg[7];
    c.bit_transition_flag = cfg[8] != 0;
    c.both_signal_components = cfg[9] != 0;
    c.caf_window_hz = (int)cfg[10];
    c.zero_padding = (int)cfg[11];
    c.item_type = (int)cfg[12];
    B200_Galileo_E5a_NoncoherentIQ_Acquisition_Caf acq(service, c, stream_id);
    if (!acq.ok()) return -1;
    AcqSynchro syn;
    acq.set_gnss_synchro(&syn);
    if (!acq.set_local_code(reinterpret_cast<const std::complex<float>*>(code_i_iq), reinterpret_cast<const std::complex<float>*>(code_q_iq), code_items)) return -1;
    acq.reset();
    int ev = 0, n = 0;
    uint64_t at = samp_count;
    for (int k = 0; k < max_runs && ev == 0; k++)
        {
            ev = acq.run(at);
            if (ev < 0) return ev;
            double* r = records + 10 * n++;
            r[0] = ev;
            r[1] = acq.code_phase_index();
            r[2] = syn.Acq_doppler_hz;
            r[3] = acq.doppler_bin();
            r[4] = acq.test_statistics();
            r[5] = acq.input_power();
            r[6] = acq.mag();
            r[7] = acq.well_count();
            r[8] = syn.Acq_delay_samples;
            r[9] = acq.updated() ? 1.0 : 0.0;
            at += acq.fft_size();
        }
    if (caf_out)
        for (size_t i = 0; i < acq.caf_vector().size(); i++) caf_out[i] = acq.caf_vector()[i];
    *n_records = n;
    return ev;
}

// the CAF filter of the E5a block alone (host arithmetic, no device): caf_q may be null
void gbh_caf_filter(const float* caf_i, const float* caf_q, float* caf_out, unsigned int num_bins, int caf_window_hz, unsigned int doppler_step)
{
    gb200::detail::caf_filter(caf_i, caf_q, caf_out, num_bins, caf_window_hz, doppler_step);
}

// ---- GLONASS L1 C/A DLL/PLL loop (gb200_glonass_tracking.hpp) on the B200 complex-code correlator ----
// conf = {fs_in, vector_length, pll_bw_hz, dll_bw_hz, early_late_space_chips, band_freq_hz, band_dfreq_hz}
// epoch = {tracking_sample_counter, prompt_i, prompt_q, carrier_doppler_hz, code_phase_samples, carrier_phase_rads, cn0_db_hz,
//          abs_e, abs_p, abs_l, valid_symbol + 2 * loss_of_lock, prn_length_samples}
struct gbh_glo
{
    B200_Multicorrelator corr;
    GlonassL1CaDllPllTracking<B200_Multicorrelator>* loop = nullptr;
};
void* gbh_glo_open(gb200_service* service, int stream_id, int item_type, const double* conf)
{
    GlonassTrkConf c;
    gbh_glo_conf(conf, &c);
    auto* h = new gbh_glo();
    if (!h->corr.init(service, stream_id, item_type, 2 * (int)c.vector_length, 3))
        {
            delete h;
            return nullptr;
        }
    h->loop = new GlonassL1CaDllPllTracking<B200_Multicorrelator>(c, &h->corr);
    return h;
}
int gbh_glo_start(void* hv, const float* code_iq, int freq_channel, double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp, uint64_t now)
{
    auto* h = static_cast<gbh_glo*>(hv);
    h->loop->set_sample_counter(now);
    return h->loop->start_tracking(reinterpret_cast<const std::complex<float>*>(code_iq), freq_channel, acq_delay_samples, acq_doppler_hz, acq_samplestamp) ? 0 : -1;
}
int gbh_glo_step(void* hv, double* epoch)
{
    auto* h = static_cast<gbh_glo*>(hv);
    GlonassEpoch e;
    const bool ok = h->loop->step(&e);
    gbh_glo_fill(e, epoch);
    return ok ? 0 : -2;
}
void gbh_glo_close(void* hv)
{
    auto* h = static_cast<gbh_glo*>(hv);
    if (!h) return;
    delete h->loop;
    delete h;
}

// ---- carrier-aided GPS L1 C/A trackers (gb200_c_aid_tracking.hpp): variant 0 = _cc on B200_Multicorrelator (32fc stream),
//      variant 1 = _sc on B200_Multicorrelator_16sc (16ic stream) ----
struct gbh_caid
{
    int variant = 0;
    B200_Multicorrelator corr_cc;
    B200_Multicorrelator_16sc corr_sc;
    GpsL1CaDllPllCAidTracking<B200_Multicorrelator>* loop_cc = nullptr;
    GpsL1CaDllPllCAidTracking<B200_Multicorrelator_16sc>* loop_sc = nullptr;
};
void* gbh_caid_open(gb200_service* service, int stream_id, int variant, const double* conf)
{
    CAidTrkConf c;
    gbh_caid_conf(conf, &c);
    auto* h = new gbh_caid();
    h->variant = variant;
    const bool ok = variant == 1 ? h->corr_sc.init(service, stream_id, 2 * (int)c.vector_length, 3)
                                 : h->corr_cc.init(service, stream_id, GB200_ITEM_32FC, 2 * (int)c.vector_length, 3);
    if (!ok)
        {
            delete h;
            return nullptr;
        }
    if (variant == 1)
        h->loop_sc = new GpsL1CaDllPllCAidTracking<B200_Multicorrelator_16sc>(c, &h->corr_sc);
    else
        h->loop_cc = new GpsL1CaDllPllCAidTracking<B200_Multicorrelator>(c, &h->corr_cc);
    return h;
}
int gbh_caid_start(void* hv, const float* code_iq, double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp, uint64_t now)
{
    auto* h = static_cast<gbh_caid*>(hv);
    const auto* code = reinterpret_cast<const std::complex<float>*>(code_iq);
    if (h->variant == 1)
        {
            h->loop_sc->set_sample_counter(now);
            return h->loop_sc->start_tracking(code, acq_delay_samples, acq_doppler_hz, acq_samplestamp) ? 0 : -1;
        }
    h->loop_cc->set_sample_counter(now);
    return h->loop_cc->start_tracking(code, acq_delay_samples, acq_doppler_hz, acq_samplestamp) ? 0 : -1;
}
void gbh_caid_set_preamble(void* hv, double preamble_timestamp_s)
{
    auto* h = static_cast<gbh_caid*>(hv);
    if (h->variant == 1)
        h->loop_sc->set_preamble_timestamp(preamble_timestamp_s);
    else
        h->loop_cc->set_preamble_timestamp(preamble_timestamp_s);
}
int gbh_caid_step(void* hv, double* epoch)
{
    auto* h = static_cast<gbh_caid*>(hv);
    CAidEpoch e;
    const bool ok = h->variant == 1 ? h->loop_sc->step(&e) : h->loop_cc->step(&e);
    gbh_caid_fill(e, epoch);
    return ok ? 0 : -2;
}
void gbh_caid_close(void* hv)
{
    auto* h = static_cast<gbh_caid*>(hv);
    if (!h) return;
    delete h->loop_cc;
    delete h->loop_sc;
    delete h;
}

// ---- Kalman GPS L1 C/A tracker (gb200_kf_tracking.hpp) on the real-code B200 correlator, any item type of the stream ----
struct gbh_kf
{
    B200_Multicorrelator_Real_Codes corr;
    GpsL1CaKfTracking<B200_Multicorrelator_Real_Codes>* loop = nullptr;
};
void* gbh_kf_open(gb200_service* service, int stream_id, int item_type, const double* conf)
{
    KfTrkConf c;
    gbh_kf_conf(conf, &c);
    auto* h = new gbh_kf();
    h->corr.set_high_dynamics_resampler(true);  // the reference correlator's constructor default, which the block keeps
    if (!h->corr.init(service, stream_id, item_type, 2 * (int)c.vector_length, 3))
        {
            delete h;
            return nullptr;
        }
    h->loop = new GpsL1CaKfTracking<B200_Multicorrelator_Real_Codes>(c, &h->corr);
    return h;
}
int gbh_kf_start(void* hv, const float* code, double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp, uint32_t acq_doppler_step, uint64_t now)
{
    auto* h = static_cast<gbh_kf*>(hv);
    h->loop->set_sample_counter(now);
    return h->loop->start_tracking(code, acq_delay_samples, acq_doppler_hz, acq_samplestamp, acq_doppler_step) ? 0 : -1;
}
int gbh_kf_step(void* hv, double* epoch)
{
    auto* h = static_cast<gbh_kf*>(hv);
    KfEpoch e;
    const bool ok = h->loop->step(&e);
    gbh_kf_fill(e, epoch);
    return ok ? 0 : -2;
}
void gbh_kf_close(void* hv)
{
    auto* h = static_cast<gbh_kf*>(hv);
    if (!h) return;
    delete h->loop;
    delete h;
}
}
