/* TEST INFRASTRUCTURE ONLY -- CPU oracle for the B200 correlator engine.
 *
 * Plain-C restatement of the reference's (zhufengGNSS/gnss-sdr-1) PCPS acquisition and
 * multicorrelator arithmetic.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this; the product library
 * (libgnss_b200.so) never links or calls it.
 *
 * Parity status: PINNED.  Tracking functions are checked bit-for-bit against the
 * reference's own _generic protokernels compiled from /root/reference (oracle/_ref,
 * tests/test_oracle_vs_ref.py, tests/golden/trk_*.npz); acquisition is checked against
 * the reference's two in-tree acquisition fixtures and their expected answers
 * (gps_l1_ca_pcps_acquisition_test.cc:285-286, galileo_e1_pcps_ambiguous_acquisition_test.cc:293-294).
 * The FFT behind acquisition is FFTW3f in the reference (un-vendored system dependency,
 * version not pinned in-tree); here it is a mixed-radix DFT written from the published
 * definition, so acquisition magnitudes agree with the reference to float32 rounding only.
 */
#ifndef GNSS_ORACLE_H
#define GNSS_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ---------------- tracking: volk_gnsssdr _generic restatements ---------------- */
void go_resampler_real(float* out /*T x n*/, const float* code, float rem_chips, float step_chips,
    const float* shifts, int code_len, int n_taps, int n);
void go_resampler_real_hd(float* out /*T x n*/, const float* code, float rem_chips, float step_chips,
    float rate_chips, const float* shifts, int code_len, int n_taps, int n);
void go_resampler_cplx(float* out_iq /*T x n x 2*/, const float* code_iq, float rem_chips, float step_chips,
    const float* shifts, int code_len, int n_taps, int n);
void go_rotator_dot_prod_real(float* result_iq, const float* in_iq, float inc_re, float inc_im,
    float* phase_iq, const float* codes /*T x n*/, int n_taps, int n);
void go_rotator_dot_prod_real_hd(float* result_iq, const float* in_iq, float inc_re, float inc_im,
    float rate_re, float rate_im, float* phase_iq, const float* codes, int n_taps, int n);
void go_rotator_dot_prod_cplx(float* result_iq, const float* in_iq, float inc_re, float inc_im,
    float* phase_iq, const float* codes_iq /*T x n x 2*/, int n_taps, int n);
/* Cpu_Multicorrelator_Real_Codes::Carrier_wipeoff_multicorrelator_resampler */
void go_multicorrelator_real(const float* in_iq, int n, const float* code, int code_len, const float* shifts,
    int n_taps, float rem_carr_phase_rad, float phase_step_rad, float phase_rate_step_rad,
    float rem_code_phase_chips, float code_phase_step_chips, float code_phase_rate_step_chips,
    int high_dyn, float* corr_out_iq);
/* Cpu_Multicorrelator::Carrier_wipeoff_multicorrelator_resampler (complex codes) */
void go_multicorrelator_cplx(const float* in_iq, int n, const float* code_iq, int code_len, const float* shifts,
    int n_taps, float rem_carr_phase_rad, float phase_step_rad, float rem_code_phase_chips,
    float code_phase_step_chips, float* corr_out_iq);
/* Cpu_Multicorrelator_16sc::Carrier_wipeoff_multicorrelator_resampler (int16 samples, complex int16 codes, saturating) */
void go_multicorrelator_16sc(const int16_t* in_iq, int n, const int16_t* code_iq, int code_len, const float* shifts, int n_taps,
    float rem_carr_phase_rad, float phase_step_rad, float rem_code_phase_chips, float code_phase_step_chips, int16_t* corr_out_iq);
/* float64 arbiter on the oracle's own chip indices */
void go_multicorrelator_real_f64(const float* in_iq, int n, const float* code, int code_len, const float* shifts,
    int n_taps, float rem_carr_phase_rad, float phase_step_rad, float rem_code_phase_chips,
    float code_phase_step_chips, double* corr_out_iq);

/* ---------------- acquisition helpers ---------------- */
void go_sincos(float* out_iq, float phase_inc, float phase0, int n);
uint32_t go_index_max(const float* src, uint32_t n);
void go_16ic_convert_32fc(float* out_iq, const int16_t* in_iq, uint32_t n);
/* un-normalised DFT, sign=-1 forward / +1 backward, any n (radix 2,3,4,5 + generic prime) */
void go_fft_f32(float* data_iq, int n, int sign);
void go_fft_f64(double* data_iq, int n, int sign);

/* ---------------- PRN codes ---------------- */
void go_gps_l1_ca_code_int(int32_t* dest, int32_t prn, uint32_t chip_shift);
void go_gps_l1_ca_code_float(float* dest, int32_t prn, uint32_t chip_shift);
void go_gps_l1_ca_code_sampled(float* dest_iq, uint32_t prn, int32_t fs, uint32_t chip_shift);
/* gnss_signal_processing.cc resampler(): nearest-neighbour code resampling used by E1/E5a/L5 set_local_code */
void go_code_resampler_f32(const float* from, float* dest, float fs_in, float fs_out, uint32_t len_in, uint32_t len_out);

/* ---------------- pcps_acquisition block ---------------- */
typedef struct go_acq_conf
{
    /* Acq_Conf (acq_conf.h:39-68) fields used by the arithmetic */
    uint32_t sampled_ms;
    uint32_t ms_per_code;
    uint32_t samples_per_chip;
    uint32_t max_dwells;
    uint32_t doppler_max;
    uint32_t doppler_step;
    uint32_t num_doppler_bins_step2;
    float doppler_step2;
    int64_t fs_in;
    float samples_per_ms;
    float samples_per_code;
    int bit_transition_flag;
    int use_CFAR_algorithm_flag;
    int make_2_steps;
    int cshort; /* it_size == sizeof(lv_16sc_t) */
    float threshold;
    int64_t freq_offset_hz; /* d_old_freq (GLONASS FDMA), pcps_acquisition.cc:277-293 */
    int fft_f64;            /* arbiter: run the DFTs in float64 */
    int analytic_wipeoff;   /* arbiter: exact phase instead of float32-accumulated table */
    int keep_grid;          /* tests: snapshot the magnitude grid before the end-of-attempt reset (:919-928) */
    int accumulator_lanes;  /* volk_32f_accumulator_s32f (upstream VOLK, not in the tree) behind d_input_power: 0 / 1 = its _generic
                             * protokernel (one sequential float32 sum), 8 = its a_avx / u_avx protokernels (eight interleaved
                             * partial sums added in lane order, then the tail) -- what VOLK dispatches on an x86 host */
} go_acq_conf;

typedef struct go_acq_result
{
    uint32_t code_phase_index; /* indext */
    int32_t doppler_hz;
    uint32_t doppler_bin;
    float peak_mag;
    float input_power;
    float test_statistic;
    int decision; /* 1 positive, 0 negative, 2 need more dwells, 3 step two armed */
    double acq_delay_samples;
    double acq_doppler_hz;
    uint64_t acq_samplestamp;
} go_acq_result;

typedef struct go_acq go_acq;
go_acq* go_acq_create(const go_acq_conf* conf);
void go_acq_destroy(go_acq* a);
uint32_t go_acq_fft_size(const go_acq* a);
uint32_t go_acq_consumed_samples(const go_acq* a);
uint32_t go_acq_num_doppler_bins(const go_acq* a);
void go_acq_set_local_code(go_acq* a, const float* code_iq);
/* samples: d_consumed_samples items (float IQ pairs, or int16 IQ pairs when conf.cshort) */
int go_acq_run(go_acq* a, const void* samples, uint64_t samp_count, go_acq_result* out);
/* kept magnitude grid, bins x effective_fft_size */
const float* go_acq_grid(const go_acq* a, int bin);

/* ---------------- sibling PCPS blocks: pcps_tong_acquisition_cc, galileo_pcps_8ms_acquisition_cc ---------------- */
#define GO_ACQV_TONG 0
#define GO_ACQV_8MS 1
typedef struct go_acqv_conf
{
    int variant;
    uint32_t fft_size;          /* d_sampled_ms * d_samples_per_ms */
    uint32_t samples_per_code;
    uint32_t doppler_max, doppler_step;
    int64_t fs_in;
    float threshold;
    uint32_t max_dwells;                                  /* 8 ms */
    uint32_t tong_init_val, tong_max_val, tong_max_dwells; /* Tong */
} go_acqv_conf;
typedef struct go_acqv_result
{
    uint32_t code_phase_index;
    int32_t doppler_hz;
    uint32_t doppler_bin;
    float mag;            /* d_mag: 8 ms: peak / N^4; Tong: accumulated test statistic of the best cell */
    float input_power;
    float test_statistic;
    int decision;         /* 1 positive, 0 negative, 2 keep dwelling */
    int32_t tong_count;
    uint32_t dwell_count;
    double acq_delay_samples;
} go_acqv_result;
typedef struct go_acqv go_acqv;
go_acqv* go_acqv_create(const go_acqv_conf* conf);
void go_acqv_destroy(go_acqv* a);
uint32_t go_acqv_num_doppler_bins(const go_acqv* a);
float go_acqv_phase_step(const go_acqv* a, uint32_t bin);
void go_acqv_set_local_code(go_acqv* a, const float* code_iq);
void go_acqv_restart(go_acqv* a);
int go_acqv_run(go_acqv* a, const float* in_iq, go_acqv_result* out);

/* ---------------- galileo_e5a_noncoherent_iq_acquisition_caf_cc (pilot + data components, A/B code pairs, CAF filter) ----------------
 * Block-level parity is UNPINNED: the reference holds no fixture for this block and it needs GNU Radio to build; the
 * arithmetic underneath (wipe-off table, FFT correlation, |.|^2, first-maximum index) is the pinned go_* code above. */
typedef struct go_e5a_conf
{
    uint32_t sampled_ms;        /* constructor argument; d_fft_size = sampled_ms * samples_per_ms */
    uint32_t samples_per_ms, samples_per_code;
    uint32_t max_dwells, doppler_max, doppler_step;
    int64_t fs_in;
    int bit_transition_flag, both_signal_components, caf_window_hz, zero_padding;
    float threshold;
} go_e5a_conf;
typedef struct go_e5a_result
{
    uint32_t code_phase_index;  /* indext of the winning bin (before % samples_per_code) */
    uint32_t doppler_bin;       /* bin of the winning cell (before the CAF stage) */
    int32_t doppler_hz;         /* Acq_doppler_hz after the CAF stage */
    float mag, input_power, test_statistic;
    int decision;               /* 1 positive, 0 negative, 2 next dwell */
    uint32_t well_count;
    double acq_delay_samples;
    int updated;                /* the dwell rewrote the Gnss_Synchro fields */
} go_e5a_result;
typedef struct go_e5a go_e5a;
go_e5a* go_e5a_create(const go_e5a_conf* conf);
void go_e5a_destroy(go_e5a* a);
uint32_t go_e5a_num_doppler_bins(const go_e5a* a);
float go_e5a_phase_step(const go_e5a* a, uint32_t bin);
void go_e5a_set_local_code(go_e5a* a, const float* code_i_iq, const float* code_q_iq);
void go_e5a_restart(go_e5a* a);   /* state 0 -> 1 */
int go_e5a_run(go_e5a* a, const float* in_iq, go_e5a_result* out);
void go_e5a_caf_filter(const float* caf_i, const float* caf_q, float* caf_out, uint32_t num_bins, int caf_window_hz, uint32_t doppler_step, int both);
const float* go_e5a_caf_vector(const go_e5a* a);  /* d_CAF_vector after the last dwell, num_bins floats (caf_window_hz > 0) */

/* ---------------- acquisition resampler of GNSSFlowgraph (gnss_flowgraph.cc:423-476) ----------------
 * gr::filter::firdes::low_pass(.., WIN_HAMMING) and gr::filter::fir_filter_ccf(decimation, taps) belong to GNU Radio
 * (>= 3.7.3 per CMakeLists.txt:398), which is NOT in the reference tree: their published algorithms are restated here and
 * parity for this row is UNPINNED (no reference output to check against). */
int go_firdes_low_pass(double gain, double sampling_freq, double cutoff_freq, double transition_width, float* taps, int capacity);
/* y[k] = sum_j taps[j] x[k D - j], x[n < 0] = 0; k = 0 .. (n_in - 1) / D; float32 accumulation in tap order. Returns the count. */
int go_fir_decimate(const float* in_iq, int n_in, const float* taps, int n_taps, int decimation, float* out_iq);
/* integer decimation the flowgraph / adapters pick (gnss_flowgraph.cc:431-437, gps_l1_ca_pcps_acquisition.cc:97-104) */
uint32_t go_acq_resampler_decimation(int64_t fs_in, uint32_t opt_acq_fs);

/* ---------------- pcps_acquisition_fine_doppler_cc (grid accumulated over dwells, first/second-peak ratio, zero-padded FFT Doppler refinement) ----------------
 * Block-level parity UNPINNED (no reference fixture; the block needs GNU Radio); the arithmetic underneath is the pinned go_* code. */
typedef struct go_fd_conf
{
    uint32_t samples_per_ms; /* d_fft_size */
    uint32_t max_dwells, doppler_max, doppler_step;
    int64_t fs_in;
    float threshold;
} go_fd_conf;
typedef struct go_fd_result
{
    uint32_t index_time, index_doppler;
    float first_peak, second_peak, test_statistic;
    int positive;             /* d_test_statistics > d_threshold */
    double acq_delay_samples, acq_doppler_hz; /* after estimate_Doppler when positive */
    uint32_t fine_index;      /* tmp_index_freq of the zero-padded FFT (positive only) */
    int fine_accepted;        /* |fftFreqBins[idx] - coarse| < 1000 */
} go_fd_result;
typedef struct go_fd go_fd;
go_fd* go_fd_create(const go_fd_conf* conf);
void go_fd_destroy(go_fd* a);
int go_fd_num_doppler_points(const go_fd* a);
float go_fd_phase_step(const go_fd* a, int bin);
void go_fd_set_local_code(go_fd* a, const float* code_iq);
/* one whole attempt: `in_iq` holds max(max_dwells, 10) * samples_per_ms items from the attempt's first sample; code_1ms_iq is
 * what gps_l1_ca_code_gen_complex_sampled(.., PRN, fs, 0) returns (the block regenerates it inside estimate_Doppler) */
int go_fd_attempt(go_fd* a, const float* in_iq, const float* code_1ms_iq, go_fd_result* out);
/* the replica estimate_Doppler builds (:412-428): 10 * samples_per_ms items, incl. its off-by-one std::rotate */
void go_fd_replica(const float* code_1ms_iq, uint32_t fft_size, int shift_index, float* replica_iq);

/* ---- pcps_quicksync_acquisition_cc: folded FFT search + time-domain disambiguation ---- */
typedef struct go_qs_conf
{
    uint32_t folding_factor, sampled_ms, max_dwells, doppler_max, doppler_step;
    int64_t fs_in;
    int32_t samples_per_ms, samples_per_code;
    int bit_transition_flag;
    float threshold;
} go_qs_conf;
typedef struct go_qs_result
{
    uint32_t well_count, folded_index, doppler_bin, best_candidate; /* of the last improvement in this dwell (0 if none) */
    int improved;                /* the dwell rewrote the Gnss_Synchro fields */
    float mag, input_power, test_statistic;
    int state;                   /* 1 more dwells, 2 positive, 3 negative (d_state after the dwell) */
    double acq_delay_samples, acq_doppler_hz;
    uint64_t acq_samplestamp;
    float corr_output[16];       /* d_corr_output_f of the last improvement */
} go_qs_result;
typedef struct go_qs go_qs;
go_qs* go_qs_create(const go_qs_conf* conf);
void go_qs_destroy(go_qs* a);
uint32_t go_qs_fft_size(const go_qs* a);
uint32_t go_qs_num_doppler_bins(const go_qs* a);
float go_qs_phase_step(const go_qs* a, uint32_t bin);
/* code_iq: samples_per_code items (the block copies that many, :181); folded_iq (may be NULL) receives the fft_size folded items */
void go_qs_set_local_code(go_qs* a, const float* code_iq, float* folded_iq);
void go_qs_restart(go_qs* a); /* set_state(1) / state 0 with d_active */
/* one pass through state 1: in_iq = one input vector (sampled_ms * samples_per_ms items, of which samples_per_code *
 * folding_factor are used); sample_counter_after = d_sample_counter after the += at :349 */
int go_qs_run(go_qs* a, const float* in_iq, uint64_t sample_counter_after, go_qs_result* out);

#ifdef __cplusplus
}
#endif
#endif
