/* libgnss_b200.so -- C ABI of the B200-native GNSS correlator engine.
 *
 * This is the drop-in boundary for GNSS-SDR's PCPS acquisition and DLL/PLL(-VEML) tracking
 * multicorrelator.  Plain C, opaque handles, int status returns (0 = ok, <0 = GB200_E_*; the
 * library never calls exit() and never throws across this boundary), caller-owned host buffers.
 * Every entry point cites the reference interface it replaces (paths relative to the
 * zhufengGNSS/gnss-sdr-1 tree).  There is NO CPU fallback: every compute entry point fails with
 * GB200_E_NO_DEVICE when no sm_100 device is present.
 */
#ifndef GNSS_B200_H
#define GNSS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
/* the library is built with -fvisibility=hidden: only the entry points below are exported */
#if defined(__GNUC__)
#define GB200_API __attribute__((visibility("default")))
#else
#define GB200_API
#endif

#define GB200_OK 0
#define GB200_E_INVALID (-1)   /* bad argument / handle */
#define GB200_E_NO_DEVICE (-2) /* no CUDA device, or not sm_100 */
#define GB200_E_CUDA (-3)      /* CUDA runtime error, see gb200_last_error() */
#define GB200_E_RANGE (-4)     /* requested samples are not (or no longer) in the device ring */
#define GB200_E_STATE (-5)     /* call order violated (e.g. run before set_local_code) */
#define GB200_E_UNSUPPORTED (-6)

/* sample item types: gr_complex, lv_16sc_t (interleaved int16 I,Q), lv_8sc_t (interleaved int8 I,Q) */
#define GB200_ITEM_32FC 0
#define GB200_ITEM_16IC 1
#define GB200_ITEM_8IC 2

typedef struct gb200_service gb200_service;
typedef struct gb200_acq gb200_acq;
typedef struct gb200_trk gb200_trk;
typedef struct gb200_trk_batch gb200_trk_batch;

/* ------------------------------------------------------------------ service ------------- */
/* One per process and GPU.  Replaces the per-channel device setup of
 * cuda_multicorrelator::init_cuda_integrated_resampler (src/algorithms/tracking/libs/cuda_multicorrelator.cu:137-241). */
typedef struct gb200_service_cfg
{
    int device;                 /* CUDA ordinal; -1 = current device */
    int max_streams;            /* number of sample streams (front-end bands); 0 -> 4 */
    uint64_t ring_capacity;     /* items kept per stream on the device; 0 -> 1<<24 */
    uint64_t ring_slack;        /* longest window a single acquisition / correlation may read; 0 -> 1<<21 */
    uint32_t batch_window_us;   /* > 0: concurrent gb200_trk_correlate callers (one GNU Radio thread per channel) are coalesced into ONE
                                 * launch; the first caller waits at most this long for the other live channels.  0: every call launches alone */
    uint32_t reserved_;
} gb200_service_cfg;

GB200_API int gb200_service_create(const gb200_service_cfg* cfg, gb200_service** out);
GB200_API int gb200_service_destroy(gb200_service* s);
GB200_API const char* gb200_last_error(const gb200_service* s); /* NULL-safe; never NULL */
GB200_API const char* gb200_version(void);
/* the cudaStream_t every launch of this service is issued on (for CUDA-event timing by the caller) */
GB200_API void* gb200_service_cuda_stream(gb200_service* s);
GB200_API int gb200_service_synchronize(gb200_service* s);
/* number of kernels launched by this library since creation (bench "gpu_launches") */
GB200_API uint64_t gb200_service_launch_count(const gb200_service* s);
/* CUDA-event stopwatch on the service stream: start records an event, stop records another, waits for it
 * and returns the device time between the two in milliseconds (what a profiler-free bench must use). */
GB200_API int gb200_service_timer_start(gb200_service* s);
GB200_API int gb200_service_timer_stop(gb200_service* s, float* elapsed_ms);
/* Optional per-kernel-class device timing (CUDA events around each launch of that class), for roofline reports.
 * kernel_timing(s, 1) zeroes and enables the accumulators; kernel_stats returns total ms and launch count. */
#define GB200_KCLASS_TRK 0         /* batched multicorrelator */
#define GB200_KCLASS_ACQ_FWD_COL 1 /* wipe-off + column DFT + twiddle */
#define GB200_KCLASS_ACQ_FWD_ROW 2 /* row DFT -> spectra */
#define GB200_KCLASS_ACQ_INV_ROW 3 /* x conj(code DFT) + row inverse DFT + twiddle */
#define GB200_KCLASS_ACQ_INV_COL 4 /* column inverse DFT + |.|^2 + arg-max */
#define GB200_KCLASS_COUNT 5
GB200_API int gb200_service_kernel_timing(gb200_service* s, int enable);
GB200_API int gb200_service_kernel_stats(gb200_service* s, int kernel_class, double* total_ms, uint64_t* launches);
/* FP32 FMA throughput of this device measured with a register-resident FMA kernel (TFLOP/s, 2 flops per FMA):
 * the roofline denominator of the multicorrelator, which is FP32-bound, not HBM- or tensor-bound */
GB200_API int gb200_service_measure_fp32(gb200_service* s, double* tflops);
/* write `bytes` of device memory on the service stream (L2 flush between timed iterations) */
GB200_API int gb200_service_flush_l2(gb200_service* s, size_t bytes);
/* how many coalesced launches served how many gb200_trk_correlate calls so far (batch_window_us > 0) */
GB200_API int gb200_service_coalescing_stats(gb200_service* s, uint64_t* batches, uint64_t* requests);
/* the same for gb200_acq_run: with batch_window_us > 0 concurrent callers (each Channel's acquisition thread,
 * pcps_acquisition.cc:1020-1043; what GNSSFlowgraph::priorize_satellites / ChannelFsm start together, gnss_flowgraph.cc:1494-1546)
 * that name the same stream and first sample are run as ONE gb200_acq_run_batch, results identical to lone calls */
GB200_API int gb200_service_acq_coalescing_stats(gb200_service* s, uint64_t* batches, uint64_t* requests);

/* Tuning switches for profiling and experiments (every default is the measured best); no counterpart in the reference.
 * Keys: "acq_lanes" (1..4 streams the inverse chunks alternate on), "acq_l2_mb" (row-buffer budget), "acq_chunk_t" (transforms
 * per chunk), "acq_fused" (one persistent launch for the inverse side), "acq_fused_lag", "acq_fused_ring", "acq_fwd_groups"
 * (1..8: forward transforms in bin groups on their own stream, bin-major inverse chunks), "host_threads", "trk_chunk_max",
 * "trace"; and, only before the first acquisition handle exists (they shape the cached FFT plans), "acq_split_n1",
 * "acq_variant" (alternative radix orders of a split), "acq_generic", "acq_fwd_generic" (run-time parametrised kernels on both
 * sides / on the forward side only).  Unknown key -> GB200_E_INVALID. */
GB200_API int gb200_service_set_tuning(gb200_service* s, const char* key, int value);

/* Hand a block of the conditioned sample stream to the device ONCE for all channels.  Replaces each
 * block's own read of input_items[0] (pcps_acquisition.cc:995-1034 state 1 buffering;
 * dll_pll_veml_tracking.cc:1544 general_work input).  `samples` is host memory (pinned or pageable).
 * Items stay packed in their wire type on the device (int8/int16 are converted in-register by the
 * kernels exactly as volk_gnsssdr_16ic_convert_32fc.h:63-72 does).
 * Buffer lifetime: the copy is asynchronous on the service's ingest stream.  A pageable `samples` buffer has been staged by the
 * time the call returns and may be reused at once; a PINNED (cudaHostAlloc / cudaHostRegister) buffer is read by the DMA engine
 * after the call returns and must stay untouched until gb200_stream_wait(s, stream_id) or gb200_service_synchronize(s) returns
 * (a GNU Radio block that hands over its input buffer calls gb200_stream_wait before consume_each()). */
GB200_API int gb200_stream_push(gb200_service* s, int stream_id, const void* samples, size_t n_items, int item_type,
    uint64_t first_sample_index);
/* The same for SEVERAL producers of one stream: in a GNU Radio flowgraph every channel's acquisition and tracking block is fed the
 * same conditioned stream (channel.cc:127-186 connects them all to the conditioner's output) and each hands over the items its
 * general_work() was given; items already in the ring are skipped, only the part beyond its newest item is copied.  A block that
 * lags more than the ring capacity behind finds its items gone (the later correlate / run call reports GB200_E_RANGE). */
GB200_API int gb200_stream_push_shared(gb200_service* s, int stream_id, const void* samples, size_t n_items, int item_type,
    uint64_t first_sample_index);
/* blocks until every push issued so far to `stream_id` has landed in the device ring (the caller's buffers are free again) */
GB200_API int gb200_stream_wait(gb200_service* s, int stream_id);
/* `inverted_spectrum` of the data-type adapters (ishort_to_complex.cc:49-59, ibyte_to_complex.cc:49-59: a conjugate_cc /
 * conjugate_sc / conjugate_ic block after the conversion): items pushed to this stream from now on are conjugated on the device
 * in their wire type as part of the push. */
GB200_API int gb200_stream_set_inverted_spectrum(gb200_service* s, int stream_id, int inverted);
/* same, `samples` already in device memory (device-to-device copy on the service stream) */
GB200_API int gb200_stream_push_device(gb200_service* s, int stream_id, const void* d_samples, size_t n_items, int item_type,
    uint64_t first_sample_index);

/* Acquisition resampler of GNSSFlowgraph (src/core/receiver/gnss_flowgraph.cc:423-476, GNSS-SDR.use_acquisition_resampler):
 * a gr::filter::fir_filter_ccf(decimation, taps) between the signal conditioner and the channel's acquisition.  Here it is a
 * property of the source stream: every later push to `src_stream` (GB200_ITEM_32FC only, as in the adapters, e.g.
 * gps_l1_ca_pcps_acquisition.cc:90-94) also appends y[k] = sum_j taps[j] * x[k * decimation - j] (x before the stream's
 * first item is 0, the block's zero history) to `dst_stream` for every k whose newest input has arrived; item k of
 * `dst_stream` is aligned with item k * decimation of `src_stream`.  The filter runs on the device right behind the copy.
 * decimation == 0 removes it.  A discontinuity of the source restarts the filter history and the destination stream. */
GB200_API int gb200_stream_set_decimator(gb200_service* s, int src_stream, int dst_stream, uint32_t decimation, const float* taps, uint32_t n_taps);
/* The taps the flowgraph asks GNU Radio for: gr::filter::firdes::low_pass(gain, sampling_freq, cutoff_freq, transition_width,
 * WIN_HAMMING) -- GNU Radio (>= 3.7.3, CMakeLists.txt:398) is not part of the reference tree; this restates its published
 * windowed-sinc design (odd tap count (int)(53 * fs / (22 * transition_width)) | 1, Hamming window, DC gain `gain`).
 * Host arithmetic only.  `*n_taps`: in = capacity of `taps`, out = number of taps (GB200_E_RANGE if it does not fit). */
GB200_API int gb200_firdes_low_pass(double gain, double sampling_freq, double cutoff_freq, double transition_width, float* taps, uint32_t* n_taps);
/* copies items [first_sample_index, first_sample_index + n_items) of a stream back to the host in their wire type
 * (what a file_sink after the conditioner / resampler would record); GB200_E_RANGE if the ring no longer holds them */
GB200_API int gb200_stream_read(gb200_service* s, int stream_id, uint64_t first_sample_index, size_t n_items, void* dst);

/* ------------------------------------------------------------------ acquisition --------- */
/* = Acq_Conf (src/algorithms/acquisition/libs/acq_conf.h:39-68) + the setters' values */
typedef struct gb200_acq_cfg
{
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
    int item_type; /* GB200_ITEM_* of the stream it will read */
    float threshold;
    int64_t freq_offset_hz; /* GLONASS FDMA d_old_freq, pcps_acquisition.cc:277-293 */
} gb200_acq_cfg;

typedef struct gb200_acq_result
{
    uint32_t code_phase_index; /* indext: arg-max index in the kept row */
    int32_t doppler_hz;
    uint32_t doppler_bin;
    float peak_mag;       /* |.|^2 grid maximum (un-normalised, as d_magnitude_grid) */
    float input_power;    /* d_input_power */
    float test_statistic; /* d_test_statistics */
    int decision;         /* 1 positive (event 1), 0 negative (event 2), 2 need more dwells, 3 step two armed */
    double acq_delay_samples;  /* Gnss_Synchro::Acq_delay_samples */
    double acq_doppler_hz;     /* Gnss_Synchro::Acq_doppler_hz */
    uint64_t acq_samplestamp;  /* Gnss_Synchro::Acq_samplestamp_samples = first_sample_index + consumed */
} gb200_acq_result;

/* = pcps_acquisition ctor + init() (pcps_acquisition.cc:63-195, 313-370) */
GB200_API int gb200_acq_create(gb200_service* s, const gb200_acq_cfg* cfg, gb200_acq** out);
/* = pcps_quicksync_acquisition_cc's geometry (src/algorithms/acquisition/gnuradio_blocks/pcps_quicksync_acquisition_cc.cc:94-95,
 * 380-398): a dwell still reads d_consumed_samples = sampled_ms * samples_per_ms items, but every Doppler bin's wiped block is
 * FOLDED `fold` times (fold = folding_factor^2: segment i is added onto segment 0 in float32, i ascending) onto an FFT of
 * consumed / fold points; gb200_acq_fft_size() returns that length, gb200_acq_set_local_code() takes the caller's folded code
 * (fft_size items, :186-196) and the input power is the mean over all consumed items (:363-365).  gb200_acq_result::peak_mag /
 * code_phase_index / doppler_bin are the folded search's maximum (first maximum in bin, then index order); the block's
 * time-domain disambiguation among the folding_factor candidate delays (:436-470) is the caller's (host/gb200_quicksync.hpp).
 * fold == 1 is gb200_acq_create.  Needs sampled_ms == ms_per_code, no bit_transition_flag / make_2_steps in `cfg`. */
GB200_API int gb200_acq_create_folded(gb200_service* s, const gb200_acq_cfg* cfg, uint32_t fold, gb200_acq** out);
GB200_API int gb200_acq_destroy(gb200_acq* a);
GB200_API uint32_t gb200_acq_fft_size(const gb200_acq* a);         /* d_fft_size */
GB200_API uint32_t gb200_acq_consumed_samples(const gb200_acq* a); /* d_consumed_samples */
GB200_API uint32_t gb200_acq_num_doppler_bins(const gb200_acq* a); /* d_num_doppler_bins */
/* = pcps_acquisition::set_local_code (pcps_acquisition.cc:239-274): zero-pad layout, FFT, conj; cached on device.
 * `code_iq`: n_items complex floats as the adapter passes them (d_consumed_samples, or d_fft_size/2 with bit_transition). */
GB200_API int gb200_acq_set_local_code(gb200_acq* a, const float* code_iq, size_t n_items);
/* = set_doppler_max / set_doppler_step (pcps_acquisition.h:226-250) + FDMA offset; re-derives the bin grid */
GB200_API int gb200_acq_set_doppler(gb200_acq* a, uint32_t doppler_max, uint32_t doppler_step, int64_t freq_offset_hz);
GB200_API int gb200_acq_set_threshold(gb200_acq* a, float threshold);
/* Sibling PCPS blocks that run the same Doppler loop with their own grid rule:
 * explicit phase step per bin [rad/sample], replacing the doppler_max / doppler_step rule until the next gb200_acq_set_doppler.
 *   pcps_tong_acquisition_cc::init            pcps_tong_acquisition_cc.cc:183-205 (inclusive bin count, double-precision step)
 *   galileo_pcps_8ms_acquisition_cc::init     galileo_pcps_8ms_acquisition_cc.cc:148-170 (float32 step, GALILEO_TWO_PI)
 * gb200_acq_result::doppler_bin indexes this list; doppler_hz is then the caller's to derive. */
GB200_API int gb200_acq_set_doppler_grid(gb200_acq* a, uint32_t n_bins, const float* phase_step_rad);
/* Tong detector arithmetic (pcps_tong_acquisition_cc.cc:330-350): every dwell adds |y|^2 / (N^4 * input_power_of_that_dwell) to
 * the kept grid; gb200_acq_result::peak_mag / test_statistic are the accumulated maximum, decision is always 2 (the
 * caller owns the Tong counter) and gb200_acq_reset starts a new attempt (= set_state(1), :226-252). */
GB200_API int gb200_acq_set_scaled_accumulation(gb200_acq* a, int enable); /* 0 off, 1 scaled (Tong), 2 unscaled sum (fine-Doppler block) */
/* = reset()/set_state(1): clears dwell counter, step-two state and the magnitude grid */
GB200_API int gb200_acq_reset(gb200_acq* a);
/* = acquisition_core (pcps_acquisition.cc:668-930) for one dwell over d_consumed_samples items of
 * `stream_id` starting at `first_sample_index`. */
GB200_API int gb200_acq_run(gb200_acq* a, int stream_id, uint64_t first_sample_index, gb200_acq_result* out);
/* Many acquisitions (channels / PRNs) over the SAME sample block in one launch sequence: the
 * Doppler-wiped forward FFTs are computed once and shared by every handle with the same geometry. */
GB200_API int gb200_acq_run_batch(gb200_acq* const* list, int n, int stream_id, uint64_t first_sample_index, gb200_acq_result* out);
/* split form of run_batch: launch is asynchronous on the service stream, collect blocks */
GB200_API int gb200_acq_launch_batch(gb200_acq* const* list, int n, int stream_id, uint64_t first_sample_index);
GB200_API int gb200_acq_collect_batch(gb200_acq* const* list, int n, gb200_acq_result* out);
/* = AcquisitionInterface::stop_acquisition() (acquisition_interface.h:71) for a launched batch: abandon it.  Nothing is waited for,
 * no result is read and no handle's dwell / step-two state advances; the next gb200_acq_launch_batch may follow at once (its
 * kernels queue behind the abandoned ones on the service stream). */
GB200_API int gb200_acq_discard_batch(gb200_acq* const* list, int n);
/* magnitude grid of the last dwell, num_bins x effective_fft_size floats (dump=true, pcps_acquisition.cc:741-744).
 * Only valid when the handle keeps a grid: max_dwells > 1, use_CFAR_algorithm_flag == 0, or after gb200_acq_keep_grid(a, 1). */
GB200_API int gb200_acq_keep_grid(gb200_acq* a, int keep);
GB200_API int gb200_acq_read_grid(gb200_acq* a, float* dst);

/* Non-coherent I/Q combination of galileo_e5a_noncoherentIQ_acquisition_caf_cc's Doppler loop body
 * (galileo_e5a_noncoherent_iq_acquisition_caf_cc.cc:466-606), over the grids the handles kept from their last batch
 * (gb200_acq_keep_grid; all four share one geometry and Doppler grid).  Codes: ia = I_A (required), qa = Q_A, ib = I_B,
 * qb = Q_B (each may be NULL as in the block: Q only with both_signal_components, B only with sampled_ms > 1).
 * Per Doppler bin the device finds each row's first maximum, picks A or B per component by magt_xA >= magt_xB after
 * dividing by `norm` (= fft_normalization_factor^2; magt_QB reads the I_B row at indext_QB, :508), adds the chosen Q row
 * onto the chosen I row and takes that row's first maximum.  `out` receives num_doppler_bins records. */
typedef struct gb200_acq_iq_bin
{
    float peak[4];     /* d_magnitudeIA[indext_IA], QA, IB, QB (not normalised; 0 when the code is absent) */
    uint32_t index[4]; /* indext_IA, _QA, _IB, _QB */
    float ib_at_qb;    /* d_magnitudeIB[indext_QB]: what the block compares as magt_QB */
    float mag;         /* maximum of the combined row (not normalised) */
    uint32_t indext;   /* its index */
    uint32_t picked;   /* bit 0: I component took B; bit 1: Q component took B */
} gb200_acq_iq_bin;
GB200_API int gb200_acq_iq_combine(gb200_acq* ia, gb200_acq* qa, gb200_acq* ib, gb200_acq* qb, float norm, gb200_acq_iq_bin* out);

/* The two device steps pcps_acquisition_fine_doppler_cc adds to the Doppler loop
 * (src/algorithms/acquisition/gnuradio_blocks/pcps_acquisition_fine_doppler_cc.cc).  Its coarse stage is the plain sum of
 * |.|^2 over max_dwells dwells (compute_and_accumulate_grid, :343-372): gb200_acq_set_scaled_accumulation(a, 2).
 * gb200_acq_second_peak = the tail of compute_CAF (:296-322): the maximum of kept grid row `bin` after zeroing
 * [idx - half_width, idx + half_width) with the block's one-sided wrap and do-while.
 * gb200_acq_wiped_spectrum_peak = estimate_Doppler's steps 2-4 (:429-441): items [first, first + n_items) of the stream
 * times `replica_iq` (n_items gr_complex: the caller's 10-period code replica, :412-428), zero-padded to fft_len, forward
 * FFT, |.|^2, first maximum -> its index (natural frequency order) and value. */
GB200_API int gb200_acq_second_peak(gb200_acq* a, uint32_t bin, uint32_t idx, uint32_t half_width, float* second_peak);
GB200_API int gb200_acq_wiped_spectrum_peak(gb200_service* s, int stream_id, uint64_t first_sample_index, const float* replica_iq, uint32_t n_items,
    uint32_t fft_len, uint32_t* peak_index, float* peak_mag);

/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same second-peak search as
 * gb200_acq_second_peak, on a row the CALLER holds in device memory: *d_second (device pointer) = maximum of d_row[0..n) after
 * zeroing [idx - half_width, idx + half_width) with the block's one-sided wrap and do-while.  Allocates nothing (no
 * cudaMalloc / cudaFree / service, no synchronisation): the launch is issued on `cuda_stream` (a cudaStream_t, may be 0) of the
 * CURRENT device and the working copy of the row lives in a fixed device array of GB200_NOALLOC_MAX_ROW floats (1 MiB per GPU
 * that loads the library, held until exit).  n > GB200_NOALLOC_MAX_ROW -> GB200_E_RANGE.  All calls share that array, which
 * keeps the last call's contents (every element read is overwritten first, nothing needs clearing): calls that overlap on the
 * GPU race -- order them on one stream or with events.  `err` (may be NULL) receives the message of a failed call.
 * The FIRST call on each device resolves the array's device address (cudaGetSymbolAddress, which may load the module): make
 * that one call per device before starting a CUDA graph capture or a no-allocation region; later calls only launch. */
#define GB200_NOALLOC_MAX_ROW 262144
GB200_API int gb200_second_peak_noalloc(const float* d_row, uint32_t n, uint32_t idx, uint32_t half_width, float* d_second, void* cuda_stream, char* err,
    size_t err_cap);

/* ------------------------------------------------------------------ tracking ------------ */
#define GB200_MAX_TAPS 8
#define GB200_CODE_REAL 0 /* Cpu_Multicorrelator_Real_Codes (float codes)        */
#define GB200_CODE_CPLX 1 /* Cpu_Multicorrelator (complex-float codes, GLONASS)  */
#define GB200_CODE_16SC 2 /* Cpu_Multicorrelator_16sc (lv_16sc_t samples and codes, saturating int16 outputs; item_type must be GB200_ITEM_16IC) */

typedef struct gb200_trk_cfg
{
    int n_taps;          /* correlators (3 = E,P,L; 5 = VE,E,P,L,VL) */
    int max_code_len;    /* longest local code table in samples (1023 ... 20460) */
    int code_kind;       /* GB200_CODE_* */
    int high_dyn;        /* set_high_dynamics_resampler(): quadratic NCO terms honoured */
    int item_type;       /* GB200_ITEM_* of the stream */
} gb200_trk_cfg;

/* = Cpu_Multicorrelator_Real_Codes::init (src/algorithms/tracking/libs/cpu_multicorrelator_real_codes.cc:62-76) */
GB200_API int gb200_trk_create(gb200_service* s, const gb200_trk_cfg* cfg, gb200_trk** out);
GB200_API int gb200_trk_destroy(gb200_trk* t);
/* = set_local_code_and_taps (cpu_multicorrelator_real_codes.cc:79-89).  `code`: code_len floats (REAL) or
 * code_len complex floats (CPLX); `shifts`: n_taps tap offsets in code-table samples. */
GB200_API int gb200_trk_set_local_code_and_taps(gb200_trk* t, const void* code, int code_len, const float* shifts, int n_taps);
/* pilot tracking: the data-component code correlated at the tap `prompt_tap` with identical NCO
 * parameters (dll_pll_veml_tracking.cc:900-910, correlator_data_cpu); adds one output after the n_taps. */
GB200_API int gb200_trk_set_data_code(gb200_trk* t, const float* code, int code_len, int prompt_tap);
/* Tap offsets only, the code tables stay: the reference's tracking block rewrites d_local_code_shift_chips in place when it
 * switches from the wide to the narrow correlator spacing (dll_pll_veml_tracking.cc:1753-1764; the multicorrelator holds the
 * pointer, cpu_multicorrelator_real_codes.cc:79-89). */
GB200_API int gb200_trk_set_taps(gb200_trk* t, const float* shifts, int n_taps);
/* = Carrier_wipeoff_multicorrelator_resampler (cpu_multicorrelator_real_codes.h:55; same argument list).
 * corr_out_iq: n_taps (+1 with a data code) complex floats. */
GB200_API int gb200_trk_correlate(gb200_trk* t, int stream_id, uint64_t first_sample_index,
    float rem_carr_phase_rad, float phase_step_rad, float phase_rate_step_rad,
    float rem_code_phase_chips, float code_phase_step_chips, float code_phase_rate_step_chips,
    int n_samples, float* corr_out_iq);

/* One epoch request of one channel, for the batched forms. */
typedef struct gb200_trk_req
{
    gb200_trk* trk;
    int stream_id;
    uint64_t first_sample_index;
    float rem_carr_phase_rad, phase_step_rad, phase_rate_step_rad;
    float rem_code_phase_chips, code_phase_step_chips, code_phase_rate_step_chips;
    int n_samples;
} gb200_trk_req;

/* All requests in ONE launch.  outs: for request i, (n_taps_i + has_data_i) complex floats, packed
 * in request order at stride GB200_MAX_TAPS+1 complex floats. */
GB200_API int gb200_trk_correlate_batch(gb200_service* s, int n, const gb200_trk_req* reqs, float* outs_iq);
/* = Cpu_Multicorrelator_16sc::Carrier_wipeoff_multicorrelator_resampler (src/algorithms/tracking/libs/cpu_multicorrelator_16sc.cc:86-105;
 * same argument list) with the exact semantics of its _generic kernels: float32 rotation, rintf to int16, complex int16
 * products narrowed to 16 bits, accumulation saturating at every sample (volk_gnsssdr_16ic_x2_rotator_dot_prod_16ic_xn.h:73-108).
 * The handle must have been created with GB200_CODE_16SC; `code` in gb200_trk_set_local_code_and_taps is then code_len
 * lv_16sc_t items.  corr_out_iq: n_taps complex int16 (I, Q interleaved).  Bit-exact against the reference. */
GB200_API int gb200_trk_correlate_16sc(gb200_trk* t, int stream_id, uint64_t first_sample_index, float rem_carr_phase_rad, float phase_step_rad,
    float rem_code_phase_chips, float code_phase_step_chips, int n_samples, int16_t* corr_out_iq);
/* all requests in one launch sequence; the rate fields of gb200_trk_req are ignored.  outs: request i at outs + 2 * GB200_MAX_TAPS * i */
GB200_API int gb200_trk_correlate_16sc_batch(gb200_service* s, int n, const gb200_trk_req* reqs, int16_t* outs_iq);
/* split form: stage -> launch (asynchronous, repeatable) -> fetch */
GB200_API int gb200_trk_batch_create(gb200_service* s, int n, const gb200_trk_req* reqs, gb200_trk_batch** out);
GB200_API int gb200_trk_batch_launch(gb200_trk_batch* b);
GB200_API int gb200_trk_batch_fetch(gb200_trk_batch* b, float* outs_iq);
GB200_API int gb200_trk_batch_destroy(gb200_trk_batch* b);

#ifdef __cplusplus
}
#endif
#endif /* GNSS_B200_H */
