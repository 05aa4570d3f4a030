// Internal declarations shared by the CUDA translation units of libgnss_b200.so (sm_100a only).
#pragma once
#include "../../include/gnss_b200.h"
#include <cuda_runtime.h>
#include <stdint.h>

#define GB200_TRK_THREADS 256
#define GB200_RING_TAIL_ITEMS 1024         // spare items allocated behind the ring: the tracking kernel reads whole groups of 4 x 256 samples
#define GB200_TRK_SMEM_MAX (96 * 1024)   // dynamic shared memory a tracking CTA may use (code window + carrier table)
#define GB200_TRK_SMEM_PREF (72 * 1024)  // preferred ceiling when choosing the chunk: 3 CTAs per SM
#define GB200_OUT_STRIDE (GB200_MAX_TAPS + 1)  // complex floats per request in the output array

// One multicorrelator request as the kernels see it (device memory, one per (channel, epoch)).
struct GbTrkWork
{
    const void* samples;     // first sample of the correlation window (contiguous: ring mirror)
    const float* code;       // local code table (float, or float2 when complex)
    const float* data_code;  // pilot tracking: data-component table correlated at tap `prompt_tap`, or nullptr
    const float* ext;        // the code repeated over ext_entries window entries (float, (pilot, data) or (re, im)) for bulk staging
    int ext_entries;
    int code_len;
    int n;                   // signal_length_samples
    int n_taps;
    int prompt_tap;
    float shifts[GB200_MAX_TAPS];
    float rem_carr, phase_step, phase_rate;
    float rem_code, code_step, code_rate;
    int out_index;           // request slot in the output / partial arrays
    float nco_mag_delta;     // |inc| - 1 of the reference's float32 phase increment (see nco_* below)
    // Reference-matched carrier NCO.  The reference rotates by phase_{n+1} = phase_n * inc in float32 with
    // phase_0 = (cosf(rem), -sinf(rem)) and inc = (cosf(-step), sinf(-step)) (cpu_multicorrelator_real_codes.cc:140-149),
    // renormalising every 256 samples.  The float32 ROUNDING OF inc shifts the NCO frequency by up to ~3e-8*step rad per
    // sample, which over 10^4..10^5 samples is visible at the 1e-4 parity budget.  The host therefore derives, in float64,
    // the angle of the rounded phase_0 and of the rounded inc; the kernel evaluates phase_0 * inc^n from those.
    double nco_a0_turns;     // arg(phase_0) / 2pi
    double nco_theta_turns;  // arg(inc) / 2pi
    double nco_log_mag;      // log|inc| (high-dynamics path: the reference never renormalises phase_doppler there)
    unsigned long long nco_a0_fix;     // frac(nco_a0_turns) * 2^64: the kernels carry the turn count in 64-bit fixed point
    unsigned long long nco_theta_fix;  // frac(nco_theta_turns) * 2^64 (two's complement for negative increments)
};

// dynamic shared memory of gb_trk_kernel: window entries (4 or 8 bytes each) + one float4 per loop iteration of the chunk
__host__ __device__ inline size_t gb_trk_smem_bytes(int win_entries, bool wide_entries, int chunk)
{
    return (((size_t)win_entries * (wide_entries ? 8 : 4) + 15) & ~(size_t)15) + (size_t)(chunk / GB200_TRK_THREADS + 4) * 16;
}

// one request of the int16 (Cpu_Multicorrelator_16sc) correlator
struct GbTrk16Work
{
    const void* samples;  // lv_16sc_t window start (contiguous)
    const void* code;     // lv_16sc_t[code_len] local code
    int code_len, n, n_taps, out_index;
    float shifts[GB200_MAX_TAPS];
    float rem_code, code_step;
    float2 phase0;  // (cosf(rem_carr), -sinf(rem_carr)), host libm like the reference (cpu_multicorrelator_16sc.cc:99)
    float2 inc;     // std::exp(lv_32fc_t(0, -phase_step)) = (cosf(-step), sinf(-step))
};

// acquisition grid-reduction record: ordering key = (mag desc, bin asc, idx asc)
struct GbPeak
{
    float mag;
    uint32_t idx;
    uint32_t bin;
    uint32_t pad_;
};

// ---- sample loads: int8 / int16 are widened exactly as volk_gnsssdr_16ic_convert_32fc does ((float)int) ----
template <int ITEM>
__device__ __forceinline__ float2 gb_load_item(const void* base, size_t i);

template <>
__device__ __forceinline__ float2 gb_load_item<GB200_ITEM_32FC>(const void* base, size_t i)
{
    return __ldg(reinterpret_cast<const float2*>(base) + i);
}
template <>
__device__ __forceinline__ float2 gb_load_item<GB200_ITEM_16IC>(const void* base, size_t i)
{
    short2 v = __ldg(reinterpret_cast<const short2*>(base) + i);
    return make_float2((float)v.x, (float)v.y);
}
template <>
__device__ __forceinline__ float2 gb_load_item<GB200_ITEM_8IC>(const void* base, size_t i)
{
    char2 v = __ldg(reinterpret_cast<const char2*>(base) + i);
    return make_float2((float)v.x, (float)v.y);
}

__host__ __device__ inline size_t gb_item_bytes(int item_type)
{
    return item_type == GB200_ITEM_32FC ? 8 : (item_type == GB200_ITEM_16IC ? 4 : 2);
}

// exp(j * 2*pi * (a + n*b)) with a, b in turns (float64), result float32
__device__ __forceinline__ float2 gb_phasor_turns(double a, double b, double n)
{
    double t = fma(n, b, a);
    t -= rint(t);
    float s, c;
    sincospif(2.0f * (float)t, &s, &c);
    return make_float2(c, s);
}

__device__ __forceinline__ float2 gb_cmul(float2 a, float2 b)
{
    return make_float2(fmaf(a.x, b.x, -a.y * b.y), fmaf(a.x, b.y, a.y * b.x));
}

// kernel launchers (defined in the .cu files, called from service.cu)
// device-side copy of `bytes` (multiple of 8) from device-mapped host memory (or anywhere) to device memory
cudaError_t gb_copy_words(void* dst, const void* src, size_t bytes, cudaStream_t st);
cudaError_t gb_conj_items(void* d_items, size_t n_items, int item_type, cudaStream_t st);
cudaError_t gb_trk_launch(const GbTrkWork* d_work, int n_work, int max_n, int n_taps, bool has_data, int prompt_tap, int item_type,
    int code_kind, bool high_dyn, float2* d_partials, unsigned* d_counters, float2* d_out, int chunk, int win_entries, cudaStream_t st,
    int* launches);

cudaError_t gb_trk16_launch(const GbTrk16Work* d_work, int n_work, int max_n, int n_taps, float2* d_ckpt, int4* d_triples, short* d_out,
    cudaStream_t st, int* launches);

// ------------------------------------------------------------------ acquisition FFT plan
#define GB200_ACQ_THREADS 256
#define GB200_MAX_RADIX_PASSES 16
#define GB200_MAX_GENERIC_RADIX 64

// Two-level ("four-step") decomposition N = N1 * N2 of the un-normalised DFT:
//   n = N2*n1 + n2,  k = k1 + N1*k2
//   forward : A[k1][n2] = W_N^(n2 k1) * sum_n1 x[N2 n1 + n2] W_N1^(n1 k1);   X[k1 + N1 k2] = sum_n2 A[k1][n2] W_N2^(n2 k2)
//   backward: B[k1][n2] = W_N^-(n2 k1) * sum_k2 Y[k1 + N1 k2] W_N2^-(n2 k2); y[N2 n1 + n2] = sum_k1 B[k1][n2] W_N1^-(n1 k1)
// Spectra live in the "[k1][k2]" layout (k1-major) everywhere, so both directions only ever touch
// whole rows (row kernels) or tiles of adjacent columns (column kernels).
struct GbFftPlan
{
    int N, N1, N2;
    int nrad1, nrad2;
    int rad1[GB200_MAX_RADIX_PASSES];
    int rad2[GB200_MAX_RADIX_PASSES];
    int tile1;  // columns (n2) per CTA in the column kernels (sub-FFT length N1)
    int tile2;  // rows (k1) per CTA in the row kernels (sub-FFT length N2)
    int inv_fast;   // 1: (N1, N2) has the compile-time in-place DIF inverse kernels (acq_dif.cuh)
    int inv_variant;  // which instantiation of the pair (GB200_ACQ_VARIANT, tuning only)
    int inv_tile1;  // columns per CTA of the inverse column kernel actually launched (sizes the per-CTA peak array)
    int fwd_generic;  // 1: forward side on the run-time parametrised kernels even where the DIF tiles exist (tuning key "acq_fwd_generic")
    const float2* tw1;  // device: W_N1^i, i < N1 (forward sign)
    const float2* tw2;  // device: W_N2^i, i < N2
    const float2* twN;  // device: W_N^i, i < max(256, 16 * max(N1, N2)): the per-column part of the four-step twiddle
    const float2* twNhi;  // device: W_N^(256 i), i < ceil(N / 256)
};

// Reference-matched Doppler wipe-off.  The reference builds each bin's table with
// volk_gnsssdr_s32f_sincos_32fc: out[n] = (cosf(ph_n), sinf(ph_n)), ph_{n+1} = fl32(ph_n + inc) (float32
// running sum, ...s32f_sincos_32fc.h:405-416).  Inside one float32 binade that running sum is exactly
// linear in n with an increment rounded to the binade's ulp, so the host simulates the sum once per bin
// and hands the kernel the piecewise-linear description; ph_n is then reproduced EXACTLY without a table.
struct GbWipeSeg
{
    int n_start;     // first sample of the segment
    float ph_start;  // ph at n_start
    float inc_eff;   // exact ph_{n+1} - ph_n inside the segment
};

struct GbAcqFwdParams
{
    const void* samples;   // consumed items
    int consumed;
    int item_type;
    const GbWipeSeg* segs;        // all bins' segments, concatenated
    const int* seg_offsets;       // [bins + 1] offsets into segs
    float2* colbuf;        // [bins][N1][N2] scratch after the column stage
    float2* spectra;       // [bins][N1][N2] forward spectra, [k1][k2] layout
    int first_bin = 0;     // this launch transforms bins first_bin .. first_bin + n_bins - 1 (the forward side runs in bin groups)
    int fold = 1;          // QuickSync: the wiped block is folded `fold` times onto the N-point FFT input (pcps_quicksync_acquisition_cc.cc:391-398)
};

struct GbAcqInvParams
{
    const float2* spectra;          // [bins][N1][N2]
    const float2* const* code_fft;  // per handle: conj(FFT(code)) in [k1][k2] layout
    float2* rowbuf;                 // [transforms in chunk][N1][N2]
    int n_bins;
    int first_transform;            // chunk offset; transform t = handle * n_bins + bin, or bin * n_handles + handle when n_handles > 0
    int n_handles = 0;              // > 0: bin-major transform order (chunks then need only the first bins' spectra: gb_acq_unit)
    int n_transforms;               // in this chunk
    int keep_offset;                // first kept output index (N/2 with bit_transition, else 0)
    int keep_len;                   // effective_fft_size
    float* const* grid;             // per handle: [bins][keep_len] magnitude grid or nullptr
    int accumulate;                 // dwell > 1: grid += mag
    GbPeak* partial;                // [handle][bin][tiles] per-CTA maxima (over the stored values where a grid is kept)
    float* const* tmpbuf;           // per handle d_tmp_buffer image or nullptr: receives the last bin's magnitudes when accumulating (:737)
    const float* scale;             // device scalar or nullptr: magnitudes are multiplied by it before they enter the grid (Tong detector)
};

#ifdef __CUDACC__
// transform number -> (handle, bin) in either order
__device__ __forceinline__ void gb_acq_unit(const GbAcqInvParams& p, int t, int& handle, int& bin)
{
    if (p.n_handles > 0)
        {
            bin = t / p.n_handles;
            handle = t - bin * p.n_handles;
        }
    else
        {
            handle = t / p.n_bins;
            bin = t - handle * p.n_bins;
        }
}
#endif

// per-kernel-class CUDA-event timing hooks (service.cu); `ctx` may be nullptr
void gb_kt_begin(void* ctx, int kernel_class);
void gb_kt_end(void* ctx);
cudaError_t gb_acq_forward(const GbFftPlan& plan, const GbAcqFwdParams& p, int n_bins, bool wipeoff, cudaStream_t st, int* launches, void* kt);
cudaError_t gb_acq_forward_dif(const GbFftPlan& plan, const GbAcqFwdParams& p, int n_bins, bool wipeoff, cudaStream_t st, int* launches, void* kt);  // acq_inv.cu
cudaError_t gb_acq_inverse(const GbFftPlan& plan, const GbAcqInvParams& p, cudaStream_t st, int* launches, void* kt);
size_t gb_acq_fused_ctl_bytes(int n_transforms);
cudaError_t gb_acq_inverse_fused(const GbFftPlan& plan, const GbAcqInvParams& p, void* d_ctl, int lag, int ring, int sm_count, cudaStream_t st, int* launches, void* kt);
bool gb_acq_inverse_tiles(int N1, int N2, int variant, int* tile1, int* tile2);
bool gb_acq_preferred_split(int N, int avoid_n1, int* n1, int* n2);  // a split of N that has the fast inverse kernels  // true when the pair has the fast inverse kernels
cudaError_t gb_acq_conj(float2* data, size_t n, cudaStream_t st, int* launches);
cudaError_t gb_acq_input_power(const void* samples, int consumed, int item_type, int N, float* d_power, float* d_mag2_tmp, cudaStream_t st, int* launches);
// reduce per-CTA maxima (all modes; over the stored grid where one is kept) -> per handle peak
cudaError_t gb_acq_reduce_peaks(const GbPeak* partial, int n_handles, int n_bins, int tiles, GbPeak* out, cudaStream_t st, int* launches);
// first_vs_second_peak_statistic's second peak with the d_tmp_buffer quirk (pcps_acquisition.cc:633-660)
// FIR + decimation of the newest items of a 32fc ring into a linear buffer (gb200_stream_set_decimator)
cudaError_t gb_fir_decimate(const float2* ring, unsigned long long cap, unsigned long long valid_begin, unsigned long long k_first, unsigned n_out,
    unsigned decimation, const float* d_taps, unsigned n_taps, float2* out, cudaStream_t st, int* launches);
cudaError_t gb_acq_iq_combine(const float* const rows[4], int n_bins, int row_len, float norm, gb200_acq_iq_bin* d_out, cudaStream_t st, int* launches);
// fine-Doppler block helpers (acq_misc.cu)
cudaError_t gb_acq_second_peak_row(const float* grid_row, float* tmp_row, int N, int idx, int half_width, float* d_second, cudaStream_t st, int* launches);
cudaError_t gb_acq_code_wipe(const void* samples, int item_type, const float2* replica, int n, float2* out, cudaStream_t st, int* launches);
cudaError_t gb_acq_spectrum_peak(const float2* spectrum, int N1, int N2, GbPeak* d_out, cudaStream_t st, int* launches);
cudaError_t gb_acq_second_peak_batch(float* const* grids, float* const* tmps, const GbPeak* d_peaks, int n_handles, int row_len, int N, int samples_per_chip,
    float* d_seconds, cudaStream_t st, int* launches);
cudaError_t gb_acq_second_peak(const float* grid_row, float* tmp_buffer, int N, int idx, int samples_per_chip, float* d_second, cudaStream_t st, int* launches);
cudaError_t gb_acq_copy_f32(float* dst, const float* src, int n, cudaStream_t st, int* launches);
size_t gb_acq_col_smem(const GbFftPlan& plan);
size_t gb_acq_row_smem(const GbFftPlan& plan);
