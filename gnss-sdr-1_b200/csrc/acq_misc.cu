// Small acquisition kernels: conjugate, input power, peak reductions, second-peak statistic.
#include "acq_fft.cuh"
#include <cstring>
#include <mutex>
#include <string>

namespace
{

__global__ void gb_conj_kernel(float2* d, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) d[i].y = -d[i].y;
}

// d_input_power = sum_{n<N} |x[n]|^2 / N with x zero-padded beyond `consumed` (pcps_acquisition.cc:703-709);
// optionally leaves |x|^2 in the block's d_tmp_buffer image (needed only to reproduce quirk 1 of SURVEY 7.3).
// Single CTA, fixed summation tree -> deterministic.
template <int ITEM>
__global__ void __launch_bounds__(1024) gb_input_power_kernel(const void* samples, int consumed, int N, float* power, float* tmp)
{
    __shared__ float s[32];
    float acc = 0.f;
    // eight loads in flight per thread; the additions stay in ascending-n order, so the sum is that of the plain strided loop
    for (int n0 = threadIdx.x; n0 < N; n0 += 8 * blockDim.x)
        {
            float m[8];
#pragma unroll
            for (int u = 0; u < 8; u++)
                {
                    const int n = n0 + u * blockDim.x;
                    m[u] = 0.f;
                    if (n < consumed && n < N)
                        {
                            const float2 x = gb_load_item<ITEM>(samples, (size_t)n);
                            m[u] = __fadd_rn(__fmul_rn(x.x, x.x), __fmul_rn(x.y, x.y));
                        }
                }
#pragma unroll
            for (int u = 0; u < 8; u++)
                {
                    const int n = n0 + u * blockDim.x;
                    if (n < N)
                        {
                            if (tmp) tmp[n] = m[u];
                            acc += m[u];
                        }
                }
        }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32)
        {
            acc = threadIdx.x < (blockDim.x >> 5) ? s[threadIdx.x] : 0.f;
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
            if (threadIdx.x == 0)
                {
                    const float p = acc / (float)N;
                    power[0] = p;
                    // Tong detector (pcps_tong_acquisition_cc.cc:333-335): per-dwell scale 1 / (N^2 * N^2 * input_power), float32
                    const float fnf = __fmul_rn((float)N, (float)N);
                    power[1] = __fdiv_rn(1.0f, __fmul_rn(__fmul_rn(fnf, fnf), p));
                }
        }
}

// ordering of the reference's two nested strict-'>' scans (bins ascending, index ascending):
// larger magnitude wins; on equal magnitude the lower bin, then the lower index
__device__ __forceinline__ bool peak_before(const GbPeak& a, const GbPeak& b)
{
    if (a.mag != b.mag) return a.mag > b.mag;
    if (a.bin != b.bin) return a.bin < b.bin;
    return a.idx < b.idx;
}

__device__ __forceinline__ GbPeak peak_block_reduce(GbPeak best)
{
    __shared__ GbPeak s[32];
    for (int o = 16; o > 0; o >>= 1)
        {
            GbPeak other;
            other.mag = __shfl_down_sync(0xffffffffu, best.mag, o);
            other.idx = __shfl_down_sync(0xffffffffu, best.idx, o);
            other.bin = __shfl_down_sync(0xffffffffu, best.bin, o);
            other.pad_ = 0;
            if (peak_before(other, best)) best = other;
        }
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x == 0)
        for (int k = 1; k < (int)(blockDim.x >> 5); k++)
            if (peak_before(s[k], best)) best = s[k];
    return best;  // valid in thread 0
}

// one CTA per handle: reduce [bins][tiles] per-CTA maxima
__global__ void __launch_bounds__(256) gb_reduce_peaks_kernel(const GbPeak* partial, int n_bins, int tiles, GbPeak* out)
{
    const GbPeak* p = partial + (size_t)blockIdx.x * n_bins * tiles;
    GbPeak best;
    best.mag = -1.f;
    best.idx = 0xffffffffu;
    best.bin = 0xffffffffu;
    best.pad_ = 0;
    for (int i = threadIdx.x; i < n_bins * tiles; i += blockDim.x)
        {
            GbPeak c = p[i];
            if (peak_before(c, best)) best = c;
        }
    best = peak_block_reduce(best);
    if (threadIdx.x == 0)
        {
            // the reference starts from grid_maximum = 0 with strict '>': an all-zero grid reports bin 0, index 0
            if (!(best.mag > 0.f))
                {
                    best.mag = 0.f;
                    best.idx = 0;
                    best.bin = 0;
                }
            out[blockIdx.x] = best;
        }
}

// first_vs_second_peak_statistic tail (pcps_acquisition.cc:633-660), single CTA:
//   memcpy(d_tmp_buffer, row, d_fft_size BYTES)  -> only the first N/4 floats are refreshed (quirk 1)
//   zero [idx - spc, idx + spc) with the one-sided wrap of :637-644 (quirk 2), do-while semantics
//   second peak = first maximum of d_tmp_buffer[0..N)
__device__ __forceinline__ void second_peak_body(const float* row, int row_len, float* tmp, int N, int idx, int spc, float* second, int ncopy_all)
{
    const int ncopy = ncopy_all ? min(N, row_len) : min(N / 4, row_len);
    for (int i = threadIdx.x; i < ncopy; i += blockDim.x) tmp[i] = row[i];
    __syncthreads();
    int ex1 = idx - spc, ex2 = idx + spc;
    if (ex1 < 0)
        ex1 = N + ex1;
    else if (ex2 >= N)
        ex2 = ex2 - N;
    // do { tmp[i] = 0; i++; if (i == N) i = 0; } while (i != ex2): at least one element, wraps the whole buffer when ex1 == ex2
    int count = ex2 - ex1;
    if (count <= 0) count += N;
    for (int k = threadIdx.x; k < count; k += blockDim.x)
        {
            int i = ex1 + k;
            if (i >= N) i -= N;
            if (i >= 0 && i < N) tmp[i] = 0.f;
        }
    __syncthreads();
    GbPeak best;
    best.mag = -INFINITY;
    best.idx = 0xffffffffu;
    best.bin = 0;
    best.pad_ = 0;
    for (int i = threadIdx.x; i < N; i += blockDim.x)
        {
            GbPeak c;
            c.mag = tmp[i];
            c.idx = (unsigned)i;
            c.bin = 0;
            c.pad_ = 0;
            if (peak_before(c, best)) best = c;
        }
    best = peak_block_reduce(best);
    if (threadIdx.x == 0) *second = best.mag;
}

__global__ void __launch_bounds__(1024) gb_second_peak_kernel(const float* row, int row_len, float* tmp, int N, int idx, int spc, float* second, int ncopy_all)
{
    second_peak_body(row, row_len, tmp, N, idx, spc, second, ncopy_all);
}

// the same for every handle of a batch group, one CTA each: the winning (bin, index) comes from the peak record the grid scan
// left on the device, so nothing crosses to the host between the scan and the second-peak search
__global__ void __launch_bounds__(1024) gb_second_peak_batch_kernel(float* const* grids, float* const* tmps, const GbPeak* peaks, int row_len, int N, int spc, float* seconds)
{
    const GbPeak pk = peaks[blockIdx.x];
    second_peak_body(grids[blockIdx.x] + (size_t)pk.bin * row_len, N, tmps[blockIdx.x], N, (int)pk.idx, spc, seconds + blockIdx.x, 0);
}

// galileo_e5a_noncoherent_iq_acquisition_caf_cc.cc:466-606, one CTA per Doppler bin (see gb200_acq_iq_combine in the header)
struct GbIqRows
{
    const float* g[4];  // I_A, Q_A, I_B, Q_B grids [bins][row_len] or nullptr
};
__global__ void __launch_bounds__(1024) gb_iq_combine_kernel(GbIqRows rows, int row_len, float norm, gb200_acq_iq_bin* out)
{
    __shared__ GbPeak s_pk[5];
    const size_t off = (size_t)blockIdx.x * row_len;
    const float* r[4];
#pragma unroll
    for (int k = 0; k < 4; k++) r[k] = rows.g[k] ? rows.g[k] + off : nullptr;
    auto row_max = [&](const float* a, const float* b, int slot) {  // first maximum of a[i] (+ b[i])
        GbPeak best;
        best.mag = -INFINITY;
        best.idx = 0xffffffffu;
        best.bin = 0;
        best.pad_ = 0;
        for (int i = threadIdx.x; i < row_len; i += blockDim.x)
            {
                GbPeak c;
                c.mag = b ? __fadd_rn(a[i], b[i]) : a[i];
                c.idx = (unsigned)i;
                c.bin = 0;
                c.pad_ = 0;
                if (peak_before(c, best)) best = c;
            }
        best = peak_block_reduce(best);
        if (threadIdx.x == 0) s_pk[slot] = best;
        __syncthreads();
    };
#pragma unroll
    for (int k = 0; k < 4; k++)
        if (r[k]) row_max(r[k], nullptr, k);
    const float m_ia = __fdiv_rn(s_pk[0].mag, norm);
    const float m_qa = r[1] ? __fdiv_rn(s_pk[1].mag, norm) : 0.f;
    const float m_ib = r[2] ? __fdiv_rn(s_pk[2].mag, norm) : 0.f;
    const float ib_at_qb = (r[2] && r[3]) ? r[2][s_pk[3].idx] : 0.f;
    const float m_qb = __fdiv_rn(ib_at_qb, norm);
    const bool i_b = r[2] && !(m_ia >= m_ib);
    const bool q_b = r[3] && !(m_qa >= m_qb);
    row_max(i_b ? r[2] : r[0], r[1] ? (q_b ? r[3] : r[1]) : nullptr, 4);
    if (threadIdx.x == 0)
        {
            gb200_acq_iq_bin o;
#pragma unroll
            for (int k = 0; k < 4; k++)
                {
                    o.peak[k] = r[k] ? s_pk[k].mag : 0.f;
                    o.index[k] = r[k] ? s_pk[k].idx : 0u;
                }
            o.ib_at_qb = ib_at_qb;
            o.mag = s_pk[4].mag;
            o.indext = s_pk[4].idx;
            o.picked = (i_b ? 1u : 0u) | (q_b ? 2u : 0u);
            out[blockIdx.x] = o;
        }
}

// last dwell's magnitudes of the LAST Doppler bin into d_tmp_buffer (pcps_acquisition.cc:737, dwell > 1)
__global__ void gb_copy_kernel(float* dst, const float* src, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[i];
}
}  // namespace

size_t gb_acq_col_smem(const GbFftPlan& plan) { return smem_for(plan.N1, plan.tile1); }
size_t gb_acq_row_smem(const GbFftPlan& plan) { return smem_for(plan.N2, plan.tile2); }

cudaError_t gb_acq_conj(float2* data, size_t n, cudaStream_t st, int* launches)
{
    gb_conj_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(data, n);
    if (launches) (*launches)++;
    return cudaGetLastError();
}

cudaError_t gb_acq_input_power(const void* samples, int consumed, int item_type, int N, float* d_power, float* d_mag2_tmp, cudaStream_t st, int* launches)
{
    switch (item_type)
        {
        case GB200_ITEM_32FC: gb_input_power_kernel<GB200_ITEM_32FC><<<1, 1024, 0, st>>>(samples, consumed, N, d_power, d_mag2_tmp); break;
        case GB200_ITEM_16IC: gb_input_power_kernel<GB200_ITEM_16IC><<<1, 1024, 0, st>>>(samples, consumed, N, d_power, d_mag2_tmp); break;
        case GB200_ITEM_8IC: gb_input_power_kernel<GB200_ITEM_8IC><<<1, 1024, 0, st>>>(samples, consumed, N, d_power, d_mag2_tmp); break;
        default: return cudaErrorInvalidValue;
        }
    if (launches) (*launches)++;
    return cudaGetLastError();
}

cudaError_t gb_acq_reduce_peaks(const GbPeak* partial, int n_handles, int n_bins, int tiles, GbPeak* out, cudaStream_t st, int* launches)
{
    gb_reduce_peaks_kernel<<<n_handles, 256, 0, st>>>(partial, n_bins, tiles, out);
    if (launches) (*launches)++;
    return cudaGetLastError();
}

cudaError_t gb_acq_iq_combine(const float* const rows[4], int n_bins, int row_len, float norm, gb200_acq_iq_bin* d_out, cudaStream_t st, int* launches)
{
    GbIqRows r;
    for (int k = 0; k < 4; k++) r.g[k] = rows[k];
    gb_iq_combine_kernel<<<n_bins, 1024, 0, st>>>(r, row_len, norm, d_out);
    if (launches) (*launches)++;
    return cudaGetLastError();
}

cudaError_t gb_acq_second_peak(const float* grid_row, float* tmp_buffer, int N, int idx, int samples_per_chip, float* d_second, cudaStream_t st, int* launches)
{
    gb_second_peak_kernel<<<1, 1024, 0, st>>>(grid_row, N, tmp_buffer, N, idx, samples_per_chip, d_second, 0);
    if (launches) (*launches)++;
    return cudaGetLastError();
}

cudaError_t gb_acq_second_peak_batch(float* const* grids, float* const* tmps, const GbPeak* d_peaks, int n_handles, int row_len, int N, int samples_per_chip,
    float* d_seconds, cudaStream_t st, int* launches)
{
    gb_second_peak_batch_kernel<<<n_handles, 1024, 0, st>>>(grids, tmps, d_peaks, row_len, N, samples_per_chip, d_seconds);
    if (launches) (*launches)++;
    return cudaGetLastError();
}

// pcps_acquisition_fine_doppler_cc.cc:296-322: the exclusion is applied to the whole row (here: to a copy of it)
cudaError_t gb_acq_second_peak_row(const float* grid_row, float* tmp_row, int N, int idx, int half_width, float* d_second, cudaStream_t st, int* launches)
{
    gb_second_peak_kernel<<<1, 1024, 0, st>>>(grid_row, N, tmp_row, N, idx, half_width, d_second, 1);
    if (launches) (*launches)++;
    return cudaGetLastError();
}

namespace
{
// volk_32fc_x2_multiply_32fc of the buffered samples with the code replica (pcps_acquisition_fine_doppler_cc.cc:430)
template <int ITEM>
__global__ void gb_code_wipe_kernel(const void* __restrict__ samples, const float2* __restrict__ rep, int n, float2* __restrict__ out)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        {
            const float2 x = gb_load_item<ITEM>(samples, i);
            const float2 c = rep[i];
            out[i] = make_float2(__fsub_rn(__fmul_rn(x.x, c.x), __fmul_rn(x.y, c.y)), __fadd_rn(__fmul_rn(x.x, c.y), __fmul_rn(x.y, c.x)));
        }
}

// first maximum of |X[k]|^2 in natural frequency order; X is stored [k1][k2] with k = k1 + N1 * k2 (four-step output order)
__global__ void __launch_bounds__(1024) gb_spectrum_peak_kernel(const float2* __restrict__ X, int N1, int N2, GbPeak* out)
{
    GbPeak best;
    best.mag = -INFINITY;
    best.idx = 0xffffffffu;
    best.bin = 0;
    best.pad_ = 0;
    const int N = N1 * N2;
    for (int pos = threadIdx.x; pos < N; pos += blockDim.x)
        {
            const int k1 = pos / N2, k2 = pos - k1 * N2;
            const float2 v = X[pos];
            GbPeak c;
            c.mag = __fadd_rn(__fmul_rn(v.x, v.x), __fmul_rn(v.y, v.y));
            c.idx = (unsigned)(k1 + N1 * k2);
            c.bin = 0;
            c.pad_ = 0;
            if (peak_before(c, best)) best = c;
        }
    best = peak_block_reduce(best);
    if (threadIdx.x == 0) *out = best;
}
}  // namespace

cudaError_t gb_acq_code_wipe(const void* samples, int item_type, const float2* replica, int n, float2* out, cudaStream_t st, int* launches)
{
    const int blocks = std::min((n + 255) / 256, 148 * 8);
    switch (item_type)
        {
        case GB200_ITEM_32FC: gb_code_wipe_kernel<GB200_ITEM_32FC><<<blocks, 256, 0, st>>>(samples, replica, n, out); break;
        case GB200_ITEM_16IC: gb_code_wipe_kernel<GB200_ITEM_16IC><<<blocks, 256, 0, st>>>(samples, replica, n, out); break;
        case GB200_ITEM_8IC: gb_code_wipe_kernel<GB200_ITEM_8IC><<<blocks, 256, 0, st>>>(samples, replica, n, out); break;
        default: return cudaErrorInvalidValue;
        }
    if (launches) (*launches)++;
    return cudaGetLastError();
}

cudaError_t gb_acq_spectrum_peak(const float2* spectrum, int N1, int N2, GbPeak* d_out, cudaStream_t st, int* launches)
{
    gb_spectrum_peak_kernel<<<1, 1024, 0, st>>>(spectrum, N1, N2, d_out);
    if (launches) (*launches)++;
    return cudaGetLastError();
}

cudaError_t gb_acq_copy_f32(float* dst, const float* src, int n, cudaStream_t st, int* launches)
{
    gb_copy_kernel<<<(n + 255) / 256, 256, 0, st>>>(dst, src, n);
    if (launches) (*launches)++;
    return cudaGetLastError();
}



// ---- ingest: acquisition resampler (gnss_flowgraph.cc:423-476: fir_filter_ccf(decimation, low-pass taps)) ----
// y[k] = sum_j taps[j] x[kD - j].  A CTA produces blockDim.x consecutive outputs: it stages the input span
// [k0 D - (T-1), (k0 + B - 1) D] once (coalesced, ring wrap and the zero history handled there) de-interleaved by phase
// (element m -> s[m % D][m / D]), so that in the dot product consecutive threads read consecutive shared-memory words for
// every tap; the taps are a warp-uniform read.  HBM traffic: 8 B per input item + 8/D B per item written.
namespace
{
__global__ void gb_fir_decimate_kernel(const float2* __restrict__ ring, unsigned long long cap, unsigned long long valid_begin, unsigned long long k_first,
    unsigned n_out, unsigned D, const float* __restrict__ taps, unsigned T, float2* __restrict__ out)
{
    extern __shared__ float2 s_x[];
    const unsigned B = blockDim.x;
    const unsigned long long k0 = k_first + (unsigned long long)blockIdx.x * B;
    const unsigned span = (B - 1) * D + T;            // inputs this CTA touches
    const unsigned rows = (span + D - 1) / D;         // elements per phase row
    const long long base = (long long)(k0 * D) - (long long)(T - 1);  // absolute index of span element 0 (may be negative)
    for (unsigned m = threadIdx.x; m < span; m += B)
        {
            const long long a = base + (long long)m;
            float2 v = make_float2(0.f, 0.f);
            if (a >= (long long)valid_begin) v = ring[(unsigned long long)a % cap];
            s_x[(m % D) * rows + m / D] = v;
        }
    __syncthreads();
    const unsigned t = threadIdx.x;
    if ((unsigned long long)blockIdx.x * B + t >= n_out) return;
    // output k0 + t reads span elements m = t D + (T - 1) - j
    float ax = 0.f, ay = 0.f;
    unsigned m = t * D + (T - 1);
    unsigned p = m % D, q = m / D;
    for (unsigned j = 0; j < T; j++)
        {
            const float w = __ldg(taps + j);
            const float2 x = s_x[p * rows + q];
            ax = fmaf(w, x.x, ax);
            ay = fmaf(w, x.y, ay);
            if (p == 0)
                {
                    p = D - 1;
                    q--;
                }
            else
                p--;
        }
    out[(unsigned long long)blockIdx.x * B + t] = make_float2(ax, ay);
}
}  // namespace

cudaError_t gb_fir_decimate(const float2* ring, unsigned long long cap, unsigned long long valid_begin, unsigned long long k_first, unsigned n_out,
    unsigned decimation, const float* d_taps, unsigned n_taps, float2* out, cudaStream_t st, int* launches)
{
    if (n_out == 0) return cudaSuccess;
    unsigned B = 256;
    auto smem = [&](unsigned b) { return sizeof(float2) * (size_t)decimation * (((size_t)(b - 1) * decimation + n_taps + decimation - 1) / decimation); };
    while (B > 32 && smem(B) > 96 * 1024) B >>= 1;
    const size_t sm = smem(B);
    if (sm > 96 * 1024) return cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute(gb_fir_decimate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
    if (e != cudaSuccess) return e;
    gb_fir_decimate_kernel<<<(n_out + B - 1) / B, B, sm, st>>>(ring, cap, valid_begin, k_first, n_out, decimation, d_taps, n_taps, out);
    if (launches) (*launches)++;
    return cudaGetLastError();
}

// ---- ingest: conjugate the pushed items in place (inverted_spectrum option of the data-type adapters) ----
// Replaces conjugate_cc / conjugate_sc / conjugate_ic (src/algorithms/libs/conjugate_{cc,sc,ic}.cc -> volk_32fc_conjugate_32fc,
// volk_gnsssdr_16ic_conjugate_16ic, volk_gnsssdr_8ic_conjugate_8ic): the imaginary part is negated in the item's own type, so
// the most negative integer stays what two's-complement negation makes of it, like the reference's generic kernels.
namespace
{
template <class T>
__global__ void gb_conj_kernel(T* __restrict__ iq, size_t n_items)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_items; i += (size_t)gridDim.x * blockDim.x)
        {
            T v = iq[2 * i + 1];
            iq[2 * i + 1] = (T)(-v);
        }
}
}  // namespace

cudaError_t gb_conj_items(void* d_items, size_t n_items, int item_type, cudaStream_t st)
{
    if (n_items == 0) return cudaSuccess;
    const int threads = 256;
    const int blocks = (int)std::min<size_t>((n_items + threads - 1) / threads, 148 * 8);
    if (item_type == GB200_ITEM_32FC)
        gb_conj_kernel<float><<<blocks, threads, 0, st>>>((float*)d_items, n_items);
    else if (item_type == GB200_ITEM_16IC)
        gb_conj_kernel<short><<<blocks, threads, 0, st>>>((short*)d_items, n_items);
    else
        gb_conj_kernel<signed char><<<blocks, threads, 0, st>>>((signed char*)d_items, n_items);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------
// gb200_second_peak_noalloc: the second-peak search of gb_acq_second_peak_row for callers that must not allocate (CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The row copy the kernel works on lives in a file-scope device array instead of service scratch:
// GB200_NOALLOC_MAX_ROW floats (1 MiB) held by every GPU that loads the library until exit.  Every call shares it and it keeps
// the last call's contents; the kernel overwrites elements [0, n) before it reads them, so nothing has to be cleared, but two
// calls that overlap on the GPU (different streams, no ordering between them) race on it.
static __device__ __align__(256) float gb_noalloc_row[GB200_NOALLOC_MAX_ROW];

namespace
{
// device address of gb_noalloc_row per device ordinal: in host code the array's name is only a host-side stand-in
float* g_noalloc_ptr[64] = {};
std::mutex g_noalloc_mu;

int noalloc_fail(char* err, size_t cap, int code, const std::string& msg)
{
    if (err && cap > 0)
        {
            strncpy(err, msg.c_str(), cap - 1);
            err[cap - 1] = 0;
        }
    return code;
}
}  // namespace

extern "C" int gb200_second_peak_noalloc(const float* d_row, uint32_t n, uint32_t idx, uint32_t half_width, float* d_second, void* cuda_stream, char* err,
    size_t err_cap)
{
    if (!d_row || !d_second) return noalloc_fail(err, err_cap, GB200_E_INVALID, "gb200_second_peak_noalloc: d_row and d_second must be device pointers");
    if (n > (uint32_t)GB200_NOALLOC_MAX_ROW)
        return noalloc_fail(err, err_cap, GB200_E_RANGE,
            "gb200_second_peak_noalloc: row of " + std::to_string(n) + " floats exceeds GB200_NOALLOC_MAX_ROW = " + std::to_string(GB200_NOALLOC_MAX_ROW) +
                " (the fixed device scratch of this entry point; use gb200_acq_second_peak for longer rows)");
    if (n == 0 || idx >= n || half_width >= n)
        return noalloc_fail(err, err_cap, GB200_E_INVALID, "gb200_second_peak_noalloc: need 0 < n, idx < n, half_width < n");
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return noalloc_fail(err, err_cap, GB200_E_NO_DEVICE, std::string("gb200_second_peak_noalloc: cudaGetDevice: ") + cudaGetErrorString(e));
    if (dev < 0 || dev >= 64)
        return noalloc_fail(err, err_cap, GB200_E_UNSUPPORTED, "gb200_second_peak_noalloc: device ordinal " + std::to_string(dev) + " outside the 64 this entry point keeps an address for");
    float* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_noalloc_mu);
        if (!g_noalloc_ptr[dev])
            {
                // once per device, with that device current
                e = cudaGetSymbolAddress((void**)&g_noalloc_ptr[dev], gb_noalloc_row);
                if (e != cudaSuccess)
                    {
                        g_noalloc_ptr[dev] = nullptr;
                        return noalloc_fail(err, err_cap, GB200_E_CUDA, std::string("gb200_second_peak_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
                    }
            }
        scratch = g_noalloc_ptr[dev];
    }
    gb_second_peak_kernel<<<1, 1024, 0, (cudaStream_t)cuda_stream>>>(d_row, (int)n, scratch, (int)n, (int)idx, (int)half_width, d_second, 1);
    e = cudaGetLastError();
    if (e != cudaSuccess) return noalloc_fail(err, err_cap, GB200_E_CUDA, std::string("gb200_second_peak_noalloc: launch: ") + cudaGetErrorString(e));
    return GB200_OK;
}
