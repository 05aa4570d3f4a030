// Batched multicorrelator for sm_100a: carrier NCO wipe-off + nearest-neighbour code resampling +
// multi-tap complex accumulation, every (channel, epoch) request of a batch in ONE launch.
//
// Replaces (semantics, not code):
//   Cpu_Multicorrelator_Real_Codes::Carrier_wipeoff_multicorrelator_resampler
//       src/algorithms/tracking/libs/cpu_multicorrelator_real_codes.cc:101-152
//   volk_gnsssdr_32f_xn_resampler_32f_xn            (chip index rule)   ...resampler_32f_xn.h:77-95
//   volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn  (wipe-off + MACs)   ...rotator_dot_prod_32fc_xn.h:81-113
//   Cpu_Multicorrelator (complex codes)             cpu_multicorrelator.cc:99-126
//
// Parity rules:
//  * chip index = floor( fl(fl(fl(step*(float)n) + shift) - rem) ) in float32, this order, NO FMA
//    (__fmul_rn/__fadd_rn/__fsub_rn), then wrapped to [0, L) -- bit-identical to the reference _generic.
//  * carrier phase: the reference recurs phase *= inc sequentially in float32.  Every thread here seeds
//    phase_0 * inc^n from float64 turn counts of the reference's own float32-rounded phase_0 and inc
//    (GbTrkWork::nco_*) and recurs 256 samples at a time, so the NCO frequency offset caused by rounding
//    inc to float32 is reproduced and only the reference's random-walk rounding (~sqrt(n)*6e-8) remains.
//
// Layout: grid = requests x chunks (1-D); a CTA owns `chunk` consecutive samples of one request.  The slice
// of the code table those samples can touch is staged once into shared memory already wrapped
// ("window"), so the hot loop does no modulo: index = float-magic floor - window base.
#include "gb200_internal.cuh"

namespace
{

constexpr float kFloorMagic = 12582912.0f;  // 1.5 * 2^23: fl_down(v + magic) has floor(v) in its low mantissa bits
constexpr int kFloorMagicBits = 0x4B400000;

__device__ __forceinline__ float chip_arg(float step, float nf, float shift, float rem)
{
    return __fsub_rn(__fadd_rn(__fmul_rn(step, nf), shift), rem);
}

__device__ __forceinline__ int wrap_index(int idx, int len)
{
    // same result as the reference's "if (<0) += L*(abs/L+1); %= L": the non-negative remainder
    int r = idx % len;
    return r < 0 ? r + len : r;
}

// shared-memory loads by 32-bit shared address: the window index is folded into ONE address computation
// (bits * 4 + (window base - 4 * wbase)) instead of a subtract plus a scaled add per tap
__device__ __forceinline__ float lds_f32(unsigned addr)
{
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ float2 lds_f32x2(unsigned addr)
{
    float2 v;
    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr));
    return v;
}

// PROMPT: compile-time index of the tap the data-component code rides on (-1: read GbTrkWork::prompt_tap at run time)
template <int ITEM, int NT, bool HAS_DATA, bool CPLX, int PROMPT>
__global__ void __launch_bounds__(GB200_TRK_THREADS)
    gb_trk_kernel(const GbTrkWork* __restrict__ work, int chunk, int max_chunks, float2* __restrict__ partials,
        unsigned* __restrict__ counters, float2* __restrict__ out, int nt_runtime)
{
    constexpr int NACC = NT + (HAS_DATA ? 1 : 0);
    constexpr int CW = CPLX ? 2 : 1;  // floats per code-table entry
    __shared__ __align__(16) float s_win[GB200_TRK_WMAX * (CPLX ? 2 : (HAS_DATA ? 2 : 1))];
    __shared__ float2 s_red[GB200_TRK_THREADS / 32][NACC];
    __shared__ bool s_last;

    const int cidx = blockIdx.x % max_chunks;  // 1-D grid: request-major, chunk-minor
    const GbTrkWork& w = work[blockIdx.x / max_chunks];
    const int n0 = cidx * chunk;
    if (n0 >= w.n) return;
    const int n1 = min(w.n, n0 + chunk);
    const int nt = (NT == GB200_MAX_TAPS) ? nt_runtime : NT;  // NT == MAX_TAPS instantiation is the generic one

    float shifts[NT];
#pragma unroll
    for (int t = 0; t < NT; t++) shifts[t] = t < nt ? w.shifts[t] : 0.0f;
    const float step = w.code_step, rem = w.rem_code;
    const int L = w.code_len;

    // ---- stage the code window ----
    float smin = shifts[0], smax = shifts[0];
#pragma unroll
    for (int t = 1; t < NT; t++)
        if (t < nt)
            {
                smin = fminf(smin, shifts[t]);
                smax = fmaxf(smax, shifts[t]);
            }
    const float vlo = chip_arg(step, (float)n0, smin, rem);
    const float vhi = chip_arg(step, (float)(n1 - 1), smax, rem);
    const bool window_ok = (step >= 0.0f) && (fabsf(vlo) < 4.0e6f) && (fabsf(vhi) < 4.0e6f) &&
                           ((int)floorf(vhi) - (int)floorf(vlo) + 1 <= GB200_TRK_WMAX);
    const int lo = window_ok ? (int)floorf(vlo) : 0;
    if (window_ok)
        {
            const int W = (int)floorf(vhi) - lo + 1;
            int src = wrap_index(lo + (int)threadIdx.x, L);
            const int adv = GB200_TRK_THREADS % L;
            for (int j = threadIdx.x; j < W; j += GB200_TRK_THREADS)
                {
                    if (CPLX)
                        {
                            s_win[2 * j] = __ldg(w.code + 2 * src);
                            s_win[2 * j + 1] = __ldg(w.code + 2 * src + 1);
                        }
                    else
                        {
                            s_win[j] = __ldg(w.code + src);
                            if (HAS_DATA) s_win[GB200_TRK_WMAX + j] = __ldg(w.data_code + src);
                        }
                    src += adv;
                    if (src >= L) src -= L;
                }
        }
    __syncthreads();

    // ---- carrier NCO: seed per thread, then advance by 256 samples per iteration ----
    int n = n0 + threadIdx.x;
    float2 ph = gb_phasor_turns(w.nco_a0_turns, w.nco_theta_turns, (double)n);
    const float2 adv256 = w.nco_adv;
    {
        // |inc| != 1 in float32: between two renormalisations (every 256 samples, chunk starts are multiples of 256)
        // the reference's phasor magnitude is |inc|^e, e = ((n-1) mod 256) + 1 -- constant per thread here
        const float m = fmaf((float)(((threadIdx.x + 255u) & 255u) + 1u), w.nco_mag_delta, 1.0f);
        ph.x *= m;
        ph.y *= m;
    }

    float2 acc[NACC];
#pragma unroll
    for (int t = 0; t < NACC; t++) acc[t] = make_float2(0.f, 0.f);

    const int wbase = kFloorMagicBits + lo;
    float nf = (float)n;  // exact: n < 2^24
    if (window_ok)
        {
            // shared address of window entry 0 minus the float-magic bias: entry(bits) = swin + bits * stride
            const unsigned swin = (unsigned)__cvta_generic_to_shared(s_win) - (unsigned)wbase * (CPLX ? 8u : 4u);
            const int prompt = PROMPT >= 0 ? PROMPT : w.prompt_tap;
#pragma unroll 4
            for (; n < n1; n += GB200_TRK_THREADS)
                {
                    const float2 x = gb_load_item<ITEM>(w.samples, (size_t)n);
                    const float wr = fmaf(x.x, ph.x, -x.y * ph.y);
                    const float wi = fmaf(x.x, ph.y, x.y * ph.x);
                    const float sn = __fmul_rn(step, nf);
                    unsigned ap = 0;  // shared address of the prompt tap's table entry
#pragma unroll
                    for (int t = 0; t < NT; t++)
                        {
                            if (t < nt)
                                {
                                    const float v = __fsub_rn(__fadd_rn(sn, shifts[t]), rem);
                                    const unsigned bits = (unsigned)__float_as_int(__fadd_rd(v, kFloorMagic));
                                    if (CPLX)
                                        {
                                            const float2 c = lds_f32x2(swin + bits * 8u);
                                            // (wr + j wi) * (c.x + j c.y), reference order: w * code
                                            acc[t].x += fmaf(wr, c.x, -wi * c.y);
                                            acc[t].y += fmaf(wr, c.y, wi * c.x);
                                        }
                                    else
                                        {
                                            const unsigned a = swin + bits * 4u;
                                            const float c = lds_f32(a);
                                            acc[t].x = fmaf(wr, c, acc[t].x);
                                            acc[t].y = fmaf(wi, c, acc[t].y);
                                            if (HAS_DATA && (PROMPT >= 0 ? t == PROMPT : t == prompt)) ap = a;
                                        }
                                }
                        }
                    if (HAS_DATA && !CPLX)
                        {
                            const float cd = lds_f32(ap + 4u * GB200_TRK_WMAX);  // the data table sits WMAX floats after the pilot's
                            acc[NT].x = fmaf(wr, cd, acc[NT].x);
                            acc[NT].y = fmaf(wi, cd, acc[NT].y);
                        }
                    ph = gb_cmul(ph, adv256);
                    nf += (float)GB200_TRK_THREADS;
                }
        }
    else
        {
            // general path (step < 0, |index| huge, or window larger than shared memory): table in L2, full modulo
            for (; n < n1; n += GB200_TRK_THREADS)
                {
                    const float2 x = gb_load_item<ITEM>(w.samples, (size_t)n);
                    const float wr = fmaf(x.x, ph.x, -x.y * ph.y);
                    const float wi = fmaf(x.x, ph.y, x.y * ph.x);
                    const float sn = __fmul_rn(step, nf);
#pragma unroll
                    for (int t = 0; t < NT; t++)
                        {
                            if (t < nt)
                                {
                                    const float v = __fsub_rn(__fadd_rn(sn, shifts[t]), rem);
                                    const int j = wrap_index((int)floor((double)v), L);
                                    if (CPLX)
                                        {
                                            const float2 c = __ldg(reinterpret_cast<const float2*>(w.code) + j);
                                            acc[t].x += fmaf(wr, c.x, -wi * c.y);
                                            acc[t].y += fmaf(wr, c.y, wi * c.x);
                                        }
                                    else
                                        {
                                            const float c = __ldg(w.code + j);
                                            acc[t].x = fmaf(wr, c, acc[t].x);
                                            acc[t].y = fmaf(wi, c, acc[t].y);
                                            if (HAS_DATA && t == w.prompt_tap)
                                                {
                                                    const float cd = __ldg(w.data_code + j);
                                                    acc[NT].x = fmaf(wr, cd, acc[NT].x);
                                                    acc[NT].y = fmaf(wi, cd, acc[NT].y);
                                                }
                                        }
                                }
                        }
                    ph = gb_cmul(ph, adv256);
                    nf += (float)GB200_TRK_THREADS;
                }
        }
    (void)CW;

    // ---- CTA reduction: warp shuffles, then one shared-memory stage ----
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int t = 0; t < NACC; t++)
        {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1)
                {
                    acc[t].x += __shfl_down_sync(0xffffffffu, acc[t].x, o);
                    acc[t].y += __shfl_down_sync(0xffffffffu, acc[t].y, o);
                }
            if (lane == 0) s_red[warp][t] = acc[t];
        }
    __syncthreads();
    const int nchunks = (w.n + chunk - 1) / chunk;
    float2* my_part = partials + ((size_t)w.out_index * max_chunks + cidx) * GB200_OUT_STRIDE;
    if (threadIdx.x < NACC)
        {
            float2 s = make_float2(0.f, 0.f);
#pragma unroll
            for (int k = 0; k < GB200_TRK_THREADS / 32; k++)
                {
                    s.x += s_red[k][threadIdx.x].x;
                    s.y += s_red[k][threadIdx.x].y;
                }
            if (nchunks == 1)
                out[(size_t)w.out_index * GB200_OUT_STRIDE + threadIdx.x] = s;
            else
                my_part[threadIdx.x] = s;
        }
    if (nchunks == 1) return;

    // ---- last CTA of the request sums the chunk partials in chunk order (deterministic) ----
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0)
        {
            unsigned prev = atomicAdd(&counters[w.out_index], 1u);
            s_last = (prev == (unsigned)nchunks - 1u);
        }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (threadIdx.x < NACC)
        {
            float2 s = make_float2(0.f, 0.f);
            const volatile float2* p = partials + (size_t)w.out_index * max_chunks * GB200_OUT_STRIDE + threadIdx.x;
            for (int c = 0; c < nchunks; c++)
                {
                    s.x += p[(size_t)c * GB200_OUT_STRIDE].x;
                    s.y += p[(size_t)c * GB200_OUT_STRIDE].y;
                }
            out[(size_t)w.out_index * GB200_OUT_STRIDE + threadIdx.x] = s;
        }
    if (threadIdx.x == 0) counters[w.out_index] = 0u;  // self-cleaning for the next launch
}

// High-dynamics variant (high_dyn=true): quadratic code and carrier terms with the reference's exact
// quirks -- unsigned n*n wrap, taps >= 1 as circular SAMPLE shifts of tap 0
// (volk_gnsssdr_32f_xn_high_dynamics_resampler_32f_xn.h:81-106) and the rate phasor applied one
// sample late, rate^((n-1)^2) (volk_gnsssdr_32fc_32f_high_dynamic_rotator_dot_prod_32fc_xn.h:82-116).
// Rare path: one accurate phasor per sample, code table read through L1/L2.
template <int ITEM>
__global__ void __launch_bounds__(GB200_TRK_THREADS)
    gb_trk_hd_kernel(const GbTrkWork* __restrict__ work, int chunk, int max_chunks, float2* __restrict__ partials,
        unsigned* __restrict__ counters, float2* __restrict__ out)
{
    __shared__ float2 s_red[GB200_TRK_THREADS / 32][GB200_OUT_STRIDE];
    __shared__ bool s_last;
    const int cidx = blockIdx.x % max_chunks;  // 1-D grid: request-major, chunk-minor
    const GbTrkWork& w = work[blockIdx.x / max_chunks];
    const int n0 = cidx * chunk;
    if (n0 >= w.n) return;
    const int n1 = min(w.n, n0 + chunk);
    const int nt = w.n_taps;
    const bool has_data = w.data_code != nullptr;
    const int nacc = nt + (has_data ? 1 : 0);
    const int L = w.code_len;
    // cumulative sample shifts of taps >= 1
    unsigned sshift[GB200_MAX_TAPS];
    sshift[0] = 0;
    for (int t = 1; t < GB200_MAX_TAPS; t++)
        sshift[t] = t < nt ? sshift[t - 1] + (unsigned)(int)round((double)__fdiv_rn(__fsub_rn(w.shifts[t], w.shifts[t - 1]), w.code_step)) : 0u;

    float2 acc[GB200_OUT_STRIDE];
    for (int t = 0; t < GB200_OUT_STRIDE; t++) acc[t] = make_float2(0.f, 0.f);
    const double inv2pi = 0.15915494309189533576888376337251;
    for (int n = n0 + threadIdx.x; n < n1; n += GB200_TRK_THREADS)
        {
            const float2 x = gb_load_item<ITEM>(w.samples, (size_t)n);
            double tt = fma((double)n, w.nco_theta_turns, w.nco_a0_turns);
            if (n >= 1)
                {
                    const unsigned m = (unsigned)(n - 1);
                    tt -= (double)(float)(m * m) * (double)w.phase_rate * inv2pi;
                }
            tt -= rint(tt);
            float s, c;
            sincospif(2.0f * (float)tt, &s, &c);
            const float g = expf((float)((double)n * w.nco_log_mag));  // phase_doppler is never renormalised in the reference
            s *= g;
            c *= g;
            const float wr = fmaf(x.x, c, -x.y * s);
            const float wi = fmaf(x.x, s, x.y * c);
            for (int t = 0; t < nt; t++)
                {
                    unsigned k = (unsigned)n + sshift[t];
                    if (k >= (unsigned)w.n) k -= (unsigned)w.n;
                    const float v = __fsub_rn(__fadd_rn(__fadd_rn(__fmul_rn(w.code_step, (float)k), __fmul_rn(w.code_rate, (float)(k * k))), w.shifts[0]), w.rem_code);
                    const int j = wrap_index((int)floor((double)v), L);
                    const float cc = __ldg(w.code + j);
                    acc[t].x = fmaf(wr, cc, acc[t].x);
                    acc[t].y = fmaf(wi, cc, acc[t].y);
                    if (has_data && t == w.prompt_tap)
                        {
                            const float cd = __ldg(w.data_code + j);
                            acc[nt].x = fmaf(wr, cd, acc[nt].x);
                            acc[nt].y = fmaf(wi, cd, acc[nt].y);
                        }
                }
        }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int t = 0; t < GB200_OUT_STRIDE; t++)
        {
            for (int o = 16; o > 0; o >>= 1)
                {
                    acc[t].x += __shfl_down_sync(0xffffffffu, acc[t].x, o);
                    acc[t].y += __shfl_down_sync(0xffffffffu, acc[t].y, o);
                }
            if (lane == 0) s_red[warp][t] = acc[t];
        }
    __syncthreads();
    const int nchunks = (w.n + chunk - 1) / chunk;
    if ((int)threadIdx.x < nacc)
        {
            float2 s = make_float2(0.f, 0.f);
            for (int k = 0; k < GB200_TRK_THREADS / 32; k++)
                {
                    s.x += s_red[k][threadIdx.x].x;
                    s.y += s_red[k][threadIdx.x].y;
                }
            if (nchunks == 1)
                out[(size_t)w.out_index * GB200_OUT_STRIDE + threadIdx.x] = s;
            else
                partials[((size_t)w.out_index * max_chunks + cidx) * GB200_OUT_STRIDE + threadIdx.x] = s;
        }
    if (nchunks == 1) return;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = (atomicAdd(&counters[w.out_index], 1u) == (unsigned)nchunks - 1u);
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if ((int)threadIdx.x < nacc)
        {
            float2 s = make_float2(0.f, 0.f);
            const volatile float2* p = partials + (size_t)w.out_index * max_chunks * GB200_OUT_STRIDE + threadIdx.x;
            for (int c = 0; c < nchunks; c++)
                {
                    s.x += p[(size_t)c * GB200_OUT_STRIDE].x;
                    s.y += p[(size_t)c * GB200_OUT_STRIDE].y;
                }
            out[(size_t)w.out_index * GB200_OUT_STRIDE + threadIdx.x] = s;
        }
    if (threadIdx.x == 0) counters[w.out_index] = 0u;
}

template <int ITEM>
cudaError_t launch_item(const GbTrkWork* d_work, dim3 grid, int n_taps, bool has_data, int prompt_tap, int code_kind, bool high_dyn,
    int chunk, int max_chunks, float2* d_partials, unsigned* d_counters, float2* d_out, cudaStream_t st)
{
    dim3 block(GB200_TRK_THREADS);
#define GB_LAUNCH(NT, HD, CP) gb_trk_kernel<ITEM, NT, HD, CP, -1><<<grid, block, 0, st>>>(d_work, chunk, max_chunks, d_partials, d_counters, d_out, n_taps)
#define GB_LAUNCH_P(NT, P) gb_trk_kernel<ITEM, NT, true, false, P><<<grid, block, 0, st>>>(d_work, chunk, max_chunks, d_partials, d_counters, d_out, n_taps)
    if (high_dyn)
        gb_trk_hd_kernel<ITEM><<<grid, block, 0, st>>>(d_work, chunk, max_chunks, d_partials, d_counters, d_out);
    else if (code_kind == GB200_CODE_CPLX)
        GB_LAUNCH(GB200_MAX_TAPS, false, true);
    else if (n_taps == 1 && !has_data)
        GB_LAUNCH(1, false, false);
    else if (n_taps == 3 && !has_data)
        GB_LAUNCH(3, false, false);
    else if (n_taps == 3 && has_data && prompt_tap == 1)
        GB_LAUNCH_P(3, 1);
    else if (n_taps == 3 && has_data)
        GB_LAUNCH(3, true, false);
    else if (n_taps == 5 && !has_data)
        GB_LAUNCH(5, false, false);
    else if (n_taps == 5 && has_data && prompt_tap == 2)
        GB_LAUNCH_P(5, 2);
    else if (n_taps == 5 && has_data)
        GB_LAUNCH(5, true, false);
    else if (has_data)
        GB_LAUNCH(GB200_MAX_TAPS, true, false);
    else
        GB_LAUNCH(GB200_MAX_TAPS, false, false);
#undef GB_LAUNCH
#undef GB_LAUNCH_P
    return cudaGetLastError();
}

}  // namespace

cudaError_t gb_trk_launch(const GbTrkWork* d_work, int n_work, int max_n, int n_taps, bool has_data, int prompt_tap, int item_type,
    int code_kind, bool high_dyn, float2* d_partials, unsigned* d_counters, float2* d_out, int chunk, cudaStream_t st,
    int* launches)
{
    if (n_work <= 0) return cudaSuccess;
    const int max_chunks = (max_n + chunk - 1) / chunk;
    dim3 grid((unsigned)max_chunks * (unsigned)n_work);
    cudaError_t e;
    switch (item_type)
        {
        case GB200_ITEM_32FC:
            e = launch_item<GB200_ITEM_32FC>(d_work, grid, n_taps, has_data, prompt_tap, code_kind, high_dyn, chunk, max_chunks, d_partials, d_counters, d_out, st);
            break;
        case GB200_ITEM_16IC:
            e = launch_item<GB200_ITEM_16IC>(d_work, grid, n_taps, has_data, prompt_tap, code_kind, high_dyn, chunk, max_chunks, d_partials, d_counters, d_out, st);
            break;
        case GB200_ITEM_8IC:
            e = launch_item<GB200_ITEM_8IC>(d_work, grid, n_taps, has_data, prompt_tap, code_kind, high_dyn, chunk, max_chunks, d_partials, d_counters, d_out, st);
            break;
        default:
            return cudaErrorInvalidValue;
        }
    if (launches) (*launches)++;
    return e;
}
