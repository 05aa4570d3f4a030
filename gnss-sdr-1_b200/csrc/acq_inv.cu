// Inverse side of the PCPS acquisition search (x conj(code DFT), inverse DFT, |.|^2, arg-max), once per (PRN, Doppler bin).
#include "acq_fft.cuh"

namespace
{

// -> * W_N^-(n2 k1) -> rowbuf[t][k1][n2].      grid = (ceil(N1 / TILE), transforms in chunk)
template <int LOGT, class FFT>
__global__ void __launch_bounds__(GB200_ACQ_THREADS)
    gb_acq_inv_row(PlanDev pl, GbAcqInvParams p)
{
    constexpr int TILE = 1 << LOGT;
    constexpr int PITCH = TILE + 1;
    extern __shared__ __align__(16) float2 smem[];
    const int M = FFT::kM ? FFT::kM : pl.N2, N1 = pl.N1;
    float2 *bufA, *bufB, *tw, *twA;
    carve<LOGT>(smem, M, bufA, bufB, tw, twA);
    for (int i = threadIdx.x; i < M; i += blockDim.x) tw[i] = __ldg(pl.tw + i);
    const int r0 = blockIdx.x * TILE;
    const int nb = min(TILE, N1 - r0);
    fill_twA(twA, M, r0, pl);
    const int t = p.first_transform + blockIdx.y;
    const int handle = t / p.n_bins, bin = t - handle * p.n_bins;
    const float2* xs = p.spectra + (size_t)bin * pl.N + (size_t)r0 * M;
    const float2* cs = p.code_fft[handle] + (size_t)r0 * M;
    const int db = (int)blockDim.x / M, dm = (int)blockDim.x % M;
    {
        int b = threadIdx.x / M, m = threadIdx.x - b * M;
        for (int idx = threadIdx.x; idx < M * TILE; idx += blockDim.x)
            {
                bufA[m * PITCH + b] = b < nb ? cmul(__ldg(xs + idx), __ldg(cs + idx)) : make_float2(0.f, 0.f);
                m += dm;
                b += db;
                if (m >= M)
                    {
                        m -= M;
                        b++;
                    }
            }
    }
    __syncthreads();
    float2* res = FFT::template run<+1, LOGT>(bufA, bufB, pl, tw);
    float2* dst = p.rowbuf + (size_t)blockIdx.y * pl.N + (size_t)r0 * M;
    {
        int b = threadIdx.x / M, m = threadIdx.x - b * M;
        for (int idx = threadIdx.x; idx < M * nb; idx += blockDim.x)
            {
                const float2 w = cmul(twA[m], __ldg(pl.twN + m * b));  // W_N^(n2 r0) * W_N^(n2 b); the inverse needs its conjugate
                dst[idx] = cmul_conj(res[m * PITCH + b], w);
                m += dm;
                b += db;
                if (m >= M)
                    {
                        m -= M;
                        b++;
                    }
            }
    }
}

__device__ __forceinline__ bool peak_better(float m, unsigned i, float bm, unsigned bi)
{
    return (m > bm) || (m == bm && i < bi);
}

// Column kernel, inverse: rowbuf[t][k1][n2] -> inverse DFT over k1 (length N1) -> y[N2*n1 + n2] -> |y|^2
// -> either per-CTA (max, first index) or the magnitude grid.   grid = (ceil(N2 / TILE), transforms in chunk)
template <int LOGT, class FFT>
__global__ void __launch_bounds__(GB200_ACQ_THREADS)
    gb_acq_inv_col(PlanDev pl, GbAcqInvParams p)
{
    constexpr int TILE = 1 << LOGT;
    constexpr int PITCH = TILE + 1;
    extern __shared__ __align__(16) float2 smem[];
    __shared__ float s_mag[GB200_ACQ_THREADS / 32];
    __shared__ unsigned s_idx[GB200_ACQ_THREADS / 32];
    const int M = FFT::kM ? FFT::kM : pl.N1, N2 = pl.N2;
    float2 *bufA, *bufB, *tw, *twA;
    carve<LOGT>(smem, M, bufA, bufB, tw, twA);
    for (int i = threadIdx.x; i < M; i += blockDim.x) tw[i] = __ldg(pl.tw + i);
    const int c0 = blockIdx.x * TILE;
    const float2* src = p.rowbuf + (size_t)blockIdx.y * pl.N;
    for (int idx = threadIdx.x; idx < (M << LOGT); idx += blockDim.x)
        {
            const int k1 = idx >> LOGT;
            const int b = idx & (TILE - 1);
            bufA[k1 * PITCH + b] = (c0 + b < N2) ? src[(size_t)k1 * N2 + c0 + b] : make_float2(0.f, 0.f);
        }
    __syncthreads();
    float2* res = FFT::template run<+1, LOGT>(bufA, bufB, pl, tw);

    const int t = p.first_transform + blockIdx.y;
    const int handle = t / p.n_bins, bin = t - handle * p.n_bins;
    float* grid = p.grid ? p.grid[handle] : nullptr;
    float best = -1.0f;
    unsigned best_i = 0xffffffffu;
    for (int idx = threadIdx.x; idx < (M << LOGT); idx += blockDim.x)
        {
            const int n1 = idx >> LOGT;
            const int b = idx & (TILE - 1);
            const int n = N2 * n1 + c0 + b - p.keep_offset;
            if (c0 + b >= N2 || n < 0 || n >= p.keep_len) continue;
            const float2 y = res[n1 * PITCH + b];
            const float mag = __fadd_rn(__fmul_rn(y.x, y.x), __fmul_rn(y.y, y.y));  // volk_32fc_magnitude_squared_32f
            if (grid)
                {
                    float* g = grid + (size_t)bin * p.keep_len + n;
                    *g = p.accumulate ? __fadd_rn(*g, mag) : mag;
                    if (p.accumulate && p.tmpbuf && p.tmpbuf[handle] && bin == p.n_bins - 1) p.tmpbuf[handle][n] = mag;
                }
            else if (peak_better(mag, (unsigned)n, best, best_i))
                {
                    best = mag;
                    best_i = (unsigned)n;
                }
        }
    if (grid) return;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        {
            const float om = __shfl_down_sync(0xffffffffu, best, o);
            const unsigned oi = __shfl_down_sync(0xffffffffu, best_i, o);
            if (peak_better(om, oi, best, best_i))
                {
                    best = om;
                    best_i = oi;
                }
        }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0)
        {
            s_mag[warp] = best;
            s_idx[warp] = best_i;
        }
    __syncthreads();
    if (threadIdx.x == 0)
        {
            for (int k = 1; k < GB200_ACQ_THREADS / 32; k++)
                if (peak_better(s_mag[k], s_idx[k], best, best_i))
                    {
                        best = s_mag[k];
                        best_i = s_idx[k];
                    }
            GbPeak pk;
            pk.mag = best;
            pk.idx = best_i;
            pk.bin = (unsigned)bin;
            pk.pad_ = 0;
            p.partial[(size_t)t * gridDim.x + blockIdx.x] = pk;
        }
}
}  // namespace

cudaError_t gb_acq_inverse(const GbFftPlan& plan, const GbAcqInvParams& p, cudaStream_t st, int* launches, void* kt)
{
    cudaError_t e;
    {
        const PlanDev pd = make_plan_dev(plan, false);
        const size_t sm = gb_acq_row_smem(plan);
        const dim3 grid((plan.N1 + plan.tile2 - 1) / plan.tile2, p.n_transforms);
        if (kt) gb_kt_begin(kt, GB200_KCLASS_ACQ_INV_ROW);
        e = gb_with_fft(plan.N2, log2_tile(plan.tile2), [&](auto fft, auto lt) -> cudaError_t {
            using FFT = decltype(fft);
            constexpr int LT = decltype(lt)::value;
            cudaError_t r = set_smem(gb_acq_inv_row<LT, FFT>, sm);
            if (r == cudaSuccess) gb_acq_inv_row<LT, FFT><<<grid, GB200_ACQ_THREADS, sm, st>>>(pd, p);
            return r;
        });
        if (kt) gb_kt_end(kt);
        if (e != cudaSuccess) return e;
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        if (launches) (*launches)++;
    }
    {
        const PlanDev pd = make_plan_dev(plan, true);
        const size_t sm = gb_acq_col_smem(plan);
        const dim3 grid((plan.N2 + plan.tile1 - 1) / plan.tile1, p.n_transforms);
        if (kt) gb_kt_begin(kt, GB200_KCLASS_ACQ_INV_COL);
        e = gb_with_fft(plan.N1, log2_tile(plan.tile1), [&](auto fft, auto lt) -> cudaError_t {
            using FFT = decltype(fft);
            constexpr int LT = decltype(lt)::value;
            cudaError_t r = set_smem(gb_acq_inv_col<LT, FFT>, sm);
            if (r == cudaSuccess) gb_acq_inv_col<LT, FFT><<<grid, GB200_ACQ_THREADS, sm, st>>>(pd, p);
            return r;
        });
        if (kt) gb_kt_end(kt);
        if (e != cudaSuccess) return e;
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        if (launches) (*launches)++;
    }
    return cudaSuccess;
}
