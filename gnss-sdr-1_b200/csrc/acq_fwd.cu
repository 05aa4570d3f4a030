// Forward side of the PCPS acquisition search (wipe-off + DFT of the sample block, once per Doppler bin).
#include "acq_fft.cuh"

namespace
{


// ---------------------------------------------------------------------------------------------
// Column kernel, forward: x[N2*n1 + n2] * wipeoff -> DFT over n1 (length N1) -> * W_N^(n2 k1) -> colbuf[k1][n2]
// grid = (ceil(N2 / TILE), bins)
template <int ITEM, int LOGT, class FFT>
__global__ void __launch_bounds__(GB200_ACQ_THREADS)
    gb_acq_fwd_col(PlanDev pl, GbAcqFwdParams p, int wipeoff)
{
    constexpr int TILE = 1 << LOGT;
    constexpr int PITCH = TILE + 1;
    extern __shared__ __align__(16) float2 smem[];
    const int M = FFT::kM ? FFT::kM : pl.N1, N2 = pl.N2;
    float2 *bufA, *bufB, *tw, *twA;
    carve<LOGT>(smem, M, bufA, bufB, tw, twA);
    for (int i = threadIdx.x; i < M; i += blockDim.x) tw[i] = __ldg(pl.tw + i);
    const int c0 = blockIdx.x * TILE;
    fill_twA(twA, M, c0, pl);
    const int bin = p.first_bin + blockIdx.y;
    const GbWipeSeg* segs = nullptr;
    int nseg = 0;
    if (wipeoff)
        {
            const int o0 = __ldg(p.seg_offsets + bin);
            nseg = __ldg(p.seg_offsets + bin + 1) - o0;
            segs = p.segs + o0;
        }
    int lo = 0;
    for (int idx = threadIdx.x; idx < (M << LOGT); idx += blockDim.x)
        {
            const int n1 = idx >> LOGT;
            const int b = idx & (TILE - 1);
            const int n2 = c0 + b;
            const int n = N2 * n1 + n2;
            float2 x = make_float2(0.f, 0.f);
            // fold == 1: one wiped sample.  QuickSync (fold > 1): the wiped samples n, n + N, n + 2N, ... added in that order onto a
            // zeroed buffer in float32, as the block's std::transform(std::plus) passes do
            for (int f = 0; f < p.fold; f++)
                {
                    const int m = n + f * pl.N;
                    if (!(n2 < N2 && m < p.consumed)) break;
                    float2 v = gb_load_item<ITEM>(p.samples, (size_t)m);
                    if (wipeoff)
                        {
                            // last segment with n_start <= m.  A thread's samples come in increasing m (3200 apart at fold == 1), a bin
                            // has ~10 segments: walking from the previous sample's segment is one comparison almost always
                            while (lo + 1 < nseg && __ldg(&segs[lo + 1].n_start) <= m) lo++;
                            while (lo > 0 && __ldg(&segs[lo].n_start) > m) lo--;
                            const GbWipeSeg sg = segs[lo];
                            const float ph = (float)fma((double)(m - sg.n_start), (double)sg.inc_eff, (double)sg.ph_start);  // exact
                            float s, c;
                            sincosf(ph, &s, &c);
                            v = cmul(v, make_float2(c, s));
                        }
                    x.x = __fadd_rn(x.x, v.x);
                    x.y = __fadd_rn(x.y, v.y);
                }
            bufA[n1 * PITCH + b] = x;
        }
    __syncthreads();
    float2* res = FFT::template run<-1, LOGT>(bufA, bufB, pl, tw);
    float2* dst = p.colbuf + (size_t)bin * pl.N;
    for (int idx = threadIdx.x; idx < (M << LOGT); idx += blockDim.x)
        {
            const int k1 = idx >> LOGT;
            const int b = idx & (TILE - 1);
            const int n2 = c0 + b;
            if (n2 < N2)
                {
                    const float2 w = cmul(twA[k1], __ldg(pl.twN + k1 * b));  // W_N^(k1 c0) * W_N^(k1 b)
                    dst[(size_t)k1 * N2 + n2] = cmul(res[k1 * PITCH + b], w);
                }
        }
}

// Row kernel, forward: colbuf[k1][n2] -> DFT over n2 (length N2) -> spectra[k1][k2]
// grid = (ceil(N1 / TILE), bins)
template <int LOGT, class FFT>
__global__ void __launch_bounds__(GB200_ACQ_THREADS)
    gb_acq_fwd_row(PlanDev pl, GbAcqFwdParams p)
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
    const float2* src = p.colbuf + (size_t)(p.first_bin + blockIdx.y) * pl.N + (size_t)r0 * M;
    const int db = (int)blockDim.x / M, dm = (int)blockDim.x % M;
    {
        int b = threadIdx.x / M, m = threadIdx.x - b * M;
        for (int idx = threadIdx.x; idx < M * TILE; idx += blockDim.x)
            {
                bufA[m * PITCH + b] = b < nb ? src[idx] : make_float2(0.f, 0.f);
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
    float2* res = FFT::template run<-1, LOGT>(bufA, bufB, pl, tw);
    float2* dst = p.spectra + (size_t)(p.first_bin + blockIdx.y) * pl.N + (size_t)r0 * M;
    {
        int b = threadIdx.x / M, m = threadIdx.x - b * M;
        for (int idx = threadIdx.x; idx < M * nb; idx += blockDim.x)
            {
                dst[idx] = res[m * PITCH + b];
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

// Row kernel, inverse: spectra[bin][k1][k2] * codefft[handle][k1][k2] -> inverse DFT over k2 (length N2)
}  // namespace

cudaError_t gb_acq_forward(const GbFftPlan& plan, const GbAcqFwdParams& p, int n_bins, bool wipeoff, cudaStream_t st, int* launches, void* kt)
{
    cudaError_t e;
    // the compile-time DIF tiles where the split has them (acq_inv.cu); the run-time parametrised kernels below otherwise
    if (plan.inv_fast && !plan.fwd_generic)
        {
            e = gb_acq_forward_dif(plan, p, n_bins, wipeoff, st, launches, kt);
            if (e != cudaErrorNotSupported) return e;
            (void)cudaGetLastError();
        }
    {
        const PlanDev pd = make_plan_dev(plan, true);
        const size_t sm = gb_acq_col_smem(plan);
        const dim3 grid((plan.N2 + plan.tile1 - 1) / plan.tile1, n_bins);
        const int wo = wipeoff ? 1 : 0;
        if (kt) gb_kt_begin(kt, GB200_KCLASS_ACQ_FWD_COL);
        e = gb_with_fft(plan.N1, log2_tile(plan.tile1), [&](auto fft, auto lt) -> cudaError_t {
            using FFT = decltype(fft);
            constexpr int LT = decltype(lt)::value;
            cudaError_t r = cudaErrorInvalidValue;
            switch (p.item_type)
                {
                case GB200_ITEM_32FC:
                    if ((r = set_smem(gb_acq_fwd_col<GB200_ITEM_32FC, LT, FFT>, sm)) == cudaSuccess)
                        gb_acq_fwd_col<GB200_ITEM_32FC, LT, FFT><<<grid, GB200_ACQ_THREADS, sm, st>>>(pd, p, wo);
                    break;
                case GB200_ITEM_16IC:
                    if ((r = set_smem(gb_acq_fwd_col<GB200_ITEM_16IC, LT, FFT>, sm)) == cudaSuccess)
                        gb_acq_fwd_col<GB200_ITEM_16IC, LT, FFT><<<grid, GB200_ACQ_THREADS, sm, st>>>(pd, p, wo);
                    break;
                case GB200_ITEM_8IC:
                    if ((r = set_smem(gb_acq_fwd_col<GB200_ITEM_8IC, LT, FFT>, sm)) == cudaSuccess)
                        gb_acq_fwd_col<GB200_ITEM_8IC, LT, FFT><<<grid, GB200_ACQ_THREADS, sm, st>>>(pd, p, wo);
                    break;
                default: break;
                }
            return r;
        });
        if (kt) gb_kt_end(kt);
        if (e != cudaSuccess) return e;
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        if (launches) (*launches)++;
    }
    {
        const PlanDev pd = make_plan_dev(plan, false);
        const size_t sm = gb_acq_row_smem(plan);
        const dim3 grid((plan.N1 + plan.tile2 - 1) / plan.tile2, n_bins);
        if (kt) gb_kt_begin(kt, GB200_KCLASS_ACQ_FWD_ROW);
        e = gb_with_fft(plan.N2, log2_tile(plan.tile2), [&](auto fft, auto lt) -> cudaError_t {
            using FFT = decltype(fft);
            constexpr int LT = decltype(lt)::value;
            cudaError_t r = set_smem(gb_acq_fwd_row<LT, FFT>, sm);
            if (r == cudaSuccess) gb_acq_fwd_row<LT, FFT><<<grid, GB200_ACQ_THREADS, sm, st>>>(pd, p);
            return r;
        });
        if (kt) gb_kt_end(kt);
        if (e != cudaSuccess) return e;
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        if (launches) (*launches)++;
    }
    return cudaSuccess;
}
