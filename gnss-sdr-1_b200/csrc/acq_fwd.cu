// Forward side of the PCPS acquisition search (wipe-off + DFT of the sample block, once per Doppler bin).
#include "acq_fft.cuh"
#include "acq_dif.cuh"
#include "acq_pairs.cuh"

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

// ---------------------------------------------------------------------------------------------------------------
// Forward side on the same in-place DIF tiles (the sample block's spectrum once per Doppler bin, and the local code's).
//
// columns: x[N2 n1 + n2] x wipe-off -> DFT over n1 -> x W_N^(n2 k1) -> colbuf[k1][n2]   (tile of columns c0 .. c0+TILE)
template <int ITEM, int LOGT, int NT, int OCC, class DCOL, int N2>
__global__ void __launch_bounds__(NT, OCC)
    gb_acq_fwd_col_dif(PlanDev pl, GbAcqFwdParams p, int wipeoff)
{
    extern __shared__ __align__(16) float2 smem[];
    constexpr int M = DCOL::kM, TILE = 1 << LOGT, N = M * N2;
    constexpr int RL = DCOL::kLastR, NIL = DCOL::kLastNI;
    using L = GbTile<LOGT, NT>;
    constexpr int IL = 8, IH = (NIL + IL - 1) / IL;  // last-pass item i = IL * ih + il
    float2* buf = smem;
    float2* tw = buf + L::size(M);
    float2* step = tw + M;          // [TILE]      W_N^(n2 NIL): ratio between consecutive outputs of one item
    float2* tlo = step + TILE;      // [TILE][IL]  W_N^(n2 il)
    float2* thi = tlo + TILE * IL;  // [TILE][IH]  W_N^(n2 IL ih)      (n2 = c0 + b)
    const int c0 = blockIdx.x << LOGT;
    const int bin = p.first_bin + blockIdx.y;
    for (int i = threadIdx.x; i < M; i += NT) tw[i] = __ldg(pl.tw + i);
    for (int q = threadIdx.x; q < TILE * (1 + IL + IH); q += NT)
        {
            const int b = q / (1 + IL + IH), c = q - b * (1 + IL + IH);
            const unsigned n2 = (unsigned)(c0 + b);
            const unsigned e = (c == 0 ? n2 * (unsigned)NIL : (c <= IL ? n2 * (unsigned)(c - 1) : n2 * (unsigned)(IL * (c - 1 - IL)))) % (unsigned)N;
            const float2 w = cmul(__ldg(pl.twN + (e & 255u)), __ldg(pl.twNhi + (e >> 8)));
            if (c == 0)
                step[b] = w;
            else if (c <= IL)
                tlo[b * IL + c - 1] = w;
            else
                thi[b * IH + c - 1 - IL] = w;
        }
    const GbWipeSeg* segs = nullptr;
    int nseg = 0;
    if (wipeoff)
        {
            const int o0 = __ldg(p.seg_offsets + bin);
            nseg = __ldg(p.seg_offsets + bin + 1) - o0;
            segs = p.segs + o0;
        }
    // Stage the wiped tile in shared memory first, every thread with all its loads in flight at once: legs fetched one by one
    // inside the first butterfly pass (each behind the previous leg's segment walk and sincos) leave the kernel waiting on
    // global-memory latency 16 times in a row.  The first pass then runs in place like a middle pass.
    {
        constexpr int TOTAL = M * TILE, ITERS = (TOTAL + NT - 1) / NT;
        float2 raw[ITERS];
#pragma unroll
        for (int it = 0; it < ITERS; it++)
            {
                const int idx = it * NT + (int)threadIdx.x;
                const int b = idx & (TILE - 1), m1 = idx >> LOGT, n2 = c0 + b, n = N2 * m1 + n2;
                raw[it] = make_float2(0.f, 0.f);
                if (idx < TOTAL && ((N2 % TILE) == 0 || n2 < N2) && n < p.consumed) raw[it] = gb_load_item<ITEM>(p.samples, (size_t)n);
            }
        int lo = 0;
#pragma unroll
        for (int it = 0; it < ITERS; it++)
            {
                const int idx = it * NT + (int)threadIdx.x;
                const int b = idx & (TILE - 1), m1 = idx >> LOGT, n2 = c0 + b, n = N2 * m1 + n2;
                if (idx >= TOTAL) break;
                float2 x = make_float2(0.f, 0.f);
                const bool inside = ((N2 % TILE) == 0 || n2 < N2);
                // fold == 1: one wiped sample.  QuickSync (fold > 1): the wiped samples n, n + N, n + 2N, ... added in that order
                // onto a zeroed buffer in float32, as the block's std::transform(std::plus) passes do
                for (int f = 0; inside && f < p.fold; f++)
                    {
                        const int m = n + f * N;
                        if (m >= p.consumed) break;
                        float2 v = f == 0 ? raw[it] : gb_load_item<ITEM>(p.samples, (size_t)m);
                        if (wipeoff)
                            {
                                // last segment with n_start <= m: a thread's samples come in increasing n, a bin has ~10 segments
                                while (lo + 1 < nseg && __ldg(&segs[lo + 1].n_start) <= m) lo++;
                                while (lo > 0 && __ldg(&segs[lo].n_start) > m) lo--;
                                const GbWipeSeg sg = segs[lo];
                                const float ph = (float)fma((double)(m - sg.n_start), (double)sg.inc_eff, (double)sg.ph_start);  // exact
                                float sn, cs;
                                sincosf(ph, &sn, &cs);
                                v = cmul(v, make_float2(cs, sn));
                            }
                        x.x = __fadd_rn(x.x, v.x);
                        x.y = __fadd_rn(x.y, v.y);
                    }
                buf[b + m1 * L::MS] = x;
            }
    }
    __syncthreads();
    dif_mid<-1, L, M, M, DCOL::kFirstR>(buf, tw);
    __syncthreads();
    float2* dst = p.colbuf + (size_t)bin * N;
    DCOL::template run_rest<-1, L, false>(buf, tw, [&](int b, int i, const float2* v) {
        const int n2 = c0 + b;
        if ((N2 % TILE) != 0 && n2 >= N2) return;
        float2 w = cmul(tlo[b * IL + (i & (IL - 1))], thi[b * IH + (i / IL)]);  // W_N^(n2 i)
        const float2 ws = step[b];
        float2* d = dst + (size_t)i * N2 + n2;
#pragma unroll
        for (int r = 0; r < RL; r++)
            {
                d[(size_t)r * NIL * N2] = cmul(v[r], w);
                if (r + 1 < RL) w = cmul(w, ws);
            }
    });
}

// rows: colbuf[k1][n2] -> DFT over n2 -> spectra[k1][k2]   (tile of rows r0 .. r0+TILE)
template <int LOGT, int NT, int OCC, class DROW, int N1>
__global__ void __launch_bounds__(NT, OCC)
    gb_acq_fwd_row_dif(PlanDev pl, GbAcqFwdParams p)
{
    extern __shared__ __align__(16) float2 smem[];
    constexpr int M = DROW::kM, TILE = 1 << LOGT, N = N1 * M;
    constexpr int RL = DROW::kLastR, NIL = DROW::kLastNI;
    using L = GbTile<LOGT, NT>;
    float2* buf = smem;
    float2* tw = buf + L::size(M);
    const int r0 = blockIdx.x << LOGT;
    const int nb = min(TILE, N1 - r0);
    const size_t off = (size_t)(p.first_bin + blockIdx.y) * N + (size_t)r0 * M;
    for (int i = threadIdx.x; i < M; i += NT) tw[i] = __ldg(pl.tw + i);
    const float2* src = p.colbuf + off;
    dif_first<-1, L, true, M, DROW::kFirstR>(buf, pl.tw, [&](int b, int m) { return b < nb ? __ldg(src + b * M + m) : make_float2(0.f, 0.f); });
    __syncthreads();
    float2* dst = p.spectra + off;
    DROW::template run_rest<-1, L, true>(buf, tw, [&](int b, int i, const float2* v) {
        if (b >= nb) return;
        float2* d = dst + b * M + i;
#pragma unroll
        for (int r = 0; r < RL; r++) d[r * NIL] = v[r];
    });
}

}  // namespace

// Forward side on the DIF tiles; cudaErrorNotSupported when the plan has no such pair (the caller then runs the run-time
// parametrised kernels of acq_fwd.cu)
cudaError_t gb_acq_forward_dif(const GbFftPlan& plan, const GbAcqFwdParams& p, int n_bins, bool wipeoff, cudaStream_t st, int* launches, void* kt)
{
    if (!plan.inv_fast) return cudaErrorNotSupported;
    cudaError_t e = cudaErrorNotSupported;
    const PlanDev pdr = make_plan_dev(plan, false), pdc = make_plan_dev(plan, true);
    gb_with_dif_pair(plan.N1, plan.N2, plan.inv_variant, [&](auto d1, auto d2, auto c1, auto c2) {
        using D1 = decltype(d1);
        using D2 = decltype(d2);
        using C1 = decltype(c1);
        using C2 = decltype(c2);
        constexpr int LT1 = C1::LOGT, LT2 = C2::LOGT, NT1 = C1::NT, NT2 = C2::NT;
        constexpr int T1 = 1 << LT1, T2 = 1 << LT2;
        {
            auto k = p.item_type == GB200_ITEM_32FC ? gb_acq_fwd_col_dif<GB200_ITEM_32FC, LT1, NT1, C1::OCC, D1, D2::kM>
                                                    : (p.item_type == GB200_ITEM_16IC ? gb_acq_fwd_col_dif<GB200_ITEM_16IC, LT1, NT1, C1::OCC, D1, D2::kM>
                                                                                      : gb_acq_fwd_col_dif<GB200_ITEM_8IC, LT1, NT1, C1::OCC, D1, D2::kM>);
            if (p.item_type != GB200_ITEM_32FC && p.item_type != GB200_ITEM_16IC && p.item_type != GB200_ITEM_8IC)
                {
                    e = cudaErrorInvalidValue;
                    return;
                }
            const size_t sm = (GbTile<LT1, NT1>::size(D1::kM) + D1::kM + T1 * (1 + 8 + (D1::kLastNI + 7) / 8)) * sizeof(float2);
            if ((e = set_smem(k, sm)) != cudaSuccess) return;
            const dim3 grid((D2::kM + T1 - 1) / T1, n_bins);
            if (kt) gb_kt_begin(kt, GB200_KCLASS_ACQ_FWD_COL);
            k<<<grid, NT1, sm, st>>>(pdc, p, wipeoff ? 1 : 0);
            if (kt) gb_kt_end(kt);
            if ((e = cudaGetLastError()) != cudaSuccess) return;
            if (launches) (*launches)++;
        }
        {
            auto k = gb_acq_fwd_row_dif<LT2, NT2, C2::OCC, D2, D1::kM>;
            const size_t sm = (GbTile<LT2, NT2>::size(D2::kM) + D2::kM) * sizeof(float2);
            if ((e = set_smem(k, sm)) != cudaSuccess) return;
            const dim3 grid((D1::kM + T2 - 1) / T2, n_bins);
            if (kt) gb_kt_begin(kt, GB200_KCLASS_ACQ_FWD_ROW);
            k<<<grid, NT2, sm, st>>>(pdr, p);
            if (kt) gb_kt_end(kt);
            if ((e = cudaGetLastError()) != cudaSuccess) return;
            if (launches) (*launches)++;
        }
    });
    return e;
}

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
