// Inverse side of the PCPS acquisition search (x conj(code DFT), inverse DFT, |.|^2, arg-max), once per (PRN, Doppler bin).
#include "acq_dif.cuh"
#include "acq_pairs.cuh"
#include <algorithm>

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
    int handle, bin;
    gb_acq_unit(p, t, handle, bin);
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
    int handle, bin;
    gb_acq_unit(p, t, handle, bin);
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
            float mag = __fadd_rn(__fmul_rn(y.x, y.x), __fmul_rn(y.y, y.y));  // volk_32fc_magnitude_squared_32f
            if (grid)
                {
                    if (p.scale) mag = __fmul_rn(mag, __ldg(p.scale));  // volk_32f_s32f_multiply_32f
                    float* g = grid + (size_t)bin * p.keep_len + n;
                    if (p.accumulate && p.tmpbuf && p.tmpbuf[handle] && bin == p.n_bins - 1) p.tmpbuf[handle][n] = mag;
                    if (p.accumulate) mag = __fadd_rn(*g, mag);
                    *g = mag;  // the search below runs over the stored (accumulated) grid, :577-600
                }
            if (peak_better(mag, (unsigned)n, best, best_i))
                {
                    best = mag;
                    best_i = (unsigned)n;
                }
        }
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
            p.partial[((size_t)handle * p.n_bins + bin) * gridDim.x + blockIdx.x] = pk;
        }
}

// ---------------------------------------------------------------------------------------------------------------
// Fast path: both sub-lengths compile-time (GbDif), in-place DIF tiles, first/last pass fused with global I/O.

// rows: spectra x code spectrum -> inverse DFT over k2 -> x W_N^+(k1 n2) -> dst[k1][n2] (one transform's row tile r0 .. r0+TILE)
__device__ __forceinline__ int gb_ld_poll(const int* p);

// `dep` / `need` (fused kernel only): a completion counter that must have reached `need` before this tile may STORE -- the
// slot of the row-buffer ring it overwrites.  Thread 0 reads it when the tile starts and looks at the value only after the
// first pass, so the round trip hides behind the tables and the first butterflies.
template <int LOGT, int NT, class DROW, int N1>
__device__ __forceinline__ void gb_inv_row_tile(float2* smem, const PlanDev& pl, const GbAcqInvParams& p, int r0, int t, float2* __restrict__ dst_transform,
    const int* dep = nullptr, int need = 0)
{
    int dep_seen = need;
    if (dep && threadIdx.x == 0) dep_seen = gb_ld_poll(dep);
    constexpr int M = DROW::kM, TILE = 1 << LOGT, N = N1 * M;
    constexpr int RL = DROW::kLastR, NIL = DROW::kLastNI;
    using L = GbTile<LOGT, NT>;
    constexpr int IL = 8, IH = (NIL + IL - 1) / IL;  // last-pass item i = IL * ih + il
    float2* buf = smem;
    float2* tw = buf + L::size(M);
    float2* step = tw + M;          // [TILE]      W_N^(k1 NIL): ratio between consecutive outputs of one item
    float2* tlo = step + TILE;      // [TILE][IL]  W_N^(k1 il)
    float2* thi = tlo + TILE * IL;  // [TILE][IH]  W_N^(k1 IL ih)      (all forward sign; k1 = r0 + b)
    const int nb = min(TILE, N1 - r0);
    for (int i = threadIdx.x; i < M; i += NT) tw[i] = __ldg(pl.tw + i);
    for (int q = threadIdx.x; q < TILE * (1 + IL + IH); q += NT)
        {
            const int b = q / (1 + IL + IH), c = q - b * (1 + IL + IH);
            const unsigned k1 = (unsigned)(r0 + b);
            const unsigned e = (c == 0 ? k1 * (unsigned)NIL : (c <= IL ? k1 * (unsigned)(c - 1) : k1 * (unsigned)(IL * (c - 1 - IL)))) % (unsigned)N;
            const float2 w = cmul(__ldg(pl.twN + (e & 255u)), __ldg(pl.twNhi + (e >> 8)));
            if (c == 0)
                step[b] = w;
            else if (c <= IL)
                tlo[b * IL + c - 1] = w;
            else
                thi[b * IH + c - 1 - IL] = w;
        }
    int handle, bin;
    gb_acq_unit(p, t, handle, bin);
    const float2* xs = p.spectra + (size_t)bin * N + (size_t)r0 * M;
    const float2* cs = p.code_fft[handle] + (size_t)r0 * M;
    dif_first<+1, L, true, M, DROW::kFirstR>(buf, pl.tw, [&](int b, int m) {
        if ((N1 % TILE) != 0 && b >= nb) return make_float2(0.f, 0.f);  // (tiles are whole when TILE divides N1: no guard compiled)
        return cmul(__ldg(xs + b * M + m), __ldg(cs + b * M + m));
    });
    if (dep && threadIdx.x == 0 && dep_seen < need)
        while (gb_ld_poll(dep) < need) __nanosleep(64);
    __syncthreads();
    float2* dst = dst_transform + (size_t)r0 * M;
    DROW::template run_rest<+1, L, true>(buf, tw, [&](int b, int i, const float2* v) {
        if ((N1 % TILE) != 0 && b >= nb) return;
        float2 w = cmul(tlo[b * IL + (i & (IL - 1))], thi[b * IH + (i / IL)]);  // W_N^(k1 i)
        const float2 ws = step[b];
        float2* d = dst + b * M + i;
#pragma unroll
        for (int r = 0; r < RL; r++)
            {
                d[r * NIL] = cmul_conj(v[r], w);  // tables hold the forward sign; the inverse needs the conjugate
                if (r + 1 < RL) w = cmul(w, ws);
            }
    });
}

template <int LOGT, int NT, int OCC, class DROW, int N1>
__global__ void __launch_bounds__(NT, OCC)
    gb_acq_inv_row_dif(PlanDev pl, GbAcqInvParams p)
{
    extern __shared__ __align__(16) float2 smem[];
    constexpr int N = N1 * DROW::kM;
    gb_inv_row_tile<LOGT, NT, DROW, N1>(smem, pl, p, blockIdx.x << LOGT, p.first_transform + blockIdx.y, p.rowbuf + (size_t)blockIdx.y * N);
}

// columns of one transform's tile c0 .. c0+TILE: src[k1][n2] -> inverse DFT over k1 -> |y|^2 -> arg-max partial / grid.
// MODE 0: arg-max over the whole row (keep_offset == 0, keep_len == N: the common single-dwell CFAR search);
// MODE 1: arg-max over a kept window (bit transition); MODE 2: magnitude grid written / accumulated, arg-max over the stored values.
// COHERENT: the source was written earlier in the SAME launch by other CTAs (fused kernel): ld.global.cg (L2) instead of
// the read-only path, see gb_ld_poll.
template <int LOGT, int NT, class DCOL, int N2, int MODE, bool COHERENT>
__device__ __forceinline__ void gb_inv_col_tile(float2* smem, float* s_mag, unsigned* s_idx, const PlanDev& pl, const GbAcqInvParams& p, int tile, int n_tiles, int t,
    const float2* __restrict__ src_transform, const int* dep = nullptr, int need = 0)
{
    constexpr int M = DCOL::kM, TILE = 1 << LOGT;
    constexpr int RL = DCOL::kLastR, NIL = DCOL::kLastNI;
    using L = GbTile<LOGT, NT>;
    float2* buf = smem;
    float2* tw = buf + L::size(M);
    const int c0 = tile * TILE;
    // `dep` / `need` (fused kernel only): the transform's row tiles must all have signalled before its row buffer is read;
    // the counter is read together with the twiddle table, one round trip for both
    int dep_seen = need;
    if (dep && threadIdx.x == 0) dep_seen = gb_ld_poll(dep);
    for (int i = threadIdx.x; i < M; i += NT) tw[i] = __ldg(pl.tw + i);
    if (dep)
        {
            if (threadIdx.x == 0 && dep_seen < need)
                while (gb_ld_poll(dep) < need) __nanosleep(64);
            __syncthreads();
        }
    const float2* src = src_transform + c0;
    dif_first<+1, L, false, M, DCOL::kFirstR>(buf, pl.tw, [&](int b, int m) {
        if (!((N2 % TILE) == 0 || c0 + b < N2)) return make_float2(0.f, 0.f);
        return COHERENT ? __ldcg(src + m * N2 + b) : __ldg(src + m * N2 + b);
    });
    __syncthreads();
    int handle, bin;
    gb_acq_unit(p, t, handle, bin);
    float* grid = MODE == 2 ? p.grid[handle] + (size_t)bin * p.keep_len : nullptr;
    float* tmp = (MODE == 2 && p.accumulate && p.tmpbuf && p.tmpbuf[handle] && bin == p.n_bins - 1) ? p.tmpbuf[handle] : nullptr;
    // Within one work item the outputs come in increasing n, so "first maximum wins" (volk_gnsssdr_32f_index_max_32u) is
    // a strict '>' there; items, lanes and warps are then merged with the full (magnitude, index) ordering.
    float best = -1.0f;
    unsigned best_i = 0xffffffffu;
    DCOL::template run_rest<+1, L, false>(buf, tw, [&](int b, int i, const float2* v) {
        if ((N2 % TILE) != 0 && c0 + b >= N2) return;
        const int nbase = N2 * i + c0 + b - (MODE == 0 ? 0 : p.keep_offset);
        float ibest = -1.0f;
        unsigned ibest_i = 0xffffffffu;
        int ibest_r = 0;  // MODE 0: the winner's leg; its index is formed once, after the loop
#pragma unroll
        for (int r = 0; r < RL; r++)
            {
                const int n = nbase + r * (NIL * N2);
                if (MODE != 0 && (n < 0 || n >= p.keep_len)) continue;
                const float2 sq = upk2(mul2x(pk2(v[r]), pk2(v[r])));  // both squares in one packed multiply (each lane rounds like __fmul_rn)
                float mag = __fadd_rn(sq.x, sq.y);                    // volk_32fc_magnitude_squared_32f
                if (MODE == 2)
                    {
                        if (p.scale) mag = __fmul_rn(mag, __ldg(p.scale));  // volk_32f_s32f_multiply_32f
                        if (tmp) tmp[n] = mag;
                        if (p.accumulate) mag = __fadd_rn(grid[n], mag);
                        grid[n] = mag;  // the search runs over the stored (accumulated) grid, :577-600
                    }
                if (mag > ibest)
                    {
                        ibest = mag;
                        if (MODE == 0)
                            ibest_r = r;
                        else
                            ibest_i = (unsigned)n;
                    }
            }
        if (MODE == 0) ibest_i = (unsigned)(nbase + ibest_r * (NIL * N2));  // every leg is inside the row: ibest >= 0 > -1 always wins once
        if (peak_better(ibest, ibest_i, best, best_i))
            {
                best = ibest;
                best_i = ibest_i;
            }
    });
        {
            // warp stage with two REDUX instructions instead of five shuffle rounds: magnitudes are >= +0 (or the -1 of "no
            // output"; never NaN -- a NaN never wins a '>'), so their bit patterns order like unsigned integers; the first
            // maximum is then the smallest index among the lanes that hold the winning pattern
            const unsigned key = best >= 0.f ? __float_as_uint(best) + 1u : 0u;
            const unsigned wkey = __reduce_max_sync(0xffffffffu, key);
            best_i = __reduce_min_sync(0xffffffffu, key == wkey ? best_i : 0xffffffffu);
            best = wkey ? __uint_as_float(wkey - 1u) : -1.0f;
            const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
            if (lane == 0)
                {
                    s_mag[warp] = best;
                    s_idx[warp] = best_i;
                }
            __syncthreads();
            if (threadIdx.x == 0)
                {
                    for (int k = 1; k < NT / 32; k++)
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
                    p.partial[((size_t)handle * p.n_bins + bin) * n_tiles + tile] = pk;
                }
        }
}

template <int LOGT, int NT, int OCC, class DCOL, int N2, int MODE>
__global__ void __launch_bounds__(NT, OCC)
    gb_acq_inv_col_dif(PlanDev pl, GbAcqInvParams p)
{
    extern __shared__ __align__(16) float2 smem[];
    __shared__ float s_mag[NT / 32];
    __shared__ unsigned s_idx[NT / 32];
    constexpr int N = N2 * DCOL::kM;
    gb_inv_col_tile<LOGT, NT, DCOL, N2, MODE, false>(smem, s_mag, s_idx, pl, p, blockIdx.x, gridDim.x, p.first_transform + blockIdx.y, p.rowbuf + (size_t)blockIdx.y * N);
}

// ---------------------------------------------------------------------------------------------------------------
// One persistent launch for the whole inverse side of a dwell (opt-in: GB200_ACQ_FUSED=1; measured 0.483 ms against the
// two-lane pipeline's ~0.49 ms at C2, profiles/r01_acq_fused_persistent.txt -- parity-tested, not yet faster, kept as the
// base of the next round's work).  Every launch of the two kernels above pays about one
// extra wave of fill and drain (profiles/r01_acq_chunk_sweep.txt) and the L2 budget keeps a launch at ~3 waves; here the
// grid is as many CTAs as fit on the device and never drains: CTAs pull tile tasks from a counter in the order
//   step s = 0, 1, ...:  the row tiles of transform s, then the column tiles of transform s - lag,
// a column tile waits until its transform's row tiles have all signalled (release / acquire on a per-transform counter),
// and a row tile of transform t waits for the column tiles of transform t - ring before it overwrites their slot of the
// row-buffer ring (which stays L2-resident: ring x N x 8 B).  A task only ever waits for tasks pulled before it, so the
// queue cannot deadlock whatever part of the grid is resident.
struct GbFusedCtl
{
    unsigned next_task;
    unsigned pad_[3];
    // followed by int row_done[T], col_done[T]
};

// Polling load of a completion counter.  Deliberately NOT ld.acquire.gpu: that compiles to LDG.STRONG.GPU + CCTL.IVALL, and
// invalidating the SM's whole L1 at the head of every task throws away the twiddle and code-spectrum lines every CTA on the
// SM keeps hitting.  The only data that needs coherence is the row-buffer ring, and that is read with ld.global.cg (L2, where
// the producer's MEMBAR.GPU + REDG made it visible before the counter moved); the loads are issued behind the poll's branch
// and a CTA barrier.
__device__ __forceinline__ int gb_ld_poll(const int* p)
{
    int v;
    asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void gb_red_release_inc(int* p)
{
    asm volatile("red.release.gpu.global.add.s32 [%0], 1;" ::"l"(p) : "memory");
}

template <int LT1, int LT2, int NT, int OCC, class D1, class D2, int MODE>
__global__ void __launch_bounds__(NT, OCC)
    gb_acq_inv_fused(PlanDev plr, PlanDev plc, GbAcqInvParams p, GbFusedCtl* ctl, int lag, int ring)
{
    extern __shared__ __align__(16) float2 smem[];
    __shared__ float s_mag[NT / 32];
    __shared__ unsigned s_idx[NT / 32];
    __shared__ unsigned s_task;
    constexpr int N = D1::kM * D2::kM;
    constexpr int TR = (D1::kM + (1 << LT2) - 1) >> LT2;  // row tiles per transform (rows of length N2, N1 of them)
    constexpr int TC = (D2::kM + (1 << LT1) - 1) >> LT1;  // column tiles per transform
    const int T = p.n_transforms;
    int* row_done = (int*)(ctl + 1);
    int* col_done = row_done + T;
    const unsigned total = (unsigned)(T + lag) * (unsigned)(TR + TC);
    unsigned pulled = 0;  // thread 0: the task id pulled one task ahead, so the atomic's round trip hides behind a tile
    if (threadIdx.x == 0) pulled = atomicAdd(&ctl->next_task, 1u);
    for (;;)
        {
            __syncthreads();  // everyone is done with s_task and with the tile of the previous task
            if (threadIdx.x == 0)
                {
                    s_task = pulled;
                    if (pulled < total) pulled = atomicAdd(&ctl->next_task, 1u);
                }
            __syncthreads();
            const unsigned u = s_task;
            if (u >= total) break;
            const int step = (int)(u / (unsigned)(TR + TC)), v = (int)(u - (unsigned)step * (unsigned)(TR + TC));
            if (v < TR)
                {
                    const int t = step;
                    if (t >= T) continue;
                    gb_inv_row_tile<LT2, NT, D2, D1::kM>(smem, plr, p, v << LT2, p.first_transform + t, p.rowbuf + (size_t)(t % ring) * N,
                        t >= ring ? col_done + t - ring : nullptr, TC);
                    __syncthreads();
                    if (threadIdx.x == 0) gb_red_release_inc(row_done + t);
                }
            else
                {
                    const int t = step - lag;
                    if (t < 0 || t >= T) continue;
                    gb_inv_col_tile<LT1, NT, D1, D2::kM, MODE, true>(smem, s_mag, s_idx, plc, p, v - TR, TC, p.first_transform + t, p.rowbuf + (size_t)(t % ring) * N,
                        row_done + t, TR);
                    __syncthreads();
                    if (threadIdx.x == 0) gb_red_release_inc(col_done + t);
                }
        }
}

}  // namespace

// (N1, N2) splits with a fast path, in order of preference for a given N = N1 * N2
bool gb_acq_preferred_split(int N, int avoid_n1, int* n1, int* n2)
{
    static const int pairs[][2] = {{50, 80}, {80, 100}, {125, 128}, {125, 160}, {125, 200}, {160, 200}, {125, 256}, {200, 200}, {200, 250}, {250, 320}, {400, 500}};
    for (const auto& p : pairs)
        if ((long long)p[0] * p[1] == N && p[0] != avoid_n1)
            {
                *n1 = p[0];
                *n2 = p[1];
                return true;
            }
    return false;
}

bool gb_acq_inverse_tiles(int N1, int N2, int variant, int* tile1, int* tile2)
{
    return gb_with_dif_pair(N1, N2, variant, [&](auto, auto, auto c1, auto c2) {
        *tile1 = 1 << decltype(c1)::LOGT;
        *tile2 = 1 << decltype(c2)::LOGT;
    });
}

size_t gb_acq_fused_ctl_bytes(int n_transforms) { return sizeof(GbFusedCtl) + 2 * sizeof(int) * (size_t)n_transforms; }

// The whole inverse side of a dwell in one persistent launch (p covers ALL transforms; p.rowbuf holds `ring` transforms).
// cudaErrorNotSupported: no such kernel for this plan / mode -- the caller falls back to gb_acq_inverse per chunk.
cudaError_t gb_acq_inverse_fused(const GbFftPlan& plan, const GbAcqInvParams& p, void* d_ctl, int lag, int ring, int sm_count, cudaStream_t st, int* launches, void* kt)
{
    if (!plan.inv_fast || p.grid || !(p.keep_offset == 0 && p.keep_len == plan.N)) return cudaErrorNotSupported;
    cudaError_t e = cudaErrorNotSupported;
    const PlanDev pdr = make_plan_dev(plan, false), pdc = make_plan_dev(plan, true);
    gb_with_dif_pair(plan.N1, plan.N2, plan.inv_variant, [&](auto d1, auto d2, auto c1, auto c2) {
        using D1 = decltype(d1);
        using D2 = decltype(d2);
        using C1 = decltype(c1);
        using C2 = decltype(c2);
        if constexpr (C1::NT == C2::NT)
            {
                constexpr int LT1 = C1::LOGT, LT2 = C2::LOGT, NT = C1::NT, OCC = C1::OCC < C2::OCC ? C1::OCC : C2::OCC;
                auto k = gb_acq_inv_fused<LT1, LT2, NT, OCC, D1, D2, 0>;
                const size_t smr = (GbTile<LT2, NT>::size(D2::kM) + D2::kM + (1 << LT2) * (1 + 8 + (D2::kLastNI + 7) / 8)) * sizeof(float2);
                const size_t smc = (GbTile<LT1, NT>::size(D1::kM) + D1::kM) * sizeof(float2);
                const size_t sm = smr > smc ? smr : smc;
                if ((e = set_smem(k, sm)) != cudaSuccess) return;
                int per_sm = 0;
                if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, NT, sm)) != cudaSuccess) return;
                constexpr int TR = (D1::kM + (1 << LT2) - 1) >> LT2, TC = (D2::kM + (1 << LT1) - 1) >> LT1;
                const long long tasks = (long long)(p.n_transforms + lag) * (TR + TC);
                const int grid = (int)std::min<long long>(tasks, (long long)std::max(1, per_sm) * sm_count);
                if ((e = cudaMemsetAsync(d_ctl, 0, gb_acq_fused_ctl_bytes(p.n_transforms), st)) != cudaSuccess) return;
                if (kt) gb_kt_begin(kt, GB200_KCLASS_ACQ_INV_ROW);
                k<<<grid, NT, sm, st>>>(pdr, pdc, p, (GbFusedCtl*)d_ctl, lag, ring);
                if (kt) gb_kt_end(kt);
                e = cudaGetLastError();
                if (e == cudaSuccess && launches) (*launches)++;
            }
    });
    return e;
}

cudaError_t gb_acq_inverse(const GbFftPlan& plan, const GbAcqInvParams& p, cudaStream_t st, int* launches, void* kt)
{
    cudaError_t e;
    if (plan.inv_fast)
        {
            e = cudaErrorInvalidValue;
            const PlanDev pdr = make_plan_dev(plan, false), pdc = make_plan_dev(plan, true);
            gb_with_dif_pair(plan.N1, plan.N2, plan.inv_variant, [&](auto d1, auto d2, auto c1, auto c2) {
                using D1 = decltype(d1);
                using D2 = decltype(d2);
                using C1 = decltype(c1);
                using C2 = decltype(c2);
                constexpr int LT1 = C1::LOGT, LT2 = C2::LOGT, NT1 = C1::NT, NT2 = C2::NT;
                constexpr int T1 = 1 << LT1, T2 = 1 << LT2;
                {
                    auto k = gb_acq_inv_row_dif<LT2, NT2, C2::OCC, D2, D1::kM>;
                    const size_t sm = (GbTile<LT2, NT2>::size(D2::kM) + D2::kM + T2 * (1 + 8 + (D2::kLastNI + 7) / 8)) * sizeof(float2);
                    if ((e = set_smem(k, sm)) != cudaSuccess) return;
                    const dim3 grid((D1::kM + T2 - 1) / T2, p.n_transforms);
                    if (kt) gb_kt_begin(kt, GB200_KCLASS_ACQ_INV_ROW);
                    k<<<grid, NT2, sm, st>>>(pdr, p);
                    if (kt) gb_kt_end(kt);
                    if ((e = cudaGetLastError()) != cudaSuccess) return;
                    if (launches) (*launches)++;
                }
                {
                    const int mode = p.grid ? 2 : ((p.keep_offset == 0 && p.keep_len == plan.N) ? 0 : 1);
                    auto k = mode == 2 ? gb_acq_inv_col_dif<LT1, NT1, C1::OCC, D1, D2::kM, 2>
                                       : (mode == 1 ? gb_acq_inv_col_dif<LT1, NT1, C1::OCC, D1, D2::kM, 1> : gb_acq_inv_col_dif<LT1, NT1, C1::OCC, D1, D2::kM, 0>);
                    const size_t sm = (GbTile<LT1, NT1>::size(D1::kM) + D1::kM) * sizeof(float2);
                    if ((e = set_smem(k, sm)) != cudaSuccess) return;
                    const dim3 grid((D2::kM + T1 - 1) / T1, p.n_transforms);
                    if (kt) gb_kt_begin(kt, GB200_KCLASS_ACQ_INV_COL);
                    k<<<grid, NT1, sm, st>>>(pdc, p);
                    if (kt) gb_kt_end(kt);
                    if ((e = cudaGetLastError()) != cudaSuccess) return;
                    if (launches) (*launches)++;
                }
            });
            return e;
        }
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
