// PCPS acquisition search for sm_100a: Doppler wipe-off -> forward DFT -> x conj(code DFT) -> inverse DFT
// -> |.|^2 -> arg-max, batched over (PRN, Doppler bin), with hand-written shared-memory mixed-radix
// Stockham FFTs (no cuFFT).
//
// Replaces (semantics, not code) src/algorithms/acquisition/gnuradio_blocks/pcps_acquisition.cc:
//   set_local_code            :239-274   (code DFT + conj, cached on device)
//   update_local_carrier      :296-310   (wipe-off phasor, generated on the fly instead of a stored table)
//   acquisition_core loop     :714-745   (per Doppler bin: multiply, FFT, multiply, IFFT, |.|^2 [+ accumulate])
//   max_to_input_power / first_vs_second_peak statistics :565-665 (grid arg-max, first maximum wins)
// The DFT lengths are the reference's own (N = d_fft_size, NOT padded to a power of two), so code-phase
// indices are directly comparable; transforms are un-normalised like FFTW's.
//
// Decomposition: see GbFftPlan in gb200_internal.cuh.  Shared-memory layout of a CTA's batch of
// sub-transforms is [m][b] (element m of sub-transform b at m*pitch + b, pitch = batch + 1 to spread banks);
// every radix pass is out-of-place between two such buffers (Stockham autosort, no bit reversal).
#include "gb200_internal.cuh"

namespace
{

__device__ __forceinline__ float2 cmul(float2 a, float2 b)
{
    return make_float2(fmaf(a.x, b.x, -a.y * b.y), fmaf(a.x, b.y, a.y * b.x));
}
__device__ __forceinline__ float2 cmul_conj(float2 a, float2 b)  // a * conj(b)
{
    return make_float2(fmaf(a.x, b.x, a.y * b.y), fmaf(a.y, b.x, -a.x * b.y));
}
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
// multiply by -j*SIGN... rot(a) = a * (SIGN>0 ? +j : -j)
template <int SIGN>
__device__ __forceinline__ float2 rotj(float2 a)
{
    return SIGN > 0 ? make_float2(-a.y, a.x) : make_float2(a.y, -a.x);
}

// W_N^(SIGN*e) for 0 <= e < N, e reduced exactly in integers, angle evaluated with sincospif
template <int SIGN>
__device__ __forceinline__ float2 twiddle_big(unsigned e, unsigned N, float two_over_N)
{
    int es = (int)e;
    if (2u * e >= N) es -= (int)N;  // [-N/2, N/2)
    float s, c;
    sincospif((float)es * two_over_N, &s, &c);  // forward sign is exp(-j...)
    return make_float2(c, SIGN > 0 ? s : -s);
}

// ---- radix-R butterflies on registers: v[q] <- sum_r v[r] W_R^(SIGN * -(-1)... ) i.e. DFT_R with exp(SIGN*j*2*pi*r*q/R) ----
template <int SIGN>
__device__ __forceinline__ void bfly2(float2* v)
{
    float2 a = v[0], b = v[1];
    v[0] = cadd(a, b);
    v[1] = csub(a, b);
}
template <int SIGN>
__device__ __forceinline__ void bfly3(float2* v)
{
    const float s60 = 0.86602540378443864676f;
    float2 t1 = cadd(v[1], v[2]);
    float2 t2 = csub(v[1], v[2]);
    float2 m = make_float2(fmaf(-0.5f, t1.x, v[0].x), fmaf(-0.5f, t1.y, v[0].y));
    float2 r = rotj<SIGN>(make_float2(s60 * t2.x, s60 * t2.y));
    v[0] = cadd(v[0], t1);
    v[1] = cadd(m, r);
    v[2] = csub(m, r);
}
template <int SIGN>
__device__ __forceinline__ void bfly4(float2* v)
{
    float2 a = cadd(v[0], v[2]), b = csub(v[0], v[2]);
    float2 c = cadd(v[1], v[3]), d = rotj<SIGN>(csub(v[1], v[3]));
    v[0] = cadd(a, c);
    v[1] = cadd(b, d);
    v[2] = csub(a, c);
    v[3] = csub(b, d);
}
template <int SIGN>
__device__ __forceinline__ void bfly5(float2* v)
{
    const float c1 = 0.30901699437494742410f;   // cos(2pi/5)
    const float c2 = -0.80901699437494742410f;  // cos(4pi/5)
    const float s1 = 0.95105651629515357212f;   // sin(2pi/5)
    const float s2 = 0.58778525229247312917f;   // sin(4pi/5)
    float2 a1 = cadd(v[1], v[4]), b1 = csub(v[1], v[4]);
    float2 a2 = cadd(v[2], v[3]), b2 = csub(v[2], v[3]);
    float2 m1 = make_float2(v[0].x + c1 * a1.x + c2 * a2.x, v[0].y + c1 * a1.y + c2 * a2.y);
    float2 m2 = make_float2(v[0].x + c2 * a1.x + c1 * a2.x, v[0].y + c2 * a1.y + c1 * a2.y);
    float2 r1 = rotj<SIGN>(make_float2(s1 * b1.x + s2 * b2.x, s1 * b1.y + s2 * b2.y));
    float2 r2 = rotj<SIGN>(make_float2(s2 * b1.x - s1 * b2.x, s2 * b1.y - s1 * b2.y));
    v[0] = make_float2(v[0].x + a1.x + a2.x, v[0].y + a1.y + a2.y);
    v[1] = cadd(m1, r1);
    v[4] = csub(m1, r1);
    v[2] = cadd(m2, r2);
    v[3] = csub(m2, r2);
}
template <int SIGN>
__device__ __forceinline__ void bfly8(float2* v)
{
    const float h = 0.70710678118654752440f;
    // three radix-2 stages (decimation in time on the 8 inputs taken in natural order)
    float2 e[4] = {v[0], v[2], v[4], v[6]};
    float2 o[4] = {v[1], v[3], v[5], v[7]};
    bfly4<SIGN>(e);
    bfly4<SIGN>(o);
    // twiddles W_8^q, q = 1,2,3
    float2 w1 = SIGN > 0 ? make_float2(h * (o[1].x - o[1].y), h * (o[1].x + o[1].y)) : make_float2(h * (o[1].x + o[1].y), h * (o[1].y - o[1].x));
    float2 w2 = rotj<SIGN>(o[2]);
    float2 w3 = SIGN > 0 ? make_float2(-h * (o[3].x + o[3].y), h * (o[3].x - o[3].y)) : make_float2(h * (o[3].y - o[3].x), -h * (o[3].x + o[3].y));
    v[0] = cadd(e[0], o[0]);
    v[4] = csub(e[0], o[0]);
    v[1] = cadd(e[1], w1);
    v[5] = csub(e[1], w1);
    v[2] = cadd(e[2], w2);
    v[6] = csub(e[2], w2);
    v[3] = cadd(e[3], w3);
    v[7] = csub(e[3], w3);
}

template <int SIGN, int R>
__device__ __forceinline__ void bfly(float2* v)
{
    if (R == 2) bfly2<SIGN>(v);
    if (R == 3) bfly3<SIGN>(v);
    if (R == 4) bfly4<SIGN>(v);
    if (R == 5) bfly5<SIGN>(v);
    if (R == 8) bfly8<SIGN>(v);
}

// One Stockham pass of radix R over a tile of TILE = 2^LOGT sub-transforms of length M (all TILE columns are
// always computed; ragged tiles are zero-filled on load and masked on store).
// tw[i] = W_M^i (forward sign); the pass twiddle for input r of butterfly position k is W_M^(r*k*M/(Ns*R)).
// k = j mod Ns through a multiply-high by magic = ceil(2^32 / Ns) (exact for j, Ns < 65536).
template <int SIGN, int R, int LOGT>
__device__ __forceinline__ void radix_pass(const float2* __restrict__ x, float2* __restrict__ y, int M, int Ns, unsigned magic,
    const float2* __restrict__ tw)
{
    constexpr int TILE = 1 << LOGT;
    constexpr int PITCH = TILE + 1;
    const int nbf = M / R;
    const int tstride = M / (Ns * R);
    const int total = nbf << LOGT;
    for (int idx = threadIdx.x; idx < total; idx += blockDim.x)
        {
            const int j = idx >> LOGT;
            const int b = idx & (TILE - 1);
            float2 v[R];
#pragma unroll
            for (int r = 0; r < R; r++) v[r] = x[(j + r * nbf) * PITCH + b];
            int k = 0;
            if (Ns > 1)
                {
                    k = j - (int)__umulhi((unsigned)j, magic) * Ns;
                    const int kt = k * tstride;
#pragma unroll
                    for (int r = 1; r < R; r++)
                        {
                            const float2 w = tw[r * kt];
                            v[r] = SIGN > 0 ? cmul_conj(v[r], w) : cmul(v[r], w);
                        }
                }
            bfly<SIGN, R>(v);
            const int j0 = (j - k) * R + k;
#pragma unroll
            for (int r = 0; r < R; r++) y[(j0 + r * Ns) * PITCH + b] = v[r];
        }
}

// generic odd-prime radix (naive O(R^2) butterfly); R <= GB200_MAX_GENERIC_RADIX
template <int SIGN, int LOGT>
__device__ void radix_pass_generic(const float2* __restrict__ x, float2* __restrict__ y, int M, int Ns, int R, const float2* __restrict__ tw)
{
    constexpr int TILE = 1 << LOGT;
    constexpr int PITCH = TILE + 1;
    const int nbf = M / R;
    const int tstride = M / (Ns * R);
    const int rstride = M / R;
    const int total = nbf << LOGT;
    for (int idx = threadIdx.x; idx < total; idx += blockDim.x)
        {
            const int j = idx >> LOGT;
            const int b = idx & (TILE - 1);
            const int k = j % Ns;
            const int j0 = (j - k) * R + k;
            for (int q = 0; q < R; q++)
                {
                    float2 acc = make_float2(0.f, 0.f);
                    for (int r = 0; r < R; r++)
                        {
                            // combined twiddle exponent: pass twiddle r*k*tstride + butterfly r*q*(M/R), mod M
                            const int e = (int)(((long long)r * k * tstride + (long long)((r * q) % R) * rstride) % M);
                            const float2 w = tw[e];
                            const float2 xv = x[(j + r * nbf) * PITCH + b];
                            const float2 p = SIGN > 0 ? cmul_conj(xv, w) : cmul(xv, w);
                            acc.x += p.x;
                            acc.y += p.y;
                        }
                    y[(j0 + q * Ns) * PITCH + b] = acc;
                }
        }
}

struct PlanDev  // by-value kernel argument
{
    int N, N1, N2, nrad, M;
    int rad[GB200_MAX_RADIX_PASSES];
    unsigned magic[GB200_MAX_RADIX_PASSES];  // ceil(2^32 / Ns) of each pass
    const float2* tw;    // W_M^i, i < M
    const float2* twN;   // W_N^i, i < max(N1, N2) * 16 (forward sign)
    float two_over_N;
};

// Full sub-transform: runs all radix passes; returns the buffer holding the result.
template <int SIGN, int LOGT>
__device__ float2* block_fft(float2* a, float2* b, const PlanDev& pl, const float2* tw)
{
    int Ns = 1;
    const int M = pl.M;
    for (int p = 0; p < pl.nrad; p++)
        {
            const int R = pl.rad[p];
            const unsigned mg = pl.magic[p];
            switch (R)
                {
                case 2: radix_pass<SIGN, 2, LOGT>(a, b, M, Ns, mg, tw); break;
                case 3: radix_pass<SIGN, 3, LOGT>(a, b, M, Ns, mg, tw); break;
                case 4: radix_pass<SIGN, 4, LOGT>(a, b, M, Ns, mg, tw); break;
                case 5: radix_pass<SIGN, 5, LOGT>(a, b, M, Ns, mg, tw); break;
                case 8: radix_pass<SIGN, 8, LOGT>(a, b, M, Ns, mg, tw); break;
                default: radix_pass_generic<SIGN, LOGT>(a, b, M, Ns, R, tw); break;
                }
            __syncthreads();
            float2* t = a;
            a = b;
            b = t;
            Ns *= R;
        }
    return a;
}

// shared-memory carve-up common to the four kernels: two [M][TILE+1] buffers, W_M table, per-CTA W_N^(m*base) table
template <int LOGT>
__device__ __forceinline__ void carve(float2* smem, int M, float2*& bufA, float2*& bufB, float2*& tw, float2*& twA)
{
    constexpr int PITCH = (1 << LOGT) + 1;
    bufA = smem;
    bufB = bufA + (size_t)M * PITCH;
    tw = bufB + (size_t)M * PITCH;
    twA = tw + M;
}

// twA[m] = W_N^(m * base) (forward sign), m < M: the part of the four-step twiddle W_N^(n2 k1) that is common to the tile
__device__ __forceinline__ void fill_twA(float2* twA, int M, int base, const PlanDev& pl)
{
    for (int m = threadIdx.x; m < M; m += blockDim.x)
        {
            const unsigned e = (unsigned)(((unsigned long long)(unsigned)m * (unsigned)base) % (unsigned)pl.N);
            twA[m] = twiddle_big<-1>(e, pl.N, pl.two_over_N);
        }
}

}  // namespace

namespace
{

// ---------------------------------------------------------------------------------------------
// Column kernel, forward: x[N2*n1 + n2] * wipeoff -> DFT over n1 (length N1) -> * W_N^(n2 k1) -> colbuf[k1][n2]
// grid = (ceil(N2 / TILE), bins)
template <int ITEM, int LOGT>
__global__ void __launch_bounds__(GB200_ACQ_THREADS)
    gb_acq_fwd_col(PlanDev pl, GbAcqFwdParams p, int wipeoff)
{
    constexpr int TILE = 1 << LOGT;
    constexpr int PITCH = TILE + 1;
    extern __shared__ __align__(16) float2 smem[];
    const int M = pl.N1, N2 = pl.N2;
    float2 *bufA, *bufB, *tw, *twA;
    carve<LOGT>(smem, M, bufA, bufB, tw, twA);
    for (int i = threadIdx.x; i < M; i += blockDim.x) tw[i] = __ldg(pl.tw + i);
    const int c0 = blockIdx.x * TILE;
    fill_twA(twA, M, c0, pl);
    const int bin = blockIdx.y;
    const GbWipeSeg* segs = nullptr;
    int nseg = 0;
    if (wipeoff)
        {
            const int o0 = __ldg(p.seg_offsets + bin);
            nseg = __ldg(p.seg_offsets + bin + 1) - o0;
            segs = p.segs + o0;
        }
    for (int idx = threadIdx.x; idx < (M << LOGT); idx += blockDim.x)
        {
            const int n1 = idx >> LOGT;
            const int b = idx & (TILE - 1);
            const int n2 = c0 + b;
            const int n = N2 * n1 + n2;
            float2 x = make_float2(0.f, 0.f);
            if (n2 < N2 && n < p.consumed)
                {
                    x = gb_load_item<ITEM>(p.samples, (size_t)n);
                    if (wipeoff)
                        {
                            // last segment with n_start <= n (binary search; a bin has ~10 segments)
                            int lo = 0, hi = nseg - 1;
                            while (lo < hi)
                                {
                                    const int mid = (lo + hi + 1) >> 1;
                                    if (__ldg(&segs[mid].n_start) <= n)
                                        lo = mid;
                                    else
                                        hi = mid - 1;
                                }
                            const GbWipeSeg sg = segs[lo];
                            const float ph = (float)fma((double)(n - sg.n_start), (double)sg.inc_eff, (double)sg.ph_start);  // exact
                            float s, c;
                            sincosf(ph, &s, &c);
                            x = cmul(x, make_float2(c, s));
                        }
                }
            bufA[n1 * PITCH + b] = x;
        }
    __syncthreads();
    float2* res = block_fft<-1, LOGT>(bufA, bufB, pl, tw);
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
template <int LOGT>
__global__ void __launch_bounds__(GB200_ACQ_THREADS)
    gb_acq_fwd_row(PlanDev pl, GbAcqFwdParams p)
{
    constexpr int TILE = 1 << LOGT;
    constexpr int PITCH = TILE + 1;
    extern __shared__ __align__(16) float2 smem[];
    const int M = pl.N2, N1 = pl.N1;
    float2 *bufA, *bufB, *tw, *twA;
    carve<LOGT>(smem, M, bufA, bufB, tw, twA);
    for (int i = threadIdx.x; i < M; i += blockDim.x) tw[i] = __ldg(pl.tw + i);
    const int r0 = blockIdx.x * TILE;
    const int nb = min(TILE, N1 - r0);
    const float2* src = p.colbuf + (size_t)blockIdx.y * pl.N + (size_t)r0 * M;
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
    float2* res = block_fft<-1, LOGT>(bufA, bufB, pl, tw);
    float2* dst = p.spectra + (size_t)blockIdx.y * pl.N + (size_t)r0 * M;
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
// -> * W_N^-(n2 k1) -> rowbuf[t][k1][n2].      grid = (ceil(N1 / TILE), transforms in chunk)
template <int LOGT>
__global__ void __launch_bounds__(GB200_ACQ_THREADS)
    gb_acq_inv_row(PlanDev pl, GbAcqInvParams p)
{
    constexpr int TILE = 1 << LOGT;
    constexpr int PITCH = TILE + 1;
    extern __shared__ __align__(16) float2 smem[];
    const int M = pl.N2, N1 = pl.N1;
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
    float2* res = block_fft<+1, LOGT>(bufA, bufB, pl, tw);
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
template <int LOGT>
__global__ void __launch_bounds__(GB200_ACQ_THREADS)
    gb_acq_inv_col(PlanDev pl, GbAcqInvParams p)
{
    constexpr int TILE = 1 << LOGT;
    constexpr int PITCH = TILE + 1;
    extern __shared__ __align__(16) float2 smem[];
    __shared__ float s_mag[GB200_ACQ_THREADS / 32];
    __shared__ unsigned s_idx[GB200_ACQ_THREADS / 32];
    const int M = pl.N1, N2 = pl.N2;
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
    float2* res = block_fft<+1, LOGT>(bufA, bufB, pl, tw);

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
    for (int n = threadIdx.x; n < N; n += blockDim.x)
        {
            float m = 0.f;
            if (n < consumed)
                {
                    const float2 x = gb_load_item<ITEM>(samples, (size_t)n);
                    m = __fadd_rn(__fmul_rn(x.x, x.x), __fmul_rn(x.y, x.y));
                }
            if (tmp) tmp[n] = m;
            acc += m;
        }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32)
        {
            acc = threadIdx.x < (blockDim.x >> 5) ? s[threadIdx.x] : 0.f;
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
            if (threadIdx.x == 0) *power = acc / (float)N;
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

// one CTA per handle: scan its stored magnitude grid [bins][row_len]
__global__ void __launch_bounds__(1024) gb_grid_peaks_kernel(float* const* grids, int n_bins, int row_len, GbPeak* out)
{
    const float* g = grids[blockIdx.x];
    GbPeak best;
    best.mag = -1.f;
    best.idx = 0xffffffffu;
    best.bin = 0xffffffffu;
    best.pad_ = 0;
    const size_t total = (size_t)n_bins * row_len;
    for (size_t i = threadIdx.x; i < total; i += blockDim.x)
        {
            GbPeak c;
            c.mag = g[i];
            c.bin = (unsigned)(i / row_len);
            c.idx = (unsigned)(i - (size_t)c.bin * row_len);
            c.pad_ = 0;
            if (peak_before(c, best)) best = c;
        }
    best = peak_block_reduce(best);
    if (threadIdx.x == 0)
        {
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
__global__ void __launch_bounds__(1024) gb_second_peak_kernel(const float* row, int row_len, float* tmp, int N, int idx, int spc, float* second)
{
    const int ncopy = min(N / 4, row_len);
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

// last dwell's magnitudes of the LAST Doppler bin into d_tmp_buffer (pcps_acquisition.cc:737, dwell > 1)
__global__ void gb_copy_kernel(float* dst, const float* src, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[i];
}

PlanDev make_plan_dev(const GbFftPlan& plan, bool columns)
{
    PlanDev d;
    d.N = plan.N;
    d.N1 = plan.N1;
    d.N2 = plan.N2;
    d.M = columns ? plan.N1 : plan.N2;
    d.nrad = columns ? plan.nrad1 : plan.nrad2;
    int Ns = 1;
    for (int i = 0; i < GB200_MAX_RADIX_PASSES; i++)
        {
            d.rad[i] = columns ? plan.rad1[i] : plan.rad2[i];
            d.magic[i] = Ns > 1 ? (unsigned)((0x100000000ull + (unsigned)Ns - 1) / (unsigned)Ns) : 0u;
            if (i < d.nrad) Ns *= d.rad[i];
        }
    d.tw = columns ? plan.tw1 : plan.tw2;
    d.twN = plan.twN;
    d.two_over_N = (float)(2.0 / (double)plan.N);
    return d;
}

// two [M][tile+1] buffers + W_M table + per-CTA W_N^(m*base) table
size_t smem_for(int M, int tile) { return ((size_t)2 * M * (tile + 1) + 2 * (size_t)M) * sizeof(float2); }

int log2_tile(int tile)
{
    int l = 0;
    while ((1 << l) < tile) l++;
    return l;
}

template <typename K>
cudaError_t set_smem(K kernel, size_t bytes)
{
    if (bytes > 48 * 1024) return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    return cudaSuccess;
}

// launch KERNEL<..., LOGT> for the run-time tile size
#define GB_TILE_DISPATCH(LOGT_VALUE, ...)      \
    switch (LOGT_VALUE)                        \
        {                                      \
        case 4: { constexpr int LT = 4; __VA_ARGS__; } break; \
        case 3: { constexpr int LT = 3; __VA_ARGS__; } break; \
        case 2: { constexpr int LT = 2; __VA_ARGS__; } break; \
        case 1: { constexpr int LT = 1; __VA_ARGS__; } break; \
        default: { constexpr int LT = 0; __VA_ARGS__; } break; \
        }

template <int ITEM>
cudaError_t launch_fwd_col(int logt, dim3 grid, size_t sm, cudaStream_t st, const PlanDev& pd, const GbAcqFwdParams& p, int wipeoff)
{
    cudaError_t e = cudaSuccess;
    GB_TILE_DISPATCH(logt, e = set_smem(gb_acq_fwd_col<ITEM, LT>, sm); if (e == cudaSuccess) gb_acq_fwd_col<ITEM, LT><<<grid, GB200_ACQ_THREADS, sm, st>>>(pd, p, wipeoff))
    return e;
}

}  // namespace

size_t gb_acq_col_smem(const GbFftPlan& plan) { return smem_for(plan.N1, plan.tile1); }
size_t gb_acq_row_smem(const GbFftPlan& plan) { return smem_for(plan.N2, plan.tile2); }

cudaError_t gb_acq_forward(const GbFftPlan& plan, const GbAcqFwdParams& p, int n_bins, bool wipeoff, cudaStream_t st, int* launches, void* kt)
{
    cudaError_t e;
    {
        PlanDev pd = make_plan_dev(plan, true);
        const size_t sm = gb_acq_col_smem(plan);
        const int logt = log2_tile(plan.tile1);
        dim3 grid((plan.N2 + plan.tile1 - 1) / plan.tile1, n_bins);
        if (kt) gb_kt_begin(kt, GB200_KCLASS_ACQ_FWD_COL);
        switch (p.item_type)
            {
            case GB200_ITEM_32FC: e = launch_fwd_col<GB200_ITEM_32FC>(logt, grid, sm, st, pd, p, wipeoff ? 1 : 0); break;
            case GB200_ITEM_16IC: e = launch_fwd_col<GB200_ITEM_16IC>(logt, grid, sm, st, pd, p, wipeoff ? 1 : 0); break;
            case GB200_ITEM_8IC: e = launch_fwd_col<GB200_ITEM_8IC>(logt, grid, sm, st, pd, p, wipeoff ? 1 : 0); break;
            default: return cudaErrorInvalidValue;
            }
        if (kt) gb_kt_end(kt);
        if (e != cudaSuccess) return e;
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        if (launches) (*launches)++;
    }
    {
        PlanDev pd = make_plan_dev(plan, false);
        const size_t sm = gb_acq_row_smem(plan);
        const int logt = log2_tile(plan.tile2);
        dim3 grid((plan.N1 + plan.tile2 - 1) / plan.tile2, n_bins);
        if (kt) gb_kt_begin(kt, GB200_KCLASS_ACQ_FWD_ROW);
        GB_TILE_DISPATCH(logt, e = set_smem(gb_acq_fwd_row<LT>, sm); if (e == cudaSuccess) gb_acq_fwd_row<LT><<<grid, GB200_ACQ_THREADS, sm, st>>>(pd, p))
        if (kt) gb_kt_end(kt);
        if (e != cudaSuccess) return e;
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        if (launches) (*launches)++;
    }
    return cudaSuccess;
}

cudaError_t gb_acq_inverse(const GbFftPlan& plan, const GbAcqInvParams& p, cudaStream_t st, int* launches, void* kt)
{
    cudaError_t e;
    {
        PlanDev pd = make_plan_dev(plan, false);
        const size_t sm = gb_acq_row_smem(plan);
        const int logt = log2_tile(plan.tile2);
        dim3 grid((plan.N1 + plan.tile2 - 1) / plan.tile2, p.n_transforms);
        if (kt) gb_kt_begin(kt, GB200_KCLASS_ACQ_INV_ROW);
        GB_TILE_DISPATCH(logt, e = set_smem(gb_acq_inv_row<LT>, sm); if (e == cudaSuccess) gb_acq_inv_row<LT><<<grid, GB200_ACQ_THREADS, sm, st>>>(pd, p))
        if (kt) gb_kt_end(kt);
        if (e != cudaSuccess) return e;
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        if (launches) (*launches)++;
    }
    {
        PlanDev pd = make_plan_dev(plan, true);
        const size_t sm = gb_acq_col_smem(plan);
        const int logt = log2_tile(plan.tile1);
        dim3 grid((plan.N2 + plan.tile1 - 1) / plan.tile1, p.n_transforms);
        if (kt) gb_kt_begin(kt, GB200_KCLASS_ACQ_INV_COL);
        GB_TILE_DISPATCH(logt, e = set_smem(gb_acq_inv_col<LT>, sm); if (e == cudaSuccess) gb_acq_inv_col<LT><<<grid, GB200_ACQ_THREADS, sm, st>>>(pd, p))
        if (kt) gb_kt_end(kt);
        if (e != cudaSuccess) return e;
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        if (launches) (*launches)++;
    }
    return cudaSuccess;
}

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

cudaError_t gb_acq_grid_peaks(float* const* grids, int n_handles, int n_bins, int row_len, GbPeak* out, cudaStream_t st, int* launches)
{
    gb_grid_peaks_kernel<<<n_handles, 1024, 0, st>>>(grids, n_bins, row_len, out);
    if (launches) (*launches)++;
    return cudaGetLastError();
}

cudaError_t gb_acq_second_peak(const float* grid_row, float* tmp_buffer, int N, int idx, int samples_per_chip, float* d_second, cudaStream_t st, int* launches)
{
    gb_second_peak_kernel<<<1, 1024, 0, st>>>(grid_row, N, tmp_buffer, N, idx, samples_per_chip, d_second);
    if (launches) (*launches)++;
    return cudaGetLastError();
}

cudaError_t gb_acq_copy_f32(float* dst, const float* src, int n, cudaStream_t st, int* launches)
{
    gb_copy_kernel<<<(n + 255) / 256, 256, 0, st>>>(dst, src, n);
    if (launches) (*launches)++;
    return cudaGetLastError();
}
