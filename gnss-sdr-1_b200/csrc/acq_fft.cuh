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
#pragma once
#include "gb200_internal.cuh"

namespace
{

// ---- complex arithmetic on packed FP32 pairs (sm_100a FADD2 / FMUL2 / FFMA2): one issue slot per complex add and two per
// complex multiply.  Every lane is the IEEE round-to-nearest operation of the scalar instruction it replaces; ptxas turns
// pk2(a, a) operands into scalar broadcasts and swapped / half-negated operands into source modifiers.
typedef unsigned long long gb_f32x2;
__device__ __forceinline__ gb_f32x2 pk2(float lo, float hi)
{
    gb_f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ gb_f32x2 pk2(float2 a) { return pk2(a.x, a.y); }
__device__ __forceinline__ float2 upk2(gb_f32x2 v)
{
    float2 r;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v));
    return r;
}
__device__ __forceinline__ gb_f32x2 fma2x(gb_f32x2 a, gb_f32x2 b, gb_f32x2 c)
{
    gb_f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ gb_f32x2 mul2x(gb_f32x2 a, gb_f32x2 b)
{
    gb_f32x2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ gb_f32x2 add2x(gb_f32x2 a, gb_f32x2 b)
{
    gb_f32x2 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ gb_f32x2 sub2x(gb_f32x2 a, gb_f32x2 b)
{
    gb_f32x2 r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ float gb_fneg(float v) { return __int_as_float(__float_as_int(v) ^ (int)0x80000000); }

__device__ __forceinline__ float2 cmul(float2 a, float2 b)
{
    // (a.x b.x - a.y b.y, a.x b.y + a.y b.x) = a.x (b.x, b.y) + a.y (-b.y, b.x)
    return upk2(fma2x(pk2(a.x, a.x), pk2(b), mul2x(pk2(a.y, a.y), pk2(gb_fneg(b.y), b.x))));
}
__device__ __forceinline__ float2 cmul_conj(float2 a, float2 b)  // a * conj(b)
{
    // (a.x b.x + a.y b.y, a.y b.x - a.x b.y) = a.y (b.y, b.x) + a.x (b.x, -b.y)
    return upk2(fma2x(pk2(a.x, a.x), pk2(b.x, gb_fneg(b.y)), mul2x(pk2(a.y, a.y), pk2(b.y, b.x))));
}
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return upk2(add2x(pk2(a), pk2(b))); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return upk2(sub2x(pk2(a), pk2(b))); }
__device__ __forceinline__ float2 cscale(float2 a, float s) { return upk2(mul2x(pk2(a), pk2(s, s))); }            // s a
__device__ __forceinline__ float2 caxpy(float s, float2 a, float2 y) { return upk2(fma2x(pk2(a), pk2(s, s), pk2(y))); }  // s a + y
// multiply by -j*SIGN... rot(a) = a * (SIGN>0 ? +j : -j)
template <int SIGN>
__device__ __forceinline__ float2 rotj(float2 a)
{
    return SIGN > 0 ? make_float2(gb_fneg(a.y), a.x) : make_float2(a.y, gb_fneg(a.x));
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
    float2 m = caxpy(-0.5f, t1, v[0]);
    float2 r = rotj<SIGN>(cscale(t2, s60));
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
    float2 m1 = caxpy(c2, a2, caxpy(c1, a1, v[0]));
    float2 m2 = caxpy(c1, a2, caxpy(c2, a1, v[0]));
    float2 r1 = rotj<SIGN>(caxpy(s2, b2, cscale(b1, s1)));
    float2 r2 = rotj<SIGN>(caxpy(-s1, b2, cscale(b1, s2)));
    v[0] = cadd(cadd(v[0], a1), a2);
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
    // o1 * h (1 + j SIGN), o3 * h (-1 + j SIGN): one packed add of the (half-negated, swapped) pair, one packed scale
    float2 w1 = cscale(cadd(o[1], rotj<SIGN>(o[1])), h);
    float2 w2 = rotj<SIGN>(o[2]);
    float2 w3 = cscale(csub(rotj<SIGN>(o[3]), o[3]), h);
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
    const float2* twN;   // W_N^i, i < max(256, max(N1, N2) * 16) (forward sign)
    const float2* twNhi; // W_N^(256 i), i < ceil(N / 256)
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
    // W_N^e = W_N^(e & 255) * W_N^(256 (e >> 8)) from two double-precision-built tables; m * base < 2^32 (plan guarantees)
    for (int m = threadIdx.x; m < M; m += blockDim.x)
        {
            const unsigned e = ((unsigned)m * (unsigned)base) % (unsigned)pl.N;
            twA[m] = cmul(__ldg(pl.twN + (e & 255u)), __ldg(pl.twNhi + (e >> 8)));
        }
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
    d.twNhi = plan.twNhi;
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


// ---------------------------------------------------------------------------------------------------------------
// Compile-time specialised sub-transforms.  For the sub-lengths the BASELINE configurations decompose into, every
// stride, pass length and twiddle step is an immediate and the pass loops are fully unrolled; the run-time
// parametrised block_fft above remains the fall-back for any other length (incl. generic prime radices).
template <int SIGN, int R, int LOGT, int M, int NS>
__device__ __forceinline__ void radix_pass_static(const float2* __restrict__ x, float2* __restrict__ y, const float2* __restrict__ tw)
{
    constexpr int TILE = 1 << LOGT;
    constexpr int PITCH = TILE + 1;
    constexpr int NBF = M / R;
    constexpr int TSTRIDE = M / (NS * R);
    constexpr int TOTAL = NBF << LOGT;
    constexpr int ITERS = (TOTAL + GB200_ACQ_THREADS - 1) / GB200_ACQ_THREADS;
#pragma unroll
    for (int it = 0; it < ITERS; it++)
        {
            const int idx = it * GB200_ACQ_THREADS + (int)threadIdx.x;
            if ((TOTAL % GB200_ACQ_THREADS) == 0 || idx < TOTAL)
                {
                    const int j = idx >> LOGT;
                    const int b = idx & (TILE - 1);
                    const float2* xp = x + j * PITCH + b;
                    float2 v[R];
#pragma unroll
                    for (int r = 0; r < R; r++) v[r] = xp[r * NBF * PITCH];
                    int k = 0;
                    if (NS > 1)
                        {
                            k = j % NS;
                            const float2* tp = tw + k * TSTRIDE;
#pragma unroll
                            for (int r = 1; r < R; r++)
                                {
                                    const float2 w = tp[(r - 1) * k * TSTRIDE];  // tw[r * k * TSTRIDE]
                                    v[r] = SIGN > 0 ? cmul_conj(v[r], w) : cmul(v[r], w);
                                }
                        }
                    bfly<SIGN, R>(v);
                    float2* yp = y + ((j - k) * R + k) * PITCH + b;
#pragma unroll
                    for (int r = 0; r < R; r++) yp[r * NS * PITCH] = v[r];
                }
        }
}

template <int SIGN, int LOGT, int M, int NS, int... RADS>
struct GbPasses;
template <int SIGN, int LOGT, int M, int NS>
struct GbPasses<SIGN, LOGT, M, NS>
{
    static_assert(NS == M, "radix list must multiply to M");
    static __device__ __forceinline__ float2* run(float2* a, float2*, const float2*) { return a; }
};
template <int SIGN, int LOGT, int M, int NS, int R, int... REST>
struct GbPasses<SIGN, LOGT, M, NS, R, REST...>
{
    static __device__ __forceinline__ float2* run(float2* a, float2* b, const float2* tw)
    {
        radix_pass_static<SIGN, R, LOGT, M, NS>(a, b, tw);
        __syncthreads();
        return GbPasses<SIGN, LOGT, M, NS * R, REST...>::run(b, a, tw);
    }
};

template <int M_, int... RADS>
struct GbFftStatic
{
    static constexpr int kM = M_;
    template <int SIGN, int LOGT>
    static __device__ __forceinline__ float2* run(float2* a, float2* b, const PlanDev&, const float2* tw)
    {
        return GbPasses<SIGN, LOGT, M_, 1, RADS...>::run(a, b, tw);
    }
};
struct GbFftGeneric
{
    static constexpr int kM = 0;
    template <int SIGN, int LOGT>
    static __device__ __forceinline__ float2* run(float2* a, float2* b, const PlanDev& pl, const float2* tw)
    {
        return block_fft<SIGN, LOGT>(a, b, pl, tw);
    }
};

// host-side dispatch: f(FftType{}, std::integral_constant<int, LOGT>{})
template <int V>
struct GbInt
{
    static constexpr int value = V;
};

template <class FFT, class F>
cudaError_t gb_with_tile(int logt, bool is_static, F&& f)
{
    switch (logt)
        {
        case 4: return f(FFT{}, GbInt<4>{});
        case 3: return f(FFT{}, GbInt<3>{});
        case 2: return f(FFT{}, GbInt<2>{});
        default: break;
        }
    if (is_static) return cudaErrorInvalidValue;  // static types are only instantiated for tiles 4, 8, 16
    if (logt == 1) return f(GbFftGeneric{}, GbInt<1>{});
    return f(GbFftGeneric{}, GbInt<0>{});
}

inline bool gb_fft_is_static(int M)
{
    switch (M)
        {
        case 50: case 80: case 125: case 128: case 160: case 200: case 250: case 256: case 320: case 400: case 500: case 800: case 1000: return true;
        default: return false;
        }
}

template <class F>
cudaError_t gb_with_fft(int M, int logt, F&& f)
{
    if (logt >= 2)
        switch (M)
            {
            case 50: return gb_with_tile<GbFftStatic<50, 5, 5, 2>>(logt, true, f);
            case 80: return gb_with_tile<GbFftStatic<80, 8, 5, 2>>(logt, true, f);
            case 125: return gb_with_tile<GbFftStatic<125, 5, 5, 5>>(logt, true, f);
            case 128: return gb_with_tile<GbFftStatic<128, 8, 8, 2>>(logt, true, f);
            case 160: return gb_with_tile<GbFftStatic<160, 8, 5, 4>>(logt, true, f);
            case 200: return gb_with_tile<GbFftStatic<200, 8, 5, 5>>(logt, true, f);
            case 250: return gb_with_tile<GbFftStatic<250, 5, 5, 5, 2>>(logt, true, f);
            case 256: return gb_with_tile<GbFftStatic<256, 8, 8, 4>>(logt, true, f);
            case 320: return gb_with_tile<GbFftStatic<320, 8, 8, 5>>(logt, true, f);
            case 400: return gb_with_tile<GbFftStatic<400, 8, 5, 5, 2>>(logt, true, f);
            case 500: return gb_with_tile<GbFftStatic<500, 5, 5, 5, 4>>(logt, true, f);
            case 800: return gb_with_tile<GbFftStatic<800, 8, 5, 5, 4>>(logt, true, f);
            case 1000: return gb_with_tile<GbFftStatic<1000, 8, 5, 5, 5>>(logt, true, f);
            default: break;
            }
    return gb_with_tile<GbFftGeneric>(logt, false, f);
}

}  // namespace
