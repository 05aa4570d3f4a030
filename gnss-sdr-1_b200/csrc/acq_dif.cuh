// In-place decimation-in-frequency sub-transforms with register-resident first and last passes.
//
// The inverse side of the acquisition search runs one (PRN, Doppler bin) transform per CTA tile, 2560 of them per
// dwell at C2, so its cost per cell decides the headline number.  Compared with the out-of-place Stockham passes in
// acq_fft.cuh this formulation
//   * loads the first pass's butterfly legs straight from global memory into registers (no staging copy),
//   * hands the last pass's outputs to the caller in registers (twiddle + store, or |.|^2 + arg-max) -- the
//     digit-reversed read order of the last pass makes its outputs consecutive frequencies k = i + r*(M/R_last),
//   * works in place on ONE shared-memory tile (thread-private element sets per pass, one barrier per pass).
//
// Index algebra for M = R_0 R_1 ... R_{P-1}: pass p works on blocks of B_p = M / (R_0..R_{p-1}) with leg stride
// S_p = B_p / R_p; item i of pass p -> block q = i / S_p, offset j = i % S_p, legs at q*B_p + j + r*S_p, output r'
// multiplied by W_M^(r' j M/B_p).  After all passes position sum_p d'_p S_p holds frequency sum_p d'_p R_0..R_{p-1}.
#pragma once
#include "acq_fft.cuh"

namespace
{

// ---- composite butterflies (Cooley-Tukey inside registers, constant inner twiddles) ----
template <int SIGN>
__device__ __forceinline__ float2 cmul_const(float2 a, float c, float s)  // a * (c + j*SIGN*s)
{
    // (a.x c - SIGN a.y s, SIGN a.x s + a.y c) = a.x (c, SIGN s) + a.y (-SIGN s, c)
    return upk2(fma2x(pk2(a.x, a.x), pk2(c, SIGN > 0 ? s : -s), mul2x(pk2(a.y, a.y), pk2(SIGN > 0 ? -s : s, c))));
}

// radix 16 = 4 x 4: x[4a+b] -> X[k1 + 4 k2]
template <int SIGN>
__device__ __forceinline__ void bfly16(float2* v)
{
    const float c1 = 0.92387953251128675613f, s1 = 0.38268343236508977173f;  // cos, sin(pi/8)
    const float h = 0.70710678118654752440f;
    float2 t[4][4];
#pragma unroll
    for (int b = 0; b < 4; b++)
        {
            float2 u[4] = {v[b], v[4 + b], v[8 + b], v[12 + b]};
            bfly4<SIGN>(u);  // over a; u[k1]
#pragma unroll
            for (int k1 = 0; k1 < 4; k1++) t[k1][b] = u[k1];
        }
    // inner twiddles W_16^(b k1)
    t[1][1] = cmul_const<SIGN>(t[1][1], c1, s1);
    t[1][2] = cmul_const<SIGN>(t[1][2], h, h);
    t[1][3] = cmul_const<SIGN>(t[1][3], s1, c1);
    t[2][1] = cmul_const<SIGN>(t[2][1], h, h);
    t[2][2] = rotj<SIGN>(t[2][2]);
    t[2][3] = cmul_const<SIGN>(t[2][3], -h, h);
    t[3][1] = cmul_const<SIGN>(t[3][1], s1, c1);
    t[3][2] = cmul_const<SIGN>(t[3][2], -h, h);
    t[3][3] = cmul_const<SIGN>(t[3][3], -c1, -s1);
#pragma unroll
    for (int k1 = 0; k1 < 4; k1++)
        {
            bfly4<SIGN>(t[k1]);  // over b; t[k1][k2]
#pragma unroll
            for (int k2 = 0; k2 < 4; k2++) v[k1 + 4 * k2] = t[k1][k2];
        }
}

// radix 10 = 2 x 5: x[5a+b] (a < 2, b < 5) -> X[k1 + 2 k2]
template <int SIGN>
__device__ __forceinline__ void bfly10(float2* v)
{
    const float c1 = 0.80901699437494742410f, s1 = 0.58778525229247312917f;  // W_10^1
    const float c2 = 0.30901699437494742410f, s2 = 0.95105651629515357212f;  // W_10^2
    float2 e[5], o[5];
#pragma unroll
    for (int b = 0; b < 5; b++)
        {
            e[b] = cadd(v[b], v[5 + b]);
            o[b] = csub(v[b], v[5 + b]);
        }
    o[1] = cmul_const<SIGN>(o[1], c1, s1);
    o[2] = cmul_const<SIGN>(o[2], c2, s2);
    o[3] = cmul_const<SIGN>(o[3], -c2, s2);
    o[4] = cmul_const<SIGN>(o[4], -c1, s1);
    bfly5<SIGN>(e);
    bfly5<SIGN>(o);
#pragma unroll
    for (int k2 = 0; k2 < 5; k2++)
        {
            v[2 * k2] = e[k2];
            v[2 * k2 + 1] = o[k2];
        }
}

// radix 25 = 5 x 5: x[5a+b] -> X[k1 + 5 k2]
template <int SIGN>
__device__ __forceinline__ void bfly25(float2* v)
{
    // W_25^m, m = b k1: (cos, sin)(2 pi m / 25)
    constexpr float C[17] = {1.f, 0.96858316112863107605f, 0.87630668004386358394f, 0.72896862742141155245f, 0.53582679497899654564f, 0.f,
        0.06279051952931352654f, 0.f, -0.42577929156507271502f, -0.63742398974868974548f, 0.f, 0.f, -0.99211470131447776488f, 0.f, 0.f, 0.f,
        -0.63742398974868952344f};
    constexpr float S[17] = {0.f, 0.24868988716485479484f, 0.48175367410171532345f, 0.68454710592868861507f, 0.84432792550201507531f, 0.f,
        0.99802672842827155897f, 0.f, 0.90482705246601946580f, 0.77051324277578925326f, 0.f, 0.f, 0.12533323356430453588f, 0.f, 0.f, 0.f,
        -0.77051324277578936428f};
#pragma unroll
    for (int b = 0; b < 5; b++)
        {
            float2 u[5] = {v[b], v[5 + b], v[10 + b], v[15 + b], v[20 + b]};
            bfly5<SIGN>(u);  // over a; u[k1]
#pragma unroll
            for (int k1 = 0; k1 < 5; k1++) v[5 * k1 + b] = (b == 0 || k1 == 0) ? u[k1] : cmul_const<SIGN>(u[k1], C[b * k1], S[b * k1]);  // t[k1][b]
        }
    float2 t[5][5];
#pragma unroll
    for (int k1 = 0; k1 < 5; k1++)
        {
#pragma unroll
            for (int b = 0; b < 5; b++) t[k1][b] = v[5 * k1 + b];
            bfly5<SIGN>(t[k1]);  // over b; t[k1][k2]
        }
#pragma unroll
    for (int k1 = 0; k1 < 5; k1++)
#pragma unroll
        for (int k2 = 0; k2 < 5; k2++) v[k1 + 5 * k2] = t[k1][k2];
}

// radix 20 = 4 x 5: x[5a+b] (a < 4, b < 5) -> X[k1 + 4 k2]
template <int SIGN>
__device__ __forceinline__ void bfly20(float2* v)
{
    // W_20^m, m = b k1: (cos, sin)(2 pi m / 20)
    constexpr float C[13] = {1.f, 0.95105651629515353118f, 0.80901699437494745126f, 0.58778525229247313710f, 0.30901699437494745126f, 0.f,
        -0.30901699437494734024f, 0.f, -0.80901699437494734024f, -0.95105651629515353118f, 0.f, 0.f, -0.80901699437494756229f};
    constexpr float S[13] = {0.f, 0.30901699437494739575f, 0.58778525229247313710f, 0.80901699437494745126f, 0.95105651629515353118f, 0.f,
        0.95105651629515364220f, 0.f, 0.58778525229247324813f, 0.30901699437494750677f, 0.f, 0.f, -0.58778525229247302608f};
    float2 t[4][5];
#pragma unroll
    for (int b = 0; b < 5; b++)
        {
            float2 u[4] = {v[b], v[5 + b], v[10 + b], v[15 + b]};
            bfly4<SIGN>(u);  // over a; u[k1]
#pragma unroll
            for (int k1 = 0; k1 < 4; k1++) t[k1][b] = (b == 0 || k1 == 0) ? u[k1] : cmul_const<SIGN>(u[k1], C[b * k1], S[b * k1]);
        }
#pragma unroll
    for (int k1 = 0; k1 < 4; k1++)
        {
            bfly5<SIGN>(t[k1]);  // over b; t[k1][k2]
#pragma unroll
            for (int k2 = 0; k2 < 5; k2++) v[k1 + 4 * k2] = t[k1][k2];
        }
}

template <int SIGN, int R>
__device__ __forceinline__ void bflyx(float2* v)
{
    static_assert(R == 2 || R == 3 || R == 4 || R == 5 || R == 8 || R == 10 || R == 16 || R == 20 || R == 25, "no butterfly for this radix");
    if (R == 10)
        bfly10<SIGN>(v);
    else if (R == 16)
        bfly16<SIGN>(v);
    else if (R == 20)
        bfly20<SIGN>(v);
    else if (R == 25)
        bfly25<SIGN>(v);
    else
        bfly<SIGN, R>(v);
}

// ---- tile layout: element m of the tile's sub-transform b lives at m * MS + b, MS = TILE + 1 (odd) ----
// With an odd element stride both access directions are bank-conflict free for 8-byte words: lanes along b (stride 1)
// and lanes along m (stride MS = 1 mod 16).  Work items (i, b) are mapped to threads either b-fastest (column kernels: a
// global row of TILE consecutive columns per instruction) or i-fastest (row kernels' first and last pass: consecutive
// elements of one contiguous sub-transform per instruction; i is padded to a multiple of GB_IFAST_PAD so a quarter-warp
// never straddles two sub-transforms).
// i-fastest maps pad the item count of one sub-transform to a multiple of this.  16 (a half-warp never straddles two
// sub-transforms) leaves 8 of 48 last-pass lanes idle at M = 200; 8 measured 2.4 % faster at C2, 1 % at C3, 0.8 % slower at C5;
// 4 and 1 (no padding) lose to shared-memory conflicts (profiles/r02_acq_fwd_sweep.txt)
#ifndef GB_IFAST_PAD
#define GB_IFAST_PAD 8
#endif
#ifndef GB_IFAST_BPERM
#define GB_IFAST_BPERM 1
#endif
template <int LOGT, int NT_>
struct GbTile
{
    static constexpr int TILE = 1 << LOGT;
    static constexpr int NT = NT_;  // threads per CTA
    static constexpr int MS = TILE + 1;
    static __host__ __device__ constexpr int size(int M) { return M * MS; }
    template <int NI, bool IFAST>
    struct Map
    {
        static constexpr int NIP = IFAST ? ((NI + GB_IFAST_PAD - 1) / GB_IFAST_PAD * GB_IFAST_PAD) : NI;
        static constexpr int TOTAL = NIP * TILE;
        static constexpr bool kExact = (NIP == NI) && (TOTAL % NT_) == 0;
        static __device__ __forceinline__ bool split(int idx, int& i, int& b)
        {
            if (IFAST)
                {
                    b = (int)((unsigned)idx / (unsigned)NIP);
                    i = idx - b * NIP;
                    // 8 items per sub-transform: a half-warp holds two sub-transforms.  With the odd pitch, neighbours b, b + 1 collide
                    // in one bank pair (items 7 and 0); b and b + TILE/2 never do -- pair those (any bijection of b is valid here)
                    if (GB_IFAST_BPERM && NIP == 8 && TILE == 16) b = ((b & 1) << 3) | (b >> 1);
                    return (kExact || idx < TOTAL) && (NIP == NI || i < NI);
                }
            b = idx & (TILE - 1);
            i = idx >> LOGT;
            return kExact || idx < TOTAL;
        }
    };
};

// ---- passes ----
// first pass: legs come from `load(b, m)` (global memory), outputs go to the tile; twiddles read through L1 (`twg`)
template <int SIGN, class L, bool IFAST, int M, int R, class LoadF>
__device__ __forceinline__ void dif_first(float2* __restrict__ buf, const float2* __restrict__ twg, LoadF&& load)
{
    constexpr int S = M / R, NI = S;
    using MAP = typename L::template Map<NI, IFAST>;
    constexpr int ITERS = (MAP::TOTAL + L::NT - 1) / L::NT;
#pragma unroll
    for (int it = 0; it < ITERS; it++)
        {
            const int idx = it * L::NT + (int)threadIdx.x;
            int i, b;
            if (MAP::split(idx, i, b))
                {
                    float2 v[R];
#pragma unroll
                    for (int r = 0; r < R; r++) v[r] = load(b, i + r * S);
                    bflyx<SIGN, R>(v);
                    float2* p = buf + b + i * L::MS;
                    p[0] = v[0];
                    if (IFAST)
                        {
                            // lanes run along i: the table reads W^(r i) would scatter over r times more cache lines per leg,
                            // and this kernel is bound by the L1 data pipe -- build the powers from W^i in registers instead
                            // (products of two earlier powers: depth log2 R)
                            float2 w[R];
                            w[1] = __ldg(twg + i);
#pragma unroll
                            for (int r = 2; r < R; r++) w[r] = cmul(w[r / 2], w[r - r / 2]);
#pragma unroll
                            for (int r = 1; r < R; r++) p[r * S * L::MS] = SIGN > 0 ? cmul_conj(v[r], w[r]) : cmul(v[r], w[r]);
                        }
                    else
                        {
#pragma unroll
                            for (int r = 1; r < R; r++)
                                {
                                    const float2 w = __ldg(twg + r * i);  // r*i < M; lanes run along b: a broadcast
                                    p[r * S * L::MS] = SIGN > 0 ? cmul_conj(v[r], w) : cmul(v[r], w);
                                }
                        }
                }
        }
}

// middle pass, in place
template <int SIGN, class L, int M, int BP, int R>
__device__ __forceinline__ void dif_mid(float2* __restrict__ buf, const float2* __restrict__ tw)
{
    constexpr int S = BP / R, NI = M / R, TW = M / BP;
    using MAP = typename L::template Map<NI, false>;
    constexpr int ITERS = (MAP::TOTAL + L::NT - 1) / L::NT;
#pragma unroll
    for (int it = 0; it < ITERS; it++)
        {
            const int idx = it * L::NT + (int)threadIdx.x;
            int i, b;
            if (MAP::split(idx, i, b))
                {
                    const int q = (int)((unsigned)i / (unsigned)S);
                    const int j = i - q * S;
                    float2* p = buf + b + (q * BP + j) * L::MS;
                    float2 v[R];
#pragma unroll
                    for (int r = 0; r < R; r++) v[r] = p[r * S * L::MS];
                    bflyx<SIGN, R>(v);
                    p[0] = v[0];
                    const float2* tp = tw + j * TW;
#pragma unroll
                    for (int r = 1; r < R; r++)
                        {
                            const float2 w = tp[(r - 1) * j * TW];  // tw[r * j * TW], r*j*TW < M
                            p[r * S * L::MS] = SIGN > 0 ? cmul_conj(v[r], w) : cmul(v[r], w);
                        }
                }
        }
}

// digit-reversed base position of last-pass item i = d_0 + R_0 (d_1 + R_1 (...)) -> sum_p d_p S_p
template <int B, int R, int... REST>
struct GbRevBase
{
    static __device__ __forceinline__ int get(unsigned i)
    {
        if constexpr (sizeof...(REST) == 0)
            return 0;
        else
            {
                constexpr int S = B / R;
                const unsigned d = i % (unsigned)R;
                return (int)d * S + GbRevBase<S, REST...>::get(i / (unsigned)R);
            }
    }
};

template <int R0, int... RS>
struct GbLastOf
{
    static constexpr int value = GbLastOf<RS...>::value;
};
template <int R0>
struct GbLastOf<R0>
{
    static constexpr int value = R0;
};

// last pass: outputs stay in registers; `store(b, i, v)` receives frequencies k = i + r * (M / R), r < R
template <int SIGN, class L, bool IFAST, int M, int R, class REV, class StoreF>
__device__ __forceinline__ void dif_last(const float2* __restrict__ buf, StoreF&& store)
{
    constexpr int NI = M / R;
    using MAP = typename L::template Map<NI, IFAST>;
    constexpr int ITERS = (MAP::TOTAL + L::NT - 1) / L::NT;
#pragma unroll
    for (int it = 0; it < ITERS; it++)
        {
            const int idx = it * L::NT + (int)threadIdx.x;
            int i, b;
            if (MAP::split(idx, i, b))
                {
                    const float2* p = buf + b + REV::get((unsigned)i) * L::MS;
                    float2 v[R];
#pragma unroll
                    for (int r = 0; r < R; r++) v[r] = p[r * L::MS];
                    bflyx<SIGN, R>(v);
                    store(b, i, v);
                }
        }
}

template <int SIGN, class L, bool IFAST, int M, int BP, class REV, int... RS>
struct GbDifRest;
template <int SIGN, class L, bool IFAST, int M, int BP, class REV, int R>
struct GbDifRest<SIGN, L, IFAST, M, BP, REV, R>
{
    static_assert(BP == R, "radix list must multiply to M");
    template <class StoreF>
    static __device__ __forceinline__ void run(float2* buf, const float2*, StoreF&& store)
    {
        dif_last<SIGN, L, IFAST, M, R, REV>(buf, store);
    }
};
template <int SIGN, class L, bool IFAST, int M, int BP, class REV, int R, int R2, int... REST>
struct GbDifRest<SIGN, L, IFAST, M, BP, REV, R, R2, REST...>
{
    template <class StoreF>
    static __device__ __forceinline__ void run(float2* buf, const float2* tw, StoreF&& store)
    {
        dif_mid<SIGN, L, M, BP, R>(buf, tw);
        __syncthreads();
        GbDifRest<SIGN, L, IFAST, M, BP / R, REV, R2, REST...>::run(buf, tw, store);
    }
};

// M = R0 * REST..., at least two passes
template <int M_, int R0, int... REST>
struct GbDif
{
    static_assert(sizeof...(REST) >= 1, "need at least two passes");
    static constexpr int kM = M_;
    static constexpr int kFirstR = R0;
    static constexpr int kLastR = GbLastOf<REST...>::value;
    static constexpr int kLastNI = M_ / kLastR;
    using Rev = GbRevBase<M_, R0, REST...>;
    // after dif_first + barrier: the remaining passes
    template <int SIGN, class L, bool IFAST, class StoreF>
    static __device__ __forceinline__ void run_rest(float2* buf, const float2* tw, StoreF&& store)
    {
        GbDifRest<SIGN, L, IFAST, M_, M_ / R0, Rev, REST...>::run(buf, tw, store);
    }
};

}  // namespace
