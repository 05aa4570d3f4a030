// The (N1, N2) splits that have compile-time in-place DIF kernels, and their tile configurations: shared by the inverse
// kernels (acq_inv.cu) and the forward kernels (acq_fwd.cu), which are separate translation units only to halve the build time.
#pragma once
#include "acq_dif.cuh"

namespace
{
// The (N1, N2) pairs with a fast path.  f(GbDif<N1...>{}, GbDif<N2...>{}, column-kernel config, row-kernel config);
// a config is GbCfg<LOGT, NT, OCC>: tile (columns / rows per CTA) = 2^LOGT, threads per CTA, CTAs per SM the register
// allocation is bounded for.  `variant` selects among alternative instantiations of one pair (tuning; 0 = default).
template <int LOGT_, int NT_, int OCC_>
struct GbCfg
{
    static constexpr int LOGT = LOGT_, NT = NT_, OCC = OCC_;
};

template <class F>
bool gb_with_dif_pair(int N1, int N2, int variant, F&& f)
{
#define GB_C ,
#define GB_DIF_PAIR(n1, n2, var, D1, D2, C1, C2)      \
    if (N1 == n1 && N2 == n2 && variant == var)       \
        {                                             \
            f(D1{}, D2{}, C1{}, C2{});                \
            return true;                              \
        }
    using D50 = GbDif<50, 10, 5>;
    using D80 = GbDif<80, 16, 5>;
    using D100 = GbDif<100, 10, 10>;
    using D125 = GbDif<125, 5, 5, 5>;
    using D128 = GbDif<128, 16, 8>;
    using D160 = GbDif<160, 16, 10>;
    using D160c = GbDif<160, 10, 16>;
    using D200 = GbDif<200, 8, 5, 5>;
    using D250 = GbDif<250, 10, 5, 5>;
    using D256 = GbDif<256, 16, 16>;
    using D200e = GbDif<200, 8, 25>;
    using D250e = GbDif<250, 10, 25>;
    using D320e = GbDif<320, 16, 20>;
    using D400e = GbDif<400, 16, 25>;
    using D500e = GbDif<500, 20, 25>;
    using D320 = GbDif<320, 8, 8, 5>;
    using D400 = GbDif<400, 16, 5, 5>;
    using D500 = GbDif<500, 10, 10, 5>;
    using S128 = GbCfg<4, 128, 8>;
    using S256 = GbCfg<4, 256, 4>;
    using E256 = GbCfg<3, 256, 4>;
    // Default (variant 0) of every pair: as few passes as the butterflies allow -- each pass is a shared-memory round trip, a
    // barrier and a table of twiddles; two passes with a radix-20 / 25 last butterfly in registers beat three passes by
    // 9 .. 24 % per step (profiles/r02_acq_fwd_sweep.txt).  The exception is the 125-point COLUMN transform: with 8 columns
    // per CTA a 25 x 5 or 5 x 25 split leaves most threads without a first- or last-pass item.
    GB_DIF_PAIR(50, 80, 0, D50, D80, S128, S128)       // 4000   GPS L1 C/A 1 ms @ 4 Msps
    GB_DIF_PAIR(80, 100, 0, D80, D100, S128, S128)     // 8000
    GB_DIF_PAIR(125, 128, 0, D125, D128, S256, S256)   // 16000
    GB_DIF_PAIR(125, 160, 0, D125, D160, S256, S256)   // 20000
    GB_DIF_PAIR(125, 200, 0, D125, D200e, E256, S256)  // 25000  GPS L5 / Galileo E5a 1 ms @ 25 Msps: rows 8 x 25
    GB_DIF_PAIR(125, 200, 1, D125, D200, E256, S256)   // rows 8 x 5 x 5 (1.7 % slower at C3)
    GB_DIF_PAIR(125, 256, 0, D125, D256, S256, S256)   // 32000, alternative split (tuning key acq_split_n1=125): 2 KB rows, 128 B aligned, but the 16x16 row passes conflict in shared memory
    GB_DIF_PAIR(160, 200, 0, D160c, D200e, S256, S256)  // 32000  Galileo E1 4 ms @ 8 Msps: columns 10 x 16, rows 8 x 25
    GB_DIF_PAIR(160, 200, 1, D160, D200, S256, S256)    // columns 16 x 10, rows 8 x 5 x 5: the default until the last session of round 2 (11 % slower at C2)
    GB_DIF_PAIR(160, 200, 2, D160c, D200, S256, S256)   // rows 8 x 5 x 5 (9.5 % slower)
    // (also measured and dropped: 32-row tiles 15 % slower, rows 10 x 20 19 %, rows 5 x 8 x 5 17 %, 3 CTAs / SM 4 %, 128-thread CTAs 5 %;
    //  125 x 200 with columns 5 x 25 30 % slower, 25 x 5 14 %; 400 x 500 with columns 20 x 20 as fast as 16 x 25)
    GB_DIF_PAIR(200, 200, 0, D200e, D200e, E256, S256) // 40000
    GB_DIF_PAIR(200, 200, 1, D200, D200, E256, S256)   // 8 x 5 x 5 both ways (15 % slower)
    GB_DIF_PAIR(200, 250, 0, D200e, D250e, E256, E256) // 50000
    GB_DIF_PAIR(200, 250, 1, D200, D250, E256, E256)   // three passes both ways (13.5 % slower)
    GB_DIF_PAIR(250, 320, 0, D250e, D320e, E256, E256) // 80000: 10 x 25, 16 x 20
    GB_DIF_PAIR(250, 320, 1, D250, D320, E256, E256)   // three passes both ways (18 % slower)
    GB_DIF_PAIR(400, 500, 0, D400e, D500e, E256, E256) // 200000: columns 16 x 25, rows 20 x 25
    GB_DIF_PAIR(400, 500, 1, D400, D500, E256, E256)   // 16 x 5 x 5, 10 x 10 x 5: the default until the last session of round 2 (24 % slower at C5)
#undef GB_DIF_PAIR
    return false;
}
}  // namespace
