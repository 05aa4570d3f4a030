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
#include <algorithm>

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

// ---- packed FP32 (sm_100a FFMA2 / FADD2 / FMUL2): two float32 lanes per issue slot, each lane IEEE round-to-nearest
// (or round-down where stated) exactly like the scalar instruction, so the index rule below stays bit-identical.
// ptxas folds a pk(a, a) operand into a scalar-broadcast source and pk(hi, lo) of an unpacked pair into a swizzle.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk(float lo, float hi)
{
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void upk(f32x2 v, float& lo, float& hi)
{
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c)
{
    f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 add2_rd(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("add.rm.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ float fneg_bits(float v) { return __int_as_float(__float_as_int(v) ^ (int)0x80000000); }  // integer pipe

// ---- raw sample words: loaded early (software prefetch), widened at use exactly as volk_gnsssdr_16ic_convert_32fc does
template <int ITEM>
struct GbRaw;
template <>
struct GbRaw<GB200_ITEM_32FC>
{
    typedef float2 T;
    static __device__ __forceinline__ T load(const void* b, size_t i) { return __ldg(reinterpret_cast<const float2*>(b) + i); }
    static __device__ __forceinline__ float2 cvt(T v) { return v; }
};
template <>
struct GbRaw<GB200_ITEM_16IC>
{
    typedef unsigned T;  // the packed (I, Q) int16 pair as one word: converted straight from its halves at use
    static __device__ __forceinline__ T load(const void* b, size_t i) { return __ldg(reinterpret_cast<const unsigned*>(b) + i); }
    static __device__ __forceinline__ float2 cvt(T v) { return make_float2((float)(short)(v & 0xffffu), (float)(short)(v >> 16)); }
};
template <>
struct GbRaw<GB200_ITEM_8IC>
{
    typedef unsigned short T;
    static __device__ __forceinline__ T load(const void* b, size_t i) { return __ldg(reinterpret_cast<const unsigned short*>(b) + i); }
    static __device__ __forceinline__ float2 cvt(T v) { return make_float2((float)(signed char)(v & 0xffu), (float)(signed char)(v >> 8)); }
};

// exp(j 2 pi t), t = fix / 2^64 turns.  The turn count is carried as a 64-bit fixed-point fraction (wraps mod 1 for
// free, integer pipe only -- no FP64 on the path); the top 32 bits, read as a signed fraction of a turn, feed sincospif.
__device__ __forceinline__ float2 phasor_fix(unsigned long long fix)
{
    const float half_turns = (float)(int)(fix >> 32) * 4.656612873077392578125e-10f;  // * 2^-31: angle / pi in [-1, 1)
    float sn, cs;
    sincospif(half_turns, &sn, &cs);
    return make_float2(cs, sn);
}

// Sum NV per-lane values over the 32 lanes of a warp with ~NV shuffles instead of 5*NV: every step halves the set of
// values a lane is responsible for (lanes with the step's bit set keep the upper half and hand the lower half to their
// partner).  Returns the index of the value whose warp total this lane holds in v[0], or NV if it holds none.
template <int NV>
__device__ __forceinline__ int warp_multi_reduce(float (&v)[NV], int lane)
{
    int id = 0, mine = NV;  // first value this lane is responsible for, and how many
    int cnt = NV;           // upper bound of `mine` over all lanes (compile time)
#pragma unroll
    for (int off = 16; off > 0; off >>= 1)
        {
            const int half = (cnt + 1) / 2;
            const bool upper = (lane & off) != 0;
#pragma unroll
            for (int j = 0; j < NV; j++)
                {
                    if (j < half)
                        {
                            const float hi = (j + half < cnt) ? v[(j + half) < NV ? (j + half) : 0] : 0.0f;
                            const float send = upper ? v[j] : hi;
                            const float keep = upper ? hi : v[j];
                            v[j] = keep + __shfl_xor_sync(0xffffffffu, send, off);
                        }
                }
            if (upper)
                {
                    id += half;
                    mine -= half;  // <= 0: this lane kept an empty upper half and owns nothing
                }
            else
                mine = min(mine, half);
            cnt = half;
        }
    return mine >= 1 ? id : NV;
}

// PROMPT: compile-time index of the tap the data-component code rides on (-1: read GbTrkWork::prompt_tap at run time)
//
// Carrier NCO: ph[n] = phase_0 * inc^n is factored per CTA as P_i * B_tid with n = n0 + tid + 256 i:
//   P_i   = exp(j 2 pi (a0 + (n0 + 256 i) theta))    one table entry per loop iteration, shared by the CTA
//   B_tid = m_tid * exp(j 2 pi tid theta)             per thread, constant over the loop
// so the loop rotates every sample by the CTA-uniform P_i only and the accumulators are multiplied by B_tid once at
// the end -- the per-sample phasor recurrence (a complex multiply per sample per thread) is gone.
template <int ITEM, int NT, bool HAS_DATA, bool CPLX, int PROMPT>
__global__ void __launch_bounds__(GB200_TRK_THREADS, 3)
    gb_trk_kernel(const GbTrkWork* __restrict__ work, int chunk, int max_chunks, int win_entries, float2* __restrict__ partials,
        unsigned* __restrict__ counters, float2* __restrict__ out, int nt_runtime, int magic_bits)
{
    constexpr int NACC = NT + (HAS_DATA ? 1 : 0);
    constexpr int EW = (CPLX || HAS_DATA) ? 2 : 1;  // floats per window entry: (re, im), (pilot, data) or one real chip
    typedef GbRaw<ITEM> RAW;
    extern __shared__ __align__(128) float s_dyn[];
    float* s_win = s_dyn;                                                          // [win_entries][EW]
    float4* s_P = reinterpret_cast<float4*>(s_dyn + (((size_t)EW * win_entries + 3) & ~(size_t)3));      // [chunk / 256 + 4]: (Px, Py, -Py, Px)
    __shared__ float s_red[GB200_TRK_THREADS / 32][2 * NACC];
    __shared__ bool s_last;
    __shared__ __align__(8) unsigned long long s_mbar;

    const int cidx = blockIdx.x % max_chunks;  // 1-D grid: request-major, chunk-minor
    const GbTrkWork& w = work[blockIdx.x / max_chunks];
    const int tid = threadIdx.x;
    if (tid < 2)
        {
            // the request record of the CTA that will follow this one on the SM: cold in L2 otherwise (one DRAM round trip
            // in front of everything else the CTA does)
            const unsigned nxt = (blockIdx.x + 3u * 148u) / (unsigned)max_chunks;
            if (nxt < gridDim.x / (unsigned)max_chunks)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char*>(work + nxt) + 128 * tid));
        }
    const int n0 = cidx * chunk;
    if (n0 >= w.n) return;
    const int n1 = min(w.n, n0 + chunk);
    const int nt = (NT == GB200_MAX_TAPS) ? nt_runtime : NT;  // NT == MAX_TAPS instantiation is the generic one

    float shifts[NT];
#pragma unroll
    for (int t = 0; t < NT; t++) shifts[t] = t < nt ? w.shifts[t] : 0.0f;
    const float step = w.code_step, rem = w.rem_code;
    const int L = w.code_len;

    // ---- stage the code window: asynchronous 4-byte copies (no register staging, all in flight at once); the carrier
    // table below is computed while they land ----
    float smin = shifts[0], smax = shifts[0];
#pragma unroll
    for (int t = 1; t < NT; t++)
        if (t < nt)
            {
                smin = fminf(smin, shifts[t]);
                smax = fmaxf(smax, shifts[t]);
            }
    // the loop runs whole groups of four samples per thread; the samples of the last group that lie beyond n1 are
    // zeroed, but their chips are still looked up, so the window covers the padded range
    const int n1_pad = n0 + ((n1 - n0 + 4 * GB200_TRK_THREADS - 1) / (4 * GB200_TRK_THREADS)) * (4 * GB200_TRK_THREADS);
    const float vlo = chip_arg(step, (float)n0, smin, rem);
    const float vhi = chip_arg(step, (float)(n1_pad - 1), smax, rem);
    constexpr int ALN = 16 / (4 * EW);  // window entries per 16 bytes (bulk-copy granularity)
    const bool window_ok = (step >= 0.0f) && (fabsf(vlo) < 4.0e6f) && (fabsf(vhi) < 4.0e6f) &&
                           ((int)floorf(vhi) - (int)floorf(vlo) + 1 + 2 * ALN <= win_entries);
    int lo = window_ok ? (int)floorf(vlo) : 0;
    bool bulk = false;
    if (window_ok)
        {
            const int W = (int)floorf(vhi) - lo + 1;
            const int start = wrap_index(lo, L);
            // Fast path: the handle's extended table (the code repeated past 2 L entries, already in window-entry layout)
            // holds [start, start + W) contiguously -> ONE bulk asynchronous copy (TMA) issued by one thread, starting at
            // the 16-byte boundary below `start`.
            const int head = start & (ALN - 1);
            const int Wb = (W + head + ALN - 1) & ~(ALN - 1);
            bulk = w.ext != nullptr && start - head + Wb <= w.ext_entries;
            if (bulk)
                {
                    lo -= head;  // window entry 0 is the aligned entry
                    if (tid == 0)
                        {
                            const unsigned mb = (unsigned)__cvta_generic_to_shared(&s_mbar);
                            const unsigned bytes = (unsigned)Wb * (4u * EW);
                            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mb) : "memory");
                            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
                            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"(bytes) : "memory");
                            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                                             (unsigned)__cvta_generic_to_shared(s_win)),
                                         "l"(w.ext + (size_t)(start - head) * EW), "r"(bytes), "r"(mb)
                                         : "memory");
                        }
                }
            else
                {
                    int src = wrap_index(lo + tid, L);
                    const int adv = GB200_TRK_THREADS % L;
                    const unsigned sw = (unsigned)__cvta_generic_to_shared(s_win);
                    for (int j = tid; j < W; j += GB200_TRK_THREADS)
                        {
                            if (CPLX)
                                asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sw + 8u * j), "l"(w.code + 2 * src) : "memory");
                            else if (HAS_DATA)
                                {
                                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sw + 8u * j), "l"(w.code + src) : "memory");
                                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sw + 8u * j + 4u), "l"(w.data_code + src) : "memory");
                                }
                            else
                                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sw + 4u * j), "l"(w.code + src) : "memory");
                            src += adv;
                            if (src >= L) src -= L;
                        }
                }
        }

    // ---- carrier table P_i for this CTA's iterations
    {
        const unsigned long long th = w.nco_theta_fix, a0 = w.nco_a0_fix;
        const int iters_cta = (n1 - n0 + GB200_TRK_THREADS - 1) / GB200_TRK_THREADS;
        for (int i = tid; i < ((iters_cta + 3) & ~3); i += GB200_TRK_THREADS)
            {
                const float2 P = phasor_fix(a0 + (unsigned long long)(n0 + GB200_TRK_THREADS * i) * th);
                s_P[i] = make_float4(P.x, P.y, -P.y, P.x);  // P and j P: the rotation needs both operand orders
            }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncthreads();
    if (bulk)
        {
            const unsigned mb = (unsigned)__cvta_generic_to_shared(&s_mbar);
            unsigned done = 0;
            while (!done)
                asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(done) : "r"(mb) : "memory");
        }

    f32x2 pacc[NACC];
#pragma unroll
    for (int t = 0; t < NACC; t++) pacc[t] = pk(0.f, 0.f);

    const int iters = (n1 - n0 - tid + GB200_TRK_THREADS - 1) / GB200_TRK_THREADS;  // samples of this thread
    const void* sbase = w.samples;
    int n = n0 + tid;
    float nf = (float)n;  // exact: n < 2^24
    if (window_ok)
        {
            // shared address of window entry 0 minus the float-magic bias: entry(bits) = swin + bits * (4 EW)
            // magic_bits == kFloorMagicBits, passed as a kernel argument so that the bias stays folded into this ONE base
            // (as a literal, ptxas re-associates it into an extra add per tap)
            const unsigned swin = (unsigned)__cvta_generic_to_shared(s_win) - (unsigned)(magic_bits + lo) * (4u * EW);
            const unsigned sP = (unsigned)__cvta_generic_to_shared(s_P);
            const int prompt = PROMPT >= 0 ? PROMPT : w.prompt_tap;
            constexpr int NP = NT / 2;  // tap pairs (packed index arithmetic); an odd last tap stays scalar
            f32x2 shp[NP > 0 ? NP : 1];
#pragma unroll
            for (int q = 0; q < NP; q++) shp[q] = pk(shifts[2 * q], shifts[2 * q + 1]);
            const f32x2 rem2 = pk(rem, rem), magic2 = pk(kFloorMagic, kFloorMagic);

            // one sample: rotate by the CTA-uniform P_i, look every tap's chip up, accumulate.
            //   rotate   (zr, zi) = xr * (Px, Py) + (-xi, xi) * (Py, Px)
            //   index    taps in pairs: ((sn, sn) + (shift_a, shift_b)) - (rem, rem), round-down add of the magic
            //   MAC      real codes (acc.x, acc.y) += (c, c) * (zr, zi); complex codes += zr * (cx, cy) + (-zi, zi) * (cy, cx)
            auto one = [&](typename RAW::T raw, int i) {
                const float2 x = RAW::cvt(raw);
                f32x2 P, Pj;
                asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];" : "=l"(P), "=l"(Pj) : "r"(sP + 16u * (unsigned)i));
                const f32x2 z = fma2(pk(x.x, x.x), P, mul2(pk(x.y, x.y), Pj));
                float zr, zi;
                upk(z, zr, zi);
                const f32x2 zs = pk(fneg_bits(zi), zi);
                const float sn = __fmul_rn(step, nf);
                const f32x2 sn2 = pk(sn, sn);
                unsigned ap = 0;  // shared address of the prompt tap's window entry
                auto tap = [&](int t, float bits_f) {
                    const unsigned a = swin + (unsigned)__float_as_int(bits_f) * (4u * EW);
                    if (CPLX)
                        {
                            const float2 c = lds_f32x2(a);
                            pacc[t] = fma2(pk(zr, zr), pk(c.x, c.y), pacc[t]);
                            pacc[t] = fma2(zs, pk(c.y, c.x), pacc[t]);
                        }
                    else if (HAS_DATA && (PROMPT >= 0 ? t == PROMPT : false))
                        {
                            const float2 c = lds_f32x2(a);  // (pilot, data) chips of the prompt tap in one load
                            pacc[t] = fma2(pk(c.x, c.x), z, pacc[t]);
                            pacc[NT] = fma2(pk(c.y, c.y), z, pacc[NT]);
                        }
                    else
                        {
                            const float c = lds_f32(a);
                            pacc[t] = fma2(pk(c, c), z, pacc[t]);
                            if (HAS_DATA && PROMPT < 0 && t == prompt) ap = a;
                        }
                };
#pragma unroll
                for (int q = 0; q < NP; q++)
                    {
                        if (2 * q < nt)
                            {
                                const f32x2 v2 = add2_rd(sub2(add2(sn2, shp[q]), rem2), magic2);
                                float b0, b1;
                                upk(v2, b0, b1);
                                tap(2 * q, b0);
                                if (2 * q + 1 < nt) tap(2 * q + 1, b1);
                            }
                    }
                if constexpr ((NT & 1) != 0)
                    {
                        if (NT - 1 < nt) tap(NT - 1, __fadd_rd(__fsub_rn(__fadd_rn(sn, shifts[NT - 1]), rem), kFloorMagic));
                    }
                if (HAS_DATA && !CPLX && PROMPT < 0)
                    {
                        const float cd = lds_f32(ap + 4u);
                        pacc[NT] = fma2(pk(cd, cd), z, pacc[NT]);
                    }
                nf += (float)GB200_TRK_THREADS;
            };

            // groups of four samples with the next group's raw words already in flight (L2 latency >> one group); the last
            // group is padded with zero samples (their chips come from the padded window)
            // Loads past n1 stay inside the ring allocation (GB200_RING_TAIL_ITEMS spare items behind every window), so only
            // the LAST group needs its out-of-range samples forced to zero.
            const int groups = (iters + 3) >> 2;
            typename RAW::T q0 = RAW::load(sbase, (size_t)n), q1 = RAW::load(sbase, (size_t)n + GB200_TRK_THREADS),
                            q2 = RAW::load(sbase, (size_t)n + 2 * GB200_TRK_THREADS), q3 = RAW::load(sbase, (size_t)n + 3 * GB200_TRK_THREADS);
            int i = 0;
            for (int g = 0; g + 1 < groups; g++, i += 4)
                {
                    const typename RAW::T r0 = q0, r1 = q1, r2 = q2, r3 = q3;
                    n += 4 * GB200_TRK_THREADS;
                    q0 = RAW::load(sbase, (size_t)n);
                    q1 = RAW::load(sbase, (size_t)n + GB200_TRK_THREADS);
                    q2 = RAW::load(sbase, (size_t)n + 2 * GB200_TRK_THREADS);
                    q3 = RAW::load(sbase, (size_t)n + 3 * GB200_TRK_THREADS);
                    one(r0, i);
                    one(r1, i + 1);
                    one(r2, i + 2);
                    one(r3, i + 3);
                }
            if (groups > 0)
                {
                    const typename RAW::T zero = typename RAW::T();
                    one(i < iters ? q0 : zero, i);
                    one(i + 1 < iters ? q1 : zero, i + 1);
                    one(i + 2 < iters ? q2 : zero, i + 2);
                    one(i + 3 < iters ? q3 : zero, i + 3);
                }
        }
    else
        {
            // general path (step < 0, |index| huge, or window larger than shared memory): table in L2, full modulo
            for (int i = 0; i < iters; i++, n += GB200_TRK_THREADS)
                {
                    const float2 x = RAW::cvt(RAW::load(sbase, (size_t)n));
                    const float4 P = s_P[i];
                    const float zr = fmaf(x.x, P.x, -x.y * P.y);
                    const float zi = fmaf(x.x, P.y, x.y * P.x);
                    const f32x2 z = pk(zr, zi), zs = pk(-zi, zi);
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
                                            pacc[t] = fma2(pk(zr, zr), pk(c.x, c.y), pacc[t]);
                                            pacc[t] = fma2(zs, pk(c.y, c.x), pacc[t]);
                                        }
                                    else
                                        {
                                            const float c = __ldg(w.code + j);
                                            pacc[t] = fma2(pk(c, c), z, pacc[t]);
                                            if (HAS_DATA && t == w.prompt_tap)
                                                {
                                                    const float cd = __ldg(w.data_code + j);
                                                    pacc[NT] = fma2(pk(cd, cd), z, pacc[NT]);
                                                }
                                        }
                                }
                        }
                    nf += (float)GB200_TRK_THREADS;
                }
        }

    // ---- per-thread carrier factor B_tid, then the CTA reduction
    float red[2 * NACC];
    {
        float2 B = phasor_fix((unsigned long long)tid * w.nco_theta_fix);
        // |inc| != 1 in float32: between two renormalisations (every 256 samples, chunk starts are multiples of 256)
        // the reference's phasor magnitude is |inc|^e, e = ((n-1) mod 256) + 1 -- constant per thread here
        const float m = fmaf((float)(((threadIdx.x + 255u) & 255u) + 1u), w.nco_mag_delta, 1.0f);
        B.x *= m;
        B.y *= m;
#pragma unroll
        for (int t = 0; t < NACC; t++)
            {
                float ar, ai;
                upk(pacc[t], ar, ai);
                red[2 * t] = fmaf(ar, B.x, -ai * B.y);
                red[2 * t + 1] = fmaf(ar, B.y, ai * B.x);
            }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    {
        const int id = warp_multi_reduce<2 * NACC>(red, lane);
        if (id < 2 * NACC) s_red[warp][id] = red[0];
    }
    __syncthreads();
    const int nchunks = (w.n + chunk - 1) / chunk;
    float* my_part = reinterpret_cast<float*>(partials + ((size_t)w.out_index * max_chunks + cidx) * GB200_OUT_STRIDE);
    if (threadIdx.x < 2 * NACC)
        {
            float sum = 0.f;
#pragma unroll
            for (int k = 0; k < GB200_TRK_THREADS / 32; k++) sum += s_red[k][threadIdx.x];
            if (nchunks == 1)
                reinterpret_cast<float*>(out + (size_t)w.out_index * GB200_OUT_STRIDE)[threadIdx.x] = sum;
            else
                my_part[threadIdx.x] = sum;
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
            float2 sum = make_float2(0.f, 0.f);
            const volatile float2* p = partials + (size_t)w.out_index * max_chunks * GB200_OUT_STRIDE + threadIdx.x;
            for (int c = 0; c < nchunks; c++)
                {
                    sum.x += p[(size_t)c * GB200_OUT_STRIDE].x;
                    sum.y += p[(size_t)c * GB200_OUT_STRIDE].y;
                }
            out[(size_t)w.out_index * GB200_OUT_STRIDE + threadIdx.x] = sum;
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
    int chunk, int max_chunks, int win_entries, float2* d_partials, unsigned* d_counters, float2* d_out, cudaStream_t st)
{
    dim3 block(GB200_TRK_THREADS);
    // dynamic shared memory: the code window (1 or 2 floats per entry) + the carrier table P_i
    const size_t smem = gb_trk_smem_bytes(win_entries, code_kind == GB200_CODE_CPLX || has_data, chunk);
#define GB_LAUNCH_K(K)                                                                                                        \
    do                                                                                                                        \
        {                                                                                                                     \
            auto kern = K;                                                                                                    \
            cudaError_t ea = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GB200_TRK_SMEM_MAX); \
            if (ea != cudaSuccess) return ea;                                                                                 \
            kern<<<grid, block, smem, st>>>(d_work, chunk, max_chunks, win_entries, d_partials, d_counters, d_out, n_taps, kFloorMagicBits); \
        }                                                                                                                     \
    while (0)
#define GB_LAUNCH(NT, HD, CP) GB_LAUNCH_K((gb_trk_kernel<ITEM, NT, HD, CP, -1>))
#define GB_LAUNCH_P(NT, P) GB_LAUNCH_K((gb_trk_kernel<ITEM, NT, true, false, P>))
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
#undef GB_LAUNCH_K
#undef GB_LAUNCH_P
    return cudaGetLastError();
}

}  // namespace

cudaError_t gb_trk_launch(const GbTrkWork* d_work, int n_work, int max_n, int n_taps, bool has_data, int prompt_tap, int item_type,
    int code_kind, bool high_dyn, float2* d_partials, unsigned* d_counters, float2* d_out, int chunk, int win_entries, cudaStream_t st,
    int* launches)
{
    if (n_work <= 0) return cudaSuccess;
    const int max_chunks = (max_n + chunk - 1) / chunk;
    dim3 grid((unsigned)max_chunks * (unsigned)n_work);
    cudaError_t e;
    switch (item_type)
        {
        case GB200_ITEM_32FC:
            e = launch_item<GB200_ITEM_32FC>(d_work, grid, n_taps, has_data, prompt_tap, code_kind, high_dyn, chunk, max_chunks, win_entries, d_partials, d_counters, d_out, st);
            break;
        case GB200_ITEM_16IC:
            e = launch_item<GB200_ITEM_16IC>(d_work, grid, n_taps, has_data, prompt_tap, code_kind, high_dyn, chunk, max_chunks, win_entries, d_partials, d_counters, d_out, st);
            break;
        case GB200_ITEM_8IC:
            e = launch_item<GB200_ITEM_8IC>(d_work, grid, n_taps, has_data, prompt_tap, code_kind, high_dyn, chunk, max_chunks, win_entries, d_partials, d_counters, d_out, st);
            break;
        default:
            return cudaErrorInvalidValue;
        }
    if (launches) (*launches)++;
    return e;
}

namespace
{
__global__ void gb_copy_words_kernel(unsigned long long* __restrict__ dst, const unsigned long long* __restrict__ src, size_t n)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}
}  // namespace

// Pulls request records out of device-mapped pinned host memory with SM loads.  (A cudaMemcpyAsync would sit in the
// host-to-device copy engine's queue behind every sample push still in flight.)
cudaError_t gb_copy_words(void* dst, const void* src, size_t bytes, cudaStream_t st)
{
    const size_t n = bytes / 8;
    if (n == 0) return cudaSuccess;
    const int threads = 256;
    const int blocks = (int)std::min<size_t>((n + threads - 1) / threads, 296);
    gb_copy_words_kernel<<<blocks, threads, 0, st>>>((unsigned long long*)dst, (const unsigned long long*)src, n);
    return cudaGetLastError();
}
