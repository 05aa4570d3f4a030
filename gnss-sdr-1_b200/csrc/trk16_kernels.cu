// int16 multicorrelator with the reference's exact fixed-point semantics, batched over (channel, epoch) requests.
//
// Replaces (semantics, not code):
//   Cpu_Multicorrelator_16sc::Carrier_wipeoff_multicorrelator_resampler   src/algorithms/tracking/libs/cpu_multicorrelator_16sc.cc:78-105
//   volk_gnsssdr_16ic_xn_resampler_16ic_xn_generic                         ..._16ic_xn_resampler_16ic_xn.h:75-94
//   volk_gnsssdr_16ic_x2_rotator_dot_prod_16ic_xn_generic                  ..._16ic_x2_rotator_dot_prod_16ic_xn.h:73-108
//
// Unlike the float correlator this one has to be BIT-exact: the rotated sample is rounded to int16 (rintf), the complex
// int16 product is narrowed to 16 bits, and the accumulation saturates at every sample, so neither an approximated NCO
// nor a re-ordered sum is admissible.  Both are made parallel without changing a bit:
//  * the carrier phasor recurrence (float32, renormalised by hypotf at n % 256 == 0) is run once per request by one
//    thread that only stores the phasor at every 256-sample boundary (gb_trk16_nco_kernel); every 256-sample block is
//    then replayed from its checkpoint with the identical float32 operation sequence;
//  * a saturating accumulation step s -> clamp(s + x) is a "clamp-add" map s -> min(max(s + a, l), h); such maps are
//    closed under composition ((a1,l1,h1) then (a2,l2,h2) = (a1+a2, clamp(l1+a2,l2,h2), clamp(h1+a2,l2,h2))), so each block
//    reduces its 256 steps to one (a, l, h) triple per tap and component and the triples are composed in block order.
#include "gb200_internal.cuh"

namespace
{
constexpr int kBlk = 256;
constexpr int kLo = -32768, kHi = 32767;
constexpr int kNegInf = -(1 << 30), kPosInf = (1 << 30);

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return min(max(v, lo), hi); }

// glibc's hypotf for finite floats: the double-precision sum of squares, one correctly rounded sqrt, one rounding to float
__device__ __forceinline__ float hypotf_like_glibc(float x, float y)
{
    return (float)sqrt((double)x * (double)x + (double)y * (double)y);
}

// one thread per request: phasor BEFORE sample 256*b, for every block b
__global__ void gb_trk16_nco_kernel(const GbTrk16Work* __restrict__ work, int n_work, int max_blocks, float2* __restrict__ ckpt)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_work) return;
    const GbTrk16Work w = work[r];
    float pr = w.phase0.x, pi = w.phase0.y;
    const float ir = w.inc.x, ii = w.inc.y;
    float2* out = ckpt + (size_t)w.out_index * max_blocks;
    for (int n = 0; n < w.n; n++)
        {
            if ((n & (kBlk - 1)) == 0)
                {
                    out[n / kBlk] = make_float2(pr, pi);
                    const float h = hypotf_like_glibc(pr, pi);
                    pr = __fdiv_rn(pr, h);
                    pi = __fdiv_rn(pi, h);
                }
            const float nr = __fsub_rn(__fmul_rn(pr, ir), __fmul_rn(pi, ii));
            const float ni = __fadd_rn(__fmul_rn(pr, ii), __fmul_rn(pi, ir));
            pr = nr;
            pi = ni;
        }
}

// one thread per (request, 256-sample block): replays the block, emits one clamp-add triple per tap and component
template <int NT>
__global__ void gb_trk16_block_kernel(const GbTrk16Work* __restrict__ work, int n_work, int max_blocks, const float2* __restrict__ ckpt,
    int4* __restrict__ triples, int nt_runtime)
{
    const int gid = blockIdx.x * blockDim.x + threadIdx.x;
    const int r = gid / max_blocks, b = gid - r * max_blocks;
    if (r >= n_work) return;
    const GbTrk16Work& w = work[r];
    const int n0 = b * kBlk;
    if (n0 >= w.n) return;
    const int n1 = min(w.n, n0 + kBlk);
    const int nt = (NT == GB200_MAX_TAPS) ? nt_runtime : NT;
    const float2 p0 = ckpt[(size_t)w.out_index * max_blocks + b];
    float pr = p0.x, pi = p0.y;
    const float ir = w.inc.x, ii = w.inc.y;
    const short2* x = reinterpret_cast<const short2*>(w.samples);
    const short2* code = reinterpret_cast<const short2*>(w.code);
    const int L = w.code_len;
    int a[NT][2], lo[NT][2], hi[NT][2];
#pragma unroll
    for (int t = 0; t < NT; t++)
#pragma unroll
        for (int c = 0; c < 2; c++)
            {
                a[t][c] = 0;
                lo[t][c] = kNegInf;
                hi[t][c] = kPosInf;
            }
    for (int n = n0; n < n1; n++)
        {
            const short2 s = __ldg(x + n);
            const float xr = (float)s.x, xi = (float)s.y;
            const float wr = __fsub_rn(__fmul_rn(xr, pr), __fmul_rn(xi, pi));
            const float wi = __fadd_rn(__fmul_rn(xr, pi), __fmul_rn(xi, pr));
            const int sr = (int)(short)__float2int_rn(wr), si = (int)(short)__float2int_rn(wi);  // rintf, narrowed like (int16_t)
            if (n == n0)
                {
                    const float h = hypotf_like_glibc(pr, pi);
                    pr = __fdiv_rn(pr, h);
                    pi = __fdiv_rn(pi, h);
                }
            {
                const float nr = __fsub_rn(__fmul_rn(pr, ir), __fmul_rn(pi, ii));
                const float ni = __fadd_rn(__fmul_rn(pr, ii), __fmul_rn(pi, ir));
                pr = nr;
                pi = ni;
            }
            const float sn = __fmul_rn(w.code_step, (float)n);
#pragma unroll
            for (int t = 0; t < NT; t++)
                {
                    if (t < nt)
                        {
                            const float v = __fsub_rn(__fadd_rn(sn, w.shifts[t]), w.rem_code);
                            int idx = (int)floor((double)v) % L;
                            if (idx < 0) idx += L;
                            const short2 c = __ldg(code + idx);
                            const int tr = (int)(short)(sr * (int)c.x - si * (int)c.y);  // complex int16 product, each part narrowed to 16 bits
                            const int ti = (int)(short)(sr * (int)c.y + si * (int)c.x);
                            a[t][0] += tr;
                            lo[t][0] = clampi(lo[t][0] + tr, kLo, kHi);
                            hi[t][0] = clampi(hi[t][0] + tr, kLo, kHi);
                            a[t][1] += ti;
                            lo[t][1] = clampi(lo[t][1] + ti, kLo, kHi);
                            hi[t][1] = clampi(hi[t][1] + ti, kLo, kHi);
                        }
                }
        }
    int4* out = triples + ((size_t)w.out_index * max_blocks + b) * (2 * GB200_MAX_TAPS);
#pragma unroll
    for (int t = 0; t < NT; t++)
        if (t < nt)
            {
                out[2 * t] = make_int4(a[t][0], lo[t][0], hi[t][0], 0);
                out[2 * t + 1] = make_int4(a[t][1], lo[t][1], hi[t][1], 0);
            }
}

// one thread per (request, tap, component): compose the block maps in order, apply to 0
__global__ void gb_trk16_compose_kernel(const GbTrk16Work* __restrict__ work, int n_work, int max_blocks, const int4* __restrict__ triples,
    short* __restrict__ out)
{
    const int gid = blockIdx.x * blockDim.x + threadIdx.x;
    const int r = gid / (2 * GB200_MAX_TAPS), k = gid - r * (2 * GB200_MAX_TAPS);
    if (r >= n_work) return;
    const GbTrk16Work& w = work[r];
    if (k >= 2 * w.n_taps) return;
    const int nblk = (w.n + kBlk - 1) / kBlk;
    const int4* tr = triples + (size_t)w.out_index * max_blocks * (2 * GB200_MAX_TAPS) + k;
    int s = 0;  // result[n_vec] = 0 before the first sample
    for (int b = 0; b < nblk; b++)
        {
            const int4 f = tr[(size_t)b * (2 * GB200_MAX_TAPS)];
            s = min(max(s + f.x, f.y), f.z);
        }
    out[(size_t)w.out_index * (2 * GB200_MAX_TAPS) + k] = (short)s;
}
}  // namespace

cudaError_t gb_trk16_launch(const GbTrk16Work* d_work, int n_work, int max_n, int n_taps, float2* d_ckpt, int4* d_triples, short* d_out,
    cudaStream_t st, int* launches)
{
    const int max_blocks = (max_n + kBlk - 1) / kBlk;
    gb_trk16_nco_kernel<<<(n_work + 63) / 64, 64, 0, st>>>(d_work, n_work, max_blocks, d_ckpt);
    const int total = n_work * max_blocks;
    const dim3 grid((total + 63) / 64);
    switch (n_taps)
        {
        case 1: gb_trk16_block_kernel<1><<<grid, 64, 0, st>>>(d_work, n_work, max_blocks, d_ckpt, d_triples, n_taps); break;
        case 3: gb_trk16_block_kernel<3><<<grid, 64, 0, st>>>(d_work, n_work, max_blocks, d_ckpt, d_triples, n_taps); break;
        case 5: gb_trk16_block_kernel<5><<<grid, 64, 0, st>>>(d_work, n_work, max_blocks, d_ckpt, d_triples, n_taps); break;
        default: gb_trk16_block_kernel<GB200_MAX_TAPS><<<grid, 64, 0, st>>>(d_work, n_work, max_blocks, d_ckpt, d_triples, n_taps); break;
        }
    const int nc = n_work * 2 * GB200_MAX_TAPS;
    gb_trk16_compose_kernel<<<(nc + 127) / 128, 128, 0, st>>>(d_work, n_work, max_blocks, d_triples, d_out);
    if (launches) *launches += 3;
    return cudaGetLastError();
}
