/* TEST INFRASTRUCTURE ONLY -- see gnss_oracle.h for scope and parity status.
 * Build: oracle/Makefile (gcc -O3 -ffp-contract=off: no FMA contraction).
 * All reference citations are relative to /root/reference/. "volk:" abbreviates
 * src/algorithms/libs/volk_gnsssdr_module/volk_gnsssdr/kernels/volk_gnsssdr/.
 */
#include "gnss_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct { float re, im; } cf;
typedef struct { double re, im; } cd;

/* C99 complex-float product as GCC evaluates `a * b` without -ffast-math / FMA */
static inline cf cmulf_(cf a, cf b)
{
    cf r;
    r.re = a.re * b.re - a.im * b.im;
    r.im = a.re * b.im + a.im * b.re;
    return r;
}

/* ======================= tracking ======================= */

/* volk: volk_gnsssdr_32f_xn_resampler_32f_xn.h:77-95 (_generic) */
static inline int chip_index_(float v, int code_len)
{
    int idx = (int)floor((double)v);
    if (idx < 0) idx += code_len * (int)((unsigned)abs(idx) / (unsigned)code_len + 1u);
    idx = (int)((unsigned)idx % (unsigned)code_len);
    return idx;
}

void go_resampler_real(float* out, const float* code, float rem_chips, float step_chips,
    const float* shifts, int code_len, int n_taps, int n)
{
    for (int t = 0; t < n_taps; t++)
        {
            float* o = out + (size_t)t * n;
            for (unsigned k = 0; k < (unsigned)n; k++)
                {
                    float v = step_chips * (float)k + shifts[t] - rem_chips; /* left to right, float32 */
                    o[k] = code[chip_index_(v, code_len)];
                }
        }
}

/* volk: volk_gnsssdr_32f_xn_high_dynamics_resampler_32f_xn.h:81-106 (_generic).
 * tap 0 gets the rate*(float)(n*n) term with n*n in unsigned arithmetic (wraps at n>=65536);
 * taps >=1 are circular SAMPLE shifts of tap 0. */
void go_resampler_real_hd(float* out, const float* code, float rem_chips, float step_chips,
    float rate_chips, const float* shifts, int code_len, int n_taps, int n)
{
    for (unsigned k = 0; k < (unsigned)n; k++)
        {
            float v = step_chips * (float)k + rate_chips * (float)(k * k) + shifts[0] - rem_chips;
            out[k] = code[chip_index_(v, code_len)];
        }
    unsigned shift_samples = 0;
    for (int t = 1; t < n_taps; t++)
        {
            shift_samples += (unsigned)(int)round((double)((shifts[t] - shifts[t - 1]) / step_chips));
            float* o = out + (size_t)t * n;
            memcpy(o, out + shift_samples, ((size_t)n - shift_samples) * sizeof(float));
            memcpy(o + ((size_t)n - shift_samples), out, (size_t)shift_samples * sizeof(float));
        }
}

/* volk: volk_gnsssdr_32fc_xn_resampler_32fc_xn.h _generic (same index maths, complex code) */
void go_resampler_cplx(float* out_iq, const float* code_iq, float rem_chips, float step_chips,
    const float* shifts, int code_len, int n_taps, int n)
{
    for (int t = 0; t < n_taps; t++)
        {
            float* o = out_iq + (size_t)t * n * 2;
            for (unsigned k = 0; k < (unsigned)n; k++)
                {
                    float v = step_chips * (float)k + shifts[t] - rem_chips;
                    int idx = chip_index_(v, code_len);
                    o[2 * k] = code_iq[2 * idx];
                    o[2 * k + 1] = code_iq[2 * idx + 1];
                }
        }
}

/* volk: volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn.h:81-113 (_generic) */
void go_rotator_dot_prod_real(float* result_iq, const float* in_iq, float inc_re, float inc_im,
    float* phase_iq, const float* codes, int n_taps, int n)
{
    cf phase = {phase_iq[0], phase_iq[1]};
    const cf inc = {inc_re, inc_im};
    for (int t = 0; t < n_taps; t++) result_iq[2 * t] = result_iq[2 * t + 1] = 0.0f;
    for (unsigned k = 0; k < (unsigned)n; k++)
        {
            cf x = {in_iq[2 * k], in_iq[2 * k + 1]};
            cf w = cmulf_(x, phase);
            if (k % 256 == 0)
                {
                    float a = hypotf(phase.re, phase.im);
                    phase.re /= a;
                    phase.im /= a;
                }
            phase = cmulf_(phase, inc);
            for (int t = 0; t < n_taps; t++)
                {
                    float c = codes[(size_t)t * n + k];
                    result_iq[2 * t] += w.re * c;
                    result_iq[2 * t + 1] += w.im * c;
                }
        }
    phase_iq[0] = phase.re;
    phase_iq[1] = phase.im;
}

/* volk: volk_gnsssdr_32fc_32f_high_dynamic_rotator_dot_prod_32fc_xn.h:120-150 (_generic_arg
 * formulation: rate phasor = (cos,sin)((float)(n*n)*arg(rate)); the cpowf form at :82-116 is the
 * same number up to libm rounding).  Sample n is rotated by phase0*inc^n*rate^((n-1)^2): the rate
 * factor is applied after the sample is used.  The n%256 renormalisation is a no-op on the result
 * because *phase is overwritten (it only changes *phase transiently). */
void go_rotator_dot_prod_real_hd(float* result_iq, const float* in_iq, float inc_re, float inc_im,
    float rate_re, float rate_im, float* phase_iq, const float* codes, int n_taps, int n)
{
    cf phase = {phase_iq[0], phase_iq[1]};
    cf phase_doppler = phase;
    const cf inc = {inc_re, inc_im};
    const float arga = atan2f(rate_im, rate_re);
    for (int t = 0; t < n_taps; t++) result_iq[2 * t] = result_iq[2 * t + 1] = 0.0f;
    for (unsigned k = 0; k < (unsigned)n; k++)
        {
            if (k % 256 == 0)
                {
                    float a = hypotf(phase.re, phase.im);
                    phase.re /= a;
                    phase.im /= a;
                }
            cf x = {in_iq[2 * k], in_iq[2 * k + 1]};
            cf w = cmulf_(x, phase);
            phase_doppler = cmulf_(phase_doppler, inc);
            float theta = (float)(k * k) * arga;
            cf pr = {cosf(theta), sinf(theta)};
            phase = cmulf_(phase_doppler, pr);
            for (int t = 0; t < n_taps; t++)
                {
                    float c = codes[(size_t)t * n + k];
                    result_iq[2 * t] += w.re * c;
                    result_iq[2 * t + 1] += w.im * c;
                }
        }
    phase_iq[0] = phase.re;
    phase_iq[1] = phase.im;
}

/* volk: volk_gnsssdr_32fc_x2_rotator_dot_prod_32fc_xn.h:81-120 (_generic) */
void go_rotator_dot_prod_cplx(float* result_iq, const float* in_iq, float inc_re, float inc_im,
    float* phase_iq, const float* codes_iq, int n_taps, int n)
{
    cf phase = {phase_iq[0], phase_iq[1]};
    const cf inc = {inc_re, inc_im};
    for (int t = 0; t < n_taps; t++) result_iq[2 * t] = result_iq[2 * t + 1] = 0.0f;
    for (unsigned k = 0; k < (unsigned)n; k++)
        {
            cf x = {in_iq[2 * k], in_iq[2 * k + 1]};
            cf w = cmulf_(x, phase);
            if (k % 256 == 0)
                {
                    float a = hypotf(phase.re, phase.im);
                    phase.re /= a;
                    phase.im /= a;
                }
            phase = cmulf_(phase, inc);
            for (int t = 0; t < n_taps; t++)
                {
                    const float* cc = codes_iq + ((size_t)t * n + k) * 2;
                    cf c = {cc[0], cc[1]};
                    cf p = cmulf_(w, c);
                    result_iq[2 * t] += p.re;
                    result_iq[2 * t + 1] += p.im;
                }
        }
    phase_iq[0] = phase.re;
    phase_iq[1] = phase.im;
}

/* src/algorithms/tracking/libs/cpu_multicorrelator_real_codes.cc:101-152 */
void go_multicorrelator_real(const float* in_iq, int n, const float* code, int code_len, const float* shifts,
    int n_taps, float rem_carr_phase_rad, float phase_step_rad, float phase_rate_step_rad,
    float rem_code_phase_chips, float code_phase_step_chips, float code_phase_rate_step_chips,
    int high_dyn, float* corr_out_iq)
{
    float* codes = (float*)malloc(sizeof(float) * (size_t)n_taps * (size_t)n);
    float phase[2] = {cosf(rem_carr_phase_rad), -sinf(rem_carr_phase_rad)}; /* :140-141 */
    /* std::exp(lv_32fc_t(0.0, -step)) == std::polar(expf(0), -step) == (cosf(-step), sinf(-step)) */
    if (high_dyn == 2)
        {
            /* the six-argument overload (:155-170) with d_use_high_dynamics_resampler left at its constructor default (true, :49):
             * high-dynamics code resampler, PLAIN rotator -- what gps_l1_ca_kf_tracking_cc runs */
            go_resampler_real_hd(codes, code, rem_code_phase_chips, code_phase_step_chips, code_phase_rate_step_chips, shifts, code_len, n_taps, n);
            go_rotator_dot_prod_real(corr_out_iq, in_iq, cosf(-phase_step_rad), sinf(-phase_step_rad), phase, codes, n_taps, n);
        }
    else if (high_dyn)
        {
            go_resampler_real_hd(codes, code, rem_code_phase_chips, code_phase_step_chips, code_phase_rate_step_chips, shifts, code_len, n_taps, n);
            go_rotator_dot_prod_real_hd(corr_out_iq, in_iq, cosf(-phase_step_rad), sinf(-phase_step_rad),
                cosf(-phase_rate_step_rad), sinf(-phase_rate_step_rad), phase, codes, n_taps, n);
        }
    else
        {
            go_resampler_real(codes, code, rem_code_phase_chips, code_phase_step_chips, shifts, code_len, n_taps, n);
            go_rotator_dot_prod_real(corr_out_iq, in_iq, cosf(-phase_step_rad), sinf(-phase_step_rad), phase, codes, n_taps, n);
        }
    free(codes);
}

/* src/algorithms/tracking/libs/cpu_multicorrelator.cc:99-126 */
void go_multicorrelator_cplx(const float* in_iq, int n, const float* code_iq, int code_len, const float* shifts,
    int n_taps, float rem_carr_phase_rad, float phase_step_rad, float rem_code_phase_chips,
    float code_phase_step_chips, float* corr_out_iq)
{
    float* codes = (float*)malloc(sizeof(float) * 2 * (size_t)n_taps * (size_t)n);
    float phase[2] = {cosf(rem_carr_phase_rad), -sinf(rem_carr_phase_rad)};
    go_resampler_cplx(codes, code_iq, rem_code_phase_chips, code_phase_step_chips, shifts, code_len, n_taps, n);
    go_rotator_dot_prod_cplx(corr_out_iq, in_iq, cosf(-phase_step_rad), sinf(-phase_step_rad), phase, codes, n_taps, n);
    free(codes);
}

/* Cpu_Multicorrelator_16sc::Carrier_wipeoff_multicorrelator_resampler (src/algorithms/tracking/libs/cpu_multicorrelator_16sc.cc:78-105)
 * = volk_gnsssdr_16ic_xn_resampler_16ic_xn_generic (..._16ic_xn_resampler_16ic_xn.h:75-94: the float32 index rule of the
 * float resampler) + volk_gnsssdr_16ic_x2_rotator_dot_prod_16ic_xn_generic (..._16ic_x2_rotator_dot_prod_16ic_xn.h:73-108):
 * rotate in float32, rintf to int16, THEN (for n % 256 == 0) renormalise the phasor by hypotf, advance it, complex int16
 * product truncated to 16 bits, saturating int16 accumulation in sample order. */
static int16_t sat16_(int32_t v) { return (int16_t)(v < -32768 ? -32768 : (v > 32767 ? 32767 : v)); }
void go_multicorrelator_16sc(const int16_t* in_iq, int n, const int16_t* code_iq, int code_len, const float* shifts, int n_taps,
    float rem_carr_phase_rad, float phase_step_rad, float rem_code_phase_chips, float code_phase_step_chips, int16_t* corr_out_iq)
{
    float pr = cosf(rem_carr_phase_rad), pi = -sinf(rem_carr_phase_rad);
    const float ir = cosf(-phase_step_rad), ii = sinf(-phase_step_rad); /* std::exp(lv_32fc_t(0, -step)) */
    for (int t = 0; t < 2 * n_taps; t++) corr_out_iq[t] = 0;
    for (unsigned k = 0; k < (unsigned)n; k++)
        {
            const float xr = (float)in_iq[2 * k], xi = (float)in_iq[2 * k + 1];
            const float wr = xr * pr - xi * pi, wi = xr * pi + xi * pr;
            const int16_t sr = (int16_t)rintf(wr), si = (int16_t)rintf(wi);
            if (k % 256 == 0)
                {
                    const float h = hypotf(pr, pi);
                    pr /= h;
                    pi /= h;
                }
            {
                const float nr = pr * ir - pi * ii, ni = pr * ii + pi * ir;
                pr = nr;
                pi = ni;
            }
            for (int t = 0; t < n_taps; t++)
                {
                    const float v = code_phase_step_chips * (float)k + shifts[t] - rem_code_phase_chips;
                    const int idx = chip_index_(v, code_len);
                    const int32_t cr = code_iq[2 * idx], ci = code_iq[2 * idx + 1];
                    const int16_t tr = (int16_t)((int32_t)sr * cr - (int32_t)si * ci); /* product narrowed to 16 bits (wraps) */
                    const int16_t ti = (int16_t)((int32_t)sr * ci + (int32_t)si * cr);
                    corr_out_iq[2 * t] = sat16_((int32_t)corr_out_iq[2 * t] + tr);
                    corr_out_iq[2 * t + 1] = sat16_((int32_t)corr_out_iq[2 * t + 1] + ti);
                }
        }
}

/* Arbiter (not a reference function): same float32 chip indices as the reference resampler, but the
 * carrier phase and the accumulation are evaluated in float64 with the exact phase
 * -(rem + n*step) of the float32 arguments. */
void go_multicorrelator_real_f64(const float* in_iq, int n, const float* code, int code_len, const float* shifts,
    int n_taps, float rem_carr_phase_rad, float phase_step_rad, float rem_code_phase_chips,
    float code_phase_step_chips, double* corr_out_iq)
{
    for (int t = 0; t < n_taps; t++) corr_out_iq[2 * t] = corr_out_iq[2 * t + 1] = 0.0;
    for (unsigned k = 0; k < (unsigned)n; k++)
        {
            double ph = -((double)rem_carr_phase_rad + (double)k * (double)phase_step_rad);
            double c = cos(ph), s = sin(ph);
            double xr = in_iq[2 * k], xi = in_iq[2 * k + 1];
            double wr = xr * c - xi * s, wi = xr * s + xi * c;
            for (int t = 0; t < n_taps; t++)
                {
                    float v = code_phase_step_chips * (float)k + shifts[t] - rem_code_phase_chips;
                    double cc = code[chip_index_(v, code_len)];
                    corr_out_iq[2 * t] += wr * cc;
                    corr_out_iq[2 * t + 1] += wi * cc;
                }
        }
}

/* ======================= acquisition helpers ======================= */

/* volk: volk_gnsssdr_s32f_sincos_32fc.h:405-416 (_generic): phase accumulated in float32 */
void go_sincos(float* out_iq, float phase_inc, float phase0, int n)
{
    float ph = phase0;
    for (int i = 0; i < n; i++)
        {
            out_iq[2 * i] = cosf(ph);
            out_iq[2 * i + 1] = sinf(ph);
            ph += phase_inc;
        }
}

/* volk: volk_gnsssdr_32f_index_max_32u.h:460-483 (_generic): strict '>', first maximum wins */
uint32_t go_index_max(const float* src, uint32_t n)
{
    uint32_t index = 0;
    if (n > 0)
        {
            float max = src[0];
            for (uint32_t i = 1; i < n; ++i)
                {
                    if (src[i] > max)
                        {
                            index = i;
                            max = src[i];
                        }
                }
        }
    return index;
}

/* volk: volk_gnsssdr_16ic_convert_32fc.h:63-72 */
void go_16ic_convert_32fc(float* out_iq, const int16_t* in_iq, uint32_t n)
{
    for (uint32_t i = 0; i < 2 * n; i++) out_iq[i] = (float)in_iq[i];
}

/* ---- un-normalised DFT of any length: X[k] = sum_n x[n] exp(sign*j*2*pi*k*n/N).
 * The reference calls FFTW3f through gr::fft::fft_complex (pcps_acquisition.cc:125-128); FFTW is an
 * un-vendored system dependency, so this is the published DFT definition evaluated by a recursive
 * decimation-in-time mixed-radix algorithm (radix = smallest prime factor, naive DFT inside each
 * butterfly), twiddles from double-precision sin/cos.  Same mathematical transform and sign/scale
 * convention as FFTW_FORWARD / FFTW_BACKWARD; results differ from FFTW by float rounding only. */
#define GO_FFT_IMPL(NAME, T, CT)                                                                          \
    static void NAME##_rec(const CT* in, CT* out, int n, int stride, const CT* tw, int tw_stride, CT* scratch) \
    {                                                                                                     \
        if (n == 1)                                                                                       \
            {                                                                                             \
                out[0] = in[0];                                                                           \
                return;                                                                                   \
            }                                                                                             \
        int p = 2;                                                                                        \
        while (n % p) p++;                                                                                \
        int m = n / p;                                                                                    \
        for (int r = 0; r < p; r++) NAME##_rec(in + (size_t)r * stride, out + (size_t)r * m, m, stride * p, tw, tw_stride * p, scratch); \
        /* out[r*m + k] = DFT_m of the r-th decimated sequence; combine */                               \
        for (int k = 0; k < m; k++)                                                                       \
            {                                                                                             \
                for (int r = 0; r < p; r++)                                                               \
                    {                                                                                     \
                        CT w = tw[(size_t)((long long)r * k % ((long long)n)) * tw_stride];               \
                        CT v = out[(size_t)r * m + k];                                                    \
                        scratch[r].re = v.re * w.re - v.im * w.im;                                        \
                        scratch[r].im = v.re * w.im + v.im * w.re;                                        \
                    }                                                                                     \
                for (int q = 0; q < p; q++)                                                               \
                    {                                                                                     \
                        T sr = scratch[0].re, si = scratch[0].im;                                         \
                        for (int r = 1; r < p; r++)                                                       \
                            {                                                                             \
                                CT w = tw[(size_t)(((long long)r * q % p) * m) * tw_stride];              \
                                sr += scratch[r].re * w.re - scratch[r].im * w.im;                        \
                                si += scratch[r].re * w.im + scratch[r].im * w.re;                        \
                            }                                                                             \
                        scratch[p + q].re = sr;                                                           \
                        scratch[p + q].im = si;                                                           \
                    }                                                                                     \
                for (int q = 0; q < p; q++) out[(size_t)q * m + k] = scratch[p + q];                      \
            }                                                                                             \
    }                                                                                                     \
    void NAME(T* data_iq, int n, int sign)                                                                \
    {                                                                                                     \
        CT* tw = (CT*)malloc(sizeof(CT) * (size_t)n);                                                     \
        CT* tmp = (CT*)malloc(sizeof(CT) * (size_t)n);                                                    \
        int maxp = 2;                                                                                     \
        for (int r = n, p = 2; r > 1;)                                                                    \
            {                                                                                             \
                if (r % p == 0) { if (p > maxp) maxp = p; r /= p; }                                       \
                else p++;                                                                                 \
            }                                                                                             \
        CT* scratch = (CT*)malloc(sizeof(CT) * 2 * (size_t)maxp);                                         \
        for (int i = 0; i < n; i++)                                                                       \
            {                                                                                             \
                double a = (double)sign * 6.283185307179586476925286766559 * (double)i / (double)n;       \
                tw[i].re = (T)cos(a);                                                                     \
                tw[i].im = (T)sin(a);                                                                     \
            }                                                                                             \
        memcpy(tmp, data_iq, sizeof(CT) * (size_t)n);                                                     \
        NAME##_rec(tmp, (CT*)data_iq, n, 1, tw, 1, scratch);                                              \
        free(scratch);                                                                                    \
        free(tmp);                                                                                        \
        free(tw);                                                                                         \
    }

GO_FFT_IMPL(go_fft_f32, float, cf)
GO_FFT_IMPL(go_fft_f64, double, cd)

/* ======================= PRN codes ======================= */

/* src/algorithms/libs/gps_sdr_signal_processing.cc:37-117: G1/G2 LFSRs of IS-GPS-200 with the
 * reference's G2 delay table (PRN 1-32, then SBAS 120-138 at table index 32..50). */
void go_gps_l1_ca_code_int(int32_t* dest, int32_t prn, uint32_t chip_shift)
{
    enum { L = 1023 };
    static const int32_t delays[51] = {5, 6, 7, 8, 17, 18, 139, 140, 141, 251, 252, 254, 255, 256, 257, 258, 469, 470, 471, 472,
        473, 474, 509, 512, 513, 514, 515, 516, 859, 860, 861, 862,
        145, 175, 52, 21, 237, 235, 886, 657, 634, 762, 355, 1012, 176, 603, 130, 359, 595, 68, 386};
    unsigned char g1[L], g2[L], r1[10], r2[10];
    int32_t prn_idx = (120 <= prn && prn <= 138) ? prn - 88 : prn - 1;
    if (prn_idx < 0 || prn_idx > 50) return;
    for (int i = 0; i < 10; i++) r1[i] = r2[i] = 1;
    for (int i = 0; i < L; i++)
        {
            g1[i] = r1[0];
            g2[i] = r2[0];
            unsigned char f1 = r1[7] ^ r1[0];
            unsigned char f2 = (unsigned char)((r2[8] + r2[7] + r2[4] + r2[2] + r2[1] + r2[0]) & 1);
            for (int j = 0; j < 9; j++)
                {
                    r1[j] = r1[j + 1];
                    r2[j] = r2[j + 1];
                }
            r1[9] = f1;
            r2[9] = f2;
        }
    uint32_t delay = (uint32_t)(L - delays[prn_idx]);
    delay += chip_shift;
    delay %= L;
    for (uint32_t i = 0; i < L; i++)
        {
            dest[i] = (g1[(i + chip_shift) % L] ^ g2[delay]) ? 1 : -1;
            delay = (delay + 1) % L;
        }
}

void go_gps_l1_ca_code_float(float* dest, int32_t prn, uint32_t chip_shift)
{
    int32_t c[1023];
    memset(c, 0, sizeof(c));
    go_gps_l1_ca_code_int(c, prn, chip_shift);
    for (int i = 0; i < 1023; i++) dest[i] = (float)c[i];
}

/* auxCeil lambda, gps_sdr_signal_processing.cc:35 */
static inline int32_t aux_ceil_(float x) { return (int32_t)(int64_t)(x + 1); }

/* gps_sdr_signal_processing.cc:151-196: float32 index maths, last sample forced to the last chip */
void go_gps_l1_ca_code_sampled(float* dest_iq, uint32_t prn, int32_t fs, uint32_t chip_shift)
{
    float code[1023];
    const int32_t code_freq_basis = 1023000, code_length = 1023;
    int32_t samples_per_code = (int32_t)((double)fs / (double)(code_freq_basis / code_length));
    float ts = (float)(1.0 / (double)(float)fs);
    float tc = (float)(1.0 / (double)(float)code_freq_basis);
    go_gps_l1_ca_code_float(code, (int32_t)prn, chip_shift);
    for (int32_t i = 0; i < samples_per_code; i++)
        {
            float aux = (ts * (float)(i + 1)) / tc;
            int32_t idx = aux_ceil_(aux) - 1;
            dest_iq[2 * i] = (i == samples_per_code - 1) ? code[code_length - 1] : code[idx];
            dest_iq[2 * i + 1] = 0.0f;
        }
}

/* src/algorithms/libs/gnss_signal_processing.cc:161-183 */
void go_code_resampler_f32(const float* from, float* dest, float fs_in, float fs_out, uint32_t len_in, uint32_t len_out)
{
    const float t_in = 1 / fs_in;
    const float t_out = 1 / fs_out;
    for (uint32_t i = 0; i < len_out - 1; i++)
        {
            float aux = (t_out * (float)(i + 1)) / t_in;
            uint32_t idx = (uint32_t)(aux_ceil_(aux) - 1);
            dest[i] = from[idx];
        }
    dest[len_out - 1] = from[len_in - 1];
}

/* ======================= pcps_acquisition block ======================= */
/* State and arithmetic of src/algorithms/acquisition/gnuradio_blocks/pcps_acquisition.cc.
 * The GNU Radio plumbing (general_work states 0/1, :976-1034) is replaced by the caller handing
 * d_consumed_samples items; everything from acquisition_core (:668) on is restated line by line,
 * including the quirks listed in SURVEY.md 7.3 (stale d_tmp_buffer, one-sided exclusion wrap,
 * N^4 normalisation, input power over the zero padding, step-two Doppler truncation). */
struct go_acq
{
    go_acq_conf c;
    uint32_t consumed, fft_size, num_bins;
    int use_cfar;
    float* tmp_buffer;   /* d_tmp_buffer, persists between calls (quirk 1) */
    float* fft_codes;    /* d_fft_codes, IQ */
    float** wipeoffs;    /* d_grid_doppler_wipeoffs[bin], IQ */
    float** wipeoffs2;   /* step two */
    float** grid;        /* d_magnitude_grid[bin][fft_size] */
    float** grid_snap;   /* tests only */
    float* work;         /* fft in/out, IQ */
    double* work64;
    float* input_signal; /* d_input_signal IQ */
    uint32_t dwell_counter; /* d_num_noncoherent_integrations_counter */
    int step_two, positive_acq;
    float doppler_center_step_two;
};

static const double GO_GPS_TWO_PI = 6.283185307179586; /* GPS_L1_CA.h:46 */

/* pcps_acquisition.cc:296-310 (+ analytic arbiter) */
static void go_acq_local_carrier_(const go_acq* a, float* carrier_iq, uint32_t n, float freq)
{
    float phase_step_rad = (float)(GO_GPS_TWO_PI * freq / (float)a->c.fs_in);
    if (a->c.analytic_wipeoff)
        {
            for (uint32_t i = 0; i < n; i++)
                {
                    double ph = -(double)phase_step_rad * (double)i;
                    carrier_iq[2 * i] = (float)cos(ph);
                    carrier_iq[2 * i + 1] = (float)sin(ph);
                }
        }
    else
        go_sincos(carrier_iq, -phase_step_rad, 0.0f, (int)n);
}

static void go_acq_update_grid_(go_acq* a)
{
    for (uint32_t i = 0; i < a->num_bins; i++)
        {
            int32_t doppler = -(int32_t)a->c.doppler_max + (int32_t)(a->c.doppler_step * i);
            go_acq_local_carrier_(a, a->wipeoffs[i], a->fft_size, (float)(a->c.freq_offset_hz + doppler));
        }
}

go_acq* go_acq_create(const go_acq_conf* conf)
{
    go_acq* a = (go_acq*)calloc(1, sizeof(go_acq));
    a->c = *conf;
    /* ctor :75-117 */
    a->consumed = (uint32_t)((float)conf->sampled_ms * conf->samples_per_ms * (float)(conf->bit_transition_flag ? 2 : 1));
    a->fft_size = (conf->sampled_ms == conf->ms_per_code) ? a->consumed : a->consumed * 2;
    if (conf->bit_transition_flag)
        {
            a->fft_size = a->consumed * 2;
            a->c.max_dwells = 1;
        }
    a->use_cfar = (a->c.max_dwells == 1) ? conf->use_CFAR_algorithm_flag : 0; /* :152-159 */
    /* init() :326 */
    a->num_bins = (uint32_t)ceil((double)((int32_t)conf->doppler_max - (int32_t)(-(int32_t)conf->doppler_max)) / (double)conf->doppler_step);
    size_t n = a->fft_size;
    a->tmp_buffer = (float*)calloc(n, sizeof(float));
    a->fft_codes = (float*)calloc(2 * n, sizeof(float));
    a->work = (float*)calloc(2 * n, sizeof(float));
    a->work64 = (double*)calloc(2 * n, sizeof(double));
    a->input_signal = (float*)calloc(2 * n, sizeof(float));
    a->wipeoffs = (float**)calloc(a->num_bins, sizeof(float*));
    a->grid = (float**)calloc(a->num_bins, sizeof(float*));
    for (uint32_t i = 0; i < a->num_bins; i++)
        {
            a->wipeoffs[i] = (float*)calloc(2 * n, sizeof(float));
            a->grid[i] = (float*)calloc(n, sizeof(float));
        }
    if (conf->keep_grid)
        {
            a->grid_snap = (float**)calloc(a->num_bins, sizeof(float*));
            for (uint32_t i = 0; i < a->num_bins; i++) a->grid_snap[i] = (float*)calloc(n, sizeof(float));
        }
    if (conf->make_2_steps)
        {
            a->wipeoffs2 = (float**)calloc(conf->num_doppler_bins_step2, sizeof(float*));
            for (uint32_t i = 0; i < conf->num_doppler_bins_step2; i++) a->wipeoffs2[i] = (float*)calloc(2 * n, sizeof(float));
        }
    go_acq_update_grid_(a);
    return a;
}

void go_acq_destroy(go_acq* a)
{
    if (!a) return;
    for (uint32_t i = 0; i < a->num_bins; i++)
        {
            free(a->wipeoffs[i]);
            free(a->grid[i]);
            if (a->grid_snap) free(a->grid_snap[i]);
        }
    free(a->grid_snap);
    if (a->wipeoffs2)
        {
            for (uint32_t i = 0; i < a->c.num_doppler_bins_step2; i++) free(a->wipeoffs2[i]);
            free(a->wipeoffs2);
        }
    free(a->wipeoffs);
    free(a->grid);
    free(a->tmp_buffer);
    free(a->fft_codes);
    free(a->work);
    free(a->work64);
    free(a->input_signal);
    free(a);
}

uint32_t go_acq_fft_size(const go_acq* a) { return a->fft_size; }
uint32_t go_acq_consumed_samples(const go_acq* a) { return a->consumed; }
uint32_t go_acq_num_doppler_bins(const go_acq* a) { return a->num_bins; }
const float* go_acq_grid(const go_acq* a, int bin) { return a->grid_snap ? a->grid_snap[bin] : a->grid[bin]; }

static void go_acq_fft_(go_acq* a, float* buf_iq, int sign)
{
    uint32_t n = a->fft_size;
    if (a->c.fft_f64)
        {
            for (uint32_t i = 0; i < 2 * n; i++) a->work64[i] = buf_iq[i];
            go_fft_f64(a->work64, (int)n, sign);
            for (uint32_t i = 0; i < 2 * n; i++) buf_iq[i] = (float)a->work64[i];
        }
    else
        go_fft_f32(buf_iq, (int)n, sign);
}

/* pcps_acquisition.cc:239-274.  `code_iq` holds what the adapter passes: d_consumed_samples items
 * (fft_size/2 with bit_transition_flag, :253-258). */
void go_acq_set_local_code(go_acq* a, const float* code_iq)
{
    uint32_t n = a->fft_size;
    memset(a->work, 0, sizeof(float) * 2 * n);
    if (a->c.bit_transition_flag)
        {
            uint32_t offset = n / 2;
            memcpy(a->work + 2 * offset, code_iq, sizeof(float) * 2 * offset);
        }
    else if (a->c.sampled_ms == a->c.ms_per_code)
        memcpy(a->work, code_iq, sizeof(float) * 2 * a->consumed);
    else
        memcpy(a->work + 2 * (size_t)a->consumed, code_iq, sizeof(float) * 2 * a->consumed);
    go_acq_fft_(a, a->work, -1);
    for (uint32_t i = 0; i < n; i++) /* volk_32fc_conjugate_32fc */
        {
            a->fft_codes[2 * i] = a->work[2 * i];
            a->fft_codes[2 * i + 1] = -a->work[2 * i + 1];
        }
}

/* pcps_acquisition.cc:565-596 */
static float go_acq_stat_cfar_(go_acq* a, uint32_t* indext, int32_t* doppler, uint32_t* bin_out, float* peak, float input_power,
    uint32_t num_bins, int32_t doppler_max, int32_t doppler_step)
{
    float grid_maximum = 0.0f;
    uint32_t index_doppler = 0, index_time = 0;
    float nf = (float)a->fft_size * (float)a->fft_size;
    for (uint32_t i = 0; i < num_bins; i++)
        {
            uint32_t t = go_index_max(a->grid[i], a->fft_size);
            if (a->grid[i][t] > grid_maximum)
                {
                    grid_maximum = a->grid[i][t];
                    index_doppler = i;
                    index_time = t;
                }
        }
    *indext = index_time;
    *bin_out = index_doppler;
    *peak = grid_maximum;
    if (!a->step_two)
        *doppler = -doppler_max + doppler_step * (int32_t)index_doppler;
    else
        *doppler = (int32_t)(a->doppler_center_step_two + ((float)index_doppler - (float)floor(a->c.num_doppler_bins_step2 / 2.0)) * a->c.doppler_step2);
    float magt = grid_maximum / (nf * nf);
    return magt / input_power;
}

/* pcps_acquisition.cc:599-665 */
static float go_acq_stat_peaks_(go_acq* a, uint32_t* indext, int32_t* doppler, uint32_t* bin_out, float* peak,
    uint32_t num_bins, int32_t doppler_max, int32_t doppler_step)
{
    float first_peak = 0.0f;
    uint32_t index_doppler = 0, index_time = 0;
    for (uint32_t i = 0; i < num_bins; i++)
        {
            uint32_t t = go_index_max(a->grid[i], a->fft_size);
            if (a->grid[i][t] > first_peak)
                {
                    first_peak = a->grid[i][t];
                    index_doppler = i;
                    index_time = t;
                }
        }
    *indext = index_time;
    *bin_out = index_doppler;
    *peak = first_peak;
    if (!a->step_two)
        *doppler = -doppler_max + doppler_step * (int32_t)index_doppler;
    else
        *doppler = (int32_t)(a->doppler_center_step_two + ((float)index_doppler - (float)floor(a->c.num_doppler_bins_step2 / 2.0)) * a->c.doppler_step2);

    int32_t ex1 = (int32_t)index_time - (int32_t)a->c.samples_per_chip;
    int32_t ex2 = (int32_t)index_time + (int32_t)a->c.samples_per_chip;
    if (ex1 < 0)
        ex1 = (int32_t)a->fft_size + ex1;
    else if (ex2 >= (int32_t)a->fft_size)
        ex2 = ex2 - (int32_t)a->fft_size;
    int32_t idx = ex1;
    memcpy(a->tmp_buffer, a->grid[index_doppler], a->fft_size); /* d_fft_size BYTES, :647 (quirk 1) */
    do
        {
            a->tmp_buffer[idx] = 0.0f;
            idx++;
            if (idx == (int32_t)a->fft_size) idx = 0;
        }
    while (idx != ex2);
    uint32_t t2 = go_index_max(a->tmp_buffer, a->fft_size);
    float second_peak = a->tmp_buffer[t2];
    return first_peak / second_peak;
}

/* pcps_acquisition.cc:668-930 */
int go_acq_run(go_acq* a, const void* samples, uint64_t samp_count, go_acq_result* out)
{
    const uint32_t n = a->fft_size;
    const uint32_t eff = a->c.bit_transition_flag ? n / 2 : n;
    int32_t doppler = 0;
    uint32_t indext = 0, bin = 0;
    float peak = 0.0f;
    memset(out, 0, sizeof(*out));
    if (a->c.cshort)
        go_16ic_convert_32fc(a->input_signal, (const int16_t*)samples, a->consumed);
    else
        memcpy(a->input_signal, samples, sizeof(float) * 2 * a->consumed);
    for (uint32_t i = 2 * a->consumed; i < 2 * n; i++) a->input_signal[i] = 0.0f;
    const float* in = a->input_signal;

    float input_power = 0.0f;
    a->dwell_counter++;
    if (a->use_cfar || a->c.bit_transition_flag) /* :703-709 */
        {
            for (uint32_t i = 0; i < n; i++) a->tmp_buffer[i] = in[2 * i] * in[2 * i] + in[2 * i + 1] * in[2 * i + 1];
            if (a->c.accumulator_lanes > 1)
                {
                    /* VOLK volk_32f_accumulator_s32f a_avx / u_avx: one vector accumulator of `lanes` floats over the whole
                     * multiples of `lanes`, lanes added in order, then the remaining points one by one */
                    const uint32_t lanes = (uint32_t)a->c.accumulator_lanes, whole = n / lanes;
                    float acc[16] = {0};
                    for (uint32_t k = 0; k < whole; k++)
                        for (uint32_t l = 0; l < lanes && l < 16; l++) acc[l] += a->tmp_buffer[k * lanes + l];
                    input_power = acc[0];
                    for (uint32_t l = 1; l < lanes && l < 16; l++) input_power += acc[l];
                    for (uint32_t i = whole * lanes; i < n; i++) input_power += a->tmp_buffer[i];
                }
            else
                for (uint32_t i = 0; i < n; i++) input_power += a->tmp_buffer[i];
            input_power /= (float)n;
        }

    uint32_t nb = a->step_two ? a->c.num_doppler_bins_step2 : a->num_bins;
    float** wip = a->step_two ? a->wipeoffs2 : a->wipeoffs;
    for (uint32_t d = 0; d < nb; d++) /* :714-745 / :773-803 */
        {
            const float* w = wip[d];
            for (uint32_t i = 0; i < n; i++)
                {
                    a->work[2 * i] = in[2 * i] * w[2 * i] - in[2 * i + 1] * w[2 * i + 1];
                    a->work[2 * i + 1] = in[2 * i] * w[2 * i + 1] + in[2 * i + 1] * w[2 * i];
                }
            go_acq_fft_(a, a->work, -1);
            for (uint32_t i = 0; i < n; i++)
                {
                    float xr = a->work[2 * i], xi = a->work[2 * i + 1];
                    float cr = a->fft_codes[2 * i], ci = a->fft_codes[2 * i + 1];
                    a->work[2 * i] = xr * cr - xi * ci;
                    a->work[2 * i + 1] = xr * ci + xi * cr;
                }
            go_acq_fft_(a, a->work, +1);
            const float* y = a->work + 2 * (size_t)(a->c.bit_transition_flag ? eff : 0);
            if (a->dwell_counter == 1)
                for (uint32_t i = 0; i < eff; i++) a->grid[d][i] = y[2 * i] * y[2 * i] + y[2 * i + 1] * y[2 * i + 1];
            else
                {
                    for (uint32_t i = 0; i < eff; i++) a->tmp_buffer[i] = y[2 * i] * y[2 * i] + y[2 * i + 1] * y[2 * i + 1];
                    for (uint32_t i = 0; i < eff; i++) a->grid[d][i] = a->grid[d][i] + a->tmp_buffer[i];
                }
        }

    float stat;
    if (!a->step_two)
        {
            if (a->use_cfar)
                stat = go_acq_stat_cfar_(a, &indext, &doppler, &bin, &peak, input_power, a->num_bins, (int32_t)a->c.doppler_max, (int32_t)a->c.doppler_step);
            else
                stat = go_acq_stat_peaks_(a, &indext, &doppler, &bin, &peak, a->num_bins, (int32_t)a->c.doppler_max, (int32_t)a->c.doppler_step);
        }
    else
        {
            int32_t dmax2 = (int32_t)(a->doppler_center_step_two - ((float)a->c.num_doppler_bins_step2 / 2.0) * a->c.doppler_step2);
            if (a->use_cfar)
                stat = go_acq_stat_cfar_(a, &indext, &doppler, &bin, &peak, input_power, a->c.num_doppler_bins_step2, dmax2, (int32_t)a->c.doppler_step2);
            else
                stat = go_acq_stat_peaks_(a, &indext, &doppler, &bin, &peak, a->c.num_doppler_bins_step2, dmax2, (int32_t)a->c.doppler_step2);
        }
    out->code_phase_index = indext;
    out->doppler_hz = doppler;
    out->doppler_bin = bin;
    out->peak_mag = peak;
    out->input_power = input_power;
    out->test_statistic = stat;
    out->acq_delay_samples = (double)fmodf((float)indext, a->c.samples_per_code); /* :766 */
    out->acq_doppler_hz = (double)doppler;
    out->acq_samplestamp = samp_count;

    /* decision logic :832-909 */
    int state = 1; /* d_state while in acquisition_core is 2; we track only what it becomes */
    int decision;
    int positive_sent = 0, negative_sent = 0;
    if (!a->c.bit_transition_flag)
        {
            if (stat > a->c.threshold)
                {
                    if (a->c.make_2_steps)
                        {
                            if (a->step_two)
                                {
                                    positive_sent = 1;
                                    a->step_two = 0;
                                    state = 0;
                                }
                            else
                                {
                                    a->step_two = 1;
                                    a->dwell_counter = 0;
                                    a->positive_acq = 0;
                                    state = 0;
                                }
                        }
                    else
                        {
                            positive_sent = 1;
                            state = 0;
                        }
                }
            else
                state = 1;
            if (a->dwell_counter == a->c.max_dwells)
                {
                    if (state != 0) negative_sent = 1;
                    state = 0;
                    a->step_two = 0;
                }
        }
    else
        {
            if (stat > a->c.threshold)
                {
                    if (a->c.make_2_steps)
                        {
                            if (a->step_two)
                                {
                                    positive_sent = 1;
                                    a->step_two = 0;
                                }
                            else
                                {
                                    a->step_two = 1;
                                    a->dwell_counter = 0;
                                }
                        }
                    else
                        positive_sent = 1;
                }
            else
                {
                    a->step_two = 0;
                    negative_sent = 1;
                }
            state = 0;
        }
    if (positive_sent) a->positive_acq = 1;
    if (negative_sent) a->positive_acq = 0;
    if (positive_sent)
        decision = 1;
    else if (negative_sent)
        decision = 0;
    else if (a->step_two)
        decision = 3;
    else
        decision = 2;
    out->decision = decision;
    if (a->step_two) /* general_work :964-970 re-centres the fine grid before the next dwell */
        {
            a->doppler_center_step_two = (float)out->acq_doppler_hz;
            for (uint32_t i = 0; i < a->c.num_doppler_bins_step2; i++)
                {
                    float dd = ((float)i - (float)floor(a->c.num_doppler_bins_step2 / 2.0)) * a->c.doppler_step2;
                    go_acq_local_carrier_(a, a->wipeoffs2[i], n, a->doppler_center_step_two + dd);
                }
        }
    if (a->grid_snap)
        for (uint32_t i = 0; i < a->num_bins; i++) memcpy(a->grid_snap[i], a->grid[i], sizeof(float) * n);
    /* :912-929 */
    if (a->dwell_counter == a->c.max_dwells || a->positive_acq == 1)
        {
            a->dwell_counter = 0;
            a->positive_acq = 0;
            for (uint32_t i = 0; i < a->num_bins; i++) memset(a->grid[i], 0, sizeof(float) * n);
        }
    (void)state;
    return 0;
}

/* ======================= sibling PCPS blocks that share the Doppler-loop arithmetic ======================= */
/* Both follow the older block layout: fft_size = sampled_ms * samples_per_ms, inclusive Doppler-bin count
 * ("doppler <= doppler_max"), wipe-off tables from volk_gnsssdr_s32f_sincos_32fc with the phase starting at 0. */
struct go_acqv
{
    go_acqv_conf c;
    uint32_t num_bins;
    float** wipeoffs;   /* [bin] IQ */
    float** grid;       /* Tong: d_grid_data[bin][fft_size] */
    float* fft_code_a;  /* conj(FFT(code)), IQ */
    float* fft_code_b;  /* 8 ms: second replica inverted */
    float* work;
    float* mag;
    uint32_t dwell_count;
    int32_t tong_count;
};

go_acqv* go_acqv_create(const go_acqv_conf* conf)
{
    go_acqv* a = (go_acqv*)calloc(1, sizeof(go_acqv));
    a->c = *conf;
    const uint32_t n = conf->fft_size;
    /* pcps_tong_acquisition_cc.cc:183-189, galileo_pcps_8ms_acquisition_cc.cc:150-158: inclusive count */
    a->num_bins = 0;
    for (int32_t d = -(int32_t)conf->doppler_max; d <= (int32_t)conf->doppler_max; d += (int32_t)conf->doppler_step) a->num_bins++;
    a->wipeoffs = (float**)calloc(a->num_bins, sizeof(float*));
    a->grid = (float**)calloc(a->num_bins, sizeof(float*));
    for (uint32_t i = 0; i < a->num_bins; i++)
        {
            a->wipeoffs[i] = (float*)calloc(2 * (size_t)n, sizeof(float));
            a->grid[i] = (float*)calloc(n, sizeof(float));
            const int32_t doppler = -(int32_t)conf->doppler_max + (int32_t)(conf->doppler_step * i);
            float phase_step_rad;
            if (conf->variant == GO_ACQV_8MS)
                {
                    /* galileo_pcps_8ms_acquisition_cc.cc:148,166: float(GALILEO_TWO_PI) * doppler / float(fs), all float32 */
                    const double GALILEO_TWO_PI = 6.283185307179600;
                    phase_step_rad = (float)GALILEO_TWO_PI * (float)doppler / (float)conf->fs_in;
                }
            else
                /* pcps_tong_acquisition_cc.cc:200: GPS_TWO_PI (double) * doppler / float(fs), rounded once to float32 */
                phase_step_rad = (float)(GO_GPS_TWO_PI * doppler / (float)conf->fs_in);
            go_sincos(a->wipeoffs[i], -phase_step_rad, 0.0f, (int)n);
        }
    a->fft_code_a = (float*)calloc(2 * (size_t)n, sizeof(float));
    a->fft_code_b = (float*)calloc(2 * (size_t)n, sizeof(float));
    a->work = (float*)calloc(2 * (size_t)n, sizeof(float));
    a->mag = (float*)calloc(n, sizeof(float));
    a->tong_count = (int32_t)conf->tong_init_val;
    return a;
}

void go_acqv_destroy(go_acqv* a)
{
    if (!a) return;
    for (uint32_t i = 0; i < a->num_bins; i++)
        {
            free(a->wipeoffs[i]);
            free(a->grid[i]);
        }
    free(a->wipeoffs);
    free(a->grid);
    free(a->fft_code_a);
    free(a->fft_code_b);
    free(a->work);
    free(a->mag);
    free(a);
}

uint32_t go_acqv_num_doppler_bins(const go_acqv* a) { return a->num_bins; }
float go_acqv_phase_step(const go_acqv* a, uint32_t bin)
{
    /* the table's increment is -phase_step_rad: entry 1 is (cos(-s), sin(-s)); tests read the step back through this */
    const int32_t doppler = -(int32_t)a->c.doppler_max + (int32_t)(a->c.doppler_step * bin);
    if (a->c.variant == GO_ACQV_8MS) return (float)6.283185307179600 * (float)doppler / (float)a->c.fs_in;
    return (float)(GO_GPS_TWO_PI * doppler / (float)a->c.fs_in);
}

static void go_acqv_code_fft_(go_acqv* a, float* dst)
{
    const uint32_t n = a->c.fft_size;
    go_fft_f32(a->work, (int)n, -1);
    for (uint32_t i = 0; i < n; i++)
        {
            dst[2 * i] = a->work[2 * i];
            dst[2 * i + 1] = -a->work[2 * i + 1];
        }
}

/* pcps_tong_acquisition_cc.cc:170-178; galileo_pcps_8ms_acquisition_cc.cc:117-138 (code B: second replica times -1) */
void go_acqv_set_local_code(go_acqv* a, const float* code_iq)
{
    const uint32_t n = a->c.fft_size;
    memcpy(a->work, code_iq, sizeof(float) * 2 * n);
    go_acqv_code_fft_(a, a->fft_code_a);
    if (a->c.variant == GO_ACQV_8MS)
        {
            memcpy(a->work, code_iq, sizeof(float) * 2 * n);
            for (uint32_t i = a->c.samples_per_code; i < 2 * a->c.samples_per_code && i < n; i++)
                {
                    /* volk_32fc_s32fc_multiply_32fc by (-1, 0): (x + jy)(-1 + j0) */
                    const float xr = code_iq[2 * i], xi = code_iq[2 * i + 1];
                    a->work[2 * i] = xr * -1.0f - xi * 0.0f;
                    a->work[2 * i + 1] = xr * 0.0f + xi * -1.0f;
                }
            go_acqv_code_fft_(a, a->fft_code_b);
        }
}

/* set_state(1) of the Tong block (:226-252): new attempt */
void go_acqv_restart(go_acqv* a)
{
    a->dwell_count = 0;
    a->tong_count = (int32_t)a->c.tong_init_val;
    for (uint32_t i = 0; i < a->num_bins; i++) memset(a->grid[i], 0, sizeof(float) * a->c.fft_size);
}

static void go_acqv_correlate_(go_acqv* a, const float* in, uint32_t bin, const float* fft_code)
{
    const uint32_t n = a->c.fft_size;
    const float* w = a->wipeoffs[bin];
    for (uint32_t i = 0; i < n; i++)
        {
            a->work[2 * i] = in[2 * i] * w[2 * i] - in[2 * i + 1] * w[2 * i + 1];
            a->work[2 * i + 1] = in[2 * i] * w[2 * i + 1] + in[2 * i + 1] * w[2 * i];
        }
    go_fft_f32(a->work, (int)n, -1);
    for (uint32_t i = 0; i < n; i++)
        {
            const float xr = a->work[2 * i], xi = a->work[2 * i + 1];
            const float cr = fft_code[2 * i], ci = fft_code[2 * i + 1];
            a->work[2 * i] = xr * cr - xi * ci;
            a->work[2 * i + 1] = xr * ci + xi * cr;
        }
    go_fft_f32(a->work, (int)n, +1);
    for (uint32_t i = 0; i < n; i++) a->mag[i] = a->work[2 * i] * a->work[2 * i] + a->work[2 * i + 1] * a->work[2 * i + 1];
}

/* one call = one pass through general_work state 1 (one dwell of fft_size gr_complex items).
 * Tong: pcps_tong_acquisition_cc.cc:290-395.  8 ms: galileo_pcps_8ms_acquisition_cc.cc:225-375. */
int go_acqv_run(go_acqv* a, const float* in_iq, go_acqv_result* out)
{
    const uint32_t n = a->c.fft_size;
    const float fnf = (float)n * (float)n;
    float input_power = 0.0f, d_mag = 0.0f;
    memset(out, 0, sizeof(*out));
    a->dwell_count++;
    for (uint32_t i = 0; i < n; i++) a->mag[i] = in_iq[2 * i] * in_iq[2 * i] + in_iq[2 * i + 1] * in_iq[2 * i + 1];
    for (uint32_t i = 0; i < n; i++) input_power += a->mag[i];
    input_power /= (float)n;
    for (uint32_t bin = 0; bin < a->num_bins; bin++)
        {
            const int32_t doppler = -(int32_t)a->c.doppler_max + (int32_t)(a->c.doppler_step * bin);
            uint32_t indext;
            float magt;
            if (a->c.variant == GO_ACQV_8MS)
                {
                    go_acqv_correlate_(a, in_iq, bin, a->fft_code_a);
                    const uint32_t ia = go_index_max(a->mag, n);
                    const float ma = a->mag[ia] / (fnf * fnf);
                    go_acqv_correlate_(a, in_iq, bin, a->fft_code_b);
                    const uint32_t ib = go_index_max(a->mag, n);
                    const float mb = a->mag[ib] / (fnf * fnf);
                    if (ma >= mb)
                        {
                            magt = ma;
                            indext = ia;
                        }
                    else
                        {
                            magt = mb;
                            indext = ib;
                        }
                }
            else
                {
                    go_acqv_correlate_(a, in_iq, bin, a->fft_code_a);
                    const float scale = 1 / (fnf * fnf * input_power);
                    for (uint32_t i = 0; i < n; i++) a->mag[i] = a->mag[i] * scale;
                    for (uint32_t i = 0; i < n; i++) a->grid[bin][i] = a->mag[i] + a->grid[bin][i];
                    indext = go_index_max(a->grid[bin], n);
                    magt = a->grid[bin][indext];
                }
            if (d_mag < magt)
                {
                    d_mag = magt;
                    out->code_phase_index = indext;
                    out->acq_delay_samples = (double)(indext % a->c.samples_per_code);
                    out->doppler_hz = doppler;
                    out->doppler_bin = bin;
                }
        }
    out->mag = d_mag;
    out->input_power = input_power;
    if (a->c.variant == GO_ACQV_8MS)
        {
            out->test_statistic = d_mag / input_power;
            if (out->test_statistic > a->c.threshold)
                out->decision = 1;
            else if (a->dwell_count == a->c.max_dwells)
                out->decision = 0;
            else
                out->decision = 2;
        }
    else
        {
            out->test_statistic = d_mag;
            int state = 1;
            if (out->test_statistic > a->c.threshold * a->dwell_count)
                {
                    a->tong_count++;
                    if (a->tong_count == (int32_t)a->c.tong_max_val) state = 2;
                }
            else
                {
                    a->tong_count--;
                    if (a->tong_count == 0) state = 3;
                }
            if (a->dwell_count >= a->c.tong_max_dwells) state = 3;
            out->decision = state == 2 ? 1 : (state == 3 ? 0 : 2);
            out->tong_count = a->tong_count;
        }
    out->dwell_count = a->dwell_count;
    return 0;
}

/* ======================= galileo_e5a_noncoherent_iq_acquisition_caf_cc =======================
 * TEST INFRASTRUCTURE.  Block-level parity unpinned (no reference fixture, block needs GNU Radio); see gnss_oracle.h. */
struct go_e5a
{
    go_e5a_conf c;
    uint32_t fft_size, num_bins, eff_sampled_ms;
    float** wipeoffs;
    float* fft_code[4]; /* I_A, Q_A, I_B, Q_B: conj(FFT(.)) */
    float* mag[4];      /* d_magnitudeIA, QA, IB, QB */
    float* work;
    float *caf, *caf_i, *caf_q;
    uint32_t well_count;
    float test_statistics; /* survives from dwell to dwell (:389-403 resets it only in state 0) */
    go_e5a_result last;    /* Gnss_Synchro image */
};

go_e5a* go_e5a_create(const go_e5a_conf* conf)
{
    go_e5a* a = (go_e5a*)calloc(1, sizeof(go_e5a));
    a->c = *conf;
    /* :95-104: Zero_padding forces d_sampled_ms to 1 while d_fft_size keeps the constructor's sampled_ms */
    a->eff_sampled_ms = conf->zero_padding > 0 ? 1u : conf->sampled_ms;
    a->fft_size = conf->sampled_ms * conf->samples_per_ms;
    const uint32_t n = a->fft_size;
    for (int32_t d = -(int32_t)conf->doppler_max; d <= (int32_t)conf->doppler_max; d += (int32_t)conf->doppler_step) a->num_bins++;
    a->wipeoffs = (float**)calloc(a->num_bins, sizeof(float*));
    for (uint32_t i = 0; i < a->num_bins; i++)
        {
            a->wipeoffs[i] = (float*)calloc(2 * (size_t)n, sizeof(float));
            go_sincos(a->wipeoffs[i], -go_e5a_phase_step(a, i), 0.0f, (int)n);
        }
    for (int k = 0; k < 4; k++)
        {
            a->fft_code[k] = (float*)calloc(2 * (size_t)n, sizeof(float));
            a->mag[k] = (float*)calloc(n, sizeof(float));
        }
    a->work = (float*)calloc(2 * (size_t)n, sizeof(float));
    a->caf = (float*)calloc(a->num_bins, sizeof(float));
    a->caf_i = (float*)calloc(a->num_bins, sizeof(float));
    a->caf_q = (float*)calloc(a->num_bins, sizeof(float));
    return a;
}

void go_e5a_destroy(go_e5a* a)
{
    if (!a) return;
    for (uint32_t i = 0; i < a->num_bins; i++) free(a->wipeoffs[i]);
    free(a->wipeoffs);
    for (int k = 0; k < 4; k++)
        {
            free(a->fft_code[k]);
            free(a->mag[k]);
        }
    free(a->work);
    free(a->caf);
    free(a->caf_i);
    free(a->caf_q);
    free(a);
}

uint32_t go_e5a_num_doppler_bins(const go_e5a* a) { return a->num_bins; }
const float* go_e5a_caf_vector(const go_e5a* a) { return a->caf; }

/* :314-316: double GALILEO_TWO_PI * int doppler / float(fs) evaluated in double, rounded once to float32 */
float go_e5a_phase_step(const go_e5a* a, uint32_t bin)
{
    const double GALILEO_TWO_PI = 6.283185307179600;
    const int doppler = -(int)a->c.doppler_max + (int)(a->c.doppler_step * bin);
    return (float)(GALILEO_TWO_PI * doppler / (float)a->c.fs_in);
}

static void go_e5a_code_fft_(go_e5a* a, float* dst)
{
    const uint32_t n = a->fft_size;
    go_fft_f32(a->work, (int)n, -1);
    for (uint32_t i = 0; i < n; i++)
        {
            dst[2 * i] = a->work[2 * i];
            dst[2 * i + 1] = -a->work[2 * i + 1];
        }
}

/* :238-272.  The B replicas overwrite only the first samples_per_code items of the FFT input, the rest still holds the
 * previous transform's INPUT in FFTW (in-buffer untouched by execute): here the previously copied code. */
void go_e5a_set_local_code(go_e5a* a, const float* code_i, const float* code_q)
{
    const uint32_t n = a->fft_size;
    const int both = a->c.both_signal_components != 0;
    float* inbuf = (float*)calloc(2 * (size_t)n, sizeof(float)); /* d_fft_if->get_inbuf() */
    memcpy(inbuf, code_i, sizeof(float) * 2 * n);
    memcpy(a->work, inbuf, sizeof(float) * 2 * n);
    go_e5a_code_fft_(a, a->fft_code[0]);
    if (both)
        {
            memcpy(inbuf, code_q, sizeof(float) * 2 * n);
            memcpy(a->work, inbuf, sizeof(float) * 2 * n);
            go_e5a_code_fft_(a, a->fft_code[1]);
        }
    if (a->eff_sampled_ms > 1)
        {
            for (int k = 0; k < (both ? 2 : 1); k++)
                {
                    const float* src = k == 0 ? code_i : code_q;
                    for (uint32_t i = 0; i < a->c.samples_per_code && i < n; i++)
                        {
                            const float xr = src[2 * i], xi = src[2 * i + 1];
                            inbuf[2 * i] = xr * -1.0f - xi * 0.0f;
                            inbuf[2 * i + 1] = xr * 0.0f + xi * -1.0f;
                        }
                    memcpy(a->work, inbuf, sizeof(float) * 2 * n);
                    go_e5a_code_fft_(a, a->fft_code[2 + k]);
                }
        }
    free(inbuf);
}

void go_e5a_restart(go_e5a* a)
{
    /* state 0 with d_active (:385-404) */
    a->well_count = 0;
    a->test_statistics = 0.0f;
    memset(&a->last, 0, sizeof(a->last));
}

static void go_e5a_correlate_(go_e5a* a, const float* spectrum, int k)
{
    const uint32_t n = a->fft_size;
    const float* fc = a->fft_code[k];
    for (uint32_t i = 0; i < n; i++)
        {
            const float xr = spectrum[2 * i], xi = spectrum[2 * i + 1];
            a->work[2 * i] = xr * fc[2 * i] - xi * fc[2 * i + 1];
            a->work[2 * i + 1] = xr * fc[2 * i + 1] + xi * fc[2 * i];
        }
    go_fft_f32(a->work, (int)n, +1);
    for (uint32_t i = 0; i < n; i++) a->mag[k][i] = a->work[2 * i] * a->work[2 * i] + a->work[2 * i + 1] * a->work[2 * i + 1];
}

/* CAF filter over the per-bin peak values, :685-768.  Types as in the block: the first and third loops count in int, the
 * second in unsigned, and the (doppler_index - i) differences are cast to unsigned BEFORE any abs() in the I sums of loops
 * one and two and in the Q sum of loop two -- bins on the far side of the centre get a huge negative weight there. */
void go_e5a_caf_filter(const float* ci, const float* cq, float* cv, uint32_t num_bins, int caf_window_hz, uint32_t doppler_step, int both)
{
    int CAF_bins_half = caf_window_hz / (2 * (int)doppler_step);
    float weighting_factor = 0.5 / (float)CAF_bins_half;
    const unsigned int nb = num_bins;
    float accum;
    for (int d = 0; d < CAF_bins_half; d++)
        {
            cv[d] = 0;
            for (int i = 0; i < CAF_bins_half + d + 1; i++) cv[d] += ci[i] * (1 - weighting_factor * (unsigned int)(d - i));
            cv[d] /= 1 + CAF_bins_half + d - weighting_factor * CAF_bins_half * (CAF_bins_half + 1) / 2 - weighting_factor * d * (d + 1) / 2;
            if (both)
                {
                    accum = 0;
                    for (int i = 0; i < CAF_bins_half + d + 1; i++) accum += cq[i] * (1 - weighting_factor * (unsigned int)(abs(d - i)));
                    accum /= 1 + CAF_bins_half + d - weighting_factor * CAF_bins_half * (CAF_bins_half + 1) / 2 - weighting_factor * d * (d + 1) / 2;
                    cv[d] += accum;
                }
        }
    for (unsigned int d = CAF_bins_half; d < nb - CAF_bins_half; d++)
        {
            cv[d] = 0;
            for (int i = d - CAF_bins_half; i < (int)(d + CAF_bins_half + 1); i++) cv[d] += ci[i] * (1 - weighting_factor * (unsigned int)(d - i));
            cv[d] /= 1 + 2 * CAF_bins_half - 2 * weighting_factor * CAF_bins_half * (CAF_bins_half + 1) / 2;
            if (both)
                {
                    accum = 0;
                    for (int i = d - CAF_bins_half; i < (int)(d + CAF_bins_half + 1); i++) accum += cq[i] * (1 - weighting_factor * (unsigned int)(d - i));
                    accum /= 1 + 2 * CAF_bins_half - 2 * weighting_factor * CAF_bins_half * (CAF_bins_half + 1) / 2;
                    cv[d] += accum;
                }
        }
    for (int d = nb - CAF_bins_half; d < (int)nb; d++)
        {
            cv[d] = 0;
            for (int i = d - CAF_bins_half; i < (int)nb; i++) cv[d] += ci[i] * (1 - weighting_factor * (abs(d - i)));
            cv[d] /= 1 + CAF_bins_half + (nb - d - 1) - weighting_factor * CAF_bins_half * (CAF_bins_half + 1) / 2 - weighting_factor * (nb - d - 1) * (nb - d) / 2;
            if (both)
                {
                    accum = 0;
                    for (int i = d - CAF_bins_half; i < (int)nb; i++) accum += cq[i] * (1 - weighting_factor * (abs(d - i)));
                    accum /= 1 + CAF_bins_half + (nb - d - 1) - weighting_factor * CAF_bins_half * (CAF_bins_half + 1) / 2 - weighting_factor * (nb - d - 1) * (nb - d) / 2;
                    cv[d] += accum;
                }
        }
}

/* one dwell = state 2 of general_work (:440-792) on d_fft_size buffered items */
int go_e5a_run(go_e5a* a, const float* in, go_e5a_result* out)
{
    const uint32_t n = a->fft_size;
    const int both = a->c.both_signal_components != 0;
    const int ab = a->eff_sampled_ms > 1;
    const float fnf = (float)n * (float)n;
    float* spectrum = (float*)malloc(sizeof(float) * 2 * n);
    float input_power = 0.0f, d_mag = 0.0f;
    a->well_count++;
    a->last.updated = 0;
    for (uint32_t i = 0; i < n; i++) a->mag[0][i] = in[2 * i] * in[2 * i] + in[2 * i + 1] * in[2 * i + 1];
    for (uint32_t i = 0; i < n; i++) input_power += a->mag[0][i];
    input_power /= (float)n;
    for (uint32_t bin = 0; bin < a->num_bins; bin++)
        {
            const int doppler = -(int)a->c.doppler_max + (int)(a->c.doppler_step * bin);
            const float* w = a->wipeoffs[bin];
            uint32_t idx[4] = {0, 0, 0, 0};
            float magt[4] = {0.0f, 0.0f, 0.0f, 0.0f}; /* magt_IA, magt_QA, magt_IB, magt_QB */
            for (uint32_t i = 0; i < n; i++)
                {
                    spectrum[2 * i] = in[2 * i] * w[2 * i] - in[2 * i + 1] * w[2 * i + 1];
                    spectrum[2 * i + 1] = in[2 * i] * w[2 * i + 1] + in[2 * i + 1] * w[2 * i];
                }
            go_fft_f32(spectrum, (int)n, -1);
            go_e5a_correlate_(a, spectrum, 0);
            idx[0] = go_index_max(a->mag[0], n);
            magt[0] = a->mag[0][idx[0]] / (fnf * fnf);
            if (both)
                {
                    go_e5a_correlate_(a, spectrum, 1);
                    idx[1] = go_index_max(a->mag[1], n);
                    magt[1] = a->mag[1][idx[1]] / (fnf * fnf);
                }
            if (ab)
                {
                    go_e5a_correlate_(a, spectrum, 2);
                    idx[2] = go_index_max(a->mag[2], n);
                    magt[2] = a->mag[2][idx[2]] / (fnf * fnf);
                    if (both)
                        {
                            go_e5a_correlate_(a, spectrum, 3);
                            idx[3] = go_index_max(a->mag[3], n);
                            magt[3] = a->mag[2][idx[3]] / (fnf * fnf); /* :508 reads d_magnitudeIB at indext_QB */
                        }
                }
            /* :512-606: pick A or B per component by its peak, add the Q row onto the I row, take the row maximum */
            const int pi = (ab && !(magt[0] >= magt[2])) ? 2 : 0;
            const int pq = (ab && !(magt[1] >= magt[3])) ? 3 : 1;
            if (a->c.caf_window_hz > 0) a->caf_i[bin] = a->mag[pi][idx[pi]];
            if (both)
                {
                    if (a->c.caf_window_hz > 0) a->caf_q[bin] = a->mag[pq][idx[pq]];
                    for (uint32_t i = 0; i < n; i++) a->mag[pi][i] += a->mag[pq][i];
                }
            const uint32_t indext = go_index_max(a->mag[pi], n);
            const float m = a->mag[pi][indext] / (fnf * fnf);
            if (d_mag < m)
                {
                    d_mag = m;
                    if (a->test_statistics < (d_mag / input_power) || !a->c.bit_transition_flag)
                        {
                            a->last.code_phase_index = indext;
                            a->last.acq_delay_samples = (double)(indext % a->c.samples_per_code);
                            a->last.doppler_hz = doppler;
                            a->last.doppler_bin = bin;
                            a->last.updated = 1;
                            a->test_statistics = d_mag / input_power;
                        }
                }
        }
    free(spectrum);
    if (a->c.caf_window_hz > 0)
        {
            const unsigned int nb = a->num_bins;
            float* cv = a->caf;
            go_e5a_caf_filter(a->caf_i, a->caf_q, cv, nb, a->c.caf_window_hz, a->c.doppler_step, both);
            const uint32_t t = go_index_max(cv, nb);
            a->last.doppler_hz = -(int)a->c.doppler_max + (int)(a->c.doppler_step * t);
        }
    a->last.mag = d_mag;
    a->last.input_power = input_power;
    a->last.test_statistic = a->test_statistics;
    a->last.well_count = a->well_count;
    if (a->well_count == a->c.max_dwells)
        a->last.decision = a->test_statistics > a->c.threshold ? 1 : 0;
    else
        a->last.decision = 2;
    *out = a->last;
    return 0;
}


/* ======================= acquisition resampler -- TEST INFRASTRUCTURE, parity unpinned (GNU Radio absent) ======================= */
#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif
static double go_hamming_max_attenuation_(void) { return 53; }

int go_firdes_low_pass(double gain, double sampling_freq, double cutoff_freq, double transition_width, float* taps, int capacity)
{
    /* firdes::compute_ntaps: (int)(attenuation * fs / (22 * transition)), made odd */
    int ntaps = (int)(go_hamming_max_attenuation_() * sampling_freq / (22.0 * transition_width));
    if ((ntaps & 1) == 0) ntaps++;
    if (ntaps > capacity) return -ntaps;
    float* w = (float*)malloc(sizeof(float) * (size_t)ntaps);
    /* gr::fft::window::hamming: 0.54 - 0.46 cos(2 pi n / (ntaps - 1)), the divisor held in float32 */
    float Mw = (float)(ntaps - 1);
    for (int n = 0; n < ntaps; n++) w[n] = (float)(0.54 - 0.46 * cos((2 * M_PI * n) / Mw));
    int M = (ntaps - 1) / 2;
    double fwT0 = 2 * M_PI * cutoff_freq / sampling_freq;
    for (int n = -M; n <= M; n++)
        {
            if (n == 0)
                taps[n + M] = (float)(fwT0 / M_PI * w[n + M]);
            else
                taps[n + M] = (float)(sin(n * fwT0) / (n * M_PI) * w[n + M]);
        }
    /* unity gain at DC: sum of the (symmetric) taps */
    double fmax = taps[0 + M];
    for (int n = 1; n <= M; n++) fmax += 2 * taps[n + M];
    gain /= fmax;
    for (int i = 0; i < ntaps; i++) taps[i] = (float)(taps[i] * gain);
    free(w);
    return ntaps;
}

int go_fir_decimate(const float* in, int n_in, const float* taps, int n_taps, int D, float* out)
{
    int count = 0;
    for (long long k = 0; k * D <= (long long)n_in - 1; k++)
        {
            float ar = 0.0f, ai = 0.0f;
            for (int j = 0; j < n_taps; j++)
                {
                    const long long m = k * D - j;
                    if (m < 0) break;
                    ar += taps[j] * in[2 * m];
                    ai += taps[j] * in[2 * m + 1];
                }
            out[2 * k] = ar;
            out[2 * k + 1] = ai;
            count++;
        }
    return count;
}

uint32_t go_acq_resampler_decimation(int64_t fs_in, uint32_t opt_acq_fs)
{
    if ((uint64_t)fs_in <= opt_acq_fs) return 1;
    uint32_t decimation = (uint32_t)(fs_in / opt_acq_fs);
    while (fs_in % decimation > 0) decimation--;
    return decimation;
}


/* ======================= pcps_acquisition_fine_doppler_cc -- TEST INFRASTRUCTURE; pinned to the block itself by
 * tests/test_ref_blocks2_cpu.py::test_fine_doppler_block_equals_restatement ======================= */
struct go_fd
{
    go_fd_conf c;
    uint32_t n;
    int num_points;
    float** wipeoffs;
    float** grid;
    float* fft_code;
    float* work;
};

go_fd* go_fd_create(const go_fd_conf* conf)
{
    go_fd* a = (go_fd*)calloc(1, sizeof(go_fd));
    a->c = *conf;
    a->n = conf->samples_per_ms;
    /* set_doppler_step (:146-160): floor(|2 doppler_max| / step) points, the last one below +doppler_max */
    a->num_points = (int)floor(abs(2 * (int)conf->doppler_max) / conf->doppler_step);
    a->wipeoffs = (float**)calloc((size_t)a->num_points, sizeof(float*));
    a->grid = (float**)calloc((size_t)a->num_points, sizeof(float*));
    for (int i = 0; i < a->num_points; i++)
        {
            a->wipeoffs[i] = (float*)calloc(2 * (size_t)a->n, sizeof(float));
            a->grid[i] = (float*)calloc(a->n, sizeof(float));
            go_sincos(a->wipeoffs[i], -go_fd_phase_step(a, i), 0.0f, (int)a->n);
        }
    a->fft_code = (float*)calloc(2 * (size_t)a->n, sizeof(float));
    a->work = (float*)calloc(2 * (size_t)a->n, sizeof(float));
    return a;
}

void go_fd_destroy(go_fd* a)
{
    if (!a) return;
    for (int i = 0; i < a->num_points; i++)
        {
            free(a->wipeoffs[i]);
            free(a->grid[i]);
        }
    free(a->wipeoffs);
    free(a->grid);
    free(a->fft_code);
    free(a->work);
    free(a);
}

int go_fd_num_doppler_points(const go_fd* a) { return a->num_points; }

/* update_carrier_wipeoff (:246-261): float(GPS_TWO_PI) * int doppler / float(fs), all float32 */
float go_fd_phase_step(const go_fd* a, int bin)
{
    const int doppler_hz = (int)a->c.doppler_step * bin - (int)a->c.doppler_max;
    return (float)GO_GPS_TWO_PI * doppler_hz / (float)a->c.fs_in;
}

void go_fd_set_local_code(go_fd* a, const float* code_iq)
{
    memcpy(a->work, code_iq, sizeof(float) * 2 * a->n);
    go_fft_f32(a->work, (int)a->n, -1);
    for (uint32_t i = 0; i < a->n; i++)
        {
            a->fft_code[2 * i] = a->work[2 * i];
            a->fft_code[2 * i + 1] = -a->work[2 * i + 1];
        }
}

void go_fd_replica(const float* code, uint32_t n, int shift_index, float* rep)
{
    memcpy(rep, code, sizeof(float) * 2 * n);
    if (shift_index != 0)
        {
            /* std::rotate(first, first + (n - shift), first + n - 1): the LAST element stays where it is (:419-423) */
            const uint32_t m = n - 1, k = (n - (uint32_t)shift_index) % (m ? m : 1);
            float* tmp = (float*)malloc(sizeof(float) * 2 * m);
            for (uint32_t i = 0; i < m; i++)
                {
                    tmp[2 * i] = rep[2 * ((i + k) % m)];
                    tmp[2 * i + 1] = rep[2 * ((i + k) % m) + 1];
                }
            memcpy(rep, tmp, sizeof(float) * 2 * m);
            free(tmp);
        }
    for (int r = 1; r < 10; r++) memcpy(rep + 2 * (size_t)r * n, rep, sizeof(float) * 2 * n);
}

int go_fd_attempt(go_fd* a, const float* in, const float* code_1ms, go_fd_result* out)
{
    const uint32_t n = a->n;
    memset(out, 0, sizeof(*out));
    for (int i = 0; i < a->num_points; i++) memset(a->grid[i], 0, sizeof(float) * n); /* reset_grid */
    /* state 1, max_dwells times: compute_and_accumulate_grid (:343-372) */
    for (uint32_t d = 0; d < a->c.max_dwells; d++)
        {
            const float* x = in + 2 * (size_t)d * n;
            for (int b = 0; b < a->num_points; b++)
                {
                    const float* w = a->wipeoffs[b];
                    for (uint32_t i = 0; i < n; i++)
                        {
                            a->work[2 * i] = x[2 * i] * w[2 * i] - x[2 * i + 1] * w[2 * i + 1];
                            a->work[2 * i + 1] = x[2 * i] * w[2 * i + 1] + x[2 * i + 1] * w[2 * i];
                        }
                    go_fft_f32(a->work, (int)n, -1);
                    for (uint32_t i = 0; i < n; i++)
                        {
                            const float xr = a->work[2 * i], xi = a->work[2 * i + 1];
                            const float cr = a->fft_code[2 * i], ci = a->fft_code[2 * i + 1];
                            a->work[2 * i] = xr * cr - xi * ci;
                            a->work[2 * i + 1] = xr * ci + xi * cr;
                        }
                    go_fft_f32(a->work, (int)n, +1);
                    for (uint32_t i = 0; i < n; i++)
                        a->grid[b][i] = a->grid[b][i] + (a->work[2 * i] * a->work[2 * i] + a->work[2 * i + 1] * a->work[2 * i + 1]);
                }
        }
    /* state 2: compute_CAF (:266-326) */
    float firstPeak = 0.0f;
    int index_doppler = 0;
    uint32_t index_time = 0;
    for (int i = 0; i < a->num_points; i++)
        {
            const uint32_t t = go_index_max(a->grid[i], n);
            if (a->grid[i][t] > firstPeak)
                {
                    firstPeak = a->grid[i][t];
                    index_doppler = i;
                    index_time = t;
                }
        }
    const double GPS_L1_CA_CHIP_PERIOD = 9.7752e-07;
    const uint32_t samplesPerChip = (uint32_t)ceil(GPS_L1_CA_CHIP_PERIOD * (float)a->c.fs_in);
    int32_t ex1 = (int32_t)(index_time - samplesPerChip), ex2 = (int32_t)(index_time + samplesPerChip);
    if (ex1 < 0)
        ex1 = (int32_t)n + ex1;
    else if (ex2 >= (int32_t)n)
        ex2 = ex2 - (int32_t)n;
    int32_t idx = ex1;
    do
        {
            a->grid[index_doppler][idx] = 0.0f;
            idx++;
            if (idx == (int32_t)n) idx = 0;
        }
    while (idx != ex2);
    const uint32_t t2 = go_index_max(a->grid[index_doppler], n);
    const float secondPeak = a->grid[index_doppler][t2];
    out->index_time = index_time;
    out->index_doppler = (uint32_t)index_doppler;
    out->first_peak = firstPeak;
    out->second_peak = secondPeak;
    out->test_statistic = firstPeak / secondPeak;
    out->acq_delay_samples = (double)index_time;
    out->acq_doppler_hz = (double)(index_doppler * (int)a->c.doppler_step - (int)a->c.doppler_max);
    out->positive = out->test_statistic > a->c.threshold;
    if (!out->positive) return 0;
    /* state 3: estimate_Doppler (:400-479) over the first 10 ms of the attempt */
    const int signal_samples = 10 * (int)n, ext = signal_samples * 8;
    float* rep = (float*)malloc(sizeof(float) * 2 * (size_t)signal_samples);
    float* buf = (float*)calloc(2 * (size_t)ext, sizeof(float));
    go_fd_replica(code_1ms, n, (int)out->acq_delay_samples, rep);
    for (int i = 0; i < signal_samples; i++)
        {
            buf[2 * i] = in[2 * i] * rep[2 * i] - in[2 * i + 1] * rep[2 * i + 1];
            buf[2 * i + 1] = in[2 * i] * rep[2 * i + 1] + in[2 * i + 1] * rep[2 * i];
        }
    go_fft_f32(buf, ext, -1);
    float* mag = (float*)malloc(sizeof(float) * (size_t)ext);
    for (int i = 0; i < ext; i++) mag[i] = buf[2 * i] * buf[2 * i] + buf[2 * i + 1] * buf[2 * i + 1];
    const uint32_t k = go_index_max(mag, (uint32_t)ext);
    /* fftFreqBins (:443-459): k < ext/2 -> +, then ext/2 .. 1 counted down with a minus sign */
    float f;
    if ((int)k < ext / 2)
        f = (((float)a->c.fs_in / 2.0) * (float)k) / ((float)ext / 2.0);
    else
        f = ((-(float)a->c.fs_in / 2.0) * (float)(ext / 2 - ((int)k - ext / 2))) / ((float)ext / 2.0);
    out->fine_index = k;
    if (fabs(f - out->acq_doppler_hz) < 1000)
        {
            out->acq_doppler_hz = (double)f;
            out->fine_accepted = 1;
        }
    free(rep);
    free(buf);
    free(mag);
    return 0;
}

/* ======================= pcps_quicksync_acquisition_cc -- TEST INFRASTRUCTURE =======================
 * src/algorithms/acquisition/gnuradio_blocks/pcps_quicksync_acquisition_cc.cc.  fft_size = samples_per_code / folding_factor
 * (:95); a dwell uses samples_per_code * folding_factor items, wiped off and folded folding_factor^2 times (:391-398).
 * The block's complex_acumulator[100] is an UNINITIALISED stack array it adds onto (:439-462); this restatement starts it at 0. */
struct go_qs
{
    go_qs_conf c;
    uint32_t fft_size, num_bins, used;
    float** wipeoffs;
    float *fft_codes, *code, *work, *in_temp, *mag;
    uint32_t well_count;
    float d_mag, test_statistics;
    double acq_delay_samples, acq_doppler_hz;
    uint64_t acq_samplestamp;
};

go_qs* go_qs_create(const go_qs_conf* conf)
{
    go_qs* a = (go_qs*)calloc(1, sizeof(go_qs));
    a->c = *conf;
    if (a->c.doppler_step == 0) a->c.doppler_step = 250; /* init() :214-217 */
    a->fft_size = (uint32_t)conf->samples_per_code / conf->folding_factor;
    a->used = (uint32_t)conf->samples_per_code * conf->folding_factor;
    for (int32_t d = -(int32_t)a->c.doppler_max; d <= (int32_t)a->c.doppler_max; d += (int32_t)a->c.doppler_step) a->num_bins++;
    a->wipeoffs = (float**)calloc(a->num_bins, sizeof(float*));
    for (uint32_t i = 0; i < a->num_bins; i++)
        {
            a->wipeoffs[i] = (float*)calloc(2 * (size_t)a->used, sizeof(float));
            go_sincos(a->wipeoffs[i], -go_qs_phase_step(a, i), 0.0f, (int)a->used);
        }
    a->fft_codes = (float*)calloc(2 * (size_t)a->fft_size, sizeof(float));
    a->code = (float*)calloc(2 * (size_t)conf->samples_per_code, sizeof(float));
    a->work = (float*)calloc(2 * (size_t)a->fft_size, sizeof(float));
    a->in_temp = (float*)calloc(2 * (size_t)a->used, sizeof(float));
    a->mag = (float*)calloc(a->fft_size, sizeof(float));
    return a;
}

void go_qs_destroy(go_qs* a)
{
    if (!a) return;
    for (uint32_t i = 0; i < a->num_bins; i++) free(a->wipeoffs[i]);
    free(a->wipeoffs);
    free(a->fft_codes);
    free(a->code);
    free(a->work);
    free(a->in_temp);
    free(a->mag);
    free(a);
}

uint32_t go_qs_fft_size(const go_qs* a) { return a->fft_size; }
uint32_t go_qs_num_doppler_bins(const go_qs* a) { return a->num_bins; }
/* :236-237: float phase_step_rad = GPS_TWO_PI * doppler / static_cast<float>(d_fs_in) -- double product rounded once */
float go_qs_phase_step(const go_qs* a, uint32_t bin)
{
    const int32_t doppler = -(int32_t)a->c.doppler_max + (int32_t)(a->c.doppler_step * bin);
    return (float)(GO_GPS_TWO_PI * doppler / (float)a->c.fs_in);
}

/* set_local_code (:177-200): keep the unfolded code, fold folding_factor segments onto a zeroed buffer, FFT, conjugate */
void go_qs_set_local_code(go_qs* a, const float* code_iq, float* folded_iq)
{
    const uint32_t n = a->fft_size;
    memcpy(a->code, code_iq, sizeof(float) * 2 * (size_t)a->c.samples_per_code);
    memset(a->work, 0, sizeof(float) * 2 * n);
    for (uint32_t i = 0; i < a->c.folding_factor; i++)
        for (uint32_t k = 0; k < 2 * n; k++) a->work[k] = code_iq[2 * (size_t)i * n + k] + a->work[k]; /* std::plus(code, inbuf) */
    if (folded_iq) memcpy(folded_iq, a->work, sizeof(float) * 2 * n);
    go_fft_f32(a->work, (int)n, -1);
    for (uint32_t i = 0; i < n; i++)
        {
            a->fft_codes[2 * i] = a->work[2 * i];
            a->fft_codes[2 * i + 1] = -a->work[2 * i + 1];
        }
}

void go_qs_restart(go_qs* a)
{
    /* state 0 with d_active (:296-310) / set_state(1) (:255-266) */
    a->acq_delay_samples = 0.0;
    a->acq_doppler_hz = 0.0;
    a->acq_samplestamp = 0;
    a->well_count = 0;
    a->d_mag = 0.0f;
    a->test_statistics = 0.0f;
}

/* state 1 (:318-532) */
int go_qs_run(go_qs* a, const float* in_iq, uint64_t sample_counter_after, go_qs_result* out)
{
    const uint32_t n = a->fft_size, F = a->c.folding_factor, spc = (uint32_t)a->c.samples_per_code, used = a->used;
    const float fnf = (float)n * (float)n;
    float input_power = 0.0f;
    memset(out, 0, sizeof(*out));
    a->d_mag = 0.0f;            /* :343-346: d_mag and d_test_statistics restart EVERY dwell */
    a->test_statistics = 0.0f;
    a->well_count++;
    for (uint32_t i = 0; i < used; i++) input_power += in_iq[2 * i] * in_iq[2 * i] + in_iq[2 * i + 1] * in_iq[2 * i + 1];
    input_power /= (float)used;
    for (uint32_t bin = 0; bin < a->num_bins; bin++)
        {
            const int32_t doppler = -(int32_t)a->c.doppler_max + (int32_t)(a->c.doppler_step * bin);
            const float* w = a->wipeoffs[bin];
            for (uint32_t i = 0; i < used; i++)
                {
                    a->in_temp[2 * i] = in_iq[2 * i] * w[2 * i] - in_iq[2 * i + 1] * w[2 * i + 1];
                    a->in_temp[2 * i + 1] = in_iq[2 * i] * w[2 * i + 1] + in_iq[2 * i + 1] * w[2 * i];
                }
            memset(a->work, 0, sizeof(float) * 2 * n);
            for (uint32_t i = 0; i < F * F; i++)
                for (uint32_t k = 0; k < 2 * n; k++) a->work[k] = a->in_temp[2 * (size_t)i * n + k] + a->work[k];
            go_fft_f32(a->work, (int)n, -1);
            for (uint32_t i = 0; i < n; i++)
                {
                    const float xr = a->work[2 * i], xi = a->work[2 * i + 1];
                    const float cr = a->fft_codes[2 * i], ci = a->fft_codes[2 * i + 1];
                    a->work[2 * i] = xr * cr - xi * ci;
                    a->work[2 * i + 1] = xr * ci + xi * cr;
                }
            go_fft_f32(a->work, (int)n, +1);
            for (uint32_t i = 0; i < n; i++) a->mag[i] = a->work[2 * i] * a->work[2 * i] + a->work[2 * i + 1] * a->work[2 * i + 1];
            const uint32_t indext = go_index_max(a->mag, n);
            const float magt = a->mag[indext] / (fnf * fnf);
            if (a->d_mag < magt)
                {
                    a->d_mag = magt;
                    if (a->test_statistics < (a->d_mag / input_power) || !a->c.bit_transition_flag)
                        {
                            const uint32_t folded = indext % spc;
                            float best = 0.0f;
                            uint32_t best_i = 0;
                            for (uint32_t i = 0; i < F; i++)
                                {
                                    const uint32_t delay = folded + i * n;
                                    float ar = 0.0f, ai = 0.0f; /* the block's accumulator is uninitialised stack memory */
                                    for (uint32_t j = 0; j < spc; j++)
                                        {
                                            const float xr = a->in_temp[2 * (size_t)(delay + j)], xi = a->in_temp[2 * (size_t)(delay + j) + 1];
                                            const float cr = a->code[2 * j], ci = a->code[2 * j + 1];
                                            ar += xr * cr - xi * ci;
                                            ai += xr * ci + xi * cr;
                                        }
                                    const float m = ar * ar + ai * ai;
                                    if (i < 16) out->corr_output[i] = m;
                                    if (i == 0 || m > best) /* volk_gnsssdr_32f_index_max_32u: first maximum */
                                        {
                                            best = m;
                                            best_i = i;
                                        }
                                }
                            a->acq_delay_samples = (double)(folded + best_i * n);
                            a->acq_doppler_hz = (double)doppler;
                            a->acq_samplestamp = sample_counter_after;
                            a->test_statistics = a->d_mag / input_power;
                            out->improved = 1;
                            out->folded_index = indext;
                            out->doppler_bin = bin;
                            out->best_candidate = best_i;
                        }
                }
        }
    int state = 1;
    if (!a->c.bit_transition_flag)
        {
            if (a->test_statistics > a->c.threshold)
                state = 2;
            else if (a->well_count == a->c.max_dwells)
                state = 3;
        }
    else if (a->well_count == a->c.max_dwells)
        state = a->test_statistics > a->c.threshold ? 2 : 3;
    out->well_count = a->well_count;
    out->mag = a->d_mag;
    out->input_power = input_power;
    out->test_statistic = a->test_statistics;
    out->state = state;
    out->acq_delay_samples = a->acq_delay_samples;
    out->acq_doppler_hz = a->acq_doppler_hz;
    out->acq_samplestamp = a->acq_samplestamp;
    return 0;
}
