/* TEST INFRASTRUCTURE ONLY (oracle/_ref build).
 * Compiles the reference's volk_gnsssdr protokernel headers AS C, from where they lie under
 * /root/reference, and defines the un-suffixed dispatcher symbols (normally produced by the
 * reference's mako-generated dispatcher) as run-time forwards to _generic or _u_avx. */
#include <volk_gnsssdr/volk_gnsssdr.h>
#include <math.h>
#include <complex.h>

int gb_ref_volk_arch = 0;
#include <volk_gnsssdr_32f_xn_resampler_32f_xn.h>
#include <volk_gnsssdr_32f_xn_high_dynamics_resampler_32f_xn.h>
#include <volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn.h>
#include <volk_gnsssdr_32fc_32f_high_dynamic_rotator_dot_prod_32fc_xn.h>
#include <volk_gnsssdr_32fc_xn_resampler_32fc_xn.h>
#include <volk_gnsssdr_32fc_x2_rotator_dot_prod_32fc_xn.h>
#include <volk_gnsssdr_16ic_xn_resampler_16ic_xn.h>
#include <volk_gnsssdr_16ic_x2_rotator_dot_prod_16ic_xn.h>
#include <volk_gnsssdr_s32f_sincos_32fc.h>
#include <volk_gnsssdr_32f_index_max_32u.h>
#include <volk_gnsssdr_16ic_convert_32fc.h>

void volk_gnsssdr_32f_xn_resampler_32f_xn(float** result, const float* local_code, float rem_code_phase_chips, float code_phase_step_chips, float* shifts_chips, unsigned int code_length_chips, int num_out_vectors, unsigned int num_points)
{
    if (gb_ref_volk_arch == 1)
        volk_gnsssdr_32f_xn_resampler_32f_xn_u_avx(result, local_code, rem_code_phase_chips, code_phase_step_chips, shifts_chips, code_length_chips, num_out_vectors, num_points);
    else
        volk_gnsssdr_32f_xn_resampler_32f_xn_generic(result, local_code, rem_code_phase_chips, code_phase_step_chips, shifts_chips, code_length_chips, num_out_vectors, num_points);
}

void volk_gnsssdr_32f_xn_high_dynamics_resampler_32f_xn(float** result, const float* local_code, float rem_code_phase_chips, float code_phase_step_chips, float code_phase_rate_step_chips, float* shifts_chips, unsigned int code_length_chips, int num_out_vectors, unsigned int num_points)
{
    if (gb_ref_volk_arch == 1)
        volk_gnsssdr_32f_xn_high_dynamics_resampler_32f_xn_u_avx(result, local_code, rem_code_phase_chips, code_phase_step_chips, code_phase_rate_step_chips, shifts_chips, code_length_chips, num_out_vectors, num_points);
    else
        volk_gnsssdr_32f_xn_high_dynamics_resampler_32f_xn_generic(result, local_code, rem_code_phase_chips, code_phase_step_chips, code_phase_rate_step_chips, shifts_chips, code_length_chips, num_out_vectors, num_points);
}

void volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn(lv_32fc_t* result, const lv_32fc_t* in_common, const lv_32fc_t phase_inc, lv_32fc_t* phase, const float** in_a, int num_a_vectors, unsigned int num_points)
{
    if (gb_ref_volk_arch == 1)
        volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn_u_avx(result, in_common, phase_inc, phase, in_a, num_a_vectors, num_points);
    else
        volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn_generic(result, in_common, phase_inc, phase, in_a, num_a_vectors, num_points);
}

/* no SIMD protokernel exists for the high-dynamics rotator */
void volk_gnsssdr_32fc_32f_high_dynamic_rotator_dot_prod_32fc_xn(lv_32fc_t* result, const lv_32fc_t* in_common, const lv_32fc_t phase_inc, const lv_32fc_t phase_inc_rate, lv_32fc_t* phase, const float** in_a, int num_a_vectors, unsigned int num_points)
{
    volk_gnsssdr_32fc_32f_high_dynamic_rotator_dot_prod_32fc_xn_generic(result, in_common, phase_inc, phase_inc_rate, phase, in_a, num_a_vectors, num_points);
}

void volk_gnsssdr_32fc_xn_resampler_32fc_xn(lv_32fc_t** result, const lv_32fc_t* local_code, float rem_code_phase_chips, float code_phase_step_chips, float* shifts_chips, unsigned int code_length_chips, int num_out_vectors, unsigned int num_points)
{
    if (gb_ref_volk_arch == 1)
        volk_gnsssdr_32fc_xn_resampler_32fc_xn_u_avx(result, local_code, rem_code_phase_chips, code_phase_step_chips, shifts_chips, code_length_chips, num_out_vectors, num_points);
    else
        volk_gnsssdr_32fc_xn_resampler_32fc_xn_generic(result, local_code, rem_code_phase_chips, code_phase_step_chips, shifts_chips, code_length_chips, num_out_vectors, num_points);
}

void volk_gnsssdr_32fc_x2_rotator_dot_prod_32fc_xn(lv_32fc_t* result, const lv_32fc_t* in_common, const lv_32fc_t phase_inc, lv_32fc_t* phase, const lv_32fc_t** in_a, int num_a_vectors, unsigned int num_points)
{
    if (gb_ref_volk_arch == 1)
        volk_gnsssdr_32fc_x2_rotator_dot_prod_32fc_xn_u_avx(result, in_common, phase_inc, phase, in_a, num_a_vectors, num_points);
    else
        volk_gnsssdr_32fc_x2_rotator_dot_prod_32fc_xn_generic(result, in_common, phase_inc, phase, in_a, num_a_vectors, num_points);
}

void volk_gnsssdr_16ic_xn_resampler_16ic_xn(lv_16sc_t** result, const lv_16sc_t* local_code, float rem_code_phase_chips, float code_phase_step_chips, float* shifts_chips, unsigned int code_length_chips, int num_out_vectors, unsigned int num_points)
{
    volk_gnsssdr_16ic_xn_resampler_16ic_xn_generic(result, local_code, rem_code_phase_chips, code_phase_step_chips, shifts_chips, code_length_chips, num_out_vectors, num_points);
}

void volk_gnsssdr_16ic_x2_rotator_dot_prod_16ic_xn(lv_16sc_t* result, const lv_16sc_t* in_common, const lv_32fc_t phase_inc, lv_32fc_t* phase, const lv_16sc_t** in_a, int num_a_vectors, unsigned int num_points)
{
    volk_gnsssdr_16ic_x2_rotator_dot_prod_16ic_xn_generic(result, in_common, phase_inc, phase, in_a, num_a_vectors, num_points);
}


/* ---- stand-alone protokernel doorways (ctypes) ---- */
void gbref_resampler_real(float* out /*T x n*/, const float* code, float rem, float step, const float* shifts, int code_len, int n_taps, int n, int avx)
{
    float* rows[32];
    float sh[32];
    int t;
    for (t = 0; t < n_taps; t++) { rows[t] = out + (size_t)t * n; sh[t] = shifts[t]; }
    if (avx)
        volk_gnsssdr_32f_xn_resampler_32f_xn_u_avx(rows, code, rem, step, sh, code_len, n_taps, n);
    else
        volk_gnsssdr_32f_xn_resampler_32f_xn_generic(rows, code, rem, step, sh, code_len, n_taps, n);
}

void gbref_rotator_dot_prod_real(float* result_iq, const float* in_iq, float inc_re, float inc_im, float* phase_iq, const float* codes /*T x n*/, int n_taps, int n, int avx)
{
    const float* rows[32];
    lv_32fc_t ph = lv_cmake(phase_iq[0], phase_iq[1]);
    int t;
    for (t = 0; t < n_taps; t++) rows[t] = codes + (size_t)t * n;
    if (avx)
        volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn_u_avx((lv_32fc_t*)result_iq, (const lv_32fc_t*)in_iq, lv_cmake(inc_re, inc_im), &ph, rows, n_taps, n);
    else
        volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn_generic((lv_32fc_t*)result_iq, (const lv_32fc_t*)in_iq, lv_cmake(inc_re, inc_im), &ph, rows, n_taps, n);
    phase_iq[0] = lv_creal(ph);
    phase_iq[1] = lv_cimag(ph);
}

void gbref_sincos(float* out_iq, float phase_inc, float phase0, int n, int variant)
{
    float ph[1];
    ph[0] = phase0;
    if (variant == 1)
        volk_gnsssdr_s32f_sincos_32fc_u_avx2((lv_32fc_t*)out_iq, phase_inc, ph, n);
    else if (variant == 2)
        volk_gnsssdr_s32f_sincos_32fc_generic_fxpt((lv_32fc_t*)out_iq, phase_inc, ph, n);
    else
        volk_gnsssdr_s32f_sincos_32fc_generic((lv_32fc_t*)out_iq, phase_inc, ph, n);
}

unsigned gbref_index_max(const float* src, unsigned n)
{
    uint32_t idx = 0;
    volk_gnsssdr_32f_index_max_32u_generic(&idx, src, n);
    return idx;
}

void gbref_16ic_convert_32fc(float* out_iq, const int16_t* in_iq, unsigned n)
{
    volk_gnsssdr_16ic_convert_32fc_generic((lv_32fc_t*)out_iq, (const lv_16sc_t*)in_iq, n);
}

/* ---- dispatcher names pcps_acquisition.cc calls: the reference's own _generic protokernels ---- */
void volk_gnsssdr_s32f_sincos_32fc(lv_32fc_t* out, const float phase_inc, float* phase, unsigned int num_points)
{
    volk_gnsssdr_s32f_sincos_32fc_generic(out, phase_inc, phase, num_points);
}
void volk_gnsssdr_32f_index_max_32u(uint32_t* target, const float* src0, uint32_t num_points)
{
    volk_gnsssdr_32f_index_max_32u_generic(target, src0, num_points);
}
void volk_gnsssdr_16ic_convert_32fc(lv_32fc_t* outputVector, const lv_16sc_t* inputVector, unsigned int num_points)
{
    volk_gnsssdr_16ic_convert_32fc_generic(outputVector, inputVector, num_points);
}

/* ---- upstream VOLK (not in the reference tree): published _generic protokernels, see shim/volk/volk.h ---- */
int gb_volk_accumulator_lanes = 0;
void volk_32fc_x2_multiply_32fc(lv_32fc_t* c, const lv_32fc_t* a, const lv_32fc_t* b, unsigned int n)
{
    const float* pa = (const float*)a;
    const float* pb = (const float*)b;
    float* pc = (float*)c;
    for (unsigned int i = 0; i < n; i++)
        {
            const float ar = pa[2 * i], ai = pa[2 * i + 1], br = pb[2 * i], bi = pb[2 * i + 1];
            pc[2 * i] = ar * br - ai * bi;
            pc[2 * i + 1] = ar * bi + ai * br;
        }
}
void volk_32fc_conjugate_32fc(lv_32fc_t* c, const lv_32fc_t* a, unsigned int n)
{
    const float* pa = (const float*)a;
    float* pc = (float*)c;
    for (unsigned int i = 0; i < n; i++)
        {
            pc[2 * i] = pa[2 * i];
            pc[2 * i + 1] = -pa[2 * i + 1];
        }
}
void volk_32fc_magnitude_squared_32f(float* m, const lv_32fc_t* a, unsigned int n)
{
    const float* pa = (const float*)a;
    for (unsigned int i = 0; i < n; i++) m[i] = pa[2 * i] * pa[2 * i] + pa[2 * i + 1] * pa[2 * i + 1];
}
void volk_32f_accumulator_s32f(float* result, const float* in, unsigned int n)
{
    float r = 0.0f;
    if (gb_volk_accumulator_lanes > 1)
        {
            const unsigned int lanes = gb_volk_accumulator_lanes > 16 ? 16 : (unsigned int)gb_volk_accumulator_lanes, whole = n / lanes;
            float acc[16] = {0};
            for (unsigned int k = 0; k < whole; k++)
                for (unsigned int l = 0; l < lanes; l++) acc[l] += in[k * lanes + l];
            r = acc[0];
            for (unsigned int l = 1; l < lanes; l++) r += acc[l];
            for (unsigned int i = whole * lanes; i < n; i++) r += in[i];
        }
    else
        for (unsigned int i = 0; i < n; i++) r += in[i];
    *result = r;
}
void volk_32f_x2_add_32f(float* c, const float* a, const float* b, unsigned int n)
{
    for (unsigned int i = 0; i < n; i++) c[i] = a[i] + b[i];
}

void volk_32f_s32f_multiply_32f(float* c, const float* a, const float scalar, unsigned int n)
{
    for (unsigned int i = 0; i < n; i++) c[i] = a[i] * scalar;
}
void volk_32fc_s32fc_multiply_32fc(lv_32fc_t* c, const lv_32fc_t* a, const lv_32fc_t scalar, unsigned int n)
{
    const float* pa = (const float*)a;
    float* pc = (float*)c;
    const float sr = crealf(scalar), si = cimagf(scalar);
    for (unsigned int i = 0; i < n; i++)
        {
            const float ar = pa[2 * i], ai = pa[2 * i + 1];
            pc[2 * i] = ar * sr - ai * si;
            pc[2 * i + 1] = ar * si + ai * sr;
        }
}
/* volk_gnsssdr_32fc_convert_16ic (gps_l1_ca_dll_pll_c_aid_tracking_sc.cc:252-255): the reference's _generic protokernel */
#include <volk_gnsssdr_32fc_convert_16ic.h>
void volk_gnsssdr_32fc_convert_16ic(lv_16sc_t* outputVector, const lv_32fc_t* inputVector, unsigned int num_points)
{
    volk_gnsssdr_32fc_convert_16ic_generic(outputVector, inputVector, num_points);
}
