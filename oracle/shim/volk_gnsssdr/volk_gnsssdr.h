/* TEST INFRASTRUCTURE ONLY (oracle/_ref build).
 *
 * Stand-in for the dispatcher header that the reference generates with mako
 * templates (gen/volk_gnsssdr_tmpl_utils.py), which cannot run in this image.
 * The reference's protokernel headers are C, so they are compiled as C in
 * oracle/ref_volk.c (included from /root/reference where they lie; nothing is
 * copied); this header only declares the un-suffixed dispatcher names that
 * cpu_multicorrelator*.cc call. ref_volk.c forwards each to a protokernel
 * chosen at run time:
 *   gb_ref_volk_arch == 0 -> _generic  (the reference QA's own oracle)
 *   gb_ref_volk_arch == 1 -> _u_avx    (what volk_gnsssdr_profile picks on x86)
 */
#ifndef GB_ORACLE_SHIM_VOLK_GNSSSDR_H
#define GB_ORACLE_SHIM_VOLK_GNSSSDR_H

#include <immintrin.h>
#include <stdlib.h>
#include <volk_gnsssdr/volk_gnsssdr_common.h>
#include <volk_gnsssdr/volk_gnsssdr_complex.h>
#include <volk_gnsssdr/volk_gnsssdr_malloc.h>

__VOLK_DECL_BEGIN
extern int gb_ref_volk_arch;
size_t volk_gnsssdr_get_alignment(void);
__VOLK_DECL_END


__VOLK_DECL_BEGIN
void volk_gnsssdr_32f_xn_resampler_32f_xn(float** result, const float* local_code, float rem_code_phase_chips, float code_phase_step_chips, float* shifts_chips, unsigned int code_length_chips, int num_out_vectors, unsigned int num_points);
void volk_gnsssdr_32f_xn_high_dynamics_resampler_32f_xn(float** result, const float* local_code, float rem_code_phase_chips, float code_phase_step_chips, float code_phase_rate_step_chips, float* shifts_chips, unsigned int code_length_chips, int num_out_vectors, unsigned int num_points);
void volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn(lv_32fc_t* result, const lv_32fc_t* in_common, const lv_32fc_t phase_inc, lv_32fc_t* phase, const float** in_a, int num_a_vectors, unsigned int num_points);
void volk_gnsssdr_32fc_32f_high_dynamic_rotator_dot_prod_32fc_xn(lv_32fc_t* result, const lv_32fc_t* in_common, const lv_32fc_t phase_inc, const lv_32fc_t phase_inc_rate, lv_32fc_t* phase, const float** in_a, int num_a_vectors, unsigned int num_points);
void volk_gnsssdr_32fc_xn_resampler_32fc_xn(lv_32fc_t** result, const lv_32fc_t* local_code, float rem_code_phase_chips, float code_phase_step_chips, float* shifts_chips, unsigned int code_length_chips, int num_out_vectors, unsigned int num_points);
void volk_gnsssdr_32fc_x2_rotator_dot_prod_32fc_xn(lv_32fc_t* result, const lv_32fc_t* in_common, const lv_32fc_t phase_inc, lv_32fc_t* phase, const lv_32fc_t** in_a, int num_a_vectors, unsigned int num_points);
void volk_gnsssdr_16ic_xn_resampler_16ic_xn(lv_16sc_t** result, const lv_16sc_t* local_code, float rem_code_phase_chips, float code_phase_step_chips, float* shifts_chips, unsigned int code_length_chips, int num_out_vectors, unsigned int num_points);
void volk_gnsssdr_16ic_x2_rotator_dot_prod_16ic_xn(lv_16sc_t* result, const lv_16sc_t* in_common, const lv_32fc_t phase_inc, lv_32fc_t* phase, const lv_16sc_t** in_a, int num_a_vectors, unsigned int num_points);
/* the three volk_gnsssdr kernels pcps_acquisition.cc calls (always the _generic protokernel) */
void volk_gnsssdr_s32f_sincos_32fc(lv_32fc_t* out, const float phase_inc, float* phase, unsigned int num_points);
void volk_gnsssdr_32f_index_max_32u(uint32_t* target, const float* src0, uint32_t num_points);
void volk_gnsssdr_16ic_convert_32fc(lv_32fc_t* outputVector, const lv_16sc_t* inputVector, unsigned int num_points);
void volk_gnsssdr_32fc_convert_16ic(lv_16sc_t* outputVector, const lv_32fc_t* inputVector, unsigned int num_points);
__VOLK_DECL_END

#endif
