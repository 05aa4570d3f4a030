/* TEST INFRASTRUCTURE ONLY (oracle/_ref build): upstream VOLK is a system dependency of the reference, not part of its tree.
 * These are the five kernels pcps_acquisition.cc calls, stated as VOLK's published _generic protokernels (plain float32 loops,
 * no fused multiply-add: the library is built -ffp-contract=off).  volk_32f_accumulator_s32f also offers VOLK's a_avx / u_avx
 * order (eight interleaved partial sums, lanes added in order, tail) through gb_volk_accumulator_lanes, because that is the
 * protokernel VOLK dispatches on an x86 host and the two orders differ beyond 1e-4 on long 8-bit blocks. */
#ifndef GB_ORACLE_SHIM_VOLK_H
#define GB_ORACLE_SHIM_VOLK_H
#include <volk/volk_complex.h>
#ifdef __cplusplus
extern "C" {
#endif
extern int gb_volk_accumulator_lanes;
void volk_32fc_x2_multiply_32fc(lv_32fc_t* c, const lv_32fc_t* a, const lv_32fc_t* b, unsigned int n);
void volk_32fc_conjugate_32fc(lv_32fc_t* c, const lv_32fc_t* a, unsigned int n);
void volk_32fc_magnitude_squared_32f(float* m, const lv_32fc_t* a, unsigned int n);
void volk_32f_accumulator_s32f(float* result, const float* in, unsigned int n);
void volk_32f_x2_add_32f(float* c, const float* a, const float* b, unsigned int n);
/* used by the sibling acquisition blocks (Tong, 8 ms, E5a CAF, fine Doppler, QuickSync) */
void volk_32f_s32f_multiply_32f(float* c, const float* a, const float scalar, unsigned int n);
void volk_32fc_s32fc_multiply_32fc(lv_32fc_t* c, const lv_32fc_t* a, const lv_32fc_t scalar, unsigned int n);
#define volk_32f_accumulator_s32f_a volk_32f_accumulator_s32f
#define volk_32f_s32f_multiply_32f_a volk_32f_s32f_multiply_32f
#ifdef __cplusplus
}
#endif
#endif
