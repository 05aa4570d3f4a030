/* TEST INFRASTRUCTURE ONLY (oracle/_ref build): shadows src/algorithms/libs/gnss_sdr_flags.h (gflags is not installed).  The
 * command-line flags the sibling tracking blocks read, as plain globals with gnss_sdr_flags.cc's defaults (:53-59); the harness
 * may set them before it creates a block. */
#ifndef GB_ORACLE_SHIM_GNSS_SDR_FLAGS_H
#define GB_ORACLE_SHIM_GNSS_SDR_FLAGS_H
#include <cstdint>
extern int32_t FLAGS_cn0_samples;
extern int32_t FLAGS_cn0_min;
extern int32_t FLAGS_max_lock_fail;
extern double FLAGS_carrier_lock_th;
extern int32_t FLAGS_doppler_max;
extern int32_t FLAGS_doppler_step;
extern double FLAGS_dll_bw_hz;
extern double FLAGS_pll_bw_hz;
#endif
