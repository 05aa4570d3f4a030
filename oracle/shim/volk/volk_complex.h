#ifndef GB_ORACLE_SHIM_VOLK_COMPLEX_H
#define GB_ORACLE_SHIM_VOLK_COMPLEX_H
#include <volk_gnsssdr/volk_gnsssdr_complex.h>
#endif
