/* TEST INFRASTRUCTURE ONLY: one-line stand-in for GNU Radio's gr_complex typedef (oracle/_ref build). */
#ifndef GB_ORACLE_SHIM_GR_COMPLEX_H
#define GB_ORACLE_SHIM_GR_COMPLEX_H
#include <complex>
typedef std::complex<float> gr_complex;
#endif
