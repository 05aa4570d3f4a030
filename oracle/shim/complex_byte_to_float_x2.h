/* TEST INFRASTRUCTURE ONLY: stands in for src/algorithms/libs/complex_byte_to_float_x2.h (a gr::sync_block) where the reference's
 * adapters are compiled only to read the Acq_Conf / Dll_Pll_Conf they derive (tests/native/adapter_conf_compare.cc). */
#ifndef GB_ORACLE_SHIM_COMPLEX_BYTE_TO_FLOAT_X2_H
#define GB_ORACLE_SHIM_COMPLEX_BYTE_TO_FLOAT_X2_H
#include <gnuradio/block.h>
#include <memory>
class complex_byte_to_float_x2 : public gr::basic_block
{
public:
    long unique_id() const { return 0; }
};
typedef std::shared_ptr<complex_byte_to_float_x2> complex_byte_to_float_x2_sptr;
inline complex_byte_to_float_x2_sptr make_complex_byte_to_float_x2() { return complex_byte_to_float_x2_sptr(new complex_byte_to_float_x2()); }
#endif
