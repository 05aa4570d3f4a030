/* TEST INFRASTRUCTURE ONLY (oracle/_ref build): Gnss_Synchro::serialize is a template nobody instantiates here. */
#ifndef GB_ORACLE_SHIM_BOOST_NVP_HPP
#define GB_ORACLE_SHIM_BOOST_NVP_HPP
#define BOOST_SERIALIZATION_NVP(x) x
#endif
