/* TEST INFRASTRUCTURE ONLY (oracle/_ref build): the blocks' .mat dumps are not exercised through the compiled reference (matio is
 * not installed); these no-op stand-ins only let save_matfile() / dump_results() compile. */
#ifndef GB_ORACLE_SHIM_MATIO_H
#define GB_ORACLE_SHIM_MATIO_H
#include <cstddef>
struct mat_t;
struct matvar_t;
enum { MAT_FT_MAT73 = 0x0200, MAT_FT_MAT5 = 0x0100 };
enum { MAT_COMPRESSION_NONE = 0, MAT_COMPRESSION_ZLIB = 1 };
enum { MAT_C_DOUBLE = 6, MAT_C_SINGLE = 7, MAT_C_INT32 = 12, MAT_C_UINT32 = 13, MAT_C_INT64 = 14, MAT_C_UINT64 = 15, MAT_C_UINT8 = 9 };
enum { MAT_T_UINT8 = 2, MAT_T_INT32 = 5, MAT_T_UINT32 = 6, MAT_T_SINGLE = 7, MAT_T_DOUBLE = 9, MAT_T_INT64 = 12, MAT_T_UINT64 = 13 };
inline mat_t* Mat_CreateVer(const char*, const char*, int) { return nullptr; }
inline matvar_t* Mat_VarCreate(const char*, int, int, int, size_t*, void*, int) { return nullptr; }
inline int Mat_VarWrite(mat_t*, matvar_t*, int) { return 0; }
inline void Mat_VarFree(matvar_t*) {}
inline int Mat_Close(mat_t*) { return 0; }
#endif
