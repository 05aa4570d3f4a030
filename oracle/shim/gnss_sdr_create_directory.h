/* TEST INFRASTRUCTURE ONLY (oracle/_ref build): shadows src/algorithms/libs/gnss_sdr_create_directory.h (Boost.Filesystem). */
#ifndef GB_ORACLE_SHIM_CREATE_DIRECTORY_H
#define GB_ORACLE_SHIM_CREATE_DIRECTORY_H
#include <string>
#include <sys/stat.h>
inline bool gnss_sdr_create_directory(const std::string& foldername)
{
    struct stat st;
    if (stat(foldername.c_str(), &st) == 0) return S_ISDIR(st.st_mode);
    return mkdir(foldername.c_str(), 0755) == 0;
}
#endif
