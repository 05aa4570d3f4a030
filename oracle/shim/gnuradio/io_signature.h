/* TEST INFRASTRUCTURE ONLY (oracle/_ref build). */
#ifndef GB_ORACLE_SHIM_GR_IO_SIGNATURE_H
#define GB_ORACLE_SHIM_GR_IO_SIGNATURE_H
#include <cstddef>
#include <memory>
namespace gr
{
struct io_signature
{
    int min_streams, max_streams;
    std::size_t item_size;
    typedef std::shared_ptr<io_signature> sptr;
    static sptr make(int mn, int mx, std::size_t sz) { return sptr(new io_signature{mn, mx, sz}); }
};
}  // namespace gr
#endif
