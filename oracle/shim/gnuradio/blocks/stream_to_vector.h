/* TEST INFRASTRUCTURE ONLY: gr::blocks::stream_to_vector as far as the reference's adapters use it. */
#ifndef GB_ORACLE_SHIM_GR_BLOCKS_STREAM_TO_VECTOR_H
#define GB_ORACLE_SHIM_GR_BLOCKS_STREAM_TO_VECTOR_H
#include <gnuradio/block.h>
#include <memory>
namespace gr
{
namespace blocks
{
class stream_to_vector : public basic_block
{
public:
    typedef std::shared_ptr<stream_to_vector> sptr;
    static sptr make(size_t itemsize, size_t nitems_per_block)
    {
        (void)itemsize;
        (void)nitems_per_block;
        return sptr(new stream_to_vector());
    }
    long unique_id() const { return 0; }
};
}  // namespace blocks
}  // namespace gr
#endif
