/* TEST INFRASTRUCTURE ONLY: gr::blocks::float_to_complex as far as the reference's adapters use it (make(), handed to top_block::connect). */
#ifndef GB_ORACLE_SHIM_GR_BLOCKS_FLOAT_TO_COMPLEX_H
#define GB_ORACLE_SHIM_GR_BLOCKS_FLOAT_TO_COMPLEX_H
#include <gnuradio/block.h>
#include <memory>
namespace gr
{
namespace blocks
{
class float_to_complex : public basic_block
{
public:
    typedef std::shared_ptr<float_to_complex> sptr;
    static sptr make(size_t vlen = 1)
    {
        (void)vlen;
        return sptr(new float_to_complex());
    }
    long unique_id() const { return 0; }
};
}  // namespace blocks
}  // namespace gr
#endif
