/* TEST INFRASTRUCTURE ONLY: gr::top_block for the GNSSBlockInterface signatures (connect / disconnect take one); the *_B200
 * adapters have nothing to connect internally, so it only needs to exist. */
#ifndef GB_ORACLE_SHIM_GR_TOP_BLOCK_H
#define GB_ORACLE_SHIM_GR_TOP_BLOCK_H
#include <gnuradio/block.h>
#include <memory>
#include <string>
namespace gr
{
class top_block
{
public:
    explicit top_block(const std::string& name = "") : d_name(name) {}
    void connect(basic_block_sptr, int, basic_block_sptr, int) {}
    void disconnect(basic_block_sptr, int, basic_block_sptr, int) {}

private:
    std::string d_name;
};
typedef std::shared_ptr<top_block> top_block_sptr;
inline top_block_sptr make_top_block(const std::string& name) { return top_block_sptr(new top_block(name)); }
}  // namespace gr
#endif
