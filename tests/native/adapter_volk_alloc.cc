// TEST INFRASTRUCTURE: the three volk_gnsssdr allocation helpers the reference's PRN generators call (oracle/shim declares them).
#include <cstdlib>
#include <cstring>
extern "C" {
size_t volk_gnsssdr_get_alignment(void) { return 64; }
void* volk_gnsssdr_malloc(size_t size, size_t alignment)
{
    const size_t a = alignment < 64 ? 64 : alignment;
    void* p = aligned_alloc(a, (size + a - 1) / a * a);
    if (p) memset(p, 0, (size + a - 1) / a * a);
    return p;
}
void volk_gnsssdr_free(void* p) { free(p); }
}
