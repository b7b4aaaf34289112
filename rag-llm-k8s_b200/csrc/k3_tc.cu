// k3_tc.cu -- placeholder until the tcgen05 kernel lands: reports "not supported" so the
// dispatch uses the exact scan kernel (K2) for every nq.
#include "k3_api.h"
namespace b2k {
struct K3State { int unused; };
bool k3_supported(int, int) { return false; }
size_t k3_scratch_bytes(int, int, int64_t, int) { return 0; }
int k3_create(K3State** out, int, int) { *out = new K3State(); return B2K_OK; }
void k3_destroy(K3State* s) { delete s; }
void k3_invalidate_maps(K3State*) {}
int k3_search(K3State*, const float*, const float*, const float*, int64_t, int64_t, int, const float*, int64_t, int,
              float*, int64_t*, unsigned char*, cudaStream_t) {
    set_error("tensor-core path not built");
    return B2K_ECUDA;
}
}  // namespace b2k
