// tcgen05 (UMMA) 3xTF32 engine — placeholder until the kernels land: reports "unsupported" so the
// C ABI routes to the fp32 FFMA paths.
#include "umma_engine.cuh"

namespace bsa {

bool umma_ce_supported(int, int, long long) { return false; }
size_t umma_ce_workspace_bytes(int, int, long long) { return 0; }
int umma_ce_fwd(int, int, long long, const float*, const float*, const long long*, float*, float*, void*, size_t, cudaStream_t) {
    set_error("tcgen05 cross-entropy path not built");
    return -1;
}
int umma_ce_bwd(int, int, long long, const float*, const float*, const long long*, const float*, const float*, float*, float*, void*, size_t, cudaStream_t) {
    set_error("tcgen05 cross-entropy path not built");
    return -1;
}
bool umma_topk_supported(int, int, int, long long) { return false; }
size_t umma_topk_workspace_bytes(int, int, int, long long) { return 0; }
int umma_score_topk(int, int, int, long long, long long, const float*, const float*, const long long*, const long long*, float*, long long*, void*, size_t, cudaStream_t) {
    set_error("tcgen05 scoring path not built");
    return -1;
}

}  // namespace bsa
