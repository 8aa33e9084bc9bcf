// K5 placeholder while the tiled ranking kernel is written (replaced in the next commit).
#include "kge_internal.h"
namespace kge {
size_t eval_workspace_bytes(const Params&, int64_t, int64_t) { return 256; }
cudaError_t launch_eval_rank(const Params&, int, const float*, const float*, const float*, const int64_t*, int64_t,
                             int64_t, int64_t, const int64_t*, const int64_t*, float*, int64_t*, int64_t*, int, void*,
                             size_t, cudaStream_t) {
    return cudaErrorNotSupported;
}
}  // namespace kge
