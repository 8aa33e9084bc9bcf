// Fused train step (host-side composition of the kernels) -- see kge_train.cu.
#pragma once
#include "kge_internal.h"

namespace kge {
size_t train_step_workspace_bytes(const Params& P, int64_t B, int64_t N, bool host_batch);
}
