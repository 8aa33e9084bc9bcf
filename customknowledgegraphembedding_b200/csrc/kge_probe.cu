// Measurement probe: L2 read bandwidth of the device this library runs on (bench.py's `roofline.l2`).
//
// The two dominant kernels of the train step stream their operands out of L2, not HBM (the 116 MB entity table and the
// staged planes are gathered once per USE: 2.1 GB per kernel and step against 0.12 GB of table), so the roofline that
// binds them is the L2 slices' throughput.  No profiler is needed at bench time: a persistent grid re-reads a buffer
// that fits L2 (default 32 MB) `iters` times with 128-bit loads, every SM touching every line (the access pattern of a
// gather from a shared table), and the caller times it with CUDA events.  No reference counterpart (housekeeping).
#include "kge_internal.h"

namespace kge {

__global__ void __launch_bounds__(512)
l2_read_kernel(const float4* __restrict__ buf, long long nvec, int iters, float* sink) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (int it = 0; it < iters; ++it) {
        // every pass starts at a different offset per CTA so that all SMs do not walk the slices in lock step
        long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x + (long long)it * 977 * blockDim.x) % nvec;
        for (long long n = 0; n < nvec; n += 4 * stride) {
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                long long j = i + u * stride;
                if (j >= nvec) j -= nvec;
                if (j >= nvec) j %= nvec;
                asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];"  // .cg: L2 only, never L1
                             : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w)
                             : "l"(buf + j));
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w;
            }
            i += 4 * stride;
            if (i >= nvec) i %= nvec;
        }
    }
    if (acc.x + acc.y + acc.z + acc.w == 123.456f) *sink = acc.x;  // keeps the loads alive
}

}  // namespace kge

using namespace kge;

extern "C" int kge_probe_l2_read(const void* buf, size_t bytes, int iters, float* sink, int64_t* bytes_read, void* stream) {
    if (!buf || !sink || bytes < 4096 || iters < 1 || (reinterpret_cast<uintptr_t>(buf) & 15))
        return fail(KGE_ERR_INVALID_ARGUMENT, "kge_probe_l2_read: bad arguments");
    const long long nvec = (long long)(bytes / 16);
    const int grid = 2 * sm_count(), block = 512;
    const long long stride = (long long)grid * block;
    const long long per_pass = ((nvec + 4 * stride - 1) / (4 * stride)) * 4 * stride;  // loads issued per pass
    l2_read_kernel<<<grid, block, 0, (cudaStream_t)stream>>>((const float4*)buf, nvec, iters, sink);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail_cuda(e, "kge_probe_l2_read");
    if (bytes_read) *bytes_read = per_pass * 16 * (long long)iters;
    return KGE_OK;
}
