// tcgen05 / TMEM / TMA (tensor-map) building blocks shared by the kernels that run 3xTF32 tiles on the 5th-gen tensor
// cores (sm_100a): kge_eval_mma.cu (K5-T, ranking) and kge_transparse.cu (the per-relation projections).
//
//   tile        128 (TMEM lanes) x 256 (TMEM columns) fp32 accumulator, K in stages of 32 fp32 = one 128-byte swizzle span
//   operands    K-major [rows, Kpad] fp32 workspaces holding the hi / lo halves of a 3xTF32 split, read by
//               cp.async.bulk.tensor.2d with SWIZZLE_128B; descriptors below match that layout
//   precision   x = hi + lo with hi = tf32(x), lo = tf32(x - hi); a tile accumulates lo*hi + hi*lo + hi*hi
#pragma once
#include <cuda.h>

#include <algorithm>

#include "kge_internal.h"
#include "kge_ptx.cuh"

namespace kge {
namespace mma {

constexpr int BM = 128;      // queries per tile (TMEM lanes)
constexpr int BN = 256;      // entities per tile (TMEM columns per accumulator)
constexpr int BK = 32;       // fp32 coordinates per stage row = 128 bytes = one swizzle span
constexpr int UMMA_K = 8;    // K of one tcgen05.mma.kind::tf32
constexpr int STAGES = 2;
constexpr uint32_t A_BYTES = BM * BK * 4;                    // 16 KB
constexpr uint32_t B_BYTES = BN * BK * 4;                    // 32 KB
constexpr uint32_t STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;  // hi + lo of both = 96 KB
constexpr uint32_t SMEM_BYTES = 1024 + STAGES * STAGE_BYTES + 128;
constexpr uint32_t TMEM_COLS = 512;

// ---- PTX wrappers ------------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
        "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* tm) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tm)) : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}

// Shared-memory matrix descriptor, K-major operand, 128-byte swizzle: rows of 128 B, 8-row groups 1024 B apart.
// bits [0,14) start >> 4 | [16,30) LBO >> 4 (unused with swizzle, 1) | [32,46) SBO >> 4 | [46,48) version 1 |
// [61,64) layout 2 = SWIZZLE_128B.  Advancing K by 8 tf32 = 32 B adds 2 to the start field.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) |
           ((uint64_t)2 << 61);
}
// Instruction descriptor kind::tf32: D fp32 (bits [4,6) = 1), A/B tf32 (bits [7,10) = [10,13) = 2), both K-major,
// N >> 3 at bit 17, M >> 4 at bit 24.
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);

__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(kIdesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// ---- CTA pair (cta_group::2): two SMs of a cluster work on ONE 256-row tile -----------------------------------------
// Each CTA stages its 128 rows of A and HALF of the B tile; the leader's tcgen05.mma.cta_group::2 reads both CTAs' shared
// memory and writes 128 accumulator rows into each CTA's TMEM.  Per SM and k-stage that is 64 KB instead of 96 KB.
constexpr int STAGES2 = 3;
constexpr uint32_t BH_BYTES = (BN / 2) * BK * 4;               // this CTA's half of the B tile (hi or lo): 16 KB
constexpr uint32_t STAGE2_BYTES = 2 * A_BYTES + 2 * BH_BYTES;  // 64 KB
constexpr uint32_t SMEM2_BYTES = 1024 + STAGES2 * STAGE2_BYTES + 128;
constexpr uint32_t kIdescPair = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((2 * BM) >> 4) << 24);

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t cluster_id_x() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t cluster_count_x() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%nclusterid.x;" : "=r"(r));
    return r;
}
// every thread of every CTA of the cluster
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `saddr` (a shared::cta address of this CTA) in the CTA of rank `rank`
__device__ __forceinline__ uint32_t map_to_rank(uint32_t saddr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_bar) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar) : "memory");
}
// TMA load of a CTA of the pair: lands in THIS CTA's shared memory, signals the barrier at `cluster_bar` (the leader's)
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* tm, uint32_t cluster_bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
        "l"(reinterpret_cast<uint64_t>(tm)), "r"(cluster_bar), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t dst_smem, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void umma_tf32_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(kIdescPair), "r"(accumulate)
        : "memory");
}
// commit of the pair's MMAs: arrives on the barrier at this offset in both CTAs
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
                 "h"((uint16_t)3)
                 : "memory");
}

// 32 lanes x 32 consecutive columns -> 32 registers per thread (thread = lane = accumulator row)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n\t"
        "tcgen05.wait::ld.sync.aligned;"  // same asm statement: no consumer can be scheduled before the wait
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// ---- operand split ------------------------------------------------------------------------------
// round-to-nearest-even onto the tf32 grid (10 explicit mantissa bits); inputs are finite table values
__device__ __forceinline__ float tf32_rne(float x) {
    uint32_t u = __float_as_uint(x);
    u += 0xFFFu + ((u >> 13) & 1u);
    return __uint_as_float(u & 0xFFFFE000u);
}
__device__ __forceinline__ void split_tf32(float x, float& hi, float& lo) {
    hi = tf32_rne(x);
    lo = tf32_rne(x - hi);  // x - hi is exact in fp32
}

// ---- tensor maps ---------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static inline EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// [rows, Kpad] fp32 row-major, box = box_rows x 32 floats (128 B), 128-byte swizzle, out-of-range rows read as 0
static inline bool make_tmap(CUtensorMap* tm, const float* ptr, long long rows, int Kpad, int box_rows) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return false;
    const cuuint64_t dims[2] = {(cuuint64_t)Kpad, (cuuint64_t)std::max<long long>(rows, 1)};
    const cuuint64_t strides[1] = {(cuuint64_t)Kpad * sizeof(float)};
    const cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    return fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(ptr), dims, strides, box, estr,
              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// CTA pairs that can be resident at once for `kernel` (1 CTA per SM at `smem` bytes): the persistent pair kernels size
// their grid with it -- a GPC with an odd number of free SMs leaves one idle, and a cluster that starts in a second wave
// would stretch the tail of a persistent loop.
template <typename K>
static inline int max_active_pairs(K kernel, int threads, size_t smem) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(2u * (unsigned)sm_count());
    cfg.blockDim = dim3((unsigned)threads);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, kernel, &cfg) != cudaSuccess || n <= 0) {
        cudaGetLastError();
        n = sm_count() / 2;
    }
    return n;
}

}  // namespace mma
}  // namespace kge
