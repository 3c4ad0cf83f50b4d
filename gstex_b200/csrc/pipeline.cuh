// pipeline.cuh -- mbarrier + TMA bulk-copy (cp.async.bulk, SASS UBLKCP) helpers of the compositing kernels.
// The forward (raster_fwd.cu) gives every warp a private double-buffered record stage: each lane issues one
// 128-byte bulk copy of a surviving splat's record, completion is counted in bytes on the stage's mbarrier, and the
// warp waits on it alone -- no producer warp, no block-wide barrier.  The recompute backward (raster_bwd.cu) keeps
// the CTA-wide ring: one producer warp stages 32-record batches into kStages slots ("full" mbarrier counted in
// bytes), the pixel warps consume them and release a slot on its "empty" mbarrier.  sm_100a.
#pragma once
#include "raster_common.cuh"

namespace gstex {

constexpr int kStageRecs = 32;  // records per stage = one per producer lane
constexpr int kStages = 8;      // ring depth (8 x 4 KB of records)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}
// global -> shared bulk copy, `bytes` a multiple of 16, both addresses 16-byte aligned; completes on `bar`.
__device__ __forceinline__ void bulk_copy_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// Which pixel rows (relative to the tile) warp w of a bw x bw tile covers: thread ids [32w, 32w+32).
__device__ __forceinline__ void warp_rows(int w, int bw, int& r0, int& r1) {
    r0 = (32 * w) / bw;
    r1 = min((32 * w + 31) / bw, bw - 1);
}

// Does the conservative support box of a splat (pixel coordinates of pixel CENTRES that may see
// alpha >= 1/255) touch the pixel centres [x0, x1] x [y0, y1]?
__device__ __forceinline__ bool box_overlaps(const float4 box, float x0, float x1, float y0, float y1) {
    return box.x <= x1 && box.z >= x0 && box.y <= y1 && box.w >= y0;
}

}  // namespace gstex
