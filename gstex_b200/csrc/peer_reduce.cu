// peer_reduce.cu -- row (e) of SURVEY.md section 8: the one exchange step of the view-sharded path, the sum of the
// per-rank gradient buckets at the end of a multi-view step, as ONE kernel over NVLink peer memory instead of an NCCL
// all-reduce.  The reference has no multi-GPU path of its own (nerfstudio wraps the model in torch DDP, which
// all-reduces the gradients bucket by bucket through NCCL, NS/pipelines/base_pipeline.py:282); this replaces the `dist.all_reduce` over the flat
// bucket of gstex_b200/multiview.py.  sm_100a, NVLink 5 / NVSwitch.
//
// Every rank maps the buckets of all ranks (torch symmetric memory is the plumbing that exchanges the handles and
// gives the barriers before and after).  Rank r owns the r-th 1/W slice of the bucket: it loads that slice from all
// W ranks (W - 1 of them over NVLink), adds them in rank order 0..W-1 and stores the sum into the slice of all W
// ranks.  Reduce-scatter and all-gather in one pass: every float crosses NVLink (W-1)/W times in and (W-1)/W times
// out per GPU, both directions busy at the same time, no intermediate buffer, and every rank ends with the same bits
// (one rank computes each element).  With an NVSwitch multicast mapping of the bucket the same slice is instead read
// with multimem.ld_reduce (the switch adds the W copies) and written with multimem.st (the switch replicates it):
// 1/W of the load and store instructions, the same bytes on the links.
//
// Algorithmic bytes: 2 * 4 * count * (W-1)/W over NVLink per GPU (half in, half out), + 2 * 4 * count / W local.
#include "common.cuh"

namespace gstex {

constexpr int kMaxPeers = 8;  // one NVSwitch domain of a B200 node
constexpr int kMulticastCtas = 16;  // default grid of the multimem variant

struct PeerBufs {
    float4* p[kMaxPeers];
};

__device__ __forceinline__ float4 add4(float4 a, float4 b) {
    return make_float4(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y), __fadd_rn(a.z, b.z), __fadd_rn(a.w, b.w));
}

// peer loads: data written by other GPUs' kernels before the barrier; nothing of it is in this SM's L1 (a new launch),
// streaming hint because every element is read once.  NVLink round trips are a few microseconds: each thread keeps
// U * W 16-byte loads in flight (U independent elements, all W copies of each) before it adds and stores anything.
template <int W, int U>
__global__ void __launch_bounds__(512)
peer_allreduce_kernel(const PeerBufs bufs, int rank, long long count4) {
    const long long per = (count4 + W - 1) / W;
    const long long begin = per * rank;
    const long long end = begin + per < count4 ? begin + per : count4;
    const long long stride = (long long)gridDim.x * blockDim.x;
    float4* mine[W];
#pragma unroll
    for (int p = 0; p < W; ++p) mine[p] = bufs.p[p];
    for (long long i = begin + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < end; i += stride * U) {
        float4 v[U][W];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long j = i + u * stride;
#pragma unroll
            for (int p = 0; p < W; ++p)
                if (j < end) v[u][p] = __ldcs(mine[p] + j);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long j = i + u * stride;
            if (j < end) {
                float4 acc = v[u][0];
#pragma unroll
                for (int p = 1; p < W; ++p) acc = add4(acc, v[u][p]);  // rank order: the same bits whoever computes
#pragma unroll
                for (int p = 0; p < W; ++p) __stcs(mine[p] + j, acc);
            }
        }
    }
}

// NVSwitch multicast variant: `mc` is the multicast mapping of the same bucket on all W ranks.
template <int U>
__global__ void __launch_bounds__(512)
multimem_allreduce_kernel(float4* __restrict__ mc, int rank, int world, long long count4) {
    const long long per = (count4 + world - 1) / world;
    const long long begin = per * rank;
    const long long end = begin + per < count4 ? begin + per : count4;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = begin + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < end; i += stride * U) {
        float4 a[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long j = i + u * stride;
            if (j < end)
                asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                             : "=f"(a[u].x), "=f"(a[u].y), "=f"(a[u].z), "=f"(a[u].w)
                             : "l"(mc + j)
                             : "memory");
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long j = i + u * stride;
            if (j < end)
                asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc + j), "f"(a[u].x),
                             "f"(a[u].y), "f"(a[u].z), "f"(a[u].w)
                             : "memory");
        }
    }
}

}  // namespace gstex

using namespace gstex;

// one wave of resident CTAs (what the occupancy calculator says fits), fewer when the slice is short
template <typename K>
static int grid_for(K kernel, long long per, int unroll, int max_ctas) {
    int resident = 1;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&resident, kernel, 512, 0) != cudaSuccess || resident < 1) resident = 1;
    long long ctas = (long long)sm_count() * resident;
    if (max_ctas > 0 && ctas > max_ctas) ctas = max_ctas;
    const long long need = (per + 512ll * unroll - 1) / (512ll * unroll);
    if (ctas > need) ctas = need;
    return ctas < 1 ? 1 : (int)ctas;
}

extern "C" int gstex_peer_allreduce_max_world(void) { return kMaxPeers; }

extern "C" int gstex_peer_allreduce_f32(int rank, int world, const void* const* peer_bufs, long long count,
                                        void* multicast_buf, int max_ctas, void* stream) {
    if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world || count < 0) return GSTEX_E_BADARG;
    if (world == 1 || count == 0) return GSTEX_OK;
    if (count % 4) return GSTEX_E_BADARG;
    const long long count4 = count / 4;
    const long long per = (count4 + world - 1) / world;
    cudaStream_t st = (cudaStream_t)stream;
    if (multicast_buf != nullptr) {
        if ((uintptr_t)multicast_buf & 15) return GSTEX_E_BADARG;
        // Few CTAs with many requests each: measured on 8 x B200 (profiles/peer_allreduce_r2.txt), 16 CTAs x 512 threads
        // x 8 ld_reduce in flight take 0.44 ms for 208 MB at 8 ranks, a CTA on every SM 0.54 ms (the switch, not the SMs,
        // is the bottleneck, and more requesters only lengthen its queues).
        multimem_allreduce_kernel<8><<<grid_for(multimem_allreduce_kernel<8>, per, 8, max_ctas > 0 ? max_ctas : kMulticastCtas),
                                       512, 0, st>>>(reinterpret_cast<float4*>(multicast_buf), rank, world, count4);
        GSTEX_LAUNCH_CHECK();
        return GSTEX_OK;
    }
    if (!peer_bufs) return GSTEX_E_BADARG;
    PeerBufs bufs;
    for (int p = 0; p < kMaxPeers; ++p) {
        bufs.p[p] = p < world ? reinterpret_cast<float4*>(const_cast<void*>(peer_bufs[p])) : nullptr;
        if (p < world && (!peer_bufs[p] || ((uintptr_t)peer_bufs[p] & 15))) return GSTEX_E_BADARG;
    }
#define GSTEX_PEER_CASE(W, U)                                                                                       \
    case W:                                                                                                         \
        peer_allreduce_kernel<W, U><<<grid_for(peer_allreduce_kernel<W, U>, per, U, max_ctas), 512, 0, st>>>(bufs, rank, \
                                                                                                           count4); \
        break;
    switch (world) {
        GSTEX_PEER_CASE(2, 8)
        GSTEX_PEER_CASE(3, 5)
        GSTEX_PEER_CASE(4, 4)
        GSTEX_PEER_CASE(5, 3)
        GSTEX_PEER_CASE(6, 2)
        GSTEX_PEER_CASE(7, 2)
        GSTEX_PEER_CASE(8, 2)
        default: return GSTEX_E_UNSUPPORTED;
    }
#undef GSTEX_PEER_CASE
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}
