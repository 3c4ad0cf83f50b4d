// scan_emit.cu -- stage (b) integer kernels: inclusive prefix sum of the per-splat tile counts, emission
// of the (tile<<32 | depth bits) keys, and extraction of the per-tile ranges from the sorted keys.
// All results are integers and bit-exact with the reference by construction.  sm_100a.
#include <atomic>

#include "common.cuh"

namespace gstex {

constexpr int kScanThreads = 1024;
constexpr int kScanItems = 4;
constexpr int kScanTile = kScanThreads * kScanItems;

// Block-wide inclusive scan of one int per thread (1024 threads): warp shuffles + one smem hop.
__device__ __forceinline__ int block_inclusive_scan(int v, int* warp_sums /*32 ints smem*/, int& block_total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    if (lane == 31) warp_sums[warp] = v;
    __syncthreads();
    if (warp == 0) {
        int s = warp_sums[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, s, o);
            if (lane >= o) s += t;
        }
        warp_sums[lane] = s;
    }
    __syncthreads();
    block_total = warp_sums[31];
    if (warp > 0) v += warp_sums[warp - 1];
    return v;
}

// Pass 1: tile-local inclusive scan -> out, tile total -> block_sums[blockIdx.x].
__global__ void __launch_bounds__(kScanThreads)
scan_local_kernel(const int32_t* __restrict__ in, int n, int32_t* __restrict__ out, int32_t* __restrict__ block_sums) {
    __shared__ int warp_sums[32];
    const int base = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
    int v[kScanItems];
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) v[k] = (base + k < n) ? in[base + k] : 0;
#pragma unroll
    for (int k = 1; k < kScanItems; ++k) v[k] += v[k - 1];
    int total;
    const int incl = block_inclusive_scan(v[kScanItems - 1], warp_sums, total);
    const int excl = incl - v[kScanItems - 1];
#pragma unroll
    for (int k = 0; k < kScanItems; ++k)
        if (base + k < n) out[base + k] = v[k] + excl;
    if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

// Pass 2: every tile adds the sum of the preceding tiles' totals (each block reduces block_sums[0..b)
// itself: at most a few thousand ints from L2, cheaper than a third launch).  The last tile publishes
// the grand total.
__global__ void __launch_bounds__(kScanThreads)
scan_add_kernel(int n, int32_t* __restrict__ out, const int32_t* __restrict__ block_sums, int32_t* __restrict__ total) {
    __shared__ int warp_sums[32];
    __shared__ int s_offset;
    int acc = 0;
    for (int j = threadIdx.x; j < (int)blockIdx.x; j += kScanThreads) acc += block_sums[j];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        int s = warp_sums[threadIdx.x];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (threadIdx.x == 0) s_offset = s;
    }
    __syncthreads();
    const int offset = s_offset;
    const int base = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
    if (offset != 0) {
#pragma unroll
        for (int k = 0; k < kScanItems; ++k)
            if (base + k < n) out[base + k] += offset;
    }
    if (total != nullptr && blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) *total = offset + block_sums[blockIdx.x];
}

// ref: forward.cu:13-71 map_gaussian_to_intersects (wrapped = false) + helpers.cuh:37-51,75-92.
// One thread per splat.  Instead of zero-filling both outputs up front (bindings.cu:99-102) every
// thread zero-fills exactly the reserved slots it does not write: the early-return splats' slots (the
// "zero-key quirk") and any slack between the reserved count and the emitted count.
__global__ void __launch_bounds__(kThreads1D)
map_gaussian_to_intersects_kernel(int n, int num_intersects, const float2* __restrict__ centers,
                                  const float2* __restrict__ extents, const float* __restrict__ depths,
                                  const int32_t* __restrict__ cum_tiles_hit, int tiles_x, int tiles_y, float fbw,
                                  int64_t* __restrict__ isect_ids, int32_t* __restrict__ gaussian_ids) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n) return;
    int cur = (idx == 0) ? 0 : cum_tiles_hit[idx - 1];
    const int end = min(cum_tiles_hit[idx], num_intersects);
    const float2 c = centers[idx], e = extents[idx];
    if (!((double)e.x <= 1e-4 && (double)e.y <= 1e-4)) {
        const float tcx = __fdiv_rn(c.x, fbw), tcy = __fdiv_rn(c.y, fbw);
        const float tex = __fdiv_rn(e.x, fbw), tey = __fdiv_rn(e.y, fbw);
        const int min_x = min(max(0, __float2int_rz(__fsub_rn(tcx, tex))), tiles_x);
        const int max_x = min(max(0, __float2int_rz(__fadd_rn(__fadd_rn(tcx, tex), 1.0f))), tiles_x);
        const int min_y = min(max(0, __float2int_rz(__fsub_rn(tcy, tey))), tiles_y);
        const int max_y = min(max(0, __float2int_rz(__fadd_rn(__fadd_rn(tcy, tey), 1.0f))), tiles_y);
        const int64_t depth_id = (int64_t)__float_as_int(depths[idx]);  // sign-extended, forward.cu:47
        for (int i = min_y; i < max_y && cur < end; ++i) {
            for (int j = min_x; j < max_x && cur < end; ++j) {
                const int64_t tile_id = (int64_t)(i * tiles_x + j);
                isect_ids[cur] = (tile_id << 32) | depth_id;
                gaussian_ids[cur] = idx;
                ++cur;
            }
        }
    }
    for (; cur < end; ++cur) {
        isect_ids[cur] = 0;
        gaussian_ids[cur] = 0;
    }
}

// ref: forward.cu:76-98 get_tile_bin_edges
__global__ void __launch_bounds__(kThreads1D)
tile_bin_edges_kernel(int num_intersects, const int64_t* __restrict__ keys, int num_tiles, int2* __restrict__ tile_bins) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= num_intersects) return;
    const int cur = (int)(keys[idx] >> 32);
    if ((unsigned)cur >= (unsigned)num_tiles) return;  // cannot happen for keys made by the emit kernel
    if (idx == 0) tile_bins[cur].x = 0;
    if (idx == num_intersects - 1) tile_bins[cur].y = num_intersects;
    if (idx == 0) return;
    const int prev = (int)(keys[idx - 1] >> 32);
    if (prev != cur) {
        if ((unsigned)prev < (unsigned)num_tiles) tile_bins[prev].y = idx;
        tile_bins[cur].x = idx;
    }
}

}  // namespace gstex

using namespace gstex;

extern "C" size_t gstex_cumsum_workspace_bytes(int n) {
    if (n <= 0) return 16;
    return (size_t)div_up(n, kScanTile) * sizeof(int32_t) + 16;
}

extern "C" int gstex_cumsum_i32(const int32_t* in, int n, int32_t* out, int32_t* total, void* workspace,
                                size_t workspace_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (n < 0) return GSTEX_E_BADARG;
    if (n == 0) {
        if (total) GSTEX_CUDA_TRY(cudaMemsetAsync(total, 0, sizeof(int32_t), st));
        return GSTEX_OK;
    }
    if (in == nullptr || out == nullptr || workspace == nullptr) return GSTEX_E_BADARG;
    if (workspace_bytes < gstex_cumsum_workspace_bytes(n)) return GSTEX_E_WORKSPACE;
    const int blocks = div_up(n, kScanTile);
    int32_t* block_sums = reinterpret_cast<int32_t*>(workspace);
    scan_local_kernel<<<blocks, kScanThreads, 0, st>>>(in, n, out, block_sums);
    GSTEX_LAUNCH_CHECK();
    scan_add_kernel<<<blocks, kScanThreads, 0, st>>>(n, out, block_sums, total);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

// ---- gstex_cumsum_i32_noalloc: the same scan with its scratch in a static device array ---------------------------
// For callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc): no workspace
// argument, no cudaMalloc / cudaFree / allocating library call.  No counterpart in the reference.
// The per-tile totals live in a file-scope __device__ array that every GPU loading this
// library holds until exit (64 KB).  In HOST code the array's name is only a stand-in, not a device address: the
// address a kernel may use is fetched once per device, with that device current, by cudaGetSymbolAddress.
constexpr int kNoAllocMaxTiles = 16384;                         // 64 KB of scratch
constexpr int kNoAllocMaxN = kNoAllocMaxTiles * kScanTile;      // 67 108 864 elements
static __device__ __align__(256) int32_t g_noalloc_block_sums[kNoAllocMaxTiles];

constexpr int kNoAllocMaxDevices = 64;
static std::atomic<int32_t*> g_noalloc_ptr[kNoAllocMaxDevices];

// Device address of the scratch array on the CURRENT device (looked up once per device, then cached).
static int noalloc_scratch(int32_t** out) {
    int dev = -1;
    GSTEX_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kNoAllocMaxDevices) return GSTEX_E_BADARG;
    int32_t* p = g_noalloc_ptr[dev].load(std::memory_order_acquire);
    if (p == nullptr) {
        void* sym = nullptr;
        GSTEX_CUDA_TRY(cudaGetSymbolAddress(&sym, g_noalloc_block_sums));
        if (sym == nullptr) return (int)cudaErrorInvalidSymbol;
        p = reinterpret_cast<int32_t*>(sym);
        g_noalloc_ptr[dev].store(p, std::memory_order_release);  // racing threads store the same address
    }
    *out = p;
    return GSTEX_OK;
}

extern "C" int gstex_cumsum_noalloc_max_n(void) { return kNoAllocMaxN; }

extern "C" int gstex_cumsum_i32_noalloc(const int32_t* in, int n, int32_t* out, int32_t* total, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (n < 0) return GSTEX_E_BADARG;
    if (n > kNoAllocMaxN) return GSTEX_E_TOO_LARGE;
    if (n == 0) {
        if (total) GSTEX_CUDA_TRY(cudaMemsetAsync(total, 0, sizeof(int32_t), st));
        return GSTEX_OK;
    }
    if (in == nullptr || out == nullptr) return GSTEX_E_BADARG;
    int32_t* block_sums = nullptr;
    const int rc = noalloc_scratch(&block_sums);
    if (rc != GSTEX_OK) return rc;
    // Nothing has to start at zero: scan_local_kernel writes the slot of every tile this call uses before
    // scan_add_kernel (next on the stream) reads it; slots beyond `blocks` keep an earlier call's values and are
    // never read.
    const int blocks = div_up(n, kScanTile);
    scan_local_kernel<<<blocks, kScanThreads, 0, st>>>(in, n, out, block_sums);
    GSTEX_LAUNCH_CHECK();
    scan_add_kernel<<<blocks, kScanThreads, 0, st>>>(n, out, block_sums, total);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

extern "C" int gstex_map_gaussian_to_intersects(int n, int num_intersects, const float* centers, const float* extents,
                                                const float* depths, const int32_t* cum_tiles_hit, int tiles_x,
                                                int tiles_y, int block_width, int64_t* isect_ids,
                                                int32_t* gaussian_ids, void* stream) {
    if (n < 0 || num_intersects < 0 || block_width < 2 || block_width > 16 || tiles_x <= 0 || tiles_y <= 0)
        return GSTEX_E_BADARG;
    if (n == 0 || num_intersects == 0) return GSTEX_OK;
    if (!centers || !extents || !depths || !cum_tiles_hit || !isect_ids || !gaussian_ids) return GSTEX_E_BADARG;
    map_gaussian_to_intersects_kernel<<<div_up(n, kThreads1D), kThreads1D, 0, (cudaStream_t)stream>>>(
        n, num_intersects, reinterpret_cast<const float2*>(centers), reinterpret_cast<const float2*>(extents), depths,
        cum_tiles_hit, tiles_x, tiles_y, (float)block_width, isect_ids, gaussian_ids);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

extern "C" int gstex_get_tile_bin_edges(int num_intersects, const int64_t* isect_ids_sorted, int num_tiles,
                                        int32_t* tile_bins, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (num_intersects < 0 || num_tiles <= 0 || tile_bins == nullptr) return GSTEX_E_BADARG;
    GSTEX_CUDA_TRY(cudaMemsetAsync(tile_bins, 0, sizeof(int32_t) * 2 * (size_t)num_tiles, st));
    if (num_intersects == 0) return GSTEX_OK;
    if (isect_ids_sorted == nullptr) return GSTEX_E_BADARG;
    tile_bin_edges_kernel<<<div_up(num_intersects, kThreads1D), kThreads1D, 0, st>>>(
        num_intersects, isect_ids_sorted, num_tiles, reinterpret_cast<int2*>(tile_bins));
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}
