// sort.cu -- stage (b): hand-written stable LSD radix sort of the (tile<<32 | depth bits) int64 keys with
// the int32 splat id as payload.  Replaces torch.sort + torch.gather of the reference
// (gstex_cuda/utils.py:159-160; ATen -> CUB 64-bit keys with a 64-bit index payload, 8 passes, then a
// gather).  Here: 8-bit digits over only the significant key bits (32 depth bits + ceil(log2 tiles),
// i.e. 6 passes instead of 8), 4-byte payload carried through (no gather), stable so that equal keys
// keep emission order = ascending splat index, exactly what the reference produces.  sm_100a.
//
// Per pass: (1) per-block digit histograms over a contiguous chunk of tiles, (2) per-digit row scans of
// the digit-major histogram matrix, (3) stable scatter with warp-level match-any ranking.
// The grid is persistent-sized (a multiple of the SM count), so the histogram matrix stays small.
#include "common.cuh"

namespace gstex {

constexpr int kSortThreads = 256;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kSortItems = 16;
constexpr int kSortTile = kSortThreads * kSortItems;  // 4096 keys per tile
constexpr int kRadix = 256;
// Upper bound of the grid AND a layout constant of the workspace (one histogram row per block), so it does not follow
// the device: 4 CTAs per SM of a 148-SM B200; a part with another SM count runs the same grid.
constexpr int kSortMaxBlocks = 592;

struct SortGeom {
    int n;                // keys
    int num_tiles;        // ceil(n / kSortTile)
    int num_blocks;       // grid size
    int tiles_per_block;  // ceil(num_tiles / num_blocks)
};

__device__ __forceinline__ int digit_of(int64_t key, int shift, int mask) { return (int)((uint64_t)key >> shift) & mask; }

// (1) block_hist[d * num_blocks + b] = number of keys with digit d in block b's chunk
__global__ void __launch_bounds__(kSortThreads)
radix_hist_kernel(const int64_t* __restrict__ keys, SortGeom g, int shift, int mask, int32_t* __restrict__ block_hist) {
    __shared__ int hist[kRadix];
    hist[threadIdx.x] = 0;
    __syncthreads();
    const long long begin = (long long)blockIdx.x * g.tiles_per_block * kSortTile;
    long long end = begin + (long long)g.tiles_per_block * kSortTile;
    if (end > g.n) end = g.n;
    for (long long base = begin; base < end; base += kSortThreads) {
        const long long i = base + threadIdx.x;
        const bool valid = i < end;
        const int d = valid ? digit_of(keys[i], shift, mask) : kRadix;
        // warp-aggregate equal digits (the high passes see long runs of equal tile digits)
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        if (valid && (int)(__ffs(peers) - 1) == (int)(threadIdx.x & 31)) atomicAdd(&hist[d], __popc(peers));
    }
    __syncthreads();
    block_hist[threadIdx.x * g.num_blocks + blockIdx.x] = hist[threadIdx.x];
}

// (2) row scan: block d turns row d of the digit-major histogram matrix (num_blocks counts of digit d, in
// block order) into exclusive offsets within the digit and publishes the digit's total.  256 blocks of
// 256 threads, coalesced; the cross-digit offset is added by the scatter kernel itself (a 256-entry scan).
__global__ void __launch_bounds__(kSortThreads)
radix_scan_rows_kernel(const int32_t* __restrict__ block_hist, int num_blocks, int32_t* __restrict__ block_off,
                       int32_t* __restrict__ digit_total) {
    __shared__ int warp_sums[kSortWarps];
    const int d = blockIdx.x;
    const int32_t* row = block_hist + (size_t)d * num_blocks;
    int32_t* orow = block_off + (size_t)d * num_blocks;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int carry = 0;
    for (int base = 0; base < num_blocks; base += kSortThreads) {
        const int i = base + threadIdx.x;
        const int c = (i < num_blocks) ? row[i] : 0;
        int v = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += t;
        }
        if (lane == 31) warp_sums[warp] = v;
        __syncthreads();
        int wprefix = 0, total = 0;
#pragma unroll
        for (int w = 0; w < kSortWarps; ++w) {
            const int sw = warp_sums[w];
            if (w < warp) wprefix += sw;
            total += sw;
        }
        if (i < num_blocks) orow[i] = carry + wprefix + v - c;
        carry += total;
        __syncthreads();
    }
    if (threadIdx.x == 0) digit_total[d] = carry;
}

// (3) stable scatter.  Within a tile, warp w owns keys [w*512, (w+1)*512) and lane l's k-th key is
// w*512 + k*32 + l, so (warp, k, lane) order == memory order; ranks are assigned in that order.
__global__ void __launch_bounds__(kSortThreads)
radix_scatter_kernel(const int64_t* __restrict__ keys_in, const int32_t* __restrict__ vals_in,
                     int64_t* __restrict__ keys_out, int32_t* __restrict__ vals_out, SortGeom g, int shift, int mask,
                     const int32_t* __restrict__ block_off, const int32_t* __restrict__ digit_total) {
    __shared__ int warp_hist[kSortWarps][kRadix];
    __shared__ int digit_base[kRadix];  // running global offset of each digit for this block
    __shared__ int scan_tmp[kSortWarps];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    {  // exclusive scan of the 256 digit totals (thread d <-> digit d) + this block's offset within the digit
        const int c = digit_total[threadIdx.x];
        int v = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += t;
        }
        if (lane == 31) scan_tmp[warp] = v;
        __syncthreads();
        int wprefix = 0;
#pragma unroll
        for (int w = 0; w < kSortWarps; ++w)
            if (w < warp) wprefix += scan_tmp[w];
        digit_base[threadIdx.x] = wprefix + v - c + block_off[threadIdx.x * g.num_blocks + blockIdx.x];
    }

    const int first_tile = blockIdx.x * g.tiles_per_block;
    const int last_tile = min(first_tile + g.tiles_per_block, g.num_tiles);
    for (int tile = first_tile; tile < last_tile; ++tile) {
#pragma unroll
        for (int w = 0; w < kSortWarps; ++w) warp_hist[w][threadIdx.x] = 0;
        __syncthreads();
        const long long tile_base = (long long)tile * kSortTile + warp * (32 * kSortItems);
        int64_t key[kSortItems];
        int rank[kSortItems];
#pragma unroll
        for (int k = 0; k < kSortItems; ++k) {
            const long long i = tile_base + k * 32 + lane;
            key[k] = (i < g.n) ? keys_in[i] : (int64_t)-1;
        }
#pragma unroll
        for (int k = 0; k < kSortItems; ++k) {
            const long long i = tile_base + k * 32 + lane;
            const bool valid = i < g.n;
            // invalid lanes take a pseudo-digit (kRadix) so they never match a real one
            const int d = valid ? digit_of(key[k], shift, mask) : kRadix;
            const unsigned peers = __match_any_sync(0xffffffffu, d);
            const int leader = __ffs(peers) - 1;
            int prev = 0;
            if (valid && lane == leader) {
                prev = warp_hist[warp][d];
                warp_hist[warp][d] = prev + __popc(peers);
            }
            prev = __shfl_sync(0xffffffffu, prev, leader);
            rank[k] = prev + __popc(peers & lt_mask);
            __syncwarp();
        }
        __syncthreads();
        {  // thread d: exclusive scan of digit d over the warps, then advance the running base
            const int d = threadIdx.x;
            int run = 0;
#pragma unroll
            for (int w = 0; w < kSortWarps; ++w) {
                const int c = warp_hist[w][d];
                warp_hist[w][d] = run + digit_base[d];
                run += c;
            }
            digit_base[d] += run;  // only thread d touches digit_base[d]
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kSortItems; ++k) {
            const long long i = tile_base + k * 32 + lane;
            if (i < g.n) {
                const int d = digit_of(key[k], shift, mask);
                const int pos = warp_hist[warp][d] + rank[k];
                keys_out[pos] = key[k];
                vals_out[pos] = vals_in[i];
            }
        }
        __syncthreads();
    }
}

static SortGeom make_geom(int n) {
    SortGeom g;
    g.n = n;
    g.num_tiles = div_up(n, kSortTile);
    g.num_blocks = g.num_tiles < kSortMaxBlocks ? g.num_tiles : kSortMaxBlocks;
    if (g.num_blocks < 1) g.num_blocks = 1;
    g.tiles_per_block = div_up(g.num_tiles, g.num_blocks);
    g.num_blocks = div_up(g.num_tiles, g.tiles_per_block);  // drop empty trailing blocks
    if (g.num_blocks < 1) g.num_blocks = 1;
    return g;
}

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

}  // namespace gstex

using namespace gstex;

extern "C" size_t gstex_sort_workspace_bytes(int num_intersects) {
    if (num_intersects <= 0) return 256;
    const SortGeom g = make_geom(num_intersects);
    return align256((size_t)num_intersects * 8) + align256((size_t)num_intersects * 4) +
           2 * align256((size_t)kRadix * g.num_blocks * 4) + align256(kRadix * 4) + 256;
}

extern "C" int gstex_sort_isects(int num_intersects, const int64_t* isect_ids, const int32_t* gaussian_ids,
                                 int64_t* isect_ids_sorted, int32_t* gaussian_ids_sorted, int key_bits,
                                 void* workspace, size_t workspace_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    const int n = num_intersects;
    if (n < 0 || key_bits < 1 || key_bits > 64) return GSTEX_E_BADARG;
    if (n == 0) return GSTEX_OK;
    if (!isect_ids || !gaussian_ids || !isect_ids_sorted || !gaussian_ids_sorted || !workspace) return GSTEX_E_BADARG;
    if (workspace_bytes < gstex_sort_workspace_bytes(n)) return GSTEX_E_WORKSPACE;
    const SortGeom g = make_geom(n);
    char* ws = reinterpret_cast<char*>(workspace);
    int64_t* tmp_keys = reinterpret_cast<int64_t*>(ws);
    ws += align256((size_t)n * 8);
    int32_t* tmp_vals = reinterpret_cast<int32_t*>(ws);
    ws += align256((size_t)n * 4);
    int32_t* block_hist = reinterpret_cast<int32_t*>(ws);
    ws += align256((size_t)kRadix * g.num_blocks * 4);
    int32_t* block_off = reinterpret_cast<int32_t*>(ws);
    ws += align256((size_t)kRadix * g.num_blocks * 4);
    int32_t* digit_total = reinterpret_cast<int32_t*>(ws);

    const int passes = (key_bits + 7) / 8;
    const int64_t* src_k = isect_ids;
    const int32_t* src_v = gaussian_ids;
    for (int p = 0; p < passes; ++p) {
        const bool to_out = ((passes - 1 - p) % 2) == 0;
        int64_t* dst_k = to_out ? isect_ids_sorted : tmp_keys;
        int32_t* dst_v = to_out ? gaussian_ids_sorted : tmp_vals;
        const int shift = 8 * p;
        const int bits = key_bits - shift < 8 ? key_bits - shift : 8;
        const int mask = (1 << bits) - 1;
        radix_hist_kernel<<<g.num_blocks, kSortThreads, 0, st>>>(src_k, g, shift, mask, block_hist);
        GSTEX_LAUNCH_CHECK();
        radix_scan_rows_kernel<<<kRadix, kSortThreads, 0, st>>>(block_hist, g.num_blocks, block_off, digit_total);
        GSTEX_LAUNCH_CHECK();
        radix_scatter_kernel<<<g.num_blocks, kSortThreads, 0, st>>>(src_k, src_v, dst_k, dst_v, g, shift, mask, block_off,
                                                                    digit_total);
        GSTEX_LAUNCH_CHECK();
        src_k = dst_k;
        src_v = dst_v;
    }
    return GSTEX_OK;
}
