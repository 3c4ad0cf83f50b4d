// tile_bucket.cu -- stage (b), fast path used by texture_gaussians: tile-bucketed binning.
// Produces exactly what map_gaussian_to_intersects -> stable sort of (tile<<32 | depth) -> get_tile_bin_edges
// produce (gstex_cuda/utils.py:106-162; cuda/csrc/forward.cu:13-98): the per-tile depth-sorted splat ids and
// the (tiles,2) ranges -- without ever sorting the whole intersection list.  sm_100a.
//
// The global list is tile-major, so its layout is known as soon as the per-tile counts are:
//   1. tile_count_kernel   296 blocks, each owning a contiguous slice of the splats, walk every splat's tile
//                          rectangle (the emit kernel's arithmetic, scan_emit.cu) and count one intersection per
//                          tile it would emit in a per-block shared-memory array (no global atomics);
//   2. tile_scan_kernel    one block: per-tile totals over the blocks (leaving each block its exclusive prefix),
//                          exclusive scan of the totals -> tile_bins (empty tiles stay (0,0)), tile starts, the
//                          total and the longest list (read back by the host with the `num_intersects` sync the
//                          reference has anyway, utils.py:58);
//   3. tile_emit_kernel    the same walk again: each intersection claims a slot of its tile's range from the
//                          block's shared-memory cursor and stores (depth bits << 32 | splat id);
//   4. tile_local_sort_kernel  one block per tile sorts its range by (depth bits, splat id) entirely in shared
//                          memory: LSD radix over the id bits then the depth bits, 8 bits per pass, passes
//                          whose digit is constant over the tile are skipped.
// Equal (tile, depth) keys keep ascending splat id in the reference (stable sort of an emission order that is
// ascending in the splat index); sorting by (depth, id) gives the same order whatever order the atomics hand
// out, so the result is deterministic and bit-identical.  Global traffic: 8 B written + 8 B read + 4 B written
// per intersection instead of six 36-byte passes.  Lists longer than the shared-memory capacity
// (gstex_tile_bucket_max_list) make the caller take the general radix-sort path (sort.cu).
#include "common.cuh"

namespace gstex {

// size classes of the per-tile sort: (warps per block, keys per thread); capacity = warps * 32 * keys
constexpr int kLsSmallW = 8, kLsSmallK = 8;    // lists <= 2048
constexpr int kLsLargeW = 16;                  // longer lists: K = 12 / 16 / 26
constexpr int kLsSmallCap = kLsSmallW * 32 * kLsSmallK;
constexpr int kLsMaxList = kLsLargeW * 32 * 26;  // 13312 keys = 104 KB of shared memory
constexpr int ls_smem_bytes(int warps, int k) { return warps * 32 * k * 8 + warps * 256 * 4 + 256; }

struct TileRect {
    int min_x, max_x, min_y, max_y;
    bool emits;
};

// tile rectangle of a splat, arithmetic of map_gaussian_to_intersects (forward.cu:13-71, helpers.cuh:37-51,75-92)
__device__ __forceinline__ TileRect tile_rect(const float2 c, const float2 e, float fbw, int tiles_x, int tiles_y) {
    TileRect r;
    r.emits = !((double)e.x <= 1e-4 && (double)e.y <= 1e-4);
    const float tcx = __fdiv_rn(c.x, fbw), tcy = __fdiv_rn(c.y, fbw);
    const float tex = __fdiv_rn(e.x, fbw), tey = __fdiv_rn(e.y, fbw);
    r.min_x = min(max(0, __float2int_rz(__fsub_rn(tcx, tex))), tiles_x);
    r.max_x = min(max(0, __float2int_rz(__fadd_rn(__fadd_rn(tcx, tex), 1.0f))), tiles_x);
    r.min_y = min(max(0, __float2int_rz(__fsub_rn(tcy, tey))), tiles_y);
    r.max_y = min(max(0, __float2int_rz(__fadd_rn(__fadd_rn(tcy, tey), 1.0f))), tiles_y);
    return r;
}

// Counting and emitting are privatised per block: kTbBlocks blocks each own a contiguous slice of the splats and
// a shared-memory array with one int per tile.  (One global atomic per intersection lands on the ~80 L2 lines
// that hold a 2500-tile count array and serialises there: measured 0.2 ms per view for 3M intersections.)
constexpr int kTbBlocks = 296;
constexpr int kTbThreads = 1024;
constexpr int kTbMaxTiles = 50 * 1024;  // 200 KB of shared memory

// block_counts[b * num_tiles + t] = intersections of block b's splats with tile t
__global__ void __launch_bounds__(kTbThreads)
tile_count_kernel(int n, int chunk, const float2* __restrict__ centers, const float2* __restrict__ extents,
                  const int32_t* __restrict__ num_tiles_hit, int tiles_x, int tiles_y, float fbw,
                  int32_t* __restrict__ block_counts) {
    extern __shared__ int tb_smem[];
    const int num_tiles = tiles_x * tiles_y;
    for (int t = threadIdx.x; t < num_tiles; t += kTbThreads) tb_smem[t] = 0;
    __syncthreads();
    const int begin = blockIdx.x * chunk, end = min(begin + chunk, n);
    for (int idx = begin + threadIdx.x; idx < end; idx += kTbThreads) {
        int left = num_tiles_hit[idx];  // slots stage a3 reserves for this splat
        if (left <= 0) continue;
        const TileRect r = tile_rect(centers[idx], extents[idx], fbw, tiles_x, tiles_y);
        if (r.emits) {
            for (int i = r.min_y; i < r.max_y && left > 0; ++i)
                for (int j = r.min_x; j < r.max_x && left > 0; ++j) {
                    atomicAdd(&tb_smem[i * tiles_x + j], 1);
                    --left;
                }
        }
        if (left > 0) atomicAdd(&tb_smem[0], left);  // reserved but not emitted: key 0 = tile 0 (the zero-key quirk)
    }
    __syncthreads();
    int32_t* out = block_counts + (size_t)blockIdx.x * num_tiles;
    for (int t = threadIdx.x; t < num_tiles; t += kTbThreads) out[t] = tb_smem[t];
}

// One block of 1024 threads: exclusive scan of the tile totals (read from tile_start, overwritten with the starts)
// -> tile_bins, stats.
__global__ void __launch_bounds__(1024)
tile_scan_kernel(int num_tiles, int2* __restrict__ tile_bins, int32_t* __restrict__ tile_start,
                 int32_t* __restrict__ stats) {
    __shared__ int warp_sums[32];
    __shared__ int warp_max[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int carry = 0, mx = 0;
    for (int base = 0; base < num_tiles; base += 1024) {
        const int t = base + threadIdx.x;
        const int c = (t < num_tiles) ? tile_start[t] : 0;
        mx = max(mx, c);
        int v = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += u;
        }
        if (lane == 31) warp_sums[warp] = v;
        __syncthreads();
        int wprefix = 0, total = 0;
        for (int w = 0; w < 32; ++w) {
            const int s = warp_sums[w];
            if (w < warp) wprefix += s;
            total += s;
        }
        if (t < num_tiles) {
            const int start = carry + wprefix + v - c;
            tile_bins[t] = c > 0 ? make_int2(start, start + c) : make_int2(0, 0);
            tile_start[t] = start;
        }
        carry += total;
        __syncthreads();
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) warp_max[warp] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        int m = 0;
        for (int w = 0; w < 32; ++w) m = max(m, warp_max[w]);
        stats[0] = carry;
        stats[1] = m;
    }
}

// Same slicing as the count kernel: block b's cursor for tile t starts at tile_start[t] + (intersections of the
// blocks before b with t), so no two blocks ever touch the same slot and all atomics stay in shared memory.
__global__ void __launch_bounds__(kTbThreads)
tile_emit_kernel(int n, int chunk, const float2* __restrict__ centers, const float2* __restrict__ extents,
                 const float* __restrict__ depths, const int32_t* __restrict__ num_tiles_hit, int tiles_x, int tiles_y,
                 float fbw, const int32_t* __restrict__ tile_start, const int32_t* __restrict__ block_counts,
                 unsigned long long* __restrict__ keys) {
    extern __shared__ int tb_smem[];
    const int num_tiles = tiles_x * tiles_y;
    const int32_t* mine = block_counts + (size_t)blockIdx.x * num_tiles;
    for (int t = threadIdx.x; t < num_tiles; t += kTbThreads) tb_smem[t] = tile_start[t] + mine[t];
    __syncthreads();
    const int begin = blockIdx.x * chunk, end = min(begin + chunk, n);
    for (int idx = begin + threadIdx.x; idx < end; idx += kTbThreads) {
        int left = num_tiles_hit[idx];
        if (left <= 0) continue;
        const TileRect r = tile_rect(centers[idx], extents[idx], fbw, tiles_x, tiles_y);
        if (r.emits) {
            const unsigned long long key = ((unsigned long long)__float_as_uint(depths[idx]) << 32) | (unsigned)idx;
            for (int i = r.min_y; i < r.max_y && left > 0; ++i)
                for (int j = r.min_x; j < r.max_x && left > 0; ++j) {
                    const int pos = atomicAdd(&tb_smem[i * tiles_x + j], 1);
                    keys[pos] = key;
                    --left;
                }
        }
        if (left > 0) {
            const int pos = atomicAdd(&tb_smem[0], left);
            for (int k = 0; k < left; ++k) keys[pos + k] = 0ull;  // key 0, splat id 0 (bindings.cu:99-102 zero-init)
        }
    }
}

// ---- per-tile sort ------------------------------------------------------------------------------------------
// One block (W warps) per tile whose list length lies in [n_lo, n_hi], n_hi <= W*32*K.  Shared memory holds ONE
// copy of the tile's 64-bit keys: in every pass each thread first takes its K keys into registers (warp w owns
// a contiguous segment, lane l's j-th key is segment[j*32 + l], so (warp, j, lane) order == list order), the
// block counts digits, scans, and only then scatters the registers back into the same array.
// Strategy: sort by the depth bits only (stable LSD, 4 passes minus those whose digit is constant over the tile);
// that is already the final order unless two entries with EQUAL depth bits arrived with descending splat ids
// (the emit cursor hands slots out in arbitrary order) -- see the fix-up in the kernel.

// Lanes of the warp holding the same 8-bit digit as this lane (invalid lanes: none).  Eight ballots instead of
// MATCH.ANY: measured here, MATCH.ANY through the MIO pipe with 32 warps per SM left the kernel at 0.02
// instructions per cycle per warp; VOTE is a few cycles.
__device__ __forceinline__ unsigned digit_peers(int d, bool valid) {
    unsigned peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
    for (int b = 0; b < 8; ++b) {
        const bool bit = (d >> b) & 1;
        const unsigned bal = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? bal : ~bal;
    }
    return valid ? peers : 0u;
}

template <int W, int K>
__device__ __forceinline__ void ls_radix_passes(unsigned long long* buf, int (*warp_hist)[256], int* scan_tmp, int n,
                                                int seg_begin, int seg_end, int first_pass, int last_pass,
                                                int id_passes, int id_bits) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    __syncthreads();  // the caller's copy into buf is complete
    for (int pass = first_pass; pass < last_pass; ++pass) {
        const int shift = pass < id_passes ? 8 * pass : 32 + 8 * (pass - id_passes);
        const int bits = pass < id_passes ? min(8, id_bits - 8 * pass) : 8;
        const int mask = (1 << bits) - 1;
        unsigned long long key[K];
        unsigned peers[K];
#pragma unroll
        for (int j = 0; j < K; ++j) {
            const int i = seg_begin + j * 32 + lane;
            key[j] = (i < seg_end) ? buf[i] : 0ull;
        }
        for (int j = tid; j < W * 256; j += W * 32) (&warp_hist[0][0])[j] = 0;
        __syncthreads();  // every key is in a register; histograms are clear
#pragma unroll
        for (int j = 0; j < K; ++j) {
            if (seg_begin + j * 32 < seg_end) {  // warp-uniform
                const bool ok = seg_begin + j * 32 + lane < seg_end;
                const int d = (int)(key[j] >> shift) & mask;
                peers[j] = digit_peers(d, ok);
                if (ok && lane == __ffs(peers[j]) - 1) warp_hist[warp][d] += __popc(peers[j]);
                __syncwarp();
            } else {
                peers[j] = 0u;
            }
        }
        __syncthreads();
        int total = 0;
        if (tid < 256) {
#pragma unroll 8
            for (int w = 0; w < W; ++w) {  // thread d: exclusive prefix of digit d over the warps
                const int c = warp_hist[w][tid];
                warp_hist[w][tid] = total;
                total += c;
            }
        }
        // a digit that holds every key: nothing to reorder in this pass (the keys in shared memory are untouched)
        if (__syncthreads_or(tid < 256 && total == n)) continue;
        if (tid < 256) {  // exclusive scan of the digit totals over the first 256 threads
            int v = total;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, v, o);
                if (lane >= o) v += u;
            }
            if (lane == 31) scan_tmp[warp] = v;
            asm volatile("bar.sync 1, 256;");  // the eight digit warps only
            int wprefix = 0;
#pragma unroll
            for (int w = 0; w < 8; ++w)
                if (w < warp) wprefix += scan_tmp[w];
            const int dbase = wprefix + v - total;
#pragma unroll 8
            for (int w = 0; w < W; ++w) warp_hist[w][tid] += dbase;
        }
        __syncthreads();
#pragma unroll
        for (int j = 0; j < K; ++j) {
            if (seg_begin + j * 32 < seg_end) {
                const unsigned p = peers[j];
                const int d = (int)(key[j] >> shift) & mask;
                const int leader = p ? __ffs(p) - 1 : lane;
                int prev = 0;
                if (p && lane == leader) {
                    prev = warp_hist[warp][d];
                    warp_hist[warp][d] = prev + __popc(p);
                }
                prev = __shfl_sync(0xffffffffu, prev, leader);
                if (p) buf[prev + __popc(p & lt_mask)] = key[j];
                __syncwarp();
            }
        }
        __syncthreads();
    }
}

template <int W, int K>
__global__ void __launch_bounds__(W * 32)
tile_local_sort_kernel(const int2* __restrict__ tile_bins, const unsigned long long* __restrict__ keys,
                       int32_t* __restrict__ ids_out, int id_bits, int n_lo, int n_hi) {
    constexpr int kThreads = W * 32, kCap = W * 32 * K;
    extern __shared__ __align__(16) unsigned char ls_smem[];
    unsigned long long* buf = reinterpret_cast<unsigned long long*>(ls_smem);
    int (*warp_hist)[256] = reinterpret_cast<int (*)[256]>(buf + kCap);
    int* scan_tmp = reinterpret_cast<int*>(warp_hist + W);

    const int2 range = tile_bins[blockIdx.x];
    const int n = range.y - range.x;
    if (n < n_lo || n > n_hi || n > kCap) return;  // n > kCap cannot happen: the host picks K from the longest list
    const int tid = threadIdx.x;
    if (n == 1) {
        if (tid == 0) ids_out[range.x] = (int32_t)(unsigned)keys[range.x];
        return;
    }
    const int warp = tid >> 5;
    const int seg = ((n + kThreads - 1) / kThreads) * 32;  // <= K*32 keys per warp, spread over all warps
    const int seg_begin = min(warp * seg, n), seg_end = min(seg_begin + seg, n);
    const int id_passes = (id_bits + 7) / 8;

    for (int i = tid; i < n; i += kThreads) buf[i] = keys[range.x + i];
    ls_radix_passes<W, K>(buf, warp_hist, scan_tmp, n, seg_begin, seg_end, id_passes, id_passes + 4, id_passes, id_bits);
    // after the last pass's barrier: equal depths with descending ids?  Isolated pairs (the usual case: two
    // splats of a tile with the same depth bits) are put right by a few odd-even exchange rounds, which only
    // ever swap equal-depth neighbours; anything longer is sorted again from the original keys, ids first.
    bool bad = false;
    for (int i = tid; i + 1 < n; i += kThreads) bad |= buf[i] > buf[i + 1];
    if (__syncthreads_or(bad)) {
        bad = true;  // uniform from here on
        for (int round = 0; round < 4 && bad; ++round) {
            bool moved = false;
#pragma unroll
            for (int parity = 0; parity < 2; ++parity) {
                for (int i = 2 * tid + parity; i + 1 < n; i += 2 * kThreads) {
                    const unsigned long long a = buf[i], b = buf[i + 1];
                    if (a > b) { buf[i] = b; buf[i + 1] = a; moved = true; }
                }
                __syncthreads();
            }
            bad = __syncthreads_or(moved);
        }
        if (bad) {  // uniform: still moving after four rounds -> a long run of equal depths
            for (int i = tid; i < n; i += kThreads) buf[i] = keys[range.x + i];
            ls_radix_passes<W, K>(buf, warp_hist, scan_tmp, n, seg_begin, seg_end, 0, id_passes + 4, id_passes, id_bits);
        }
    }
    for (int i = tid; i < n; i += kThreads) ids_out[range.x + i] = (int32_t)(unsigned)buf[i];
}

// tile totals over the blocks: one warp per tile; block_counts[b][t] becomes the exclusive prefix over b
__global__ void __launch_bounds__(256)
tile_colscan_kernel(int num_tiles, int num_blocks, int32_t* __restrict__ block_counts, int32_t* __restrict__ tile_total) {
    const int t = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (t >= num_tiles) return;
    int carry = 0;
    for (int b0 = 0; b0 < num_blocks; b0 += 32) {
        const int b = b0 + lane;
        int32_t* q = block_counts + (size_t)b * num_tiles + t;
        const int c = (b < num_blocks) ? *q : 0;
        int v = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += u;
        }
        if (b < num_blocks) *q = carry + v - c;
        carry += __shfl_sync(0xffffffffu, v, 31);
    }
    if (lane == 0) tile_total[t] = carry;
}


}  // namespace gstex

using namespace gstex;

extern "C" int gstex_tile_bucket_max_list(void) { return kLsMaxList; }
extern "C" int gstex_tile_bucket_max_tiles(void) { return kTbMaxTiles; }
extern "C" size_t gstex_tile_bucket_scratch_bytes(int num_tiles) {
    return num_tiles > 0 ? (size_t)kTbBlocks * (size_t)num_tiles * sizeof(int32_t) : 16;
}

static inline int tb_chunk(int n) { return (n + kTbBlocks - 1) / kTbBlocks; }

extern "C" int gstex_tile_bucket_count(int n, const float* centers, const float* extents, const int32_t* num_tiles_hit,
                                       int tiles_x, int tiles_y, int block_width, int32_t* block_counts,
                                       int32_t* tile_bins, int32_t* tile_start, int32_t* stats, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (n < 0 || block_width < 2 || block_width > 16 || tiles_x <= 0 || tiles_y <= 0) return GSTEX_E_BADARG;
    if (!block_counts || !tile_bins || !tile_start || !stats) return GSTEX_E_BADARG;
    if (n > 0 && (!centers || !extents || !num_tiles_hit)) return GSTEX_E_BADARG;
    const long long num_tiles = (long long)tiles_x * tiles_y;
    if (num_tiles > kTbMaxTiles) return GSTEX_E_UNSUPPORTED;
    const size_t smem = (size_t)num_tiles * sizeof(int);
    GSTEX_CUDA_TRY(cudaFuncSetAttribute(tile_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTbMaxTiles * 4));
    tile_count_kernel<<<kTbBlocks, kTbThreads, smem, st>>>(
        n, tb_chunk(n), reinterpret_cast<const float2*>(centers), reinterpret_cast<const float2*>(extents), num_tiles_hit,
        tiles_x, tiles_y, (float)block_width, block_counts);
    GSTEX_LAUNCH_CHECK();
    tile_colscan_kernel<<<div_up(num_tiles * 32, 256), 256, 0, st>>>((int)num_tiles, kTbBlocks, block_counts, tile_start);
    GSTEX_LAUNCH_CHECK();
    tile_scan_kernel<<<1, 1024, 0, st>>>((int)num_tiles, reinterpret_cast<int2*>(tile_bins), tile_start, stats);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

extern "C" int gstex_tile_bucket_sort(int n, int num_intersects, int max_list, const float* centers,
                                      const float* extents, const float* depths, const int32_t* num_tiles_hit,
                                      int tiles_x, int tiles_y, int block_width, const int32_t* block_counts,
                                      const int32_t* tile_start, const int32_t* tile_bins, void* keys,
                                      int32_t* gaussian_ids_sorted, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (n < 0 || num_intersects < 0 || max_list < 0 || block_width < 2 || block_width > 16 || tiles_x <= 0 || tiles_y <= 0)
        return GSTEX_E_BADARG;
    if (num_intersects == 0 || n == 0) return GSTEX_OK;
    const long long num_tiles = (long long)tiles_x * tiles_y;
    if (max_list > kLsMaxList || num_tiles > kTbMaxTiles) return GSTEX_E_UNSUPPORTED;
    if (!centers || !extents || !depths || !num_tiles_hit || !block_counts || !tile_start || !tile_bins || !keys ||
        !gaussian_ids_sorted)
        return GSTEX_E_BADARG;
    GSTEX_CUDA_TRY(cudaFuncSetAttribute(tile_emit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTbMaxTiles * 4));
    tile_emit_kernel<<<kTbBlocks, kTbThreads, (size_t)num_tiles * sizeof(int), st>>>(
        n, tb_chunk(n), reinterpret_cast<const float2*>(centers), reinterpret_cast<const float2*>(extents), depths,
        num_tiles_hit, tiles_x, tiles_y, (float)block_width, tile_start, block_counts,
        reinterpret_cast<unsigned long long*>(keys));
    GSTEX_LAUNCH_CHECK();
    int id_bits = 1;
    while (id_bits < 32 && (1ll << id_bits) < (long long)n) ++id_bits;
    const int2* bins = reinterpret_cast<const int2*>(tile_bins);
    const unsigned long long* k64 = reinterpret_cast<const unsigned long long*>(keys);
    // size classes: a view whose lists are all short takes the 8-warp kernel (24 KB of shared memory per block); a
    // view with long lists takes one 16-warp launch for all tiles -- the short lists then fill the tail the long
    // ones leave (two launches, one per class, measured 20 us slower at the headline size)
    if (max_list <= kLsSmallCap) {
        tile_local_sort_kernel<kLsSmallW, kLsSmallK>
            <<<(int)num_tiles, kLsSmallW * 32, ls_smem_bytes(kLsSmallW, kLsSmallK), st>>>(bins, k64, gaussian_ids_sorted,
                                                                                          id_bits, 1, max_list);
        GSTEX_LAUNCH_CHECK();
    }
    if (max_list > kLsSmallCap) {
        // cudaFuncSetAttribute is per device and cheap: set every time rather than caching a flag across devices
#define GSTEX_LS_LAUNCH(KK)                                                                                            \
    do {                                                                                                               \
        GSTEX_CUDA_TRY(cudaFuncSetAttribute(tile_local_sort_kernel<kLsLargeW, KK>,                                     \
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, ls_smem_bytes(kLsLargeW, KK))); \
        tile_local_sort_kernel<kLsLargeW, KK><<<(int)num_tiles, kLsLargeW * 32, ls_smem_bytes(kLsLargeW, KK), st>>>(   \
            bins, k64, gaussian_ids_sorted, id_bits, 1, max_list);                                       \
    } while (0)
        if (max_list <= kLsLargeW * 32 * 12) GSTEX_LS_LAUNCH(12);
        else if (max_list <= kLsLargeW * 32 * 16) GSTEX_LS_LAUNCH(16);
        else GSTEX_LS_LAUNCH(26);
#undef GSTEX_LS_LAUNCH
        GSTEX_LAUNCH_CHECK();
    }
    return GSTEX_OK;
}
