// pair_stream.cuh -- the "pair stream": what the forward compositing kernel records for the backward.
//
// The reference backward (texture.cu:331-760) and our recompute backward (raster_bwd.cu) re-walk every tile
// list, re-evaluate every (pixel, splat) pair and do the gradient math for the few pairs that contributed
// with ~7 of 32 lanes active.  B200 has 180 GB of HBM3e and this path uses ~2 % of its bandwidth, so the
// forward instead WRITES one 32-byte item per contributing pair (~10 M per 800x800 view = 320 MB) and the
// stream backward (raster_bwd_stream.cu) reads them back fully lane-dense: no culling, no re-evaluation of
// non-contributing pairs, no texel re-fetch.
//
// Layout of the caller-provided buffer (`pair_stream`, `pair_stream_bytes`):
//   [ StreamHeader 256 B | int2 units[tiles * 8] | int32 chunk_prev[num_chunks] | pad to 128 | chunks ]
// A "unit" is one pixel warp of one tile (32 pixels); its items are appended in compositing order
// (sorted-list position major, lane minor) into a chain of 4 KB chunks of 128 items taken from the pool
// with one atomicAdd per chunk.  units[u] = {last chunk, item count}; chunk_prev links a chunk to the one
// before it in its unit (-1 for the first).  If the pool runs dry the forward raises `overflow` and the
// backward falls back to the recompute kernel -- results never depend on the pool size.
#pragma once
#include "common.cuh"

namespace gstex {

constexpr int kChunkItems = 128;
constexpr int kUnitsPerTile = 8;  // pixel warps of a 16x16 tile

struct __align__(16) PairItem {
    uint32_t gid_lane;  // splat id << 5 | pixel lane within the unit
    int32_t idx;        // position in the global sorted intersection list
    float T;            // transmittance in front of this pair
    float u, v;         // clamped chart coordinates
    float val[3];       // bilinear (or nearest) texture sample, channels 0..2
};
static_assert(sizeof(PairItem) == 32, "PairItem must be 32 bytes");

struct StreamHeader {
    uint32_t next_chunk;   // pool allocation cursor (forward)
    uint32_t overflow;     // set by the forward when the pool ran dry
    uint32_t next_unit;    // work counter of the stream backward
    uint32_t reserved;
    uint32_t pad[60];
};
static_assert(sizeof(StreamHeader) == 256, "StreamHeader must be 256 bytes");

struct StreamView {
    StreamHeader* hdr;
    int2* units;
    int32_t* chunk_prev;
    PairItem* items;
    int num_chunks;
    int num_units;
};

static inline size_t stream_fixed_bytes(int num_units) { return sizeof(StreamHeader) + (size_t)num_units * sizeof(int2); }

// Carve the caller's buffer.  Returns false if it cannot even hold one chunk per 64 units.
static inline bool make_stream_view(void* buf, size_t bytes, int num_tiles, StreamView& sv) {
    sv.num_units = num_tiles * kUnitsPerTile;
    const size_t fixed = stream_fixed_bytes(sv.num_units) + 256;
    if (!buf || bytes <= fixed) return false;
    const size_t per_chunk = sizeof(PairItem) * kChunkItems + sizeof(int32_t);
    const size_t nc = (bytes - fixed) / per_chunk;
    if (nc < 1 || nc > 0x7fffffffull) return false;
    char* p = reinterpret_cast<char*>(buf);
    sv.hdr = reinterpret_cast<StreamHeader*>(p);
    p += sizeof(StreamHeader);
    sv.units = reinterpret_cast<int2*>(p);
    p += (size_t)sv.num_units * sizeof(int2);
    sv.chunk_prev = reinterpret_cast<int32_t*>(p);
    p += nc * sizeof(int32_t);
    p = reinterpret_cast<char*>(((uintptr_t)p + 127) & ~(uintptr_t)127);
    sv.items = reinterpret_cast<PairItem*>(p);
    sv.num_chunks = (int)nc;
    // the alignment pad may eat into the last chunk
    while (sv.num_chunks > 0 &&
           reinterpret_cast<char*>(sv.items + (size_t)sv.num_chunks * kChunkItems) > reinterpret_cast<char*>(buf) + bytes)
        --sv.num_chunks;
    return sv.num_chunks > 0;
}

}  // namespace gstex
