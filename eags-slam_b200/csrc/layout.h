// layout.h -- how the three opaque state buffers are carved (host + device).
//
// HBM layout (all arrays 256-byte aligned inside their buffer, SoA so that per-Gaussian kernels
// stream with unit stride and 16-byte vectors):
//
//   geom state   (per Gaussian, written by preprocess_fwd)
//     splat_rec     float4[3P]  the splat record of every visible Gaussian, ALREADY in the form the blend
//                               kernels stream (blend.cu): {x, y, bits(index), depth} {A, B, C, opacity}
//                               {r, g, b, log2(255 opacity)} with (A, B, C) the conic pre-scaled by -log2(e)/2,
//                               -log2(e), -log2(e)/2 -- written once per Gaussian by preprocess_fwd, so that
//                               sort_pack copies 48 contiguous bytes per key instead of gathering from three
//                               arrays and redoing the scaling K / P times
//     tiles_touched u32[P]
//     clamped       u8[P]       bit c: colour channel c was clamped at 0 (gates dL/dsh)
//
//   image state  (per tile / per pixel)
//     ctrl          Ctrl        counters, zeroed at the start of every forward
//     tile_count    u32[G*32]   keys per tile, one counter per 128-byte line (word 0; zeroed with ctrl; RED
//                               atomics in preprocess_fwd); tile_scan leaves the tile's scatter cursor
//                               in word 1 of the same line, which duplicate advances with ATOM
//     ranges        uint2[G]    [start, end) of each tile in the sorted list
//     big_list      u32[G]      tiles whose list does not fit the small sort CTA
//     tile_max      u32[G]      max n_contrib over the tile's pixels (bounds the backward walk)
//     final_T       f32[N]
//     n_contrib     u32[N]
//
//   binning state (per (tile, Gaussian) key; `cap` = capacity)
//     unsorted      u64[max(cap, G * tile_cap)]  (depth_bits << 32 | gaussian) bucketed by tile, unsorted;
//                               last in the buffer
//     point_list    u32[cap]    sorted Gaussian indices (upstream: binningState.point_list)
//     records       float4[3*cap]  the sorted, packed splat stream the blend kernels read:
//                                  {x, y, idx, depth} {conic.xx, conic.xy, conic.yy, opacity} {r, g, b, 0}
//     keys          u64[cap]    sorted (tile << 32 | depth_bits) keys, only when with_keys
#pragma once
#include <stddef.h>
#include <stdint.h>

#define EAGS_TILE 16
#define EAGS_CTR_CURSOR 1           // word of a tile's counter line that holds its scatter cursor (exact path)
#define EAGS_CTR_STRIDE 32          // per-tile counters live one per 128-byte line: the L2 atomic unit
                                   // serialises RED/ATOM traffic per line, 38 hot lines cost 60 us at c2
#define EAGS_SORT_SMALL_CAP 2048   // keys a 256-thread sort CTA holds in shared memory
#define EAGS_SORT_BIG_CAP 16384    // keys a 1024-thread sort CTA holds in shared memory

struct EagsCtrl {
    unsigned long long num_rendered;   // K = sum tiles_touched (atomics in preprocess_fwd)
    unsigned long long num_scanned;    // K as seen by tile_scan (debug cross-check)
    unsigned int n_big;                // entries in big_list
    unsigned int work_sort;            // dynamic work queues of the persistent kernels
    unsigned int work_fwd;
    unsigned int max_tile;             // largest per-tile key count (tile_scan): sizes the next call's buckets
    unsigned int bucket_overflow;      // bucketed forward: some tile outgrew its bucket, nothing was rendered
    unsigned int pad0;
    // "kernel_stats" option only (bench.py's executed-work roofline): work blend_fwd really did
    unsigned long long stat_fwd_warp_iters;   // (8x4 block, splat) pairs that survived the cull
    unsigned long long stat_fwd_lane_pairs;   // (pixel, splat) pairs that were blended
    unsigned long long pad1;
};
static_assert(sizeof(EagsCtrl) == 64, "ctrl block is 64 bytes");

static inline size_t eags_align256(size_t x) { return (x + 255) & ~(size_t)255; }

struct GeomLayout {
    size_t splat_rec, tiles_touched, clamped, total;
};
static inline GeomLayout geom_layout(int64_t P) {
    GeomLayout L;
    size_t o = 0;
    L.splat_rec = o;     o += eags_align256((size_t)P * 48);
    L.tiles_touched = o; o += eags_align256((size_t)P * 4);
    L.clamped = o;       o += eags_align256((size_t)P);
    L.total = o + 256;
    return L;
}

struct ImageLayout {
    size_t ctrl, tile_count, ranges, big_list, tile_max, final_T, n_contrib, total;
    int gx, gy, G;
    int64_t N;
};
static inline ImageLayout image_layout(int W, int H) {
    ImageLayout L;
    L.gx = (W + EAGS_TILE - 1) / EAGS_TILE;
    L.gy = (H + EAGS_TILE - 1) / EAGS_TILE;
    L.G = L.gx * L.gy;
    L.N = (int64_t)W * H;
    size_t o = 0;
    L.ctrl = o;        o += 64;
    L.tile_count = o;  o += eags_align256((size_t)L.G * 4 * EAGS_CTR_STRIDE + 64) - 64;   // ctrl + counters: one memset
    L.ranges = o;      o += eags_align256((size_t)L.G * 8);
    L.big_list = o;    o += eags_align256((size_t)L.G * 4);
    L.tile_max = o;    o += eags_align256((size_t)L.G * 4);
    L.final_T = o;     o += eags_align256((size_t)L.N * 4);
    L.n_contrib = o;   o += eags_align256((size_t)L.N * 4);
    L.total = o + 256;
    return L;
}

struct BinningLayout {
    size_t unsorted, point_list, records, keys, total;
};
// `unsorted` comes last so that the other offsets do not depend on its size: it holds either `cap`
// keys laid out by the tile scan (exact path) or G fixed-stride buckets of `bucket_slots / G` keys
// (bucketed path, where preprocess_fwd scatters the keys itself and there is no duplicate kernel).
static inline BinningLayout binning_layout(int64_t cap, int with_keys, int64_t bucket_slots = 0) {
    BinningLayout L;
    if (cap < 1) cap = 1;
    size_t o = 0;
    L.point_list = o; o += eags_align256((size_t)cap * 4);
    L.records = o;    o += eags_align256((size_t)cap * 48);
    L.keys = o;       o += with_keys ? eags_align256((size_t)cap * 8) : 0;
    L.unsorted = o;   o += eags_align256((size_t)(bucket_slots > cap ? bucket_slots : cap) * 8);
    L.total = o + 256;
    return L;
}

// backward scratch: grad_acc float[P][12] | pose accumulators double[16] | work queue u32 (at pose + 128)
//                   | "kernel_stats": blend_bwd warp iterations, lane pairs (2 x u64 at pose + 144)
#define EAGS_SCRATCH_WORK_OFF 128
#define EAGS_SCRATCH_STATS_OFF 144
struct ScratchLayout {
    size_t grad_acc, pose, total;
};
static inline ScratchLayout scratch_layout(int64_t P) {
    ScratchLayout L;
    size_t o = 0;
    L.grad_acc = o; o += eags_align256((size_t)P * 48);
    L.pose = o;     o += 256;
    L.total = o + 256;
    return L;
}
