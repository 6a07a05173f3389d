// knn.cu -- neighbour queries that bracket the rasterizer when a keyframe seeds new Gaussians
// (sm_100a).
//
//   eags_knn_mean_dist2       : mean squared distance of every point to its 3 nearest neighbours --
//                               the `distCUDA2(points)` of the un-vendored `simple_knn` pip package
//                               (reference call sites src/entities/gaussian_model.py:143,146,253,257,
//                               298,301,616: the initial log-scale of new Gaussians).
//   eags_radius_has_neighbor  : "does a query point have a reference point with squared distance
//                               below `radius`" -- the faiss IndexFlatL2 search of
//                               src/utils/mapper_utils.py:193-226 (compute_new_points_ids keeps the
//                               query points for which the answer is no).
//
// Upstream simple-knn Morton-sorts the cloud (CUB radix sort), cuts it into 1 024-point boxes and
// lets every point scan every box its current 3rd-best distance cannot reject: exact, but
// O(P * boxes) box tests plus whole-box scans.  Here both queries run on one *hashed uniform grid*:
// points are counting-sorted into cells (cell -> bucket by a multiplicative hash, so the table
// size depends on P only, never on the data's extent), a query walks Chebyshev rings of cells and
// stops as soon as its 3rd-best distance is inside the searched cube.  Results are exact: ring
// termination and cell assignment carry explicit rounding margins; queries a 4-ring search cannot
// settle (points in sparse regions) are handed to two coarser grids (5 h, 25 h cells) together
// with the bound they already have, which prunes the coarse cells; only points isolated beyond
// 100 h scan the whole cloud.  The finest cell size adapts to the data in two counting passes:
// clouds that sit on surfaces fill far fewer cells than a volume estimate predicts.  Everything
// is L2-latency-bound integer + fp32 work; no tensor cores.
#include <float.h>

#include "kernels.h"

namespace {

constexpr int NB_THREADS = 256;
constexpr int MAX_DIM = 1024;            // cells per axis (10 bits each in the packed cell key)
constexpr int MAX_RING = 4;
constexpr unsigned FULLMASK = 0xffffffffu;

struct GridHdr {
    unsigned lo[3], hi[3];    // bounding box, order-preserving uint encoding of the floats
    float o[3];               // grid origin
    float h, inv_h;
    int n[3];                 // cells per axis
    unsigned n_occ;           // non-empty buckets seen by the last counting pass
    unsigned log2_table;
    unsigned pad[16];
};
static_assert(sizeof(GridHdr) <= 256, "grid header fits its 256-byte slot");

struct GridPtrs {
    GridHdr* hdr;
    unsigned* start;       // [T + 1] bucket counts, then (after the scan) bucket starts
    unsigned* cursor;      // [T]
    unsigned* block_sums;  // [scan blocks]
    float4* sorted;        // [n] (x, y, z, bits(cell key)) in bucket order
    unsigned* sorted_idx;  // [n] original index
    unsigned T, log2T;
};

__host__ __device__ inline unsigned table_log2(long long n) {
    unsigned l = 10;
    while ((1ll << l) < 2 * n && l < 26) l++;
    return l;
}
constexpr int SCAN_ITEMS = 2048;   // elements per scan block (256 threads x 8)

size_t grid_bytes(long long n) {
    const unsigned l = table_log2(n);
    const size_t T = (size_t)1 << l;
    size_t o = 256;
    o += eags_align256((T + 1) * 4);
    o += eags_align256(T * 4);
    o += eags_align256(((T + 1 + SCAN_ITEMS - 1) / SCAN_ITEMS + 1) * 4);
    o += eags_align256((size_t)(n > 0 ? n : 1) * 16);
    o += eags_align256((size_t)(n > 0 ? n : 1) * 4);
    return o + 256;
}
GridPtrs grid_ptrs(void* ws, long long n) {
    GridPtrs g;
    g.log2T = table_log2(n);
    g.T = 1u << g.log2T;
    char* b = (char*)ws;
    size_t o = 0;
    g.hdr = (GridHdr*)(b + o);        o += 256;
    g.start = (unsigned*)(b + o);     o += eags_align256(((size_t)g.T + 1) * 4);
    g.cursor = (unsigned*)(b + o);    o += eags_align256((size_t)g.T * 4);
    g.block_sums = (unsigned*)(b + o); o += eags_align256((((size_t)g.T + 1 + SCAN_ITEMS - 1) / SCAN_ITEMS + 1) * 4);
    g.sorted = (float4*)(b + o);      o += eags_align256((size_t)(n > 0 ? n : 1) * 16);
    g.sorted_idx = (unsigned*)(b + o);
    return g;
}

__device__ __forceinline__ unsigned enc(float f) {   // order-preserving float -> uint
    const unsigned b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float dec(unsigned u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}
__device__ __forceinline__ unsigned pack_cell(int x, int y, int z) {
    return (unsigned)x | ((unsigned)y << 10) | ((unsigned)z << 20);
}
__device__ __forceinline__ unsigned bucket_of(unsigned key, unsigned log2T) {
    return (key * 2654435761u) >> (32 - log2T);
}
__device__ __forceinline__ int cell_coord(float p, float o, float inv_h, int n) {
    float c = floorf((p - o) * inv_h);
    c = fminf(fmaxf(c, 0.f), (float)(n - 1));   // NaN -> 0
    return (int)c;
}
// squared distance in upstream's operation order (simple_knn.cu::updateKBest, contracted by nvcc)
__device__ __forceinline__ float dist2(float ax, float ay, float az, float bx, float by, float bz) {
    const float dx = ax - bx, dy = ay - by, dz = az - bz;
    return fmaf(dz, dz, fmaf(dy, dy, dx * dx));
}

// ---- grid construction ----------------------------------------------------------------------------
__global__ void grid_init_kernel(GridHdr* hdr, unsigned log2T) {
    if (threadIdx.x < 3) {
        hdr->lo[threadIdx.x] = 0xffffffffu;
        hdr->hi[threadIdx.x] = 0u;
    }
    if (threadIdx.x == 0) {
        hdr->n_occ = 0;
        hdr->log2_table = log2T;
    }
}

__global__ void __launch_bounds__(NB_THREADS)
bbox_kernel(const float* __restrict__ pts, int n, GridHdr* hdr) {
    unsigned lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0u, 0u, 0u};
    for (int i = blockIdx.x * NB_THREADS + threadIdx.x; i < n; i += gridDim.x * NB_THREADS) {
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const float v = pts[(size_t)3 * i + k];
            if (isfinite(v)) {
                const unsigned e = enc(v);
                lo[k] = min(lo[k], e);
                hi[k] = max(hi[k], e);
            }
        }
    }
#pragma unroll
    for (int k = 0; k < 3; k++) {
        lo[k] = __reduce_min_sync(FULLMASK, lo[k]);
        hi[k] = __reduce_max_sync(FULLMASK, hi[k]);
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int k = 0; k < 3; k++) {
            if (lo[k] != 0xffffffffu) atomicMin(&hdr->lo[k], lo[k]);
            if (hi[k] != 0u) atomicMax(&hdr->hi[k], hi[k]);
        }
    }
}

// One thread decides the cell size.  mode 0: first guess for the k-NN grid (one point per cell if the
// cloud filled its bounding box); mode 1: refine from the occupancy the first counting pass saw;
// mode 2: radius query, h = sqrt(radius) (+0.1 %: cell indices are computed in fp32, see
// cell_coord), so that every neighbour within the radius sits in the 27 surrounding cells.
__global__ void grid_setup_kernel(GridHdr* hdr, int n, int mode, float radius2) {
    float e[3], emax = 0.f;
    for (int k = 0; k < 3; k++) {
        const bool empty = hdr->lo[k] == 0xffffffffu;
        const float lo = empty ? 0.f : dec(hdr->lo[k]), hi = empty ? 0.f : dec(hdr->hi[k]);
        hdr->o[k] = lo;
        e[k] = hi - lo;
        emax = fmaxf(emax, e[k]);
    }
    float h;
    if (mode == 2) {
        h = sqrtf(fmaxf(radius2, 0.f)) * 1.001f;
    } else if (mode == 0) {
        const float floor_e = fmaxf(emax * 1e-3f, 1e-12f);
        const float vol = fmaxf(e[0], floor_e) * fmaxf(e[1], floor_e) * fmaxf(e[2], floor_e);
        h = cbrtf(vol / (float)(n > 0 ? n : 1));
    } else {
        h = hdr->h;
        const float avg = (float)n / (float)(hdr->n_occ > 0 ? hdr->n_occ : 1);
        if (avg > 6.f) h *= sqrtf(3.f / avg);   // surface-like: occupancy scales with h^2
    }
    h = fmaxf(h, emax / (float)(MAX_DIM - 4));   // at most MAX_DIM cells per axis
    if (!(h > 0.f) || !isfinite(h)) h = 1.f;     // all points coincide (or no finite point)
    hdr->h = h;
    hdr->inv_h = 1.f / h;
    for (int k = 0; k < 3; k++) {
        int c = (int)floorf(e[k] / h) + 2;
        hdr->n[k] = min(MAX_DIM, max(1, c));
    }
    hdr->n_occ = 0;
}

__global__ void __launch_bounds__(NB_THREADS)
cell_count_kernel(const float* __restrict__ pts, int n, GridPtrs g) {
    const int i = blockIdx.x * NB_THREADS + threadIdx.x;
    if (i >= n) return;
    const GridHdr* h = g.hdr;
    const int cx = cell_coord(pts[(size_t)3 * i], h->o[0], h->inv_h, h->n[0]);
    const int cy = cell_coord(pts[(size_t)3 * i + 1], h->o[1], h->inv_h, h->n[1]);
    const int cz = cell_coord(pts[(size_t)3 * i + 2], h->o[2], h->inv_h, h->n[2]);
    const unsigned b = bucket_of(pack_cell(cx, cy, cz), g.log2T);
    if (atomicAdd(&g.start[b], 1u) == 0u) atomicAdd(&g.hdr->n_occ, 1u);
}

// block-wide exclusive scan of 8 items per thread (256 threads); returns the block total
__device__ __forceinline__ unsigned block_scan8(unsigned v[8], unsigned* s_warp) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned t = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) { const unsigned x = v[k]; v[k] = t; t += x; }
    unsigned incl = t;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned u = __shfl_up_sync(FULLMASK, incl, o);
        if (lane >= o) incl += u;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        const unsigned w = lane < NB_THREADS / 32 ? s_warp[lane] : 0u;
        unsigned wi = w;
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
            const unsigned u = __shfl_up_sync(FULLMASK, wi, o);
            if (lane >= o) wi += u;
        }
        if (lane < NB_THREADS / 32) s_warp[lane] = wi - w;
        if (lane == NB_THREADS / 32 - 1) s_warp[NB_THREADS / 32] = wi;
    }
    __syncthreads();
    const unsigned base = s_warp[warp] + incl - t;
#pragma unroll
    for (int k = 0; k < 8; k++) v[k] += base;
    return s_warp[NB_THREADS / 32];
}

__global__ void __launch_bounds__(NB_THREADS)
scan_block_sums_kernel(const unsigned* __restrict__ a, unsigned n, unsigned* __restrict__ block_sums) {
    __shared__ unsigned s_warp[NB_THREADS / 32 + 1];
    const unsigned base = blockIdx.x * SCAN_ITEMS + threadIdx.x * 8;
    unsigned v[8];
#pragma unroll
    for (int k = 0; k < 8; k++) v[k] = base + k < n ? a[base + k] : 0u;
    const unsigned tot = block_scan8(v, s_warp);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = tot;
}
__global__ void __launch_bounds__(NB_THREADS)
scan_top_kernel(unsigned* __restrict__ block_sums, unsigned nb) {
    __shared__ unsigned s_warp[NB_THREADS / 32 + 1];
    __shared__ unsigned s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (unsigned c = 0; c < nb; c += SCAN_ITEMS) {
        const unsigned base = c + threadIdx.x * 8;
        unsigned v[8];
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = base + k < nb ? block_sums[base + k] : 0u;
        const unsigned tot = block_scan8(v, s_warp);
        const unsigned carry = s_carry;
#pragma unroll
        for (int k = 0; k < 8; k++)
            if (base + k < nb) block_sums[base + k] = v[k] + carry;
        __syncthreads();
        if (threadIdx.x == 0) s_carry = carry + tot;
        __syncthreads();
    }
}
__global__ void __launch_bounds__(NB_THREADS)
scan_apply_kernel(unsigned* __restrict__ a, unsigned n, const unsigned* __restrict__ block_sums) {
    __shared__ unsigned s_warp[NB_THREADS / 32 + 1];
    const unsigned base = blockIdx.x * SCAN_ITEMS + threadIdx.x * 8;
    unsigned v[8];
#pragma unroll
    for (int k = 0; k < 8; k++) v[k] = base + k < n ? a[base + k] : 0u;
    block_scan8(v, s_warp);
    const unsigned off = block_sums[blockIdx.x];
#pragma unroll
    for (int k = 0; k < 8; k++)
        if (base + k < n) a[base + k] = v[k] + off;
}

__global__ void __launch_bounds__(NB_THREADS)
cell_scatter_kernel(const float* __restrict__ pts, int n, GridPtrs g) {
    const int i = blockIdx.x * NB_THREADS + threadIdx.x;
    if (i >= n) return;
    const GridHdr* h = g.hdr;
    const float x = pts[(size_t)3 * i], y = pts[(size_t)3 * i + 1], z = pts[(size_t)3 * i + 2];
    const unsigned key = pack_cell(cell_coord(x, h->o[0], h->inv_h, h->n[0]),
                                   cell_coord(y, h->o[1], h->inv_h, h->n[1]),
                                   cell_coord(z, h->o[2], h->inv_h, h->n[2]));
    const unsigned b = bucket_of(key, g.log2T);
    const unsigned pos = g.start[b] + atomicAdd(&g.cursor[b], 1u);
    g.sorted[pos] = make_float4(x, y, z, __uint_as_float(key));
    g.sorted_idx[pos] = (unsigned)i;
}

// ---- queries ------------------------------------------------------------------------------------------
__device__ __forceinline__ void insert3(float d, float best[3]) {   // upstream updateKBest<3>
#pragma unroll
    for (int j = 0; j < 3; j++) {
        if (best[j] > d) { const float t = best[j]; best[j] = d; d = t; }
    }
}

// Ring search of one grid level around p.  `self_pos` (level 0: position in this grid's bucket order)
// or `self_idx` (coarser levels: original index, checked only for zero distances) excludes the query
// itself exactly once.  `bound` is an upper bound of the 3rd-nearest squared distance (FLT_MAX if
// unknown): cells farther away are skipped.  Returns true when `best` is final.
template <bool LEVEL0>
__device__ __forceinline__ bool ring_search(const GridPtrs& g, float px, float py, float pz, unsigned self_pos,
                                            unsigned self_idx, float bound, float best[3]) {
    const GridHdr* hd = g.hdr;
    const int nx = hd->n[0], ny = hd->n[1], nz = hd->n[2];
    const float h = hd->h, inv_h = hd->inv_h, ox = hd->o[0], oy = hd->o[1], oz = hd->o[2];
    const int cx = cell_coord(px, ox, inv_h, nx), cy = cell_coord(py, oy, inv_h, ny),
              cz = cell_coord(pz, oz, inv_h, nz);
    const float slack = 1e-3f * h;   // fp32 rounding of cell_coord
    for (int r = 0; r <= MAX_RING; r++) {
        for (int dz = -r; dz <= r; dz++) {
            const int z = cz + dz;
            if (z < 0 || z >= nz) continue;
            for (int dy = -r; dy <= r; dy++) {
                const int y = cy + dy;
                if (y < 0 || y >= ny) continue;
                const bool shell_yz = (dz == -r || dz == r || dy == -r || dy == r);
                for (int dx = -r; dx <= r; dx += (shell_yz || r == 0) ? 1 : 2 * r) {
                    const int x = cx + dx;
                    if (x < 0 || x >= nx) continue;
                    if (!LEVEL0) {   // sphere / cell-box rejection with the bound carried from finer levels
                        const float lx = ox + (float)x * h, ly = oy + (float)y * h, lz = oz + (float)z * h;
                        const float ex = fmaxf(0.f, fmaxf(lx - px, px - (lx + h)) - slack);
                        const float ey = fmaxf(0.f, fmaxf(ly - py, py - (ly + h)) - slack);
                        const float ez = fmaxf(0.f, fmaxf(lz - pz, pz - (lz + h)) - slack);
                        if (ex * ex + ey * ey + ez * ez > fminf(bound, best[2])) continue;
                    }
                    const unsigned key = pack_cell(x, y, z);
                    const unsigned b = bucket_of(key, g.log2T);
                    const unsigned beg = g.start[b], end = g.start[b + 1];
                    for (unsigned j = beg; j < end; j++) {
                        const float4 q = g.sorted[j];
                        if (__float_as_uint(q.w) != key) continue;
                        if (LEVEL0 && j == self_pos) continue;
                        const float d = dist2(q.x, q.y, q.z, px, py, pz);
                        if (!LEVEL0 && d == 0.f && g.sorted_idx[j] == self_idx) continue;
                        insert3(d, best);
                    }
                }
            }
        }
        // everything outside the searched cube is at least d_min away (faces beyond the grid hold no points)
        float dmin = FLT_MAX;
        if (cx - r > 0) dmin = fminf(dmin, px - (ox + (float)(cx - r) * h));
        if (cx + r < nx - 1) dmin = fminf(dmin, (ox + (float)(cx + r + 1) * h) - px);
        if (cy - r > 0) dmin = fminf(dmin, py - (oy + (float)(cy - r) * h));
        if (cy + r < ny - 1) dmin = fminf(dmin, (oy + (float)(cy + r + 1) * h) - py);
        if (cz - r > 0) dmin = fminf(dmin, pz - (oz + (float)(cz - r) * h));
        if (cz + r < nz - 1) dmin = fminf(dmin, (oz + (float)(cz + r + 1) * h) - pz);
        if (dmin == FLT_MAX) return true;   // the cube covers the whole grid
        dmin = fmaxf(0.f, dmin - slack);
        if (fminf(best[2], bound) <= dmin * dmin) return true;
    }
    return false;
}

struct Pending {          // queries a level could not settle, handed to the next coarser level
    unsigned* count;      // [1]
    uint2* items;         // (position in level-0 bucket order, bits(bound))
};

// level 0: every point, in bucket order (a warp's queries walk the same buckets)
__global__ void __launch_bounds__(NB_THREADS)
knn3_query_kernel(int n, GridPtrs g, Pending next, float* __restrict__ out) {
    const int s = blockIdx.x * NB_THREADS + threadIdx.x;
    if (s >= n) return;
    const float4 p = g.sorted[s];
    float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    if (ring_search<true>(g, p.x, p.y, p.z, (unsigned)s, 0u, FLT_MAX, best)) {
        out[g.sorted_idx[s]] = (best[0] + best[1] + best[2]) / 3.0f;
    } else {
        next.items[atomicAdd(next.count, 1u)] = make_uint2((unsigned)s, __float_as_uint(best[2]));
    }
}

// coarser levels: only the queries still pending; LAST scans the whole cloud for what remains
template <bool LAST>
__global__ void __launch_bounds__(NB_THREADS)
knn3_pending_kernel(int n, GridPtrs g0, GridPtrs g, Pending cur, Pending next, float* __restrict__ out) {
    const unsigned cnt = *cur.count;
    for (unsigned t = blockIdx.x * NB_THREADS + threadIdx.x; t < cnt; t += gridDim.x * NB_THREADS) {
        const uint2 it = cur.items[t];
        const float4 p = g0.sorted[it.x];
        const unsigned self_idx = g0.sorted_idx[it.x];
        float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
        bool settled = ring_search<false>(g, p.x, p.y, p.z, 0u, self_idx, __uint_as_float(it.y), best);
        if (!settled) {
            if (LAST) {   // isolated point: exact scan of the whole cloud
                best[0] = best[1] = best[2] = FLT_MAX;
                for (int j = 0; j < n; j++) {
                    if (j == (int)it.x) continue;
                    const float4 q = g0.sorted[j];
                    insert3(dist2(q.x, q.y, q.z, p.x, p.y, p.z), best);
                }
                settled = true;
            } else {
                next.items[atomicAdd(next.count, 1u)] = make_uint2(it.x, __float_as_uint(fminf(best[2], __uint_as_float(it.y))));
            }
        }
        if (settled) out[self_idx] = (best[0] + best[1] + best[2]) / 3.0f;
    }
}

// header of a coarser level: same box and origin, `factor` x the cell size
__global__ void grid_derive_kernel(const GridHdr* base, GridHdr* hdr, float factor, unsigned log2T) {
    float emax = 0.f;
    for (int k = 0; k < 3; k++) {
        hdr->lo[k] = base->lo[k];
        hdr->hi[k] = base->hi[k];
        hdr->o[k] = base->o[k];
    }
    const float h = base->h * factor;
    hdr->h = h;
    hdr->inv_h = 1.f / h;
    for (int k = 0; k < 3; k++) {
        const bool empty = base->lo[k] == 0xffffffffu;
        const float e = empty ? 0.f : dec(base->hi[k]) - dec(base->lo[k]);
        emax = fmaxf(emax, e);
        hdr->n[k] = min(MAX_DIM, max(1, (int)floorf(e / h) + 2));
    }
    hdr->n_occ = 0;
    hdr->log2_table = log2T;
    (void)emax;
}

__global__ void __launch_bounds__(NB_THREADS)
radius_query_kernel(const float* __restrict__ qpts, int nq, GridPtrs g, float radius2,
                    uint8_t* __restrict__ has_neighbor) {
    const int i = blockIdx.x * NB_THREADS + threadIdx.x;
    if (i >= nq) return;
    const GridHdr* hd = g.hdr;
    const float px = qpts[(size_t)3 * i], py = qpts[(size_t)3 * i + 1], pz = qpts[(size_t)3 * i + 2];
    const int nx = hd->n[0], ny = hd->n[1], nz = hd->n[2];
    // unclamped cell of the query (it may lie outside the reference cloud's box)
    const float fx = fminf(fmaxf(floorf((px - hd->o[0]) * hd->inv_h), -2.f), (float)MAX_DIM + 1.f);
    const float fy = fminf(fmaxf(floorf((py - hd->o[1]) * hd->inv_h), -2.f), (float)MAX_DIM + 1.f);
    const float fz = fminf(fmaxf(floorf((pz - hd->o[2]) * hd->inv_h), -2.f), (float)MAX_DIM + 1.f);
    const int cx = (int)fx, cy = (int)fy, cz = (int)fz;
    bool found = false;
    for (int dz = -1; dz <= 1 && !found; dz++) {
        const int z = cz + dz;
        if (z < 0 || z >= nz) continue;
        for (int dy = -1; dy <= 1 && !found; dy++) {
            const int y = cy + dy;
            if (y < 0 || y >= ny) continue;
            for (int dx = -1; dx <= 1 && !found; dx++) {
                const int x = cx + dx;
                if (x < 0 || x >= nx) continue;
                const unsigned key = pack_cell(x, y, z);
                const unsigned b = bucket_of(key, g.log2T);
                const unsigned beg = g.start[b], end = g.start[b + 1];
                for (unsigned j = beg; j < end; j++) {
                    const float4 q = g.sorted[j];
                    if (__float_as_uint(q.w) != key) continue;
                    if (dist2(q.x, q.y, q.z, px, py, pz) < radius2) { found = true; break; }
                }
            }
        }
    }
    has_neighbor[i] = found ? 1 : 0;
}

#define LAUNCH_OK() (cudaPeekAtLastError() == cudaSuccess)

// count -> scan -> scatter of `n` points on the grid described by the header; returns launches or -1
int build_cells(const float* pts, int n, GridPtrs g, bool scatter, cudaStream_t s) {
    int launches = 0;
    const int blocks = (n + NB_THREADS - 1) / NB_THREADS;
    if (cudaMemsetAsync(g.start, 0, ((size_t)g.T + 1) * 4, s) != cudaSuccess) return -1;
    cell_count_kernel<<<blocks, NB_THREADS, 0, s>>>(pts, n, g);
    if (!LAUNCH_OK()) return -1;
    launches++;
    if (!scatter) return launches;
    const unsigned ne = g.T + 1, nb = (ne + SCAN_ITEMS - 1) / SCAN_ITEMS;
    scan_block_sums_kernel<<<nb, NB_THREADS, 0, s>>>(g.start, ne, g.block_sums);
    scan_top_kernel<<<1, NB_THREADS, 0, s>>>(g.block_sums, nb);
    scan_apply_kernel<<<nb, NB_THREADS, 0, s>>>(g.start, ne, g.block_sums);
    if (!LAUNCH_OK()) return -1;
    if (cudaMemsetAsync(g.cursor, 0, (size_t)g.T * 4, s) != cudaSuccess) return -1;
    cell_scatter_kernel<<<blocks, NB_THREADS, 0, s>>>(pts, n, g);
    if (!LAUNCH_OK()) return -1;
    return launches + 4;
}

}  // namespace

constexpr int KNN_LEVELS = 3;
constexpr float KNN_LEVEL_FACTOR = 5.f;   // 4 rings reach 4 h, 20 h, 100 h
static size_t pending_bytes(long long n) { return 256 + eags_align256((size_t)(n > 0 ? n : 1) * 8); }

size_t neighbors_workspace_bytes(long long n_ref) {
    return KNN_LEVELS * grid_bytes(n_ref) + (KNN_LEVELS - 1) * pending_bytes(n_ref);
}

int launch_knn_mean_dist2(const float* pts, int n, float* out, void* ws, cudaStream_t s) {
    if (n <= 0) return 0;
    GridPtrs g[KNN_LEVELS];
    Pending pend[KNN_LEVELS];   // pend[l]: input of level l (l >= 1)
    char* b = (char*)ws;
    for (int l = 0; l < KNN_LEVELS; l++) { g[l] = grid_ptrs(b, n); b += grid_bytes(n); }
    pend[0].count = nullptr; pend[0].items = nullptr;
    for (int l = 1; l < KNN_LEVELS; l++) {
        pend[l].count = (unsigned*)b;
        pend[l].items = (uint2*)(b + 256);
        b += pending_bytes(n);
    }
    int launches = 0, k;
    const int blocks = (n + NB_THREADS - 1) / NB_THREADS;
    grid_init_kernel<<<1, 32, 0, s>>>(g[0].hdr, g[0].log2T);
    bbox_kernel<<<min(blocks, 4 * 148), NB_THREADS, 0, s>>>(pts, n, g[0].hdr);
    grid_setup_kernel<<<1, 1, 0, s>>>(g[0].hdr, n, 0, 0.f);
    if (!LAUNCH_OK()) return -1;
    launches += 3;
    if ((k = build_cells(pts, n, g[0], false, s)) < 0) return -1;   // occupancy probe
    launches += k;
    grid_setup_kernel<<<1, 1, 0, s>>>(g[0].hdr, n, 1, 0.f);
    if (!LAUNCH_OK()) return -1;
    launches++;
    float factor = 1.f;
    for (int l = 0; l < KNN_LEVELS; l++) {
        if (l > 0) {
            factor *= KNN_LEVEL_FACTOR;
            grid_derive_kernel<<<1, 1, 0, s>>>(g[0].hdr, g[l].hdr, factor, g[l].log2T);
            if (cudaMemsetAsync(pend[l].count, 0, 4, s) != cudaSuccess) return -1;
            launches++;
        }
        if ((k = build_cells(pts, n, g[l], true, s)) < 0) return -1;
        launches += k;
    }
    knn3_query_kernel<<<blocks, NB_THREADS, 0, s>>>(n, g[0], pend[1], out);
    const int pblocks = min(blocks, 8 * 148);
    for (int l = 1; l < KNN_LEVELS; l++) {
        if (l == KNN_LEVELS - 1)
            knn3_pending_kernel<true><<<pblocks, NB_THREADS, 0, s>>>(n, g[0], g[l], pend[l], pend[l], out);
        else
            knn3_pending_kernel<false><<<pblocks, NB_THREADS, 0, s>>>(n, g[0], g[l], pend[l], pend[l + 1], out);
    }
    if (!LAUNCH_OK()) return -1;
    return launches + KNN_LEVELS;
}

int launch_radius_has_neighbor(const float* ref, int n_ref, const float* query, int n_query,
                               float radius2, uint8_t* has_neighbor, void* ws, cudaStream_t s) {
    if (n_query <= 0) return 0;
    if (n_ref <= 0) {
        return cudaMemsetAsync(has_neighbor, 0, (size_t)n_query, s) == cudaSuccess ? 0 : -1;
    }
    GridPtrs g = grid_ptrs(ws, n_ref);
    int launches = 0, k;
    grid_init_kernel<<<1, 32, 0, s>>>(g.hdr, g.log2T);
    bbox_kernel<<<min((n_ref + NB_THREADS - 1) / NB_THREADS, 4 * 148), NB_THREADS, 0, s>>>(ref, n_ref, g.hdr);
    grid_setup_kernel<<<1, 1, 0, s>>>(g.hdr, n_ref, 2, radius2);
    if (!LAUNCH_OK()) return -1;
    launches += 3;
    if ((k = build_cells(ref, n_ref, g, true, s)) < 0) return -1;
    launches += k;
    radius_query_kernel<<<(n_query + NB_THREADS - 1) / NB_THREADS, NB_THREADS, 0, s>>>(
        query, n_query, g, radius2, has_neighbor);
    if (!LAUNCH_OK()) return -1;
    return launches + 1;
}
