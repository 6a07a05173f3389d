// blend.cu -- per-16x16-tile alpha blending, forward and backward (sm_100a).
//
//   K6 blend_fwd : front-to-back blend of colour + depth + alpha/opacity (+ n_touched in variant B);
//                  upstream forward.cu::renderCUDA, SURVEY.md A.3.
//   K7 blend_bwd : back-to-front replay producing dL/d{mean2D, conic, opacity, colour, depth} per
//                  Gaussian; upstream backward.cu::renderCUDA, SURVEY.md A.4.
//
// Both kernels read the *sorted, packed splat stream* written by sort_pack (binning.cu): a tile's
// Gaussians are 48-byte records laid out contiguously in blend order, so a batch is one linear
// cp.async.bulk (TMA bulk copy, UBLKCP) into shared memory, completion-tracked by an mbarrier and
// multi-buffered so the copy of batch b+2 overlaps the blending of batch b.  Upstream instead gathers
// id -> xy / conic / colour / depth from four arrays behind a __syncthreads per batch.
//
// Each warp owns an 8x4 pixel block.  Per batch it culls one splat per lane against the block (exact-arithmetic-
// independent, conservative), and blend_fwd then walks a compacted list of the survivors' shared-memory record
// addresses (one broadcast LDS per survivor instead of a bitmask walk); blend_bwd walks the ballot masks.
//
// blend_bwd replaces upstream's 10 global atomics per (pixel, Gaussian) pair by
//   (1) a pixel-per-lane pass that forms only two scalars per pair (gG = dL/dG * G and w = alpha * T;
//       dL/dalpha in suffix-sum form needs one running scalar instead of five accumulators),
//   (2) a transposed, splat-per-lane pass that sums the ten gradient moments of a splat over the
//       warp's 32 pixels in registers (flush_slots), and
//   (3) three vector atomics per (8x4 block, Gaussian) into a 48-byte accumulator row.
// The walk starts at the tile's max n_contrib instead of the end of the list.
//
// Tensor cores are not used: blending is a per-pixel ordered recurrence, not a contraction; the
// bound is the FP32/issue pipe.
#include "kernels.h"

namespace {

constexpr int BLK = 128;          // 4 warps = one 16x8 half tile per work unit
constexpr int WARPS = BLK / 32;
constexpr int UNITS_PER_TILE = 8 / WARPS;
constexpr unsigned FULL = 0xffffffffu;

// ---- mbarrier / TMA bulk-copy helpers (PTX) ------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    } while (!done);
}

// ---- splat records (written by sort_pack, binning.cu) ---------------------------------------------
//   R0 = {x, y, bits(gaussian index), depth}
//   R1 = {A, B, C, opacity}   with  log2(G) = A dx^2 + B dx dy + C dy^2,
//                                   A = -0.5 log2(e) conic.xx, B = -log2(e) conic.xy, C = -0.5 log2(e) conic.yy
//   R2 = {r, g, b, tau2}      tau2 = log2(255 opacity): the splat reaches alpha >= 1/255 iff -log2(G) <= tau2
constexpr float LN2 = 0.6931471805599453f;
__device__ __forceinline__ float fast_rcp(float x) {    // MUFU.RCP, 1 ulp
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float fast_exp2(float x) {   // MUFU.EX2, 2 ulp, flushes to 0 below 2^-126
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// 16-byte shared-memory load from a 32-bit shared address (+ immediate): the survivor lists below hold the
// records' shared addresses, so a record is read without any address arithmetic
template <int OFF>
__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4+%5];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr), "n"(OFF));
    return v;
}
constexpr float CULL_MARGIN = 0.02f;   // log2 units (1.4 %) >> any fp32 rounding of the pixel test

// A 16x16 tile is eight 8x4 pixel blocks; block s sits at (8 (s & 1), 4 (s >> 1)); lane l -> (l & 7, l >> 3).
// A work unit is half a tile: warp w of the CTA owns block s = 4 * half + w.
__device__ __forceinline__ void warp_region(int tile, int gx, int sub, int lane, int& px, int& py,
                                            float& rx0, float& ry0) {
    const int tx = tile % gx, ty = tile / gx;
    const int ox = tx * EAGS_TILE + (sub & 1) * 8, oy = ty * EAGS_TILE + (sub >> 1) * 4;
    px = ox + (lane & 7);
    py = oy + (lane >> 3);
    rx0 = (float)ox;
    ry0 = (float)oy;
}

// Conservative, exact-arithmetic-independent cull: can the splat reach alpha >= 1/255 anywhere in the
// continuous rectangle [rx0, rx0+7] x [ry0, ry0+3]?  q(d) = -log2 G(d) is a convex quadratic; its
// minimum over the rectangle is 0 if the mean is inside, else it lies on an edge facing the mean,
// where the 1-D minimiser is a clamp.  Non-positive-definite conics are never culled.
__device__ __forceinline__ bool splat_may_touch(const float4 r0, const float4 r1, float tau2, float rx0,
                                                float ry0) {
    const float a2 = -2.f * r1.x, b2 = -r1.y, c2 = -2.f * r1.z;
    const float lx = rx0 - r0.x, ux = lx + 7.f, ly = ry0 - r0.y, uy = ly + 3.f;
    const float ex = fminf(fmaxf(0.f, lx), ux), ey = fminf(fmaxf(0.f, ly), uy);
    const float dy1 = fminf(fmaxf(__fdividef(-b2 * ex, c2), ly), uy);
    const float q1 = 0.5f * a2 * ex * ex + b2 * ex * dy1 + 0.5f * c2 * dy1 * dy1;
    const float dx2 = fminf(fmaxf(__fdividef(-b2 * ey, a2), lx), ux);
    const float q2 = 0.5f * a2 * dx2 * dx2 + b2 * dx2 * ey + 0.5f * c2 * ey * ey;
    const bool pd = a2 > 0.f && c2 > 0.f && a2 * c2 - b2 * b2 > 0.f;
    return !(pd && fminf(q1, q2) > tau2 + CULL_MARGIN);
}

// ================================================================================================
// K6  blend forward
// ================================================================================================
// Persistent CTAs pull half-tile work units from an atomic queue: no wave quantisation (1 200 tiles
// on 148 SMs x 6 resident 256-thread CTAs was 1.35 waves = 2 rounds) and the tail is at most one
// half tile long.  The mbarrier ring keeps running across units (`it` = batches issued so far).
constexpr int FWD_BATCH = 128;   // records per stage (6 KB)
constexpr int FWD_STAGES = 3;
constexpr int FWD_ROUNDS = FWD_BATCH / 32;

struct FwdArgs {
    int W, H, gx, n_units;
    const EagsCtrl* ctrl;
    EagsCtrl* stats;
    unsigned* work;
    long long capacity;
    unsigned tile_cap;
    const uint2* ranges;
    const float4* records;
    const float* bg;
    float* out_color;
    float* out_depth;
    float* out_third;
    int32_t* n_touched;
    float* final_T;
    uint32_t* n_contrib;
    uint32_t* tile_max;
};

// STATS ("kernel_stats" option, bench.py only): count the (block, splat) iterations of pass 2 and the
// (pixel, splat) pairs that were blended into the control block.
template <bool VARIANT_B, bool TMA, bool STATS>
__global__ void __launch_bounds__(BLK)
blend_fwd_kernel(FwdArgs a) {
    __shared__ __align__(128) float4 s_rec[FWD_STAGES][FWD_BATCH * 3];
    __shared__ __align__(8) uint64_t s_bar[FWD_STAGES];
    __shared__ int s_touched[VARIANT_B ? FWD_BATCH : 1];
    __shared__ int s_unit;
    __shared__ uint32_t s_list[WARPS][FWD_BATCH];   // shared addresses of the batch's survivors, in blend order

    if (a.ctrl->num_rendered > (unsigned long long)a.capacity) return;
    if (a.tile_cap && a.ctrl->bucket_overflow) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool count_touched = VARIANT_B && a.n_touched != nullptr;
    if (TMA && tid == 0) {
        for (int s = 0; s < FWD_STAGES; s++) mbar_init(&s_bar[s], 1);
        mbar_fence_init();
    }
    if (VARIANT_B) s_touched[tid] = 0;   // FWD_BATCH == BLK
    unsigned it = 0;   // batches issued to the mbarrier ring so far (identical in every thread)
    unsigned st_iters = 0, st_pairs = 0;

    for (;;) {
        if (tid == 0) s_unit = (int)atomicAdd(a.work, 1u);
        __syncthreads();
        const int unit = s_unit;
        __syncthreads();
        if (unit >= a.n_units) break;
        const int tile = unit / UNITS_PER_TILE, half = unit % UNITS_PER_TILE;
        const uint2 rg = a.ranges[tile];
        const int n = (int)(rg.y - rg.x);
        EAGS_CHECK(rg.y >= rg.x && (long long)rg.y <= a.capacity);
        int px, py;
        float rx0, ry0;
        warp_region(tile, a.gx, half * WARPS + warp, lane, px, py, rx0, ry0);
        const bool inside = px < a.W && py < a.H;
        const float pxf = (float)px, pyf = (float)py;
        const float4* src = a.records + (size_t)3 * rg.x;
        const int nb = (n + FWD_BATCH - 1) / FWD_BATCH;

        if (TMA && tid == 0) {
            for (int b = 0; b < FWD_STAGES && b < nb; b++) {
                const int cnt = min(FWD_BATCH, n - b * FWD_BATCH);
                const int st = (int)((it + b) % FWD_STAGES);
                mbar_expect_tx(&s_bar[st], (uint32_t)cnt * 48u);
                tma_bulk_g2s(&s_rec[st][0], src + (size_t)3 * b * FWD_BATCH, (uint32_t)cnt * 48u, &s_bar[st]);
            }
        }

        // T carries the "done" flag in its sign: a live pixel has T >= 1e-4 > 0; termination (or a pixel
        // outside the image) flips the sign and keeps the magnitude, so the saturation test below
        // (T (1 - alpha) < 1e-4) also holds for finished pixels and no separate predicate is needed.
        float T = inside ? 1.f : -1.f, C0 = 0.f, C1 = 0.f, C2 = 0.f, D = 0.f, A = 0.f;
        unsigned last = 0;       // 1 + list position of the pixel's last contributor
        int issued = nb;

        for (int b = 0; b < nb; b++) {
            const int base = b * FWD_BATCH;
            const int cnt = min(FWD_BATCH, n - base);
            int stage;
            if (TMA) {
                stage = (int)((it + b) % FWD_STAGES);
                mbar_wait(&s_bar[stage], (uint32_t)(((it + b) / FWD_STAGES) & 1));
                __syncwarp();   // the try_wait loop exits per thread: reconverge before warp-wide code
            } else {
                stage = 0;
                for (int k = tid; k < cnt * 3; k += BLK) s_rec[0][k] = src[(size_t)3 * base + k];
                __syncthreads();
            }
            const float4* rec = s_rec[stage];
            if (!__all_sync(FULL, T < 0.f)) {
                // pass 1: one splat per lane; the survivors' record addresses are compacted into a per-warp list
                // (ballot + popc prefix).  Pass 2 then costs one 4-byte broadcast load per survivor instead of
                // the seven uniform-datapath / integer instructions of walking a bitmask (brev, flo, add, and,
                // setp, mov, imad): the loop is issue-bound, so they were 1/5 of it.
                uint32_t* list = s_list[warp];
                const uint32_t rec_u32 = smem_u32(rec);
                const unsigned lt = (1u << lane) - 1u;
                int nsurv = 0;
#pragma unroll
                for (int r = 0; r < FWD_ROUNDS; r++) {
                    const int jj = r * 32 + lane;
                    bool keep = false;
                    if (jj < cnt) keep = splat_may_touch(rec[3 * jj], rec[3 * jj + 1], rec[3 * jj + 2].w, rx0, ry0);
                    const unsigned bal = __ballot_sync(FULL, keep);
                    if (keep) list[nsurv + __popc(bal & lt)] = rec_u32 + 48u * (unsigned)jj;
                    nsurv += __popc(bal);
                }
                __syncwarp();
                uint32_t last_ad = 0;
                auto blend_one = [&](const uint32_t ad) {
                    EAGS_CHECK(ad >= rec_u32 && ad < rec_u32 + 48u * (unsigned)cnt && (ad - rec_u32) % 48u == 0);
                    const float4 r0 = lds128<0>(ad), r1 = lds128<16>(ad);
                    const float dx = r0.x - pxf, dy = r0.y - pyf;
                    const float power2 = (r1.x * dx + r1.y * dy) * dx + (r1.z * dy) * dy;
                    const float alpha = fminf(0.99f, r1.w * fast_exp2(power2));
                    if (STATS) st_iters += (lane == 0);
                    const bool hit = !(power2 > 0.f || alpha < 1.0f / 255.0f);
                    const float test_T = T * (1.f - alpha);
                    const bool blend = hit && !(test_T < 0.0001f);
                    if (hit && !blend) T = -fabsf(T);
                    if (blend) {
                        if (STATS) st_pairs++;
                        const float4 r2 = lds128<32>(ad);
                        const float w = alpha * T;
                        C0 += r2.x * w; C1 += r2.y * w; C2 += r2.z * w;
                        D += r0.w * w;
                        A += w;
                        if (VARIANT_B) {
                            if (count_touched && test_T > 0.5f) atomicAdd(&s_touched[(ad - rec_u32) / 48u], 1);   // Q10
                        }
                        T = test_T;
                        last_ad = ad;
                    }
                };
                int k = 0;
                for (const int head = nsurv & 3; k < head; k++) blend_one(list[k]);
                for (; k < nsurv; k += 4) {
                    blend_one(list[k]);
                    blend_one(list[k + 1]);
                    blend_one(list[k + 2]);
                    blend_one(list[k + 3]);
                }
                if (last_ad) last = (unsigned)base + (last_ad - rec_u32) / 48u + 1u;
            }
            const int ndone = __syncthreads_count(T < 0.f ? 1 : 0);
            if (VARIANT_B && count_touched) {
                if (tid < cnt) {
                    const int v = s_touched[tid];
                    if (v) {
                        atomicAdd(&a.n_touched[__float_as_uint(rec[3 * tid].z)], v);
                        s_touched[tid] = 0;
                    }
                }
                __syncthreads();
            }
            if (ndone == BLK) {
                issued = min(nb, b + FWD_STAGES);
                if (TMA)   // consume the copies still in flight: their stages are reused by the next unit
                    for (int bb = b + 1; bb < issued; bb++)
                        mbar_wait(&s_bar[(it + bb) % FWD_STAGES], (uint32_t)(((it + bb) / FWD_STAGES) & 1));
                break;
            }
            if (TMA && tid == 0 && b + FWD_STAGES < nb) {
                const int nbase = (b + FWD_STAGES) * FWD_BATCH;
                const int ncnt = min(FWD_BATCH, n - nbase);
                mbar_expect_tx(&s_bar[stage], (uint32_t)ncnt * 48u);
                tma_bulk_g2s(&s_rec[stage][0], src + (size_t)3 * nbase, (uint32_t)ncnt * 48u, &s_bar[stage]);
            }
        }
        it += (unsigned)issued;
        __syncwarp();

        T = fabsf(T);
        if (inside) {
            const size_t pix = (size_t)py * a.W + px, HW = (size_t)a.H * a.W;
            a.final_T[pix] = T;
            a.n_contrib[pix] = last;
            a.out_color[pix] = C0 + T * __ldg(a.bg);
            a.out_color[HW + pix] = C1 + T * __ldg(a.bg + 1);
            a.out_color[2 * HW + pix] = C2 + T * __ldg(a.bg + 2);
            a.out_depth[pix] = D;
            a.out_third[pix] = VARIANT_B ? 1.f - T : A;
        }
        const unsigned m = __reduce_max_sync(FULL, last);
        if (lane == 0 && m) atomicMax(&a.tile_max[tile], m);   // tile_scan zeroed it
    }
    if (STATS) {
        st_iters = __reduce_add_sync(FULL, st_iters);
        st_pairs = __reduce_add_sync(FULL, st_pairs);
        if (lane == 0) {
            atomicAdd(&a.stats->stat_fwd_warp_iters, (unsigned long long)st_iters);
            atomicAdd(&a.stats->stat_fwd_lane_pairs, (unsigned long long)st_pairs);
        }
    }
}

// ================================================================================================
// K7  blend backward
// ================================================================================================
constexpr int BWD_BATCH = 64;    // records per stage (3 KB); also the width of the per-warp bitmask
constexpr int BWD_STAGES = 3;
constexpr int BWD_ROUNDS = BWD_BATCH / 32;
constexpr int SLOTS = 16;        // (warp, splat) survivors buffered before a transposed accumulation
constexpr int SLOT_STRIDE = 33;  // float2 row stride: conflict-free row writes and column reads

struct BwdArgs {
    int W, H, gx, n_units;
    unsigned* work;
    const uint2* ranges;
    const float4* records;
    const uint32_t* tile_max;
    const float* final_T;
    const uint32_t* n_contrib;
    const float* bg;
    const float* dL_dcolor;
    const float* dL_ddepth;
    const float* dL_dthird;   // nullable
    float4* grad_acc;         // [P][3]
    unsigned long long* stats;   // STATS: {warp iterations, lane pairs}
};

// Transposed accumulation ("phase 2").  Phase 1 runs pixel-per-lane and leaves, for each of up to 16
// buffered (warp, splat) survivors, the two per-pixel scalars gG = dL/dG * G and w = alpha * T in
// shared memory.  Here the roles flip: lane (s, h) owns survivor s and sums its ten gradient moments
// over the 16 pixels of block half h in registers -- no cross-lane reduction except one xor-16 fold --
// then applies the per-splat constant factors (conic, 1/opacity, pixel scale) once and commits
// with three vector atomics.  Replaces a 12-shuffle / 20-select butterfly per (warp, splat).
__device__ __forceinline__ void flush_slots(const float2 (*gw)[SLOT_STRIDE], const float4* pixg,
                                            const float4* meta, int count, int lane, float rx0,
                                            float ry0, float ddelx_dx, float ddely_dy, float4* grad_acc) {
    __syncwarp();
    const int s = lane & 15, h = lane >> 4;
    float v[10];
#pragma unroll
    for (int k = 0; k < 10; k++) v[k] = 0.f;
    const bool mine = s < count;
    float4 q0 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (mine) {
        q0 = meta[2 * s];
        float dxc[8], dyr[2];
#pragma unroll
        for (int c = 0; c < 8; c++) dxc[c] = q0.x - (rx0 + (float)c);
#pragma unroll
        for (int r = 0; r < 2; r++) dyr[r] = q0.y - (ry0 + (float)(2 * h + r));
        const float2* row = gw[s] + 16 * h;
        const float4* pg = pixg + 16 * h;
        // dy is constant along a pixel row: the x-moments (1, dx, dx^2) are summed per row of 8 pixels and the
        // row's dy factors applied once -- 3 instead of 8 FP instructions per pixel for the six geometric moments
#pragma unroll
        for (int r = 0; r < 2; r++) {
            float R0 = 0.f, R1 = 0.f, R2 = 0.f;
#pragma unroll
            for (int c = 0; c < 8; c++) {
                const float2 e = row[8 * r + c];
                const float4 pc = pg[8 * r + c];
                const float m1 = e.x * dxc[c];
                R0 += e.x; R1 += m1; R2 = fmaf(m1, dxc[c], R2);
                v[6] = fmaf(e.y, pc.x, v[6]); v[7] = fmaf(e.y, pc.y, v[7]);
                v[8] = fmaf(e.y, pc.z, v[8]); v[9] = fmaf(e.y, pc.w, v[9]);
            }
            const float dy = dyr[r], t = R0 * dy;
            v[0] += R0; v[1] += R1; v[3] += R2;
            v[2] += t; v[4] = fmaf(R1, dy, v[4]); v[5] = fmaf(t, dy, v[5]);
        }
    }
#pragma unroll
    for (int k = 0; k < 10; k++) v[k] += __shfl_xor_sync(FULL, v[k], 16);
    if (mine) {
        const float4 q1 = meta[2 * s + 1];
        float4* dst = grad_acc + (size_t)3 * __float_as_uint(q0.z);
        if (h == 0) {
            // moments -> gradients: true conic from the log2-scaled record
            const float cxx = -2.f * LN2 * q1.x, cxy = -LN2 * q1.y, cyy = -2.f * LN2 * q1.z;
            const float dmx = -ddelx_dx * (cxx * v[1] + cxy * v[2]);
            const float dmy = -ddely_dy * (cyy * v[2] + cxy * v[1]);
            // {dmean2D.x, dmean2D.y, dconic.xx, dconic.xy (half, Q8)}
            atomicAdd(dst + 0, make_float4(dmx, dmy, -0.5f * v[3], -0.5f * v[4]));
        } else {
            // {dconic.yy, dopacity, dr, dg} {db, ddepth}
            atomicAdd(dst + 1, make_float4(-0.5f * v[5], __fdividef(v[0], q1.w), v[6], v[7]));
            atomicAdd(reinterpret_cast<float2*>(dst + 2), make_float2(v[8], v[9]));
        }
    }
    __syncwarp();
}

#ifndef BWD_MIN_CTAS
#define BWD_MIN_CTAS 7
#endif
template <bool TMA, bool STATS>
__global__ void __launch_bounds__(BLK, BWD_MIN_CTAS)
blend_bwd_kernel(BwdArgs a) {
    __shared__ __align__(128) float4 s_rec[BWD_STAGES][BWD_BATCH * 3];
    __shared__ __align__(8) uint64_t s_bar[BWD_STAGES];
    __shared__ __align__(16) float2 s_gw[WARPS][SLOTS][SLOT_STRIDE];
    __shared__ __align__(16) float4 s_pix[WARPS][32];
    __shared__ __align__(16) float4 s_meta[WARPS][SLOTS * 2];   // records R0, R1 of the buffered survivors
    __shared__ int s_unit;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (TMA && tid == 0) {
        for (int s = 0; s < BWD_STAGES; s++) mbar_init(&s_bar[s], 1);
        mbar_fence_init();
    }
    unsigned it = 0;
    unsigned st_iters = 0, st_pairs = 0;
    const float ddelx_dx = 0.5f * (float)a.W, ddely_dy = 0.5f * (float)a.H;

    for (;;) {
        if (tid == 0) s_unit = (int)atomicAdd(a.work, 1u);
        __syncthreads();
        const int unit = s_unit;
        __syncthreads();
        if (unit >= a.n_units) break;
        const int tile = unit / UNITS_PER_TILE, half = unit % UNITS_PER_TILE;
        const uint2 rg = a.ranges[tile];
        const int n = min((int)(rg.y - rg.x), (int)a.tile_max[tile]);   // nothing past max n_contrib matters
        if (n <= 0) continue;
        int px, py;
        float rx0, ry0;
        warp_region(tile, a.gx, half * WARPS + warp, lane, px, py, rx0, ry0);
        const bool inside = px < a.W && py < a.H;
        const float pxf = (float)px, pyf = (float)py;
        const float4* src = a.records + (size_t)3 * rg.x;
        const int nb = (n + BWD_BATCH - 1) / BWD_BATCH;

        if (TMA && tid == 0) {
            for (int bi = 0; bi < BWD_STAGES && bi < nb; bi++) {
                const int base = (nb - 1 - bi) * BWD_BATCH;
                const int cnt = min(BWD_BATCH, n - base);
                const int st = (int)((it + bi) % BWD_STAGES);
                mbar_expect_tx(&s_bar[st], (uint32_t)cnt * 48u);
                tma_bulk_g2s(&s_rec[st][0], src + (size_t)3 * base, (uint32_t)cnt * 48u, &s_bar[st]);
            }
        }

        // per-pixel state
        float T_final = 1.f, dC0 = 0.f, dC1 = 0.f, dC2 = 0.f, dD = 0.f, dA = 0.f;
        int last_contributor = 0;
        if (inside) {
            const size_t pix = (size_t)py * a.W + px, HW = (size_t)a.H * a.W;
            T_final = a.final_T[pix];
            last_contributor = (int)a.n_contrib[pix];
            dC0 = a.dL_dcolor[pix]; dC1 = a.dL_dcolor[HW + pix]; dC2 = a.dL_dcolor[2 * HW + pix];
            dD = a.dL_ddepth[pix];
            dA = a.dL_dthird ? a.dL_dthird[pix] : 0.f;
        }
        s_pix[warp][lane] = make_float4(dC0, dC1, dC2, dD);   // phase 2 reads it after a __syncwarp
        const int warp_last = __reduce_max_sync(FULL, last_contributor);
        float T = T_final;
        // A.4 in suffix-sum form.  With e_j = rgb_j . dL/dC + z_j dL/dD + dL/dA and
        // S_i = sum_{k > i} e_k w_k + T_final (bg . dL/dC):   dL/dalpha_i = e_i T_i - S_i / (1 - alpha_i)
        // (identical to upstream's accum_rec recurrence; one scalar instead of five accumulators).
        float S = T_final * (__ldg(a.bg) * dC0 + __ldg(a.bg + 1) * dC1 + __ldg(a.bg + 2) * dC2);
        int nslot = 0;   // survivors are carried across batches and flushed 16 at a time

        for (int bi = 0; bi < nb; bi++) {
            const int base = (nb - 1 - bi) * BWD_BATCH;
            const int cnt = min(BWD_BATCH, n - base);
            int stage;
            if (TMA) {
                stage = (int)((it + bi) % BWD_STAGES);
                mbar_wait(&s_bar[stage], (uint32_t)(((it + bi) / BWD_STAGES) & 1));
                __syncwarp();
            } else {
                stage = 0;
                for (int k = tid; k < cnt * 3; k += BLK) s_rec[0][k] = src[(size_t)3 * base + k];
                __syncthreads();
            }
            const float4* rec = s_rec[stage];
            if (base < warp_last) {
                // pass 1: cull against this warp's 8x4 block and its deepest contributor
                unsigned mask[BWD_ROUNDS];
#pragma unroll
                for (int r = 0; r < BWD_ROUNDS; r++) {
                    const int jj = r * 32 + lane;
                    bool keep = false;
                    if (jj < cnt && base + jj < warp_last)
                        keep = splat_may_touch(rec[3 * jj], rec[3 * jj + 1], rec[3 * jj + 2].w, rx0, ry0);
                    mask[r] = __ballot_sync(FULL, keep);
                }
                // pass 2: back-to-front over the survivors, pixel per lane
#pragma unroll
                for (int r = BWD_ROUNDS - 1; r >= 0; r--) {
                    unsigned m = mask[r];
                    while (m) {
                        const int bit = 31 - __clz(m);
                        m &= ~(1u << bit);
                        const int j = r * 32 + bit;
                        const float4 r0 = rec[3 * j], r1 = rec[3 * j + 1];
                        const float dx = r0.x - pxf, dy = r0.y - pyf;
                        const float power2 = (r1.x * dx + r1.y * dy) * dx + (r1.z * dy) * dy;
                        const float G = fast_exp2(power2);
                        const float alpha = fminf(0.99f, r1.w * G);
                        const bool active = (base + j < last_contributor) && (power2 <= 0.f) &&
                                            (alpha >= 1.0f / 255.0f);
                        if (STATS) { st_iters += (lane == 0); st_pairs += active; }
#ifdef EAGS_BWD_ANY_SKIP
                        // (skipping the 7 % of iterations in which no lane is active costs more in vote + branch
                        // than it saves: 205.0 -> 202.9 us at c2, 1025 -> 1009 us at c4 without it)
                        if (!__any_sync(FULL, active)) continue;   // warp-uniform
#endif
                        float gG = 0.f, w = 0.f;
                        if (active) {
                            const float4 r2 = rec[3 * j + 2];
                            const float inv1ma = fast_rcp(1.f - alpha);
                            T = T * inv1ma;                      // transmittance in front of splat j
                            w = alpha * T;
                            float e = fmaf(r2.x, dC0, dA);
                            e = fmaf(r2.y, dC1, e);
                            e = fmaf(r2.z, dC2, e);
                            e = fmaf(r0.w, dD, e);
                            const float d = fmaf(e, T, -S * inv1ma);
                            S = fmaf(e, w, S);
                            gG = (r1.w * d) * G;                 // Q1: the 0.99 clamp is straight-through
                        }
                        s_gw[warp][nslot][lane] = make_float2(gG, w);
                        if (lane == 0) {
                            s_meta[warp][2 * nslot] = r0;
                            s_meta[warp][2 * nslot + 1] = r1;
                        }
                        if (++nslot == SLOTS) {
                            flush_slots(s_gw[warp], s_pix[warp], s_meta[warp], SLOTS, lane, rx0, ry0, ddelx_dx,
                                        ddely_dy, a.grad_acc);
                            nslot = 0;
                        }
                    }
                }
            }
            __syncthreads();   // every warp is done with this stage
            if (TMA && tid == 0 && bi + BWD_STAGES < nb) {
                const int nbase = (nb - 1 - (bi + BWD_STAGES)) * BWD_BATCH;
                const int ncnt = min(BWD_BATCH, n - nbase);
                mbar_expect_tx(&s_bar[stage], (uint32_t)ncnt * 48u);
                tma_bulk_g2s(&s_rec[stage][0], src + (size_t)3 * nbase, (uint32_t)ncnt * 48u, &s_bar[stage]);
            }
        }
        it += (unsigned)nb;
        if (nslot)
            flush_slots(s_gw[warp], s_pix[warp], s_meta[warp], nslot, lane, rx0, ry0, ddelx_dx, ddely_dy,
                        a.grad_acc);
    }
    if (STATS) {
        st_iters = __reduce_add_sync(FULL, st_iters);
        st_pairs = __reduce_add_sync(FULL, st_pairs);
        if (lane == 0) {
            atomicAdd(&a.stats[0], (unsigned long long)st_iters);
            atomicAdd(&a.stats[1], (unsigned long long)st_pairs);
        }
    }
}

// resident CTAs of a persistent kernel on the current device (occupancy x SM count), cached
template <typename K>
int persistent_grid(K kernel, int units) {
    int dev = 0, sms = 148, occ = 1;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, BLK, 0) != cudaSuccess || occ < 1) occ = 1;
    const int slots = occ * sms;
    return units < slots ? (units < 1 ? 1 : units) : slots;
}

}  // namespace

int launch_blend_fwd(const EagsView& v, ImagePtrs img, BinningPtrs bin, const EagsImages& out,
                     int use_tma, int stats, unsigned tile_cap, cudaStream_t s) {
    FwdArgs a;
    a.tile_cap = tile_cap;
    a.W = v.image_width; a.H = v.image_height; a.gx = img.gx; a.n_units = img.G * UNITS_PER_TILE;
    a.ctrl = img.ctrl; a.stats = img.ctrl; a.work = &img.ctrl->work_fwd;
    a.capacity = bin.capacity; a.ranges = img.ranges; a.records = bin.records;
    a.bg = v.bg; a.out_color = out.color; a.out_depth = out.depth; a.out_third = out.third;
    a.n_touched = out.n_touched; a.final_T = img.final_T; a.n_contrib = img.n_contrib;
    a.tile_max = img.tile_max;
#define EAGS_FWD(B, T, S) blend_fwd_kernel<B, T, S><<<persistent_grid(blend_fwd_kernel<B, T, S>, a.n_units), BLK, 0, s>>>(a)
    const bool vb = v.variant == EAGS_VARIANT_B;
    if (stats) {
        if (vb) { if (use_tma) EAGS_FWD(true, true, true); else EAGS_FWD(true, false, true); }
        else { if (use_tma) EAGS_FWD(false, true, true); else EAGS_FWD(false, false, true); }
    } else {
        if (vb) { if (use_tma) EAGS_FWD(true, true, false); else EAGS_FWD(true, false, false); }
        else { if (use_tma) EAGS_FWD(false, true, false); else EAGS_FWD(false, false, false); }
    }
#undef EAGS_FWD
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

int launch_blend_bwd(const EagsView& v, ImagePtrs img, BinningPtrs bin, const EagsImageGrads& gin,
                     float* grad_acc, unsigned* work, unsigned long long* stats, int use_tma, cudaStream_t s) {
    BwdArgs a;
    a.W = v.image_width; a.H = v.image_height; a.gx = img.gx; a.n_units = img.G * UNITS_PER_TILE;
    a.work = work;
    a.ranges = img.ranges; a.records = bin.records; a.tile_max = img.tile_max;
    a.final_T = img.final_T; a.n_contrib = img.n_contrib; a.bg = v.bg;
    a.dL_dcolor = gin.dL_dcolor; a.dL_ddepth = gin.dL_ddepth; a.dL_dthird = gin.dL_dthird;
    a.grad_acc = (float4*)grad_acc;
    a.stats = stats;
#define EAGS_BWD(T, S) blend_bwd_kernel<T, S><<<persistent_grid(blend_bwd_kernel<T, S>, a.n_units), BLK, 0, s>>>(a)
    if (stats) { if (use_tma) EAGS_BWD(true, true); else EAGS_BWD(false, true); }
    else { if (use_tma) EAGS_BWD(true, false); else EAGS_BWD(false, false); }
#undef EAGS_BWD
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}
