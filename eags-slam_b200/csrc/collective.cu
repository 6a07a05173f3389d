// collective.cu -- the gradient exchange of sharded mapping, fused with the kernel
// that produces the gradients, over NVLink peer memory (sm_100a).
//
// Every rank renders its own keyframe view of a replicated submap; the per-Gaussian parameter
// gradients (11 floats: xyz 3, scale 3, rotation 4, opacity 1) must be summed over ranks.  The
// baseline is preprocess_bwd -> local buffer -> ncclAllReduce (reduce-scatter + all-gather).  Here:
//
//   1. reduce-scatter, fused into preprocess_bwd (preprocess.cu, `route`): Gaussian i is owned by rank
//      o = i / chunk; preprocess_bwd on rank q stores i's gradient row straight into slot q of rank
//      o's receive area through the peer mapping (posted NVLink stores, overlapped with the rest of
//      the kernel) instead of writing it locally and having a collective read it back.
//   2. peer_reduce_kernel (below), one launch: cross-rank barrier (release/acquire flags in peer
//      memory, monotonically increasing epochs, nothing to reset), owner sums its `world` slots in
//      rank order -- every rank receives the SAME bits, so replicas cannot drift --, applies the
//      1/views scale, and pushes the result into every rank's result buffer (all-gather by peer
//      stores); the last CTA signals completion to all peers and waits for theirs.
//
// NVLink traffic per GPU: 2 (N-1)/N x 44 B per Gaussian, the same as a ring/NVLS all-reduce, but the
// first half is hidden behind preprocess_bwd and there is no staging copy.  Spin waits are bounded:
// a peer that never arrives sets an error word instead of hanging the GPU.
#include "kernels.h"

namespace {

constexpr int RED_THREADS = 256;

struct PeerFlags {                      // one per rank, in peer-mapped memory, zeroed once at setup
    unsigned arrive[EAGS_MAX_PEERS];    // arrive[q] = last epoch rank q's gradients are complete for
    unsigned done[EAGS_MAX_PEERS];      // done[q]   = last epoch rank q finished pushing its slice
    unsigned cta_done;                  // local: CTAs of the running reduce kernel that finished
    unsigned error;                     // local: a bounded wait expired
    unsigned epoch_done;                // local: reductions completed so far (device-side epoch counter:
                                        // a CUDA-graph replay cannot pass a fresh epoch as an argument)
};

__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// Bounded wait for a peer's flag (epochs wrap around safely: compared as a signed difference).  The
// bound is wall time, EAGS_PEER_TIMEOUT_NS (30 s: first-step skew, a GC pause or a data-loader stall on a
// peer are all far below it); when it expires the error word is set and the CALLER MUST NOT USE THE
// SLOTS: peer_reduce poisons its result with NaN, peer_adam skips the update (and the host raises the
// next time it looks at the error word, eags_raster.peer).
constexpr unsigned long long EAGS_PEER_TIMEOUT_NS = 30ull * 1000ull * 1000ull * 1000ull;
__device__ __forceinline__ bool wait_epoch(const unsigned* p, unsigned epoch, unsigned* error) {
    if ((int)(ld_acquire_sys(p) - epoch) >= 0) return true;
    const unsigned long long t0 = global_ns();
    for (;;) {
        for (int spin = 0; spin < 256; spin++) {
            if ((int)(ld_acquire_sys(p) - epoch) >= 0) return true;
            __nanosleep(64);
        }
        if (global_ns() - t0 > EAGS_PEER_TIMEOUT_NS) break;
    }
    atomicMax(error, epoch ? epoch : 1u);   // sticky: the host reads it (error_flag())
    return false;
}

__global__ void __launch_bounds__(RED_THREADS)
peer_reduce_kernel(EagsPeerRoute r, unsigned epoch, float scale, int wait) {
    PeerFlags* mine = (PeerFlags*)r.flags[r.rank];
    // epoch == 0: use the device-side counter (identical on all ranks: every rank launches the same
    // sequence of reductions); it is advanced by the last CTA, after every CTA of this launch read it
    if (epoch == 0) epoch = *(volatile unsigned*)&mine->epoch_done + 1u;
    // ---- barrier A: every rank's scatter (its preprocess_bwd, earlier on its stream) is complete
    if (blockIdx.x == 0 && threadIdx.x < r.world) {
        __threadfence_system();
        st_release_sys(&((PeerFlags*)r.flags[threadIdx.x])->arrive[r.rank], epoch);
    }
    bool ok = true;
    if (wait && threadIdx.x < r.world) ok = wait_epoch(&mine->arrive[threadIdx.x], epoch, &mine->error);
    // a peer never arrived: its slot is incomplete -> the result of this step is poisoned (NaN) on every
    // replica instead of silently being a partial sum
    const float poison = __syncthreads_and(ok) ? 0.f : __int_as_float(0x7fc00000);

    // ---- owner reduces its Gaussians' rows over the `world` slots (rank order) and pushes the result
    const long long P = (long long)r.P, stride = (P + 3) & ~3ll, flat = 11 * stride;
    const long long chunk = (((P + r.world - 1) / r.world) + 3) & ~3ll;
    const long long g0 = min(P, (long long)r.rank * chunk), g1 = min(P, g0 + chunk);
    const float* slots = r.recv[r.rank];
    const long long seg_off[4] = {0, 3 * stride, 6 * stride, 10 * stride};
    const int seg_k[4] = {3, 3, 4, 1};
#pragma unroll
    for (int sgm = 0; sgm < 4; sgm++) {
        // lo is a multiple of 4 floats (stride and chunk are), so the body moves 16-byte vectors --
        // the pushes to the peers are 128-byte coalesced NVLink stores -- and only the cloud's last
        // rows (P not a multiple of 4) take the scalar tail
        const long long lo = seg_off[sgm] + seg_k[sgm] * g0, hi = seg_off[sgm] + seg_k[sgm] * g1;
        const long long nvec = (hi - lo) >> 2;
        for (long long v = (long long)blockIdx.x * RED_THREADS + threadIdx.x; v < nvec;
             v += (long long)gridDim.x * RED_THREADS) {
            const long long e = lo + 4 * v;
            float4 acc = *reinterpret_cast<const float4*>(slots + e);
            for (int q = 1; q < r.world; q++) {
                const float4 x = *reinterpret_cast<const float4*>(slots + q * flat + e);
                acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w;
            }
            acc.x = acc.x * scale + poison; acc.y = acc.y * scale + poison;
            acc.z = acc.z * scale + poison; acc.w = acc.w * scale + poison;
            for (int p = 0; p < r.world; p++) *reinterpret_cast<float4*>(r.reduced[p] + e) = acc;
        }
        if (blockIdx.x == 0) {
            for (long long e = lo + 4 * nvec + threadIdx.x; e < hi; e += RED_THREADS) {
                float acc = slots[e];
                for (int q = 1; q < r.world; q++) acc += slots[q * flat + e];
                acc = acc * scale + poison;
                for (int p = 0; p < r.world; p++) r.reduced[p][e] = acc;
            }
        }
    }

    // ---- barrier B: my slice has landed everywhere; leave only when every peer's slice is here
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned ticket = atomicAdd(&mine->cta_done, 1u);
        if (ticket == gridDim.x - 1) {
            mine->cta_done = 0;
            mine->epoch_done = epoch;
            __threadfence_system();
            for (int p = 0; p < r.world; p++) st_release_sys(&((PeerFlags*)r.flags[p])->done[r.rank], epoch);
            if (wait)
                for (int q = 0; q < r.world; q++)
                    if (!wait_epoch(&mine->done[q], epoch, &mine->error)) break;
        }
    }
}

// ================================================================================================
// peer_adam: reduce-scatter epilogue + activation backward + isotropic regulariser + Adam + parameter
// all-gather in ONE kernel ("owner computes", ZeRO-1 style).
// ================================================================================================
// After preprocess_bwd scattered every rank's gradient rows to the owners, the baseline finishes the step
// with: sum at the owner, push the summed GRADIENTS to every rank (peer_reduce above), autograd through the
// model's activations (exp / normalize / sigmoid, reference gaussian_model.py:77-81), the isotropic
// regulariser's backward (losses.py:115-126, mapper.py:166), and an Adam step over all 11 P floats on
// every rank (gaussian_model.py:341-356: per-group lr, eps 1e-15).  Here the owner of Gaussian i does all
// of that for i alone -- the sum over the `world` slots in rank order, the chain rule through the
// activations (recomputed from the raw parameters it reads from its own replica), the regulariser's
// gradient, the Adam update with moments that exist ONLY on the owner (optimizer state / world per GPU)
// -- and pushes the new raw PARAMETERS (11 floats, the same bytes a gradient all-gather would move) into
// every rank's replica through the peer mapping.  Every replica receives the owner's bits, so replicas
// cannot drift.  One barrier before (all scatters complete), one after (all slices landed): the second
// one now also covers the optimiser.
constexpr int ADAM_THREADS = 256;

__global__ void __launch_bounds__(ADAM_THREADS)
peer_adam_kernel(EagsPeerRoute r, EagsPeerAdam o, unsigned epoch, float scale, int wait) {
    __shared__ __align__(16) float s_out[11 * ADAM_THREADS];   // new parameters of the CTA's Gaussians
    PeerFlags* mine = (PeerFlags*)r.flags[r.rank];
    if (epoch == 0) epoch = *(volatile unsigned*)&mine->epoch_done + 1u;
    if (blockIdx.x == 0 && threadIdx.x < r.world) {
        __threadfence_system();
        st_release_sys(&((PeerFlags*)r.flags[threadIdx.x])->arrive[r.rank], epoch);
    }
    bool ok = true;
    if (wait && threadIdx.x < r.world) ok = wait_epoch(&mine->arrive[threadIdx.x], epoch, &mine->error);
    const bool go = __syncthreads_and(ok);   // a peer never arrived: no update at all (error word is set)

    const long long P = (long long)r.P, stride = (P + 3) & ~3ll, flat = 11 * stride;
    const long long chunk = (((P + r.world - 1) / r.world) + 3) & ~3ll;
    const long long g0 = min(P, (long long)r.rank * chunk), g1 = min(P, g0 + chunk);
    const float* slots = r.recv[r.rank];
    const float* par = o.params[r.rank];
    // bias corrections of step t = steps taken so far + 1 (device-side counter: valid under graph replay)
    const unsigned t = *(volatile unsigned*)o.step + 1u;
    const float bc1 = 1.f - powf(o.beta1, (float)t), bc2_rsqrt = rsqrtf(1.f - powf(o.beta2, (float)t));
    const float iso = o.isotropic_weight / (3.f * (float)P);

    for (long long base = g0 + (long long)blockIdx.x * ADAM_THREADS; go && base < g1;
         base += (long long)gridDim.x * ADAM_THREADS) {
        const long long i = base + threadIdx.x;
        const bool live = i < g1;
        float g[11], p[11];
#pragma unroll
        for (int k = 0; k < 11; k++) { g[k] = 0.f; p[k] = 0.f; }
        if (live) {
            // ---- sum over ranks, in rank order (activated-input gradients) x 1 / views
            for (int q = 0; q < r.world; q++) {
                const float* sl = slots + q * flat;
#pragma unroll
                for (int k = 0; k < 3; k++) { g[k] += sl[3 * i + k]; g[3 + k] += sl[3 * stride + 3 * i + k]; }
                const float4 gr = *reinterpret_cast<const float4*>(sl + 6 * stride + 4 * i);
                g[6] += gr.x; g[7] += gr.y; g[8] += gr.z; g[9] += gr.w;
                g[10] += sl[10 * stride + i];
            }
#pragma unroll
            for (int k = 0; k < 11; k++) g[k] *= scale;
            // ---- raw parameters of this replica (identical on all ranks)
#pragma unroll
            for (int k = 0; k < 3; k++) { p[k] = par[3 * i + k]; p[3 + k] = par[3 * stride + 3 * i + k]; }
            const float4 pr = *reinterpret_cast<const float4*>(par + 6 * stride + 4 * i);
            p[6] = pr.x; p[7] = pr.y; p[8] = pr.z; p[9] = pr.w;
            p[10] = par[10 * stride + i];
            // ---- chain rule through the activations
            // scaling = exp(raw) (+ isotropic regulariser mean |s_k - mean_j s_j| over (P, 3), weight 1)
            const float s0 = __expf(p[3]), s1 = __expf(p[4]), s2 = __expf(p[5]);
            const float sm = (s0 + s1 + s2) * (1.f / 3.f);
            const float e0 = (s0 > sm) - (s0 < sm), e1 = (s1 > sm) - (s1 < sm), e2 = (s2 > sm) - (s2 < sm);
            const float em = (e0 + e1 + e2) * (1.f / 3.f);
            g[3] = (g[3] + iso * (e0 - em)) * s0;
            g[4] = (g[4] + iso * (e1 - em)) * s1;
            g[5] = (g[5] + iso * (e2 - em)) * s2;
            // rotation = raw / max(|raw|, 1e-12)  (torch.nn.functional.normalize)
            const float nq = fmaxf(sqrtf(p[6] * p[6] + p[7] * p[7] + p[8] * p[8] + p[9] * p[9]), 1e-12f);
            const float inq = 1.f / nq;
            const float qh0 = p[6] * inq, qh1 = p[7] * inq, qh2 = p[8] * inq, qh3 = p[9] * inq;
            const float dotq = g[6] * qh0 + g[7] * qh1 + g[8] * qh2 + g[9] * qh3;
            g[6] = (g[6] - dotq * qh0) * inq; g[7] = (g[7] - dotq * qh1) * inq;
            g[8] = (g[8] - dotq * qh2) * inq; g[9] = (g[9] - dotq * qh3) * inq;
            // opacity = sigmoid(raw)
            const float op = 1.f / (1.f + __expf(-p[10]));
            g[10] *= op * (1.f - op);
            // ---- Adam (torch.optim.Adam semantics, amsgrad off, no weight decay); moments live here only
            float* m = o.m + (i - g0);      // moments: 11 planes of `chunk` floats (coalesced)
            float* v = o.v + (i - g0);
#pragma unroll
            for (int k = 0; k < 11; k++) {
                const float lr = o.lr[k < 3 ? 0 : (k < 6 ? 1 : (k < 10 ? 2 : 3))];
                const float mk = o.beta1 * m[k * chunk] + (1.f - o.beta1) * g[k];
                const float vk = o.beta2 * v[k * chunk] + (1.f - o.beta2) * g[k] * g[k];
                m[k * chunk] = mk; v[k * chunk] = vk;
                p[k] -= (lr / bc1) * mk / (sqrtf(vk) * bc2_rsqrt + o.eps);
            }
        }
        // ---- stage, then push the CTA's rows to every replica as 16-byte vectors
#pragma unroll
        for (int k = 0; k < 3; k++) {
            s_out[3 * threadIdx.x + k] = p[k];
            s_out[3 * ADAM_THREADS + 3 * threadIdx.x + k] = p[3 + k];
        }
#pragma unroll
        for (int k = 0; k < 4; k++) s_out[6 * ADAM_THREADS + 4 * threadIdx.x + k] = p[6 + k];
        s_out[10 * ADAM_THREADS + threadIdx.x] = p[10];
        __syncthreads();
        // base and g1 are multiples of 4 Gaussians except at the cloud's end (P): vectors never straddle it
        // because the parameter buffers are padded to `stride`
        constexpr int V3 = 3 * ADAM_THREADS / 4, V4 = ADAM_THREADS, V1 = ADAM_THREADS / 4;
        for (int tt = threadIdx.x; tt < 2 * V3 + V4 + V1; tt += ADAM_THREADS) {
            int seg, vv;
            if (tt < V3) { seg = 0; vv = tt; } else if (tt < 2 * V3) { seg = 1; vv = tt - V3; }
            else if (tt < 2 * V3 + V4) { seg = 2; vv = tt - 2 * V3; } else { seg = 3; vv = tt - 2 * V3 - V4; }
            const int k = seg < 2 ? 3 : (seg == 2 ? 4 : 1);
            const long long first = base + (4 * vv) / k;          // first Gaussian the vector touches
            if (first >= g1) continue;
            const long long off = (seg == 0 ? 0 : (seg == 1 ? 3 : (seg == 2 ? 6 : 10))) * stride + k * base + 4 * vv;
            const float4 val = *reinterpret_cast<const float4*>(
                &s_out[(seg == 0 ? 0 : (seg == 1 ? 3 : (seg == 2 ? 6 : 10))) * ADAM_THREADS + 4 * vv]);
            for (int pr = 0; pr < r.world; pr++) *reinterpret_cast<float4*>(o.params[pr] + off) = val;
        }
        __syncthreads();
    }

    // ---- barrier B: my slice (and my optimiser step) has landed everywhere
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned ticket = atomicAdd(&mine->cta_done, 1u);
        if (ticket == gridDim.x - 1) {
            mine->cta_done = 0;
            mine->epoch_done = epoch;
            if (go) *o.step = t;
            __threadfence_system();
            for (int p = 0; p < r.world; p++) st_release_sys(&((PeerFlags*)r.flags[p])->done[r.rank], epoch);
            if (wait)
                for (int q = 0; q < r.world; q++)
                    if (!wait_epoch(&mine->done[q], epoch, &mine->error)) break;
        }
    }
}

// The activations the rasterizer's inputs go through (reference src/entities/gaussian_model.py:103-122:
// get_scaling = exp, get_rotation = F.normalize, get_opacity = sigmoid), one pass instead of torch's six
// kernels (exp, norm reduction, clamp, div, sigmoid): per step, not per view, so it only shows once the views
// are spread over GPUs.  Their chain rule is the one inside peer_adam_kernel above.
__global__ void __launch_bounds__(256)
activations_kernel(int P, const float* __restrict__ raw_s, const float* __restrict__ raw_q,
                   const float* __restrict__ raw_o, float* __restrict__ s, float* __restrict__ q,
                   float* __restrict__ o) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= P) return;
    const float a = raw_s[3 * (size_t)i], b = raw_s[3 * (size_t)i + 1], c = raw_s[3 * (size_t)i + 2];
    const float4 r = *reinterpret_cast<const float4*>(raw_q + 4 * (size_t)i);
    const float op = raw_o[i];
    s[3 * (size_t)i] = expf(a); s[3 * (size_t)i + 1] = expf(b); s[3 * (size_t)i + 2] = expf(c);
    const float n = fmaxf(sqrtf(r.x * r.x + r.y * r.y + r.z * r.z + r.w * r.w), 1e-12f);   // F.normalize eps
    *reinterpret_cast<float4*>(q + 4 * (size_t)i) = make_float4(r.x / n, r.y / n, r.z / n, r.w / n);
    o[i] = 1.f / (1.f + expf(-op));
}

}  // namespace

int launch_activations(int P, const float* raw_s, const float* raw_q, const float* raw_o, float* s, float* q,
                       float* o, cudaStream_t st) {
    activations_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, raw_s, raw_q, raw_o, s, q, o);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

size_t peer_flags_bytes() { return eags_align256(sizeof(PeerFlags)); }

int launch_peer_reduce(const EagsPeerRoute& r, unsigned epoch, float scale, int wait, cudaStream_t s) {
    const long long chunk = ((((long long)r.P + r.world - 1) / r.world) + 3) & ~3ll;
    long long blocks = (11 * chunk / 4 + RED_THREADS * 2 - 1) / (RED_THREADS * 2);   // ~2 vectors per thread
    if (blocks < 1) blocks = 1;
    if (blocks > 148 * 4) blocks = 148 * 4;
    peer_reduce_kernel<<<(int)blocks, RED_THREADS, 0, s>>>(r, epoch, scale, wait);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

int launch_peer_adam(const EagsPeerRoute& r, const EagsPeerAdam& o, unsigned epoch, float scale, int wait,
                     cudaStream_t s) {
    const long long chunk = ((((long long)r.P + r.world - 1) / r.world) + 3) & ~3ll;
    long long blocks = (chunk + ADAM_THREADS - 1) / ADAM_THREADS;
    if (blocks < 1) blocks = 1;
    if (blocks > 148 * 6) blocks = 148 * 6;
    peer_adam_kernel<<<(int)blocks, ADAM_THREADS, 0, s>>>(r, o, epoch, scale, wait);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}
