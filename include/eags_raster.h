/*
 * eags_raster.h -- C ABI of the B200-native (sm_100a) differentiable Gaussian-splatting rasterizer
 * that replaces the two un-vendored CUDA extensions on EAGS-SLAM's hot path:
 *
 *   gaussian_rasterizer._C          (variant A; reference requirements.txt:21,
 *                                    call sites src/utils/utils.py:105-117,149,174-184)
 *   diff_gaussian_rasterization._C  (variant B "w-pose"; reference requirements.txt:22,
 *                                    call sites src/gsr/renderer.py:59-73,75,131-142)
 *
 * Upstream binds three pybind entry points (`rasterize_gaussians`, `rasterize_gaussians_backward`,
 * `mark_visible`); the reference never calls the third.  This header is what a Python/ctypes,
 * pybind or cgo binding of that operator links against instead.  Plain C: pointers, sizes and a
 * stream handle -- no torch types.
 *
 * Ownership: every buffer (inputs, outputs, the three state buffers, scratch) is allocated by the
 * caller (PyTorch's caching allocator in the shipped binding) on the CURRENT device and passed as
 * raw device pointers.  The library keeps no global state, never allocates device memory, never
 * synchronises the host with the device and launches only on the stream it is given, so it is
 * re-entrant: the reference's loop-closure thread (src/entities/lc.py:625) can render on another
 * device while the main thread tracks/maps.
 *
 * Errors: every entry point returns 0 on success or a negative EAGS_ERR_* / positive cudaError_t
 * value; eags_last_error() gives a thread-local message.  There is no CPU fallback.
 */
#ifndef EAGS_RASTER_H
#define EAGS_RASTER_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EAGS_ABI_VERSION 2

#define EAGS_VARIANT_A 0 /* third image = alpha = sum_i alpha_i T_i          (gaussian_rasterizer)          */
#define EAGS_VARIANT_B 1 /* third image = opacity = 1 - T_final, + n_touched (diff_gaussian_rasterization)  */

#define EAGS_ERR_INVALID -1   /* bad argument (null pointer, negative size, ...)  */
#define EAGS_ERR_TOO_SMALL -2 /* a state buffer is smaller than eags_*_state_bytes */

/* Per-view configuration == the reference's GaussianRasterizationSettings NamedTuple
 * (src/utils/utils.py:105-117; variant B adds projmatrix_raw, src/gsr/renderer.py:59-73).
 * Matrices are DEVICE pointers to 16 row-major floats exactly as the reference passes them:
 * viewmatrix = w2c^T, projmatrix = viewmatrix @ P^T (utils.py:95-104, camera.py:97-107). */
typedef struct EagsView {
    int32_t image_height, image_width;
    float tanfovx, tanfovy;
    float scale_modifier;
    int32_t sh_degree;  /* active SH degree 0..3 */
    int32_t variant;    /* EAGS_VARIANT_A / EAGS_VARIANT_B */
    int32_t prefiltered;
    int32_t debug;
    const float* bg;             /* device (3)  */
    const float* viewmatrix;     /* device (16) */
    const float* projmatrix;     /* device (16) */
    const float* projmatrix_raw; /* device (16) or NULL; carried for ABI parity with variant B */
    const float* campos;         /* device (3)  */
} EagsView;

/* The Gaussians handed to GaussianRasterizer.forward (utils.py:162-184, renderer.py:131-142).
 * All device pointers, contiguous fp32.  Exactly one of shs / colors_precomp and exactly one of
 * (scales, rotations) / cov3D_precomp is non-NULL. */
typedef struct EagsGaussians {
    int32_t P;                   /* number of Gaussians                              */
    int32_t M;                   /* SH coefficients per Gaussian (stride of shs)     */
    const float* means3D;        /* (P,3)                                            */
    const float* opacities;      /* (P,1) post-sigmoid                               */
    const float* shs;            /* (P,M,3) or NULL                                  */
    const float* colors_precomp; /* (P,3)   or NULL                                  */
    const float* scales;         /* (P,3) post-exp or NULL                           */
    const float* rotations;      /* (P,4) wxyz, normalised, or NULL                  */
    const float* cov3D_precomp;  /* (P,6) xx,xy,xz,yy,yz,zz or NULL                  */
} EagsGaussians;

/* Images produced by the forward pass; all device, caller-allocated. */
typedef struct EagsImages {
    float* color;       /* (3,H,W)                                                     */
    float* depth;       /* (1,H,W)  sum_i z_i alpha_i T_i (not normalised)             */
    float* third;       /* (1,H,W)  alpha (variant A) or opacity (variant B)           */
    int32_t* n_touched; /* (P) or NULL; variant B: #pixels where the Gaussian was blended with T > 0.5 */
} EagsImages;

/* Gradients w.r.t. the images (inputs of the backward pass). */
typedef struct EagsImageGrads {
    const float* dL_dcolor; /* (3,H,W)          */
    const float* dL_ddepth; /* (1,H,W)          */
    const float* dL_dthird; /* (1,H,W) or NULL  */
} EagsImageGrads;

/* Gradients w.r.t. the operator inputs (outputs of the backward pass).  Any pointer may be NULL
 * (that gradient is skipped; with only dL_dtau set the per-Gaussian scale / rotation / SH backward
 * is not even computed).  Non-NULL arrays are fully written (zeros for culled Gaussians). */
typedef struct EagsInputGrads {
    float* dL_dmeans3D;        /* (P,3) */
    float* dL_dmeans2D;        /* (P,3) screen-space gradient (x,y,0): lands in means2D.grad   */
    float* dL_dopacities;      /* (P,1) */
    float* dL_dshs;            /* (P,M,3) */
    float* dL_dcolors_precomp; /* (P,3) */
    float* dL_dscales;         /* (P,3) */
    float* dL_drotations;      /* (P,4) */
    float* dL_dcov3D_precomp;  /* (P,6) */
    float* dL_dtau;            /* (6) = [rho(3), theta(3)]: camera-pose gradient of the left
                                  perturbation T_CW <- Exp(tau) T_CW (pose_utils.py:76-93)     */
} EagsInputGrads;

int eags_abi_version(void);
const char* eags_last_error(void);

/* Sizes of the three opaque state buffers (upstream: geomBuffer / binningBuffer / imgBuffer)
 * and of the backward scratch.  `capacity` is the number of (tile, Gaussian) keys the binning
 * buffer can hold; `with_keys` != 0 also reserves the sorted 64-bit (tile|depth) key list. */
size_t eags_geom_state_bytes(int32_t P);
size_t eags_image_state_bytes(int32_t width, int32_t height);
size_t eags_binning_state_bytes(int64_t capacity, int32_t with_keys);
size_t eags_backward_scratch_bytes(int32_t P);
/* byte offset, inside the backward scratch, of two uint64 counters the "kernel_stats" option makes
 * blend_bwd fill: (8x4 pixel block, splat) pairs walked, (pixel, splat) pairs with a gradient */
size_t eags_backward_stats_offset(int32_t P);

/* Forward, stage 1: per-Gaussian preprocess (cull, EWA projection, conic, radius, tile rectangle,
 * SH->RGB), per-tile counts and the tile scan.  Writes `radii` (P) and, when
 * `num_rendered_pinned` is non-NULL, enqueues an async copy of K = sum(tiles_touched) (uint64)
 * into that pinned host slot -- the caller decides when (if ever) to wait for it. */
int eags_forward_geom(const EagsView* view, const EagsGaussians* g, void* geom_state,
                      void* image_state, int32_t* radii, uint64_t* num_rendered_pinned,
                      void* stream);

/* Forward, stage 2: key duplication, per-tile (depth, index) sort, packing of the sorted splat
 * stream, and front-to-back blending.  If K > capacity nothing is rendered (all kernels exit);
 * the caller re-issues this stage with a larger binning buffer. */
int eags_forward_render(const EagsView* view, const EagsGaussians* g, const int32_t* radii,
                        void* geom_state, void* image_state, void* binning_state,
                        int64_t capacity, int32_t with_keys, const EagsImages* out, void* stream);

/* Bucketed forward (optional fast path; same results).  When the caller knows an upper bound
 * `tile_cap` of the per-tile key count (e.g. 1.25 x the largest count of the previous frame --
 * `max_tile` below), stage 1 scatters every (depth | gaussian) key straight into its tile's
 * fixed-stride bucket inside `binning_state` (which must be eags_binning_state_bytes_bucketed(...)
 * bytes and is needed BEFORE stage 1) and stage 2 launches no duplicate kernel.  The whole 64-byte
 * status block is copied to pinned memory.  If `bucket_overflow` comes back non-zero (or
 * num_rendered > capacity) nothing was rendered: re-issue stage 2 with eags_forward_render (exact
 * path; the tile scan of stage 1 is still valid).  tile_cap == 0 degenerates to the exact path.
 * `sticky_status_dev` (optional, device, 64 bytes, zeroed ONCE by the caller and never by the library): when
 * given, num_rendered / max_tile are folded into it with atomicMax and bucket_overflow with atomicOr, and
 * THAT block is what reaches `status_pinned` (stored by the tile scan's last CTA when the block is pinned host
 * memory the device can address, copied by a cudaMemcpyAsync otherwise) -- for CUDA-graph replays, where the host looks at the
 * status once per many replays and an overflow of any of them must still be seen. */
typedef struct EagsForwardStatus {
    uint64_t num_rendered;    /* K = sum(tiles_touched)                                   */
    uint64_t num_scanned;
    uint32_t n_big, work_sort, work_fwd;
    uint32_t max_tile;        /* largest per-tile key count of this view                  */
    uint32_t bucket_overflow; /* != 0: a tile outgrew tile_cap                            */
    uint32_t pad0;
    uint64_t stat_fwd_warp_iters; /* "kernel_stats" option: (8x4 pixel block, splat) pairs blend_fwd walked */
    uint64_t stat_fwd_lane_pairs; /*                        (pixel, splat) pairs blend_fwd blended          */
    uint64_t pad1;
} EagsForwardStatus;
size_t eags_binning_state_bytes_bucketed(int64_t capacity, int32_t with_keys, int32_t width,
                                         int32_t height, int32_t tile_cap);
int eags_forward_geom_bucketed(const EagsView* view, const EagsGaussians* g, void* geom_state,
                               void* image_state, int32_t* radii, EagsForwardStatus* status_pinned,
                               void* binning_state, int64_t capacity, int32_t with_keys,
                               int32_t tile_cap, EagsForwardStatus* sticky_status_dev, void* stream);
int eags_forward_render_bucketed(const EagsView* view, const EagsGaussians* g, const int32_t* radii,
                                 void* geom_state, void* image_state, void* binning_state,
                                 int64_t capacity, int32_t with_keys, int32_t tile_cap,
                                 const EagsImages* out, void* stream);

/* Backward: blend backward (warp-reduced, <= 1 vector atomic per (tile, Gaussian)) and preprocess
 * backward (+ closed-form camera-pose gradient when dL_dtau is non-NULL). */
int eags_backward(const EagsView* view, const EagsGaussians* g, const int32_t* radii,
                  const void* geom_state, const void* image_state, const void* binning_state,
                  int64_t capacity, const EagsImageGrads* gin, const EagsInputGrads* gout,
                  void* scratch, void* stream);

/* ---- Sharded mapping: gradient exchange over NVLink peer memory -------------------
 * One process per GPU, every rank renders its own keyframe view of a replicated submap
 * (the reference optimises one view per step on one GPU, src/entities/mapper.py:143-170; it has no
 * collective).  Instead of "write gradients locally, then all-reduce", eags_backward_routed makes
 * preprocess_bwd store the 11 parameter-gradient floats of Gaussian i (xyz 3, scale 3, rotation 4,
 * opacity 1) directly into slot `rank` of the receive area of the rank that owns i
 * (owner = i / chunk, chunk = ceil(P / world) rounded up to 4), through peer-mapped pointers;
 * eags_peer_reduce then runs ONE kernel per rank: barrier, sum of the `world` slots in rank order
 * (bit-identical result on every rank), x `scale`, push of the owned slice into every rank's
 * `reduced` buffer, barrier.  All buffers are caller-owned symmetric (peer-mapped) device memory:
 *   recv[o]    : rank o's receive area, world x 11 x stride floats, stride = (P + 3) & ~3
 *                (slot q at q * 11 * stride; inside a slot: xyz at 0, scale at 3 stride, rotation at
 *                6 stride, opacity at 10 stride)
 *   reduced[p] : rank p's result, 11 x stride floats, same layout
 *   flags[p]   : rank p's signal words, eags_peer_flags_bytes() bytes, zeroed once before first use
 * `epoch` = 0 uses a device-side counter kept in `flags` (valid inside CUDA-graph replays); a non-zero
 * value must increase by one per eags_peer_reduce call (start at 1), identically on all ranks.
 * `wait` = 0 skips the cross-rank waits (single-GPU emulation of several ranks in tests).
 * The cross-rank waits are bounded (30 s of wall time): if a peer never arrives the error word in `flags`
 * is set (sticky) and the result of that step is poisoned with NaN on every replica -- never a partial sum. */
#define EAGS_MAX_PEERS 8
typedef struct EagsPeerRoute {
    int32_t rank, world;
    int64_t P;
    float* recv[EAGS_MAX_PEERS];
    float* reduced[EAGS_MAX_PEERS];
    uint32_t* flags[EAGS_MAX_PEERS];
    /* optional (all four or none): this rank's partial sums over the views it already rendered this
     * step -- (P,3) xyz, (P,3) scale, (P,4) rotation, (P,1) opacity, local memory -- added to the routed
     * rows of this backward pass, so a rank that renders several views per step (8-view keyframe batch
     * on fewer than 8 GPUs) still sends each row over NVLink once */
    const float* accum[4];
} EagsPeerRoute;
size_t eags_peer_flags_bytes(void);
int eags_backward_routed(const EagsView* view, const EagsGaussians* g, const int32_t* radii,
                         const void* geom_state, const void* image_state, const void* binning_state,
                         int64_t capacity, const EagsImageGrads* gin, const EagsInputGrads* gout,
                         void* scratch, const EagsPeerRoute* route, void* stream);
int eags_peer_reduce(const EagsPeerRoute* route, uint32_t epoch, float scale, int32_t wait, void* stream);

/* Fused optimiser step of sharded mapping ("owner computes"): replaces eags_peer_reduce + the autograd of
 * the model's activations + the isotropic regulariser's backward + torch.optim.Adam by ONE kernel per
 * rank.  After eags_backward_routed scattered the gradient rows, the owner of Gaussian i sums its `world`
 * slots in rank order (x `scale` = 1 / views), applies the chain rule of the reference's parametrisation
 * (src/entities/gaussian_model.py:77-81: scaling = exp(raw), rotation = normalize(raw), opacity =
 * sigmoid(raw), xyz raw), adds `isotropic_weight` x d isotropic_loss / d scaling
 * (src/entities/losses.py:128-138, mapper.py:166), takes the Adam step of
 * gaussian_model.py:341-356 (per-group lr, eps 1e-15; torch.optim.Adam semantics, amsgrad off) with moments
 * that live ONLY on the owner, and stores the new raw parameters into EVERY rank's replica through the peer
 * mapping (the all-gather moves parameters instead of gradients: same bytes, and replicas hold the owner's
 * bits by construction).  Barriers, `epoch` and `wait` as eags_peer_reduce.
 *   params[p] : rank p's raw parameter replica, peer-mapped, 11 x stride floats in the segment layout of
 *               `reduced` (xyz at 0, log-scale at 3 stride, raw quaternion at 6 stride, logit opacity at
 *               10 stride); the rasterizer reads xyz from it directly
 *   m, v      : local, 11 x chunk floats each (plane k of coordinate k: m[k * chunk + (i - first owned)]),
 *               zeroed by the caller before the first step; chunk = ceil(P / world) rounded up to 4
 *   step      : local uint32 device counter of optimiser steps taken (zeroed once; read for the bias
 *               corrections, advanced by the kernel -- valid under CUDA-graph replay)
 * If a peer never arrives (bounded wait, 30 s) NOTHING is updated and the error word in `flags` is set. */
typedef struct EagsPeerAdam {
    float* params[EAGS_MAX_PEERS];
    float* m;
    float* v;
    uint32_t* step;
    float lr[4]; /* xyz, scaling, rotation, opacity (gaussian_model.py:342-347) */
    float beta1, beta2, eps;
    float isotropic_weight; /* 1 in the reference (mapper.py:166-167); 0 disables the term */
} EagsPeerAdam;
int eags_peer_adam_step(const EagsPeerRoute* route, const EagsPeerAdam* opt, uint32_t epoch, float scale,
                        int32_t wait, void* stream);

/* The three activations between the optimiser's raw parameters and the rasterizer's inputs
 * (src/entities/gaussian_model.py:103-122: get_scaling = exp, get_rotation = torch.nn.functional.normalize
 * (eps 1e-12), get_opacity = sigmoid) in one pass; xyz has none.  raw_rotation / rotations must be 16-byte
 * aligned.  No gradient: the chain rule through them is part of eags_peer_adam_step. */
int eags_gaussian_activations(int32_t P, const float* raw_scaling, const float* raw_rotation,
                              const float* raw_opacity, float* scales, float* rotations, float* opacities,
                              void* stream);

/* ---- Fused image-space losses ---------------------------------------------
 * The per-iteration losses that bracket the rasterizer in the reference: the tracking loss
 * (src/entities/tracker.py:99-128: |x - y| + eps, depth-valid mask, alpha^3 weighting, each sum
 * divided by its number of positive elements) and the masked L1 means of mapping
 * (src/entities/mapper.py:161-165 with src/entities/losses.py:8-27).  One pass produces
 * sums = {colour sum, colour normaliser, depth sum, depth normaliser, colour loss, depth loss} (6
 * floats, loss = sum / normaliser) and the raw gradient images;
 * eags_image_l1_backward scales them by upstream_grad / normaliser (device scalars, no host sync). */
typedef struct EagsImageL1 {
    int32_t width, height;
    int32_t soft_alpha;    /* 1: weight both losses by alpha^3 (tracker.py:118-121); 2: registration weighting of
                              src/gsr/loss.py:66-95 -- colour weight alpha x mask, depth weight mask_depth x
                              (alpha > alpha_min), no power                                              */
    int32_t mask_color;    /* mask the colour loss as well (tracker.py:122-125)                    */
    int32_t norm_mode;     /* 0: divide by #(element > 0) (tracker.py:127-128); 1: by #masked;
                              2: by the element count (a plain .mean(), src/gsr/loss.py:77,95)        */
    float eps;             /* 1e-8 in the tracker, 0 for a plain L1                                */
    const float* color;    /* (3,H,W) */
    const float* depth;    /* (1,H,W) */
    const float* alpha;    /* (1,H,W) or NULL */
    const float* gt_color; /* (3,H,W) */
    const float* gt_depth; /* (1,H,W) */
    const uint8_t* mask;   /* (H,W) 0/1 or NULL */
    const uint8_t* mask_depth; /* (H,W) 0/1 or NULL (= mask): separate validity mask of the depth term  */
    float alpha_min;       /* soft_alpha == 2: opacity threshold of the depth term (0.95, loss.py:89)  */
} EagsImageL1;
size_t eags_image_l1_scratch_bytes(int32_t width, int32_t height);
/* g_color (3,H,W), g_depth (1,H,W), g_alpha (2,H,W) or NULL, sums4 (SIX floats): device, caller-owned */
int eags_image_l1_forward(const EagsImageL1* in, float* g_color, float* g_depth, float* g_alpha,
                          void* scratch, float* sums4, void* stream);
int eags_image_l1_backward(int32_t width, int32_t height, const float* sums4, const float* up_color,
                           const float* up_depth, float* g_color, float* g_depth, float* g_alpha,
                           void* stream);

/* eags_image_l1_forward without ANY allocation and without a caller scratch buffer.  This entry point has no
 * counterpart in the reference; it was added for callers that cannot
 * call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  Same
 * arguments, outputs and arithmetic as eags_image_l1_forward (bit-identical results), minus `scratch`: the
 * kernel's ticket line and per-CTA partial sums live in a static device array of the library (38 144 bytes,
 * held by every GPU that loads the library until the process exits; its address is resolved once per device
 * with cudaGetSymbolAddress).  Limit: width x height <= EAGS_IMAGE_L1_STATIC_MAX_PIXELS (2048 x 2048; the
 * reference's largest image is 1752 x 1168) -- larger images are rejected with EAGS_ERR_INVALID and a message.
 * All calls share the array and it keeps the last call's contents (each call clears the ticket line, which is
 * all that must start at zero): calls on one stream are safe back to back; calls that can OVERLAP on the GPU
 * (different streams, several host threads on one device) race -- use eags_image_l1_forward with a scratch
 * buffer per stream for those.  The call itself performs no cudaMalloc / cudaFree and uses no allocating
 * library routine.  The FIRST call on a device resolves the array's address (and, under CUDA's lazy module
 * loading, loads the kernel): make that one call before a stream capture begins, not inside it. */
#define EAGS_IMAGE_L1_STATIC_MAX_PIXELS 4194304
int eags_image_l1_forward_static(const EagsImageL1* in, float* g_color, float* g_depth, float* g_alpha,
                                 float* sums4, void* stream);

/* Fused SSIM (11x11 Gaussian window, sigma 1.5, zero padding, mean over C*H*W) -- the `ssim(img1, img2)`
 * of src/entities/losses.py:65-113 as used by the mapping colour loss (src/entities/mapper.py:162-163).
 * Forward writes the mean SSIM to out_mean (device float) and keeps three per-pixel partial maps in
 * `scratch` (eags_ssim_scratch_bytes, caller-owned, must survive until the backward);
 * backward writes d(mean ssim)/d img1 x upstream[0] (device scalar, no host sync) to grad_img1 (C,H,W). */
size_t eags_ssim_scratch_bytes(int32_t C, int32_t height, int32_t width);
int eags_ssim_forward(int32_t C, int32_t height, int32_t width, const float* img1, const float* img2,
                      void* scratch, float* out_mean, void* stream);
int eags_ssim_backward(int32_t C, int32_t height, int32_t width, const float* img1, const float* img2,
                       const void* scratch, const float* upstream, float* grad_img1, void* stream);

/* The whole per-view mapping loss in two kernels (src/entities/mapper.py:161-165):
 *   colour loss = (1 - lambda) l1_loss(image[:, mask], gt_image[:, mask]) + lambda (1 - ssim(image, gt_image))
 *   depth  loss = l1_loss(depth[:, mask], gt_depth[mask])
 * The masked L1 sums ride on the SSIM forward pass, the sign gradients on the SSIM backward pass: no raw
 * gradient images, no scaling pass, no gradient adds.  out8 (device, 8 floats) = {colour |d| sum, n_colour,
 * depth |d| sum, n_depth, colour L1, depth L1, SSIM, colour loss}; the reduction order is fixed (same bits
 * every run).  backward: grad_image (3,H,W) = up_color[0] d(colour loss)/d image, grad_depth (1,H,W) =
 * up_depth[0] d(depth loss)/d depth; up_* are device scalars (no host sync).  `scratch`
 * (eags_mapping_loss_scratch_bytes) and out8 must survive until the backward. */
typedef struct EagsMappingLoss {
    int32_t width, height;
    float lambda_dssim;    /* opt.lambda_dssim (0.2) */
    const float* image;    /* (3,H,W) rendered */
    const float* depth;    /* (1,H,W) rendered */
    const float* gt_image; /* (3,H,W) */
    const float* gt_depth; /* (1,H,W) */
    const uint8_t* mask;   /* (H,W) 0/1 or NULL (all valid) */
    int32_t valid_depth_mask; /* 1: the mask of mapper.py:159, (gt_depth > 0) & ~isnan(depth), formed per pixel
                                 inside the kernels (AND-ed with `mask` when that is given); 0: `mask` only */
} EagsMappingLoss;
size_t eags_mapping_loss_scratch_bytes(int32_t width, int32_t height);
int eags_mapping_loss_forward(const EagsMappingLoss* in, void* scratch, float* out8, void* stream);
int eags_mapping_loss_backward(const EagsMappingLoss* in, const void* scratch, const float* out8,
                               const float* up_color, const float* up_depth, float* grad_image,
                               float* grad_depth, void* stream);

/* Isotropic regulariser of the mapping loss (src/entities/losses.py:115-126, mapper.py:166):
 * out_mean = mean |s_ij - mean_j s_ij| over scaling (P,3); raw_grad (P,3) = d(sum)/d scaling (the caller
 * scales it by upstream / (3 P)); sum_scratch8: 8 bytes of device scratch (fp64 accumulator). */
int eags_isotropic_loss(int32_t P, const float* scaling, float* raw_grad, void* sum_scratch8,
                        float* out_mean, void* stream);

/* ---- Tracking fast path ----------------------------------------
 * Replaces the P-sized pose transform + its autograd of src/entities/tracker.py:85-97 by folding the
 * optimisable relative pose (quaternion q wxyz, un-normalised; translation t) into the view matrices:
 * rendering R(q) p + t, q (x) r from camera L equals rendering p, r from camera L * Rel.
 * eags_pose_compose writes viewmatrix' / projmatrix' / campos' (device, 16/16/3 floats);
 * eags_pose_chain turns the rasterizer's dL_dtau (6) into dL/dq (4) and dL/dt (3).  `viewmatrix`,
 * `projmatrix`, `campos` are the ORIGINAL settings (the previous camera, utils.py:95-104). */
int eags_pose_compose(const float* q, const float* t, const float* viewmatrix, const float* projmatrix,
                      const float* campos, float* viewmatrix_out, float* projmatrix_out,
                      float* campos_out, void* stream);
int eags_pose_chain(const float* q, const float* t, const float* viewmatrix, const float* dL_dtau,
                    float* dL_dq, float* dL_dt, void* stream);

/* Camera update of loop-closure registration (variant B call site): the reference's update_pose
 * (src/utils/pose_utils.py:76-93) followed by the view / projection / centre properties of
 * src/gsr/camera.py:97-111, on the device and in place:
 *   w2c <- SE3_exp([rho, theta]) w2c  (pose_utils.py:40-74), viewmatrix = w2c^T (updated in place),
 *   projmatrix_out = viewmatrix x projmatrix_raw, campos_out = camera centre, theta = rho = 0;
 * tau_norm (nullable, device float) receives |[rho, theta]| -- the reference's convergence measure
 * (`tau.norm() < 1e-4`), readable whenever the host chooses to look.  All pointers device fp32. */
int eags_pose_se3_update(float* theta, float* rho, float* viewmatrix, const float* projmatrix_raw,
                         float* projmatrix_out, float* campos_out, float* tau_norm, void* stream);

/* The optimiser side of one registration iteration in one launch (reference src/gsr/solver.py:83-91:
 * `optimizer.step()` on cam_rot_delta / cam_trans_delta -- torch.optim.Adam, lr 3 x base for the rotation, base
 * for the translation, betas (0.9, 0.999), eps 1e-8 -- then `update_pose`).  The deltas are zero when a step
 * starts (update_pose zeroes them), so the Adam step is the delta itself; `state13` holds the moments of
 * (theta, rho) and the step count as 13 device floats, zeroed once by the caller.  Camera outputs as
 * eags_pose_se3_update. */
int eags_pose_adam_se3_update(const float* grad_theta, const float* grad_rho, float* state13, float lr_rot,
                              float lr_trans, float beta1, float beta2, float eps, float* viewmatrix,
                              const float* projmatrix_raw, float* projmatrix_out, float* campos_out,
                              float* tau_norm, void* stream);

/* ---- Neighbour queries around keyframe seeding -------------------------------
 * eags_knn_mean_dist2 replaces `simple_knn._C.distCUDA2(points)` (un-vendored pip dependency
 * `simple-knn`; reference call sites src/entities/gaussian_model.py:143,146,253,257,298,301,616):
 * out[i] = mean of the squared distances from point i to its 3 nearest other points (FLT_MAX
 * stand-ins when fewer than 4 points exist, as upstream).  points (n,3) fp32, out (n,) fp32.
 * eags_radius_has_neighbor replaces the faiss IndexFlatL2 search of
 * src/utils/mapper_utils.py:193-226 (compute_new_points_ids): has_neighbor[q] = 1 iff some reference
 * point has SQUARED distance < radius (the reference compares faiss' squared distances with
 * `radius` directly).  Both build a hashed uniform grid in `workspace` (device, caller-owned,
 * eags_neighbors_workspace_bytes(n_ref) bytes; n_ref = n for the k-NN query); exact results. */
size_t eags_neighbors_workspace_bytes(int64_t n_ref);
int eags_knn_mean_dist2(const float* points, int32_t n, float* out, void* workspace,
                        size_t workspace_bytes, void* stream);
int eags_radius_has_neighbor(const float* ref_points, int32_t n_ref, const float* query_points,
                             int32_t n_query, float radius, uint8_t* has_neighbor, void* workspace,
                             size_t workspace_bytes, void* stream);

/* Introspection of the state buffers (parity tests, debugging).  `what` selects the array; the
 * function returns its device address, element count and element size.  Names:
 *   "splat_rec" float4[3P] ({x, y, bits(index), depth} {A, B, C, opacity} {r, g, b, log2(255 opacity)} per
 *   Gaussian, the conic pre-scaled by -log2(e)/2, -log2(e), -log2(e)/2; valid where radii > 0),
 *   "tiles_touched" u32[P], "clamped" u8[P] (bit c = channel c clamped),
 *   "ranges" uint2[G], "tile_count" u32[G], "final_T" f32[N], "n_contrib" u32[N],
 *   "tile_max_contrib" u32[G], "point_list" u32[cap], "point_list_keys" u64[cap] (with_keys),
 *   "records" float4[3*cap]. */
int eags_state_view(const char* what, int32_t P, int32_t width, int32_t height, int64_t capacity,
                    int32_t with_keys, const void* geom_state, const void* image_state,
                    const void* binning_state, const void** ptr, int64_t* count,
                    int32_t* elem_bytes);

/* Tuning knobs for A/B measurements (process-wide, read at launch time):
 *   "tma"      1 (default): blend kernels stage splat batches with cp.async.bulk + mbarrier;
 *              0: plain cooperative loads.                                                     */
/*   "preprocess_agg" 1 (default): CTA-aggregated preprocess_fwd (shared-memory tile histogram, one global
 *              atomic per (CTA, touched tile)); 0: one global atomic per (Gaussian, tile) key.
 *   "radix_sort" per-tile sort: 2 (default) one-pass bucket sort over the tile's depth range + thread-local
 *              insertion finish (clustered depths fall back to 1); 1: LSD radix sort on the varying depth
 *              bits; 0: bitonic network on the 64-bit (depth | index) composite.  Bit-identical results.
 *   "kernel_stats" 1: the blend kernels count the work they really executed (post-cull) into
 *              EagsForwardStatus.stat_* / the backward scratch (bench.py's executed-work roofline);
 *              default 0 (the counting instantiations are never launched).                       */
/*   "kernel_timing" 1: record CUDA events on the caller's stream around every kernel launcher;
 *              read them with eags_kernel_times().  Default 0 (no events are ever recorded).    */
int eags_set_option(const char* name, int32_t value);
int eags_get_option(const char* name, int32_t* value);

/* Measurement hook (bench.py's live per-kernel roofline): waits for and returns the spans recorded
 * by ANY thread of the process since the last call (autograd runs the backward on its own thread) -- names[i] is a static string ("preprocess_fwd",
 * "tile_scan", "duplicate", "sort_pack", "blend_fwd", "blend_bwd", "preprocess_bwd"), ms[i] the
 * device time of that launcher.  Returns the number of spans written (<= max_spans). */
int eags_kernel_times(int32_t max_spans, const char** names, float* ms);

/* Number of kernel launches issued by this library in the calling process since load (all
 * threads); bench.py reports the per-step delta as `gpu_launches`. */
uint64_t eags_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* EAGS_RASTER_H */
