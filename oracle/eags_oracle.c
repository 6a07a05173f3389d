/*
 * eags_oracle.c -- CPU ORACLE for the EAGS-SLAM Gaussian-splatting rasterizer hot path.
 *
 * THIS FILE IS TEST INFRASTRUCTURE.  It is a plain-C restatement of the algorithm that the
 * reference's two un-vendored rasterizer dependencies implement
 *   - gaussian_rasterizer          @ 9c40173 (reference requirements.txt:21; call sites
 *                                   src/utils/utils.py:105-117,149,174-184)       "variant A"
 *   - diff_gaussian_rasterization  @ 43e21bf (reference requirements.txt:22; call sites
 *                                   src/gsr/renderer.py:59-73,75,131-142)         "variant B"
 * following the math spec in SURVEY.md Appendix A (A.1 preprocess, A.2 binning, A.3 blend
 * forward, A.4 blend backward, A.5 preprocess backward + pose gradient, A.6 quirks Q1-Q10)
 * and the in-tree conventions: matrices src/utils/utils.py:93-104 and
 * src/utils/graphics_utils.py:72-93, SH basis src/utils/gaussian_model_utils.py:28-122,
 * quaternion->R src/utils/gaussian_model_utils.py:182-211, covariance 6-vector order
 * src/utils/gaussian_model_utils.py:165-179, pose perturbation src/utils/pose_utils.py:61-93.
 *
 * PARITY UNPINNED: the CUDA sources of both dependencies are absent from /root/reference and the
 * reference ships no tests / golden vectors for this path.  The
 * oracle is pinned instead by (i) golden vectors generated from the reference's importable
 * helpers (tests/golden/make_golden.py) and (ii) float64 PyTorch autograd of an independent
 * dense restatement (oracle/torch_ref.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product path (eags-slam_b200/) never does.
 *
 * Canonical floating-point order: every expression below is evaluated in IEEE binary32 with no
 * FMA contraction (compile with -ffp-contract=off), left to right as written.  The CUDA
 * preprocess kernel is compiled with -fmad=false and follows the same order so that depth keys,
 * radii and tile rectangles are bit-exact.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define TILE 16
#define SH_C0 0.28209479177387814f
#define SH_C1 0.4886025119029199f

typedef struct {
    int32_t P;              /* number of Gaussians */
    int32_t M;              /* SH coefficients per Gaussian (stride), >= (degree+1)^2 */
    int32_t degree;         /* active SH degree (0..3) */
    int32_t W, H;           /* image size */
    int32_t variant;        /* 0 = A: third image is alpha = sum(alpha_i T_i); 1 = B: opacity = 1 - T */
    float tanfovx, tanfovy;
    float scale_modifier;
    float view[16];         /* viewmatrix  = w2c^T, row-major flat (utils.py:95-97) */
    float proj[16];         /* projmatrix  = viewmatrix @ P^T (utils.py:103-104) */
    float campos[3];
    float bg[3];
} OracleSettings;

static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                               -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                               0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                               -0.5900435899266435f};

/* CUDA's float->int conversion (cvt.rzi.s32.f32) saturates and maps NaN to 0; C leaves that
 * undefined.  Make it explicit so both sides agree. */
static inline int f2i_rz(float f) {
    if (f != f) return 0;
    if (f >= 2147483648.0f) return 2147483647;
    if (f <= -2147483648.0f) return (int)0x80000000;
    return (int)f;
}
static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }
static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void oracle_set_num_threads(int n) {
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* quaternion (w,x,y,z) -> rotation, gaussian_model_utils.py:182-211 (no normalisation here: the
 * operator receives the already-normalised quaternion). */
static void quat_to_R(const float* q, float R[3][3]) {
    float r = q[0], x = q[1], y = q[2], z = q[3];
    R[0][0] = 1.f - 2.f * (y * y + z * z);
    R[0][1] = 2.f * (x * y - r * z);
    R[0][2] = 2.f * (x * z + r * y);
    R[1][0] = 2.f * (x * y + r * z);
    R[1][1] = 1.f - 2.f * (x * x + z * z);
    R[1][2] = 2.f * (y * z - r * x);
    R[2][0] = 2.f * (x * z - r * y);
    R[2][1] = 2.f * (y * z + r * x);
    R[2][2] = 1.f - 2.f * (x * x + y * y);
}

/* Sigma = (R S)(R S)^T, 6-vector (xx,xy,xz,yy,yz,zz): gaussian_model.py:70-74 */
static void cov3d_from_scale_rot(const float* s, float mod, const float* q, float* c6) {
    float R[3][3], Mx[3][3];
    quat_to_R(q, R);
    float sx = mod * s[0], sy = mod * s[1], sz = mod * s[2];
    for (int i = 0; i < 3; i++) {
        Mx[i][0] = R[i][0] * sx;
        Mx[i][1] = R[i][1] * sy;
        Mx[i][2] = R[i][2] * sz;
    }
    c6[0] = Mx[0][0] * Mx[0][0] + Mx[0][1] * Mx[0][1] + Mx[0][2] * Mx[0][2];
    c6[1] = Mx[0][0] * Mx[1][0] + Mx[0][1] * Mx[1][1] + Mx[0][2] * Mx[1][2];
    c6[2] = Mx[0][0] * Mx[2][0] + Mx[0][1] * Mx[2][1] + Mx[0][2] * Mx[2][2];
    c6[3] = Mx[1][0] * Mx[1][0] + Mx[1][1] * Mx[1][1] + Mx[1][2] * Mx[1][2];
    c6[4] = Mx[1][0] * Mx[2][0] + Mx[1][1] * Mx[2][1] + Mx[1][2] * Mx[2][2];
    c6[5] = Mx[2][0] * Mx[2][0] + Mx[2][1] * Mx[2][1] + Mx[2][2] * Mx[2][2];
}

/* Everything the EWA projection needs again in the backward pass. */
typedef struct {
    float t[3];          /* clamped view-space point */
    float xmul, ymul;    /* 0 where the 1.3x frustum clamp was active (Q4) */
    float J[2][3];       /* only J00,J02,J11,J12 non-zero */
    float Wm[3][3];      /* R_w2c */
    float Mt[2][3];      /* J * R_w2c */
    float cov[3];        /* a,b,c (with +0.3 on the diagonal) */
} Ewa;

static void ewa_project(const OracleSettings* S, const float* pview, const float* c6, Ewa* e) {
    const float* V = S->view;
    float fx = (float)S->W / (2.f * S->tanfovx);
    float fy = (float)S->H / (2.f * S->tanfovy);
    float limx = 1.3f * S->tanfovx, limy = 1.3f * S->tanfovy;
    float tz = pview[2];
    float txtz = pview[0] / tz, tytz = pview[1] / tz;
    float cx = fminf(limx, fmaxf(-limx, txtz));
    float cy = fminf(limy, fmaxf(-limy, tytz));
    e->xmul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
    e->ymul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
    e->t[0] = cx * tz;
    e->t[1] = cy * tz;
    e->t[2] = tz;
    memset(e->J, 0, sizeof(e->J));
    e->J[0][0] = fx / tz;
    e->J[0][2] = -(fx * e->t[0]) / (tz * tz);
    e->J[1][1] = fy / tz;
    e->J[1][2] = -(fy * e->t[1]) / (tz * tz);
    for (int i = 0; i < 3; i++)
        for (int k = 0; k < 3; k++) e->Wm[i][k] = V[k * 4 + i];
    for (int k = 0; k < 3; k++) {
        e->Mt[0][k] = e->J[0][0] * e->Wm[0][k] + e->J[0][2] * e->Wm[2][k];
        e->Mt[1][k] = e->J[1][1] * e->Wm[1][k] + e->J[1][2] * e->Wm[2][k];
    }
    float Sg[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
    float N[2][3];
    for (int i = 0; i < 2; i++)
        for (int k = 0; k < 3; k++)
            N[i][k] = e->Mt[i][0] * Sg[0][k] + e->Mt[i][1] * Sg[1][k] + e->Mt[i][2] * Sg[2][k];
    float c00 = N[0][0] * e->Mt[0][0] + N[0][1] * e->Mt[0][1] + N[0][2] * e->Mt[0][2];
    float c01 = N[0][0] * e->Mt[1][0] + N[0][1] * e->Mt[1][1] + N[0][2] * e->Mt[1][2];
    float c11 = N[1][0] * e->Mt[1][0] + N[1][1] * e->Mt[1][1] + N[1][2] * e->Mt[1][2];
    e->cov[0] = c00 + 0.3f;
    e->cov[1] = c01;
    e->cov[2] = c11 + 0.3f;
}

static void sh_to_rgb(const OracleSettings* S, const float* p, const float* sh /* M x 3 */,
                      float* rgb, uint8_t* clamped) {
    /* src/utils/gaussian_model_utils.py:59-114; shs layout (P, M, 3) gaussian_model.py:116-119 */
    float res[3];
    for (int c = 0; c < 3; c++) res[c] = SH_C0 * sh[0 * 3 + c];
    if (S->degree > 0) {
        float dx = p[0] - S->campos[0], dy = p[1] - S->campos[1], dz = p[2] - S->campos[2];
        float len = sqrtf(dx * dx + dy * dy + dz * dz);
        float x = dx / len, y = dy / len, z = dz / len;
        for (int c = 0; c < 3; c++)
            res[c] = res[c] - SH_C1 * y * sh[1 * 3 + c] + SH_C1 * z * sh[2 * 3 + c] -
                     SH_C1 * x * sh[3 * 3 + c];
        if (S->degree > 1) {
            float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            for (int c = 0; c < 3; c++)
                res[c] = res[c] + SH_C2[0] * xy * sh[4 * 3 + c] + SH_C2[1] * yz * sh[5 * 3 + c] +
                         SH_C2[2] * (2.f * zz - xx - yy) * sh[6 * 3 + c] +
                         SH_C2[3] * xz * sh[7 * 3 + c] + SH_C2[4] * (xx - yy) * sh[8 * 3 + c];
            if (S->degree > 2) {
                for (int c = 0; c < 3; c++)
                    res[c] = res[c] + SH_C3[0] * y * (3.f * xx - yy) * sh[9 * 3 + c] +
                             SH_C3[1] * xy * z * sh[10 * 3 + c] +
                             SH_C3[2] * y * (4.f * zz - xx - yy) * sh[11 * 3 + c] +
                             SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy) * sh[12 * 3 + c] +
                             SH_C3[4] * x * (4.f * zz - xx - yy) * sh[13 * 3 + c] +
                             SH_C3[5] * z * (xx - yy) * sh[14 * 3 + c] +
                             SH_C3[6] * x * (xx - 3.f * yy) * sh[15 * 3 + c];
            }
        }
    }
    for (int c = 0; c < 3; c++) {
        res[c] = res[c] + 0.5f;
        clamped[c] = res[c] < 0.f;
        rgb[c] = fmaxf(res[c], 0.f);
    }
}

/* ------------------------------------------------------------------------------------------
 * A.1  preprocess.  Outputs are zero/cleared for culled Gaussians.
 * colors_precomp (P,3) may be NULL; if given it replaces the SH evaluation (no clamp).
 * rect: (P,4) = min.x, min.y, max.x, max.y in tiles.
 * ------------------------------------------------------------------------------------------ */
int oracle_preprocess(const OracleSettings* S, const float* means3D, const float* scales,
                      const float* rots, const float* opacities, const float* shs,
                      const float* colors_precomp, float* depths, int32_t* radii, float* xy,
                      float* conic_opacity, float* rgb, uint8_t* clamped, int32_t* rect,
                      uint32_t* tiles_touched, float* cov3D) {
    const int gx = (S->W + TILE - 1) / TILE, gy = (S->H + TILE - 1) / TILE;
    const float* V = S->view;
    const float* F = S->proj;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < S->P; i++) {
        radii[i] = 0;
        tiles_touched[i] = 0;
        depths[i] = 0.f;
        xy[2 * i] = xy[2 * i + 1] = 0.f;
        for (int k = 0; k < 4; k++) conic_opacity[4 * i + k] = 0.f, rect[4 * i + k] = 0;
        for (int k = 0; k < 3; k++) rgb[3 * i + k] = 0.f, clamped[3 * i + k] = 0;
        for (int k = 0; k < 6; k++) cov3D[6 * i + k] = 0.f;
        const float* p = means3D + 3 * i;
        float pv[3];
        pv[0] = V[0] * p[0] + V[4] * p[1] + V[8] * p[2] + V[12];
        pv[1] = V[1] * p[0] + V[5] * p[1] + V[9] * p[2] + V[13];
        pv[2] = V[2] * p[0] + V[6] * p[1] + V[10] * p[2] + V[14];
        if (pv[2] <= 0.2f) continue;
        float hx = F[0] * p[0] + F[4] * p[1] + F[8] * p[2] + F[12];
        float hy = F[1] * p[0] + F[5] * p[1] + F[9] * p[2] + F[13];
        float hw = F[3] * p[0] + F[7] * p[1] + F[11] * p[2] + F[15];
        float pw = 1.0f / (hw + 0.0000001f);
        float ppx = hx * pw, ppy = hy * pw;
        float c6[6];
        cov3d_from_scale_rot(scales + 3 * i, S->scale_modifier, rots + 4 * i, c6);
        Ewa e;
        ewa_project(S, pv, c6, &e);
        float a = e.cov[0], b = e.cov[1], c = e.cov[2];
        float det = a * c - b * b;
        if (det == 0.0f) continue;
        float det_inv = 1.f / det;
        float mid = 0.5f * (a + c);
        float sq = sqrtf(fmaxf(0.1f, mid * mid - det));
        float lambda1 = mid + sq, lambda2 = mid - sq;
        float my_radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
        float px = ((ppx + 1.0f) * (float)S->W - 1.0f) * 0.5f;
        float py = ((ppy + 1.0f) * (float)S->H - 1.0f) * 0.5f;
        int r = f2i_rz(my_radius);
        float rf = (float)r;
        int rminx = imin(gx, imax(0, f2i_rz((px - rf) / 16.0f)));
        int rminy = imin(gy, imax(0, f2i_rz((py - rf) / 16.0f)));
        int rmaxx = imin(gx, imax(0, f2i_rz((px + rf + 16.0f - 1.0f) / 16.0f)));
        int rmaxy = imin(gy, imax(0, f2i_rz((py + rf + 16.0f - 1.0f) / 16.0f)));
        if ((rmaxx - rminx) * (rmaxy - rminy) == 0) continue;
        if (colors_precomp) {
            for (int k = 0; k < 3; k++) rgb[3 * i + k] = colors_precomp[3 * i + k];
        } else {
            sh_to_rgb(S, p, shs + (size_t)i * S->M * 3, rgb + 3 * i, clamped + 3 * i);
        }
        for (int k = 0; k < 6; k++) cov3D[6 * i + k] = c6[k];
        depths[i] = pv[2];
        radii[i] = r;
        xy[2 * i] = px;
        xy[2 * i + 1] = py;
        conic_opacity[4 * i + 0] = c * det_inv;
        conic_opacity[4 * i + 1] = -b * det_inv;
        conic_opacity[4 * i + 2] = a * det_inv;
        conic_opacity[4 * i + 3] = opacities[i];
        rect[4 * i + 0] = rminx;
        rect[4 * i + 1] = rminy;
        rect[4 * i + 2] = rmaxx;
        rect[4 * i + 3] = rmaxy;
        tiles_touched[i] = (uint32_t)((rmaxx - rminx) * (rmaxy - rminy));
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * A.2  binning: inclusive scan, key duplication, stable (tile|depth) sort, tile ranges.
 * ------------------------------------------------------------------------------------------ */
int64_t oracle_scan(int32_t P, const uint32_t* tiles_touched, uint32_t* offsets) {
    uint64_t acc = 0;
    for (int i = 0; i < P; i++) {
        acc += tiles_touched[i];
        offsets[i] = (uint32_t)acc;
    }
    return (int64_t)acc;
}

typedef struct { uint64_t key; uint32_t val; uint32_t pos; } KV;
static int kv_cmp(const void* a, const void* b) {
    const KV* x = (const KV*)a; const KV* y = (const KV*)b;
    if (x->key != y->key) return x->key < y->key ? -1 : 1;
    return x->pos < y->pos ? -1 : (x->pos > y->pos ? 1 : 0);   /* stable: emission order (Q9) */
}

int oracle_binning(const OracleSettings* S, const float* depths, const int32_t* rect,
                   const uint32_t* offsets, int64_t K, uint64_t* keys_sorted,
                   uint32_t* vals_sorted, uint32_t* ranges /* G x 2 */) {
    const int gx = (S->W + TILE - 1) / TILE, gy = (S->H + TILE - 1) / TILE;
    KV* kv = (KV*)malloc(sizeof(KV) * (size_t)(K > 0 ? K : 1));
    if (!kv) return -1;
    for (int i = 0; i < S->P; i++) {
        uint32_t off = (i == 0) ? 0 : offsets[i - 1];
        if (offsets[i] == off) continue;
        const int32_t* r = rect + 4 * i;
        for (int y = r[1]; y < r[3]; y++)
            for (int x = r[0]; x < r[2]; x++) {
                uint64_t key = (uint64_t)(uint32_t)(y * gx + x);
                key <<= 32;
                key |= f2u(depths[i]);
                kv[off].key = key;
                kv[off].val = (uint32_t)i;
                kv[off].pos = off;
                off++;
            }
    }
    qsort(kv, (size_t)K, sizeof(KV), kv_cmp);
    memset(ranges, 0, sizeof(uint32_t) * 2 * (size_t)gx * gy);
    for (int64_t k = 0; k < K; k++) {
        keys_sorted[k] = kv[k].key;
        vals_sorted[k] = kv[k].val;
        uint32_t tile = (uint32_t)(kv[k].key >> 32);
        if (k == 0 || tile != (uint32_t)(kv[k - 1].key >> 32)) ranges[2 * tile] = (uint32_t)k;
        if (k == K - 1 || tile != (uint32_t)(kv[k + 1].key >> 32))
            ranges[2 * tile + 1] = (uint32_t)(k + 1);
    }
    free(kv);
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * A.3  blend forward.  color (3,H,W), depth/alpha/final_T (H,W), n_contrib (H,W).
 * third = sum(alpha_i T_i) for variant A, 1 - T for variant B.
 * fragile (H,W) flags pixels where one of the hard tests (Q6, Q10) is within a 1e-5 relative
 * band of its threshold: any two correct fp32 implementations may disagree there.
 * n_touched_fragile (P) likewise for the n_touched counter.  pair_count = S of SURVEY 8(d).
 * ------------------------------------------------------------------------------------------ */
int oracle_blend_forward(const OracleSettings* S, const uint32_t* ranges, const uint32_t* vals,
                         const float* xy, const float* conic_opacity, const float* rgb,
                         const float* depths, float* out_color, float* out_depth,
                         float* out_third, float* final_T, uint32_t* n_contrib,
                         int32_t* n_touched, uint8_t* fragile, uint8_t* n_touched_fragile,
                         int64_t* pair_count) {
    const int W = S->W, H = S->H;
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    const float band = 1e-5f;
    int64_t pairs = 0;
    if (n_touched) memset(n_touched, 0, sizeof(int32_t) * (size_t)S->P);
    if (n_touched_fragile) memset(n_touched_fragile, 0, (size_t)S->P);
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : pairs)
    for (int tile = 0; tile < gx * gy; tile++) {
        int tx = tile % gx, ty = tile / gx;
        uint32_t lo = ranges[2 * tile], hi = ranges[2 * tile + 1];
        for (int ly = 0; ly < TILE; ly++)
            for (int lx = 0; lx < TILE; lx++) {
                int px = tx * TILE + lx, py = ty * TILE + ly;
                if (px >= W || py >= H) continue;
                float pxf = (float)px, pyf = (float)py;
                float T = 1.f, C[3] = {0, 0, 0}, D = 0.f, A = 0.f;
                uint32_t contributor = 0, last = 0;
                uint8_t frag = 0;
                for (uint32_t k = lo; k < hi; k++) {
                    contributor++;
                    pairs++;
                    uint32_t g = vals[k];
                    float dx = xy[2 * g] - pxf, dy = xy[2 * g + 1] - pyf;
                    const float* co = conic_opacity + 4 * g;
                    float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
                    if (power > 0.f) continue;
                    float alpha = fminf(0.99f, co[3] * expf(power));
                    float thr = 1.0f / 255.0f;
                    if ((alpha * (1.f - band) < thr) != (alpha * (1.f + band) < thr)) frag = 1;
                    if (alpha < thr) continue;
                    float test_T = T * (1.f - alpha);
                    if ((test_T * (1.f - band) < 0.0001f) != (test_T * (1.f + band) < 0.0001f))
                        frag = 1;
                    if (test_T < 0.0001f) break;
                    float w = alpha * T;
                    for (int c = 0; c < 3; c++) C[c] += rgb[3 * g + c] * w;
                    D += depths[g] * w;
                    A += w;
                    if (n_touched) {
                        if ((test_T * (1.f - band) > 0.5f) != (test_T * (1.f + band) > 0.5f)) {
                            if (n_touched_fragile) {
#pragma omp atomic write
                                n_touched_fragile[g] = 1;
                            }
                        }
                        if (test_T > 0.5f) {
#pragma omp atomic
                            n_touched[g] += 1;
                        }
                    }
                    T = test_T;
                    last = contributor;
                }
                size_t pix = (size_t)py * W + px;
                final_T[pix] = T;
                n_contrib[pix] = last;
                for (int c = 0; c < 3; c++)
                    out_color[(size_t)c * H * W + pix] = C[c] + T * S->bg[c];
                out_depth[pix] = D;
                out_third[pix] = S->variant == 0 ? A : 1.f - T;
                if (fragile) fragile[pix] = frag;
            }
    }
    if (pair_count) *pair_count = pairs;
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * A.4  blend backward.  Per-Gaussian sums are accumulated in double (the reference uses fp32
 * atomics in arbitrary order, Q: order-of-summation) and rounded once at the end.
 * dL_dmean2D (P,2), dL_dconic (P,3) [y is the half off-diagonal convention, Q8],
 * dL_dopacity (P), dL_dcolor (P,3), dL_ddepth (P).
 * dL_dthird may be NULL (variant B upstream ignores the opacity-image gradient).
 * ------------------------------------------------------------------------------------------ */
int oracle_blend_backward(const OracleSettings* S, const uint32_t* ranges, const uint32_t* vals,
                          const float* xy, const float* conic_opacity, const float* rgb,
                          const float* depths, const float* final_T, const uint32_t* n_contrib,
                          const float* dL_dcolor_pix, const float* dL_ddepth_pix,
                          const float* dL_dthird_pix, float* dL_dmean2D, float* dL_dconic,
                          float* dL_dopacity, float* dL_dcolor, float* dL_ddepth) {
    const int W = S->W, H = S->H, P = S->P;
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    double* acc = (double*)calloc((size_t)P * 10, sizeof(double));
    if (!acc) return -1;
    const float ddelx_dx = 0.5f * (float)W, ddely_dy = 0.5f * (float)H;
#pragma omp parallel
    {
        double* loc = (double*)calloc((size_t)P * 10, sizeof(double));
#pragma omp for schedule(dynamic, 1)
        for (int tile = 0; tile < gx * gy; tile++) {
            int tx = tile % gx, ty = tile / gx;
            uint32_t lo = ranges[2 * tile], hi = ranges[2 * tile + 1];
            for (int ly = 0; ly < TILE; ly++)
                for (int lx = 0; lx < TILE; lx++) {
                    int px = tx * TILE + lx, py = ty * TILE + ly;
                    if (px >= W || py >= H) continue;
                    size_t pix = (size_t)py * W + px;
                    float pxf = (float)px, pyf = (float)py;
                    const float T_final = final_T[pix];
                    float T = T_final;
                    uint32_t last_contributor = n_contrib[pix];
                    float accum_rec[3] = {0, 0, 0}, accum_d = 0.f, accum_a = 0.f;
                    float dpix[3], dD = dL_ddepth_pix[pix];
                    float dA = dL_dthird_pix ? dL_dthird_pix[pix] : 0.f;
                    for (int c = 0; c < 3; c++) dpix[c] = dL_dcolor_pix[(size_t)c * H * W + pix];
                    float last_alpha = 0.f, last_color[3] = {0, 0, 0}, last_depth = 0.f;
                    float bg_dot = S->bg[0] * dpix[0] + S->bg[1] * dpix[1] + S->bg[2] * dpix[2];
                    /* walk back-to-front over the entries this pixel actually iterated */
                    for (uint32_t n = last_contributor; n > 0; n--) {
                        uint32_t k = lo + n - 1;
                        if (k >= hi) continue;
                        uint32_t g = vals[k];
                        float dx = xy[2 * g] - pxf, dy = xy[2 * g + 1] - pyf;
                        const float* co = conic_opacity + 4 * g;
                        float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
                        if (power > 0.f) continue;
                        float G = expf(power);
                        float alpha = fminf(0.99f, co[3] * G);
                        if (alpha < 1.0f / 255.0f) continue;
                        T = T / (1.f - alpha);
                        float w = alpha * T;
                        float dL_dalpha = 0.f;
                        double* a = loc + (size_t)g * 10;
                        for (int c = 0; c < 3; c++) {
                            float col = rgb[3 * g + c];
                            accum_rec[c] = last_alpha * last_color[c] + (1.f - last_alpha) * accum_rec[c];
                            last_color[c] = col;
                            dL_dalpha += (col - accum_rec[c]) * dpix[c];
                            a[6 + c] += (double)(w * dpix[c]);
                        }
                        float dep = depths[g];
                        accum_d = last_alpha * last_depth + (1.f - last_alpha) * accum_d;
                        last_depth = dep;
                        dL_dalpha += (dep - accum_d) * dD;
                        a[9] += (double)(w * dD);
                        accum_a = last_alpha + (1.f - last_alpha) * accum_a;
                        dL_dalpha += (1.f - accum_a) * dA;
                        dL_dalpha *= T;
                        last_alpha = alpha;
                        dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot;
                        float dL_dG = co[3] * dL_dalpha;          /* Q1: clamp is straight-through */
                        float gdx = G * dx, gdy = G * dy;
                        float dG_ddelx = -gdx * co[0] - gdy * co[1];
                        float dG_ddely = -gdy * co[2] - gdx * co[1];
                        a[0] += (double)(dL_dG * dG_ddelx * ddelx_dx);
                        a[1] += (double)(dL_dG * dG_ddely * ddely_dy);
                        a[2] += (double)(-0.5f * gdx * dx * dL_dG);
                        a[3] += (double)(-0.5f * gdx * dy * dL_dG);
                        a[4] += (double)(-0.5f * gdy * dy * dL_dG);
                        a[5] += (double)(G * dL_dalpha);
                    }
                }
        }
#pragma omp critical
        {
            for (size_t i = 0; i < (size_t)P * 10; i++) acc[i] += loc[i];
        }
        free(loc);
    }
    for (int g = 0; g < P; g++) {
        const double* a = acc + (size_t)g * 10;
        dL_dmean2D[2 * g] = (float)a[0];
        dL_dmean2D[2 * g + 1] = (float)a[1];
        dL_dconic[3 * g] = (float)a[2];
        dL_dconic[3 * g + 1] = (float)a[3];
        dL_dconic[3 * g + 2] = (float)a[4];
        dL_dopacity[g] = (float)a[5];
        dL_dcolor[3 * g] = (float)a[6];
        dL_dcolor[3 * g + 1] = (float)a[7];
        dL_dcolor[3 * g + 2] = (float)a[8];
        dL_ddepth[g] = (float)a[9];
    }
    free(acc);
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * A.5  preprocess backward (+ closed-form camera-pose gradient for variant B).
 * dL_dtau: 6 doubles-as-floats [rho(3), theta(3)] (may be NULL).  Only degree-0 SH backward
 * contributes no mean gradient; degrees 1..3 add the view-direction term.
 * dL_dcolors_precomp (P,3) may be NULL, dL_dshs (P,M,3) may be NULL.
 * ------------------------------------------------------------------------------------------ */
int oracle_preprocess_backward(const OracleSettings* S, const float* means3D, const float* scales,
                               const float* rots, const float* shs, const int32_t* radii,
                               const uint8_t* clamped, const float* dL_dmean2D,
                               const float* dL_dconic, const float* dL_dcolor,
                               const float* dL_ddepth, float* dL_dmeans3D, float* dL_dscales,
                               float* dL_drots, float* dL_dshs, float* dL_dcolors_precomp,
                               float* dL_dtau) {
    const float* V = S->view;
    const float* F = S->proj;
    const float fx = (float)S->W / (2.f * S->tanfovx);
    const float fy = (float)S->H / (2.f * S->tanfovy);
    double pose[15];
    memset(pose, 0, sizeof(pose));
    for (int i = 0; i < S->P; i++) {
        float* dm = dL_dmeans3D + 3 * i;
        dm[0] = dm[1] = dm[2] = 0.f;
        for (int k = 0; k < 3; k++) dL_dscales[3 * i + k] = 0.f;
        for (int k = 0; k < 4; k++) dL_drots[4 * i + k] = 0.f;
        if (dL_dshs) memset(dL_dshs + (size_t)i * S->M * 3, 0, sizeof(float) * S->M * 3);
        if (dL_dcolors_precomp) for (int k = 0; k < 3; k++) dL_dcolors_precomp[3 * i + k] = 0.f;
        if (radii[i] <= 0) continue;
        const float* p = means3D + 3 * i;
        float pv[3];
        pv[0] = V[0] * p[0] + V[4] * p[1] + V[8] * p[2] + V[12];
        pv[1] = V[1] * p[0] + V[5] * p[1] + V[9] * p[2] + V[13];
        pv[2] = V[2] * p[0] + V[6] * p[1] + V[10] * p[2] + V[14];
        float c6[6];
        cov3d_from_scale_rot(scales + 3 * i, S->scale_modifier, rots + 4 * i, c6);
        Ewa e;
        ewa_project(S, pv, c6, &e);
        float a = e.cov[0], b = e.cov[1], c = e.cov[2];
        float denom = a * c - b * b;
        float denom2inv = 1.0f / (denom * denom + 0.0000001f);      /* Q2 */
        float gxx = dL_dconic[3 * i], gxy = dL_dconic[3 * i + 1], gyy = dL_dconic[3 * i + 2];
        float dLa = 0.f, dLb = 0.f, dLc = 0.f;
        float dS[6] = {0, 0, 0, 0, 0, 0};
        float dMt[2][3] = {{0, 0, 0}, {0, 0, 0}};
        if (denom2inv != 0.f) {
            dLa = denom2inv * (-c * c * gxx + 2.f * b * c * gxy + (denom - a * c) * gyy);
            dLc = denom2inv * (-a * a * gyy + 2.f * a * b * gxy + (denom - a * c) * gxx);
            dLb = denom2inv * 2.f * (b * c * gxx - (denom + 2.f * b * b) * gxy + a * b * gyy);
            const float(*Mt)[3] = e.Mt;
            dS[0] = Mt[0][0] * Mt[0][0] * dLa + Mt[0][0] * Mt[1][0] * dLb + Mt[1][0] * Mt[1][0] * dLc;
            dS[3] = Mt[0][1] * Mt[0][1] * dLa + Mt[0][1] * Mt[1][1] * dLb + Mt[1][1] * Mt[1][1] * dLc;
            dS[5] = Mt[0][2] * Mt[0][2] * dLa + Mt[0][2] * Mt[1][2] * dLb + Mt[1][2] * Mt[1][2] * dLc;
            dS[1] = 2.f * Mt[0][0] * Mt[0][1] * dLa + (Mt[0][0] * Mt[1][1] + Mt[0][1] * Mt[1][0]) * dLb +
                    2.f * Mt[1][0] * Mt[1][1] * dLc;
            dS[2] = 2.f * Mt[0][0] * Mt[0][2] * dLa + (Mt[0][0] * Mt[1][2] + Mt[0][2] * Mt[1][0]) * dLb +
                    2.f * Mt[1][0] * Mt[1][2] * dLc;
            dS[4] = 2.f * Mt[0][2] * Mt[0][1] * dLa + (Mt[0][1] * Mt[1][2] + Mt[0][2] * Mt[1][1]) * dLb +
                    2.f * Mt[1][1] * Mt[1][2] * dLc;
            float Sg[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
            for (int k = 0; k < 3; k++) {
                float m0 = Mt[0][0] * Sg[0][k] + Mt[0][1] * Sg[1][k] + Mt[0][2] * Sg[2][k];
                float m1 = Mt[1][0] * Sg[0][k] + Mt[1][1] * Sg[1][k] + Mt[1][2] * Sg[2][k];
                dMt[0][k] = 2.f * m0 * dLa + m1 * dLb;
                dMt[1][k] = 2.f * m1 * dLc + m0 * dLb;
            }
        }
        /* dL/dJ = dL/dMt * W^T */
        float dJ00 = dMt[0][0] * e.Wm[0][0] + dMt[0][1] * e.Wm[0][1] + dMt[0][2] * e.Wm[0][2];
        float dJ02 = dMt[0][0] * e.Wm[2][0] + dMt[0][1] * e.Wm[2][1] + dMt[0][2] * e.Wm[2][2];
        float dJ11 = dMt[1][0] * e.Wm[1][0] + dMt[1][1] * e.Wm[1][1] + dMt[1][2] * e.Wm[1][2];
        float dJ12 = dMt[1][0] * e.Wm[2][0] + dMt[1][1] * e.Wm[2][1] + dMt[1][2] * e.Wm[2][2];
        float tz = 1.f / e.t[2], tz2 = tz * tz, tz3 = tz2 * tz;
        float dt[3];
        dt[0] = e.xmul * (-fx * tz2 * dJ02);
        dt[1] = e.ymul * (-fy * tz2 * dJ12);
        dt[2] = -fx * tz2 * dJ00 - fy * tz2 * dJ11 + (2.f * fx * e.t[0]) * tz3 * dJ02 +
                (2.f * fy * e.t[1]) * tz3 * dJ12;
        /* world mean via cov path: W^T dt */
        float dmean_cov[3];
        for (int k = 0; k < 3; k++)
            dmean_cov[k] = e.Wm[0][k] * dt[0] + e.Wm[1][k] * dt[1] + e.Wm[2][k] * dt[2];
        /* projection path */
        float hx = F[0] * p[0] + F[4] * p[1] + F[8] * p[2] + F[12];
        float hy = F[1] * p[0] + F[5] * p[1] + F[9] * p[2] + F[13];
        float hw = F[3] * p[0] + F[7] * p[1] + F[11] * p[2] + F[15];
        float mw = 1.0f / (hw + 0.0000001f);                           /* Q3 */
        float mul1 = hx * mw * mw, mul2 = hy * mw * mw;
        float g2x = dL_dmean2D[2 * i], g2y = dL_dmean2D[2 * i + 1];
        float dmean_proj[3];
        dmean_proj[0] = (F[0] * mw - F[3] * mul1) * g2x + (F[1] * mw - F[3] * mul2) * g2y;
        dmean_proj[1] = (F[4] * mw - F[7] * mul1) * g2x + (F[5] * mw - F[7] * mul2) * g2y;
        dmean_proj[2] = (F[8] * mw - F[11] * mul1) * g2x + (F[9] * mw - F[11] * mul2) * g2y;
        /* depth path: depth = row 2 of the view transform */
        float dd = dL_ddepth[i];
        float dmean_dep[3] = {V[2] * dd, V[6] * dd, V[10] * dd};
        float dmean_sh[3] = {0, 0, 0};
        /* colour path */
        if (dL_dcolors_precomp) {
            for (int k = 0; k < 3; k++) dL_dcolors_precomp[3 * i + k] = dL_dcolor[3 * i + k];
        } else if (dL_dshs) {
            float dRGB[3];
            for (int k = 0; k < 3; k++) dRGB[k] = clamped[3 * i + k] ? 0.f : dL_dcolor[3 * i + k];
            float* dsh = dL_dshs + (size_t)i * S->M * 3;
            const float* sh = shs + (size_t)i * S->M * 3;
            for (int k = 0; k < 3; k++) dsh[0 * 3 + k] = SH_C0 * dRGB[k];
            if (S->degree > 0) {
                float ox = p[0] - S->campos[0], oy = p[1] - S->campos[1], oz = p[2] - S->campos[2];
                float len = sqrtf(ox * ox + oy * oy + oz * oz);
                float x = ox / len, y = oy / len, z = oz / len;
                float dRGBdx[3] = {0, 0, 0}, dRGBdy[3] = {0, 0, 0}, dRGBdz[3] = {0, 0, 0};
                for (int k = 0; k < 3; k++) {
                    dsh[1 * 3 + k] = -SH_C1 * y * dRGB[k];
                    dsh[2 * 3 + k] = SH_C1 * z * dRGB[k];
                    dsh[3 * 3 + k] = -SH_C1 * x * dRGB[k];
                    dRGBdx[k] = -SH_C1 * sh[3 * 3 + k];
                    dRGBdy[k] = -SH_C1 * sh[1 * 3 + k];
                    dRGBdz[k] = SH_C1 * sh[2 * 3 + k];
                }
                if (S->degree > 1) {
                    float xx = x * x, yy = y * y, zz = z * z, xy_ = x * y, yz = y * z, xz = x * z;
                    for (int k = 0; k < 3; k++) {
                        dsh[4 * 3 + k] = SH_C2[0] * xy_ * dRGB[k];
                        dsh[5 * 3 + k] = SH_C2[1] * yz * dRGB[k];
                        dsh[6 * 3 + k] = SH_C2[2] * (2.f * zz - xx - yy) * dRGB[k];
                        dsh[7 * 3 + k] = SH_C2[3] * xz * dRGB[k];
                        dsh[8 * 3 + k] = SH_C2[4] * (xx - yy) * dRGB[k];
                        dRGBdx[k] += SH_C2[0] * y * sh[4 * 3 + k] + SH_C2[2] * 2.f * -x * sh[6 * 3 + k] +
                                     SH_C2[3] * z * sh[7 * 3 + k] + SH_C2[4] * 2.f * x * sh[8 * 3 + k];
                        dRGBdy[k] += SH_C2[0] * x * sh[4 * 3 + k] + SH_C2[1] * z * sh[5 * 3 + k] +
                                     SH_C2[2] * 2.f * -y * sh[6 * 3 + k] + SH_C2[4] * 2.f * -y * sh[8 * 3 + k];
                        dRGBdz[k] += SH_C2[1] * y * sh[5 * 3 + k] + SH_C2[2] * 2.f * 2.f * z * sh[6 * 3 + k] +
                                     SH_C2[3] * x * sh[7 * 3 + k];
                    }
                    if (S->degree > 2) {
                        for (int k = 0; k < 3; k++) {
                            dsh[9 * 3 + k] = SH_C3[0] * y * (3.f * xx - yy) * dRGB[k];
                            dsh[10 * 3 + k] = SH_C3[1] * xy_ * z * dRGB[k];
                            dsh[11 * 3 + k] = SH_C3[2] * y * (4.f * zz - xx - yy) * dRGB[k];
                            dsh[12 * 3 + k] = SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy) * dRGB[k];
                            dsh[13 * 3 + k] = SH_C3[4] * x * (4.f * zz - xx - yy) * dRGB[k];
                            dsh[14 * 3 + k] = SH_C3[5] * z * (xx - yy) * dRGB[k];
                            dsh[15 * 3 + k] = SH_C3[6] * x * (xx - 3.f * yy) * dRGB[k];
                            dRGBdx[k] += SH_C3[0] * sh[9 * 3 + k] * 3.f * 2.f * xy_ +
                                         SH_C3[1] * sh[10 * 3 + k] * yz +
                                         SH_C3[2] * sh[11 * 3 + k] * -2.f * xy_ +
                                         SH_C3[3] * sh[12 * 3 + k] * -3.f * 2.f * xz +
                                         SH_C3[4] * sh[13 * 3 + k] * (-3.f * xx + 4.f * zz - yy) +
                                         SH_C3[5] * sh[14 * 3 + k] * 2.f * xz +
                                         SH_C3[6] * sh[15 * 3 + k] * 3.f * (xx - yy);
                            dRGBdy[k] += SH_C3[0] * sh[9 * 3 + k] * 3.f * (xx - yy) +
                                         SH_C3[1] * sh[10 * 3 + k] * xz +
                                         SH_C3[2] * sh[11 * 3 + k] * (-3.f * yy + 4.f * zz - xx) +
                                         SH_C3[3] * sh[12 * 3 + k] * -3.f * 2.f * yz +
                                         SH_C3[4] * sh[13 * 3 + k] * -2.f * xy_ +
                                         SH_C3[5] * sh[14 * 3 + k] * -2.f * yz +
                                         SH_C3[6] * sh[15 * 3 + k] * -3.f * 2.f * xy_;
                            dRGBdz[k] += SH_C3[1] * sh[10 * 3 + k] * xy_ +
                                         SH_C3[2] * sh[11 * 3 + k] * 4.f * 2.f * yz +
                                         SH_C3[3] * sh[12 * 3 + k] * 3.f * (2.f * zz - xx - yy) +
                                         SH_C3[4] * sh[13 * 3 + k] * 4.f * 2.f * xz +
                                         SH_C3[5] * sh[14 * 3 + k] * (xx - yy);
                        }
                    }
                }
                float ddir[3] = {0, 0, 0};
                for (int k = 0; k < 3; k++) {
                    ddir[0] += dRGBdx[k] * dRGB[k];
                    ddir[1] += dRGBdy[k] * dRGB[k];
                    ddir[2] += dRGBdz[k] * dRGB[k];
                }
                /* backward of normalisation dir = o/|o| */
                float inv = 1.f / len;
                float dot = (x * ddir[0] + y * ddir[1] + z * ddir[2]);
                dmean_sh[0] = (ddir[0] - x * dot) * inv;
                dmean_sh[1] = (ddir[1] - y * dot) * inv;
                dmean_sh[2] = (ddir[2] - z * dot) * inv;
            }
        }
        for (int k = 0; k < 3; k++) dm[k] = dmean_cov[k] + dmean_proj[k] + dmean_dep[k] + dmean_sh[k];
        /* Sigma -> scale, rotation */
        {
            float R[3][3];
            quat_to_R(rots + 4 * i, R);
            float s[3] = {S->scale_modifier * scales[3 * i], S->scale_modifier * scales[3 * i + 1],
                          S->scale_modifier * scales[3 * i + 2]};
            float Dm[3][3] = {{dS[0], 0.5f * dS[1], 0.5f * dS[2]},
                              {0.5f * dS[1], dS[3], 0.5f * dS[4]},
                              {0.5f * dS[2], 0.5f * dS[4], dS[5]}};
            float Mx[3][3], dM[3][3], dR[3][3];
            for (int r = 0; r < 3; r++)
                for (int k = 0; k < 3; k++) Mx[r][k] = R[r][k] * s[k];
            for (int r = 0; r < 3; r++)
                for (int k = 0; k < 3; k++)
                    dM[r][k] = 2.f * (Dm[r][0] * Mx[0][k] + Dm[r][1] * Mx[1][k] + Dm[r][2] * Mx[2][k]);
            for (int k = 0; k < 3; k++) {
                dL_dscales[3 * i + k] =
                    S->scale_modifier * (dM[0][k] * R[0][k] + dM[1][k] * R[1][k] + dM[2][k] * R[2][k]);
                for (int r = 0; r < 3; r++) dR[r][k] = dM[r][k] * s[k];
            }
            float qr = rots[4 * i], qx = rots[4 * i + 1], qy = rots[4 * i + 2], qz = rots[4 * i + 3];
            dL_drots[4 * i + 0] = 2.f * (-qz * dR[0][1] + qy * dR[0][2] + qz * dR[1][0] - qx * dR[1][2] -
                                         qy * dR[2][0] + qx * dR[2][1]);
            dL_drots[4 * i + 1] = 2.f * (qy * dR[0][1] + qz * dR[0][2] + qy * dR[1][0] - 2.f * qx * dR[1][1] -
                                         qr * dR[1][2] + qz * dR[2][0] + qr * dR[2][1] - 2.f * qx * dR[2][2]);
            dL_drots[4 * i + 2] = 2.f * (-2.f * qy * dR[0][0] + qx * dR[0][1] + qr * dR[0][2] + qx * dR[1][0] +
                                         qz * dR[1][2] - qr * dR[2][0] + qz * dR[2][1] - 2.f * qy * dR[2][2]);
            dL_drots[4 * i + 3] = 2.f * (-2.f * qz * dR[0][0] - qr * dR[0][1] + qx * dR[0][2] + qr * dR[1][0] -
                                         2.f * qz * dR[1][1] + qy * dR[1][2] + qx * dR[2][0] + qy * dR[2][1]);
        }
        /* camera-pose pieces (closed form): g = dL/dp_C with R held fixed,
         * G = dL/dR_w2c from the cov2D path.  The SH view-direction path does not go through
         * p_C (campos is held constant), so it is excluded. */
        if (dL_dtau) {
            float dm_pc[3] = {dmean_cov[0] + dmean_proj[0] + dmean_dep[0],
                              dmean_cov[1] + dmean_proj[1] + dmean_dep[1],
                              dmean_cov[2] + dmean_proj[2] + dmean_dep[2]};
            double g[3];
            for (int r = 0; r < 3; r++)
                g[r] = (double)e.Wm[r][0] * dm_pc[0] + (double)e.Wm[r][1] * dm_pc[1] +
                       (double)e.Wm[r][2] * dm_pc[2];
            double pc[3] = {pv[0], pv[1], pv[2]};
            pose[0] += g[0]; pose[1] += g[1]; pose[2] += g[2];
            pose[3] += pc[1] * g[2] - pc[2] * g[1];
            pose[4] += pc[2] * g[0] - pc[0] * g[2];
            pose[5] += pc[0] * g[1] - pc[1] * g[0];
            for (int m = 0; m < 3; m++)
                for (int k = 0; k < 3; k++)
                    pose[6 + 3 * m + k] += (double)e.J[0][m] * dMt[0][k] + (double)e.J[1][m] * dMt[1][k];
        }
    }
    if (dL_dtau) {
        /* A = (sum G) R_w2c^T ; theta += vee(A - A^T) */
        double Wm[3][3], A[3][3];
        for (int i = 0; i < 3; i++)
            for (int k = 0; k < 3; k++) Wm[i][k] = V[k * 4 + i];
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) {
                A[i][j] = 0;
                for (int k = 0; k < 3; k++) A[i][j] += pose[6 + 3 * i + k] * Wm[j][k];
            }
        dL_dtau[0] = (float)pose[0];
        dL_dtau[1] = (float)pose[1];
        dL_dtau[2] = (float)pose[2];
        dL_dtau[3] = (float)(pose[3] + (A[2][1] - A[1][2]));
        dL_dtau[4] = (float)(pose[4] + (A[0][2] - A[2][0]));
        dL_dtau[5] = (float)(pose[5] + (A[1][0] - A[0][1]));
    }
    return 0;
}
