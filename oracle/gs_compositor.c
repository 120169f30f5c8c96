/* CPU oracle of the tile compositor (rasterize_to_pixels forward + backward) in plain C + OpenMP.
 *
 * TEST INFRASTRUCTURE, NOT A PRODUCT PATH: only tests/, __graft_entry__.smoke() and bench.py's CPU legs
 * (cpu_baseline / --impl reference) load this library (through oracle/c_oracle.py). Nothing under
 * gsplat-triton_b200/ links or calls it.
 *
 * What it restates (paths relative to the reference checkout, hieu1999210/gsplat-triton @ 46beece):
 *   forward   gsplat/cuda/csrc/RasterizeToPixels3DGSFwd.cu:105-187 (== triton_impl/rasterize_to_pixels_fwd.py:105-166)
 *             math of one pair: gsplat/cuda/_torch_impl.py:490-501 (accumulate)
 *   backward  gsplat/cuda/csrc/RasterizeToPixels3DGSBwd.cu:156-243 (== triton_impl/rasterize_to_pixels_bwd.py:170-337),
 *             v_backgrounds: triton_impl/_wrapper.py:159-162
 * The reference has no CPU implementation of this stage that can run (its `_rasterize_to_pixels` needs the CUDA op
 * rasterize_to_indices_3dgs and the un-vendored third-party nerfacc), so this is a port ("kind": "port").
 * Pinning: tests/test_oracle.py checks it against tests/golden/raster.npz and tests/golden/e2e.npz — outputs and
 * autograd gradients of the reference's own `_rasterize_to_pixels` + `accumulate` frozen by oracle/make_golden.py —
 * and against the torch-autograd restatement in oracle/gs_oracle.py.
 *
 * Arithmetic: every per-pair quantity in float32 with expf (the reference's fast-math __expf differs in the last
 * ulps; the tests' tolerances cover it); gradient sums over pixels are accumulated in double (the reference sums
 * them with float atomics in an unspecified order, so any summation order is a valid reference result).
 *
 * Beyond the reference outputs, the forward pass reports per pixel how close any decision came to flipping
 * (`margin`): the tests use it to tell an alpha >= 1/255 / T <= 1e-4 / 0.999-clamp / sigma < 0 threshold flip
 * (SURVEY 8a hazard 8) from a real error.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ALPHA_MIN (1.0f / 255.0f)
#define ALPHA_MAX 0.999f
#define T_MIN 1e-4f

static inline float rel_gap(float v, float thr) { return fabsf(v - thr) / thr; }

/* Forward. Layout as the reference: means2d [P,2], conics [P,3], colors [P,D], opacities [P] (P = C*N or nnz,
 * flatten_ids index into it), offsets [C,th,tw], masks uint8 [C,th,tw] or NULL (1 = render), backgrounds [C,D] or
 * NULL. Outputs: out_colors [C,H,W,D], out_alphas [C,H,W], last_ids [C,H,W], margin [C,H,W] or NULL. */
int gso_rasterize_fwd(int C, int64_t n_isects, int D, const float *means2d, const float *conics, const float *colors,
                      const float *opacities, const float *backgrounds, const uint8_t *masks, int W, int H,
                      int tile, int tw, int th, const int32_t *offsets, const int32_t *flatten_ids,
                      float *out_colors, float *out_alphas, int32_t *last_ids, float *margin) {
    const int64_t n_tiles = (int64_t)C * th * tw;
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t t = 0; t < n_tiles; ++t) {
        const int cam = (int)(t / ((int64_t)th * tw));
        const int ty = (int)((t / tw) % th), tx = (int)(t % tw);
        const int lo = offsets[t];
        const int hi = (t + 1 < n_tiles) ? offsets[t + 1] : (int)n_isects;
        const int render = (masks == NULL) || masks[t];
        float pix[64];
        float *pixd = D <= 64 ? pix : (float *)malloc(sizeof(float) * D);
        for (int iy = 0; iy < tile; ++iy) {
            const int y = ty * tile + iy;
            if (y >= H) break;
            for (int ix = 0; ix < tile; ++ix) {
                const int x = tx * tile + ix;
                if (x >= W) break;
                const float px = (float)x + 0.5f, py = (float)y + 0.5f;
                float T = 1.0f, m = INFINITY;
                int cur = 0;
                for (int k = 0; k < D; ++k) pixd[k] = 0.f;
                for (int i = lo; render && i < hi; ++i) {
                    const int32_t g = flatten_ids[i];
                    const float dx = means2d[2 * (int64_t)g] - px, dy = means2d[2 * (int64_t)g + 1] - py;
                    const float a = conics[3 * (int64_t)g], b = conics[3 * (int64_t)g + 1], c = conics[3 * (int64_t)g + 2];
                    const float o = opacities[g];
                    const float sigma = 0.5f * (a * dx * dx + c * dy * dy) + b * dx * dy;
                    const float ov = o * expf(-sigma);
                    const float alpha = fminf(ALPHA_MAX, ov);
                    {   /* sigma < 0 can only come from rounding: how many "relative roundings" away from 0 is it */
                        const float scale = 0.5f * (fabsf(a * dx * dx) + fabsf(c * dy * dy)) + fabsf(b * dx * dy);
                        if (scale > 0.f) m = fminf(m, fabsf(sigma) / scale);
                    }
                    if (!(sigma < 0.f)) m = fminf(m, rel_gap(ov, ALPHA_MIN));
                    if (sigma < 0.f || alpha < ALPHA_MIN) continue;
                    m = fminf(m, rel_gap(ov, ALPHA_MAX));
                    const float next_T = T * (1.0f - alpha);
                    m = fminf(m, rel_gap(next_T, T_MIN));
                    if (next_T <= T_MIN) break; /* exclusive: this Gaussian is not composited */
                    const float w = alpha * T;
                    const float *col = colors + (int64_t)D * g;
                    for (int k = 0; k < D; ++k) pixd[k] += col[k] * w;
                    cur = i;
                    T = next_T;
                }
                const int64_t p = ((int64_t)cam * H + y) * W + x;
                for (int k = 0; k < D; ++k)
                    out_colors[p * D + k] = backgrounds ? pixd[k] + T * backgrounds[(int64_t)cam * D + k] : pixd[k];
                out_alphas[p] = 1.0f - T;
                last_ids[p] = cur;
                if (margin) margin[p] = m;
            }
        }
        if (pixd != pix) free(pixd);
    }
    return 0;
}

static inline void atomic_add_d(double *p, double v) {
#pragma omp atomic
    *p += v;
}

/* Backward. v_* outputs are float arrays shaped like their primals and are OVERWRITTEN; v_means2d_abs may be NULL.
 * pixel_weight (uint8 [C,H,W] or NULL): pixels with 0 are left out of the loss (their cotangents are taken as zero);
 * the tests use it to exclude pixels whose forward decisions sit on a threshold. */
int gso_rasterize_bwd(int C, int64_t P, int64_t n_isects, int D, const float *means2d, const float *conics,
                      const float *colors, const float *opacities, const float *backgrounds, const uint8_t *masks,
                      int W, int H, int tile, int tw, int th, const int32_t *offsets, const int32_t *flatten_ids,
                      const float *render_alphas, const int32_t *last_ids, const float *v_render_colors,
                      const float *v_render_alphas, const uint8_t *pixel_weight, float *v_means2d, float *v_conics,
                      float *v_colors, float *v_opacities, float *v_means2d_abs, float *v_backgrounds) {
    const int64_t n_tiles = (int64_t)C * th * tw;
    const int S = 6 + D + 2; /* per-Gaussian slots: xy(2) conic(3) opacity(1) colour(D) |xy|(2) */
    double *acc = (double *)calloc((size_t)P * S, sizeof(double));
    double *bgacc = (double *)calloc((size_t)C * D + 1, sizeof(double));
    if (!acc || !bgacc) return -1;
#pragma omp parallel
    {
        double *local = NULL;
        size_t local_cap = 0;
        double *S_buf = (double *)malloc(sizeof(double) * D);
        float *vout = (float *)malloc(sizeof(float) * D);
        double *bgl = (double *)malloc(sizeof(double) * D);
#pragma omp for schedule(dynamic, 1)
        for (int64_t t = 0; t < n_tiles; ++t) {
            const int cam = (int)(t / ((int64_t)th * tw));
            const int ty = (int)((t / tw) % th), tx = (int)(t % tw);
            const int lo = offsets[t];
            const int hi = (t + 1 < n_tiles) ? offsets[t + 1] : (int)n_isects;
            /* a masked-out tile shows the background only (alpha 0): it still feeds v_backgrounds */
            const int n = (masks != NULL && !masks[t]) ? 0 : hi - lo;
            for (int k = 0; k < D; ++k) bgl[k] = 0.0;
            if (n > 0) {
                if ((size_t)n * S > local_cap) {
                    free(local);
                    local_cap = (size_t)n * S;
                    local = (double *)malloc(sizeof(double) * local_cap);
                }
                memset(local, 0, sizeof(double) * (size_t)n * S);
            }
            for (int iy = 0; iy < tile; ++iy) {
                const int y = ty * tile + iy;
                if (y >= H) break;
                for (int ix = 0; ix < tile; ++ix) {
                    const int x = tx * tile + ix;
                    if (x >= W) break;
                    const int64_t p = ((int64_t)cam * H + y) * W + x;
                    if (pixel_weight != NULL && !pixel_weight[p]) continue;
                    const float px = (float)x + 0.5f, py = (float)y + 0.5f;
                    const float T_final = 1.0f - render_alphas[p];
                    const float v_a = v_render_alphas[p];
                    float bg_dot = 0.f;
                    for (int k = 0; k < D; ++k) {
                        vout[k] = v_render_colors[p * D + k];
                        S_buf[k] = 0.0;
                        if (backgrounds) {
                            bg_dot += backgrounds[(int64_t)cam * D + k] * vout[k];
                            bgl[k] += (double)vout[k] * (double)T_final; /* v_bg = sum v_col (1 - alpha) */
                        }
                    }
                    if (n == 0) continue;
                    float T = T_final;
                    const int last = last_ids[p];
                    for (int i = (last < hi - 1 ? last : hi - 1); i >= lo; --i) {
                        const int32_t g = flatten_ids[i];
                        const float dx = means2d[2 * (int64_t)g] - px, dy = means2d[2 * (int64_t)g + 1] - py;
                        const float a = conics[3 * (int64_t)g], b = conics[3 * (int64_t)g + 1],
                                    c = conics[3 * (int64_t)g + 2];
                        const float o = opacities[g];
                        const float sigma = 0.5f * (a * dx * dx + c * dy * dy) + b * dx * dy;
                        const float vis = expf(-sigma);
                        const float alpha = fminf(ALPHA_MAX, o * vis);
                        if (sigma < 0.f || alpha < ALPHA_MIN) continue;
                        const float ra = 1.0f / (1.0f - alpha);
                        T *= ra; /* transmittance in front of this Gaussian */
                        const float fac = alpha * T;
                        const float *col = colors + (int64_t)D * g;
                        double *L = local + (size_t)(i - lo) * S;
                        float v_alpha = 0.f;
                        for (int k = 0; k < D; ++k) {
                            L[6 + k] += (double)(fac * vout[k]);
                            v_alpha += (col[k] * T - (float)S_buf[k] * ra) * vout[k];
                        }
                        v_alpha += T_final * ra * v_a;
                        if (backgrounds) v_alpha += -T_final * ra * bg_dot;
                        if (o * vis <= ALPHA_MAX) { /* no gradient through the clamp */
                            const float v_sigma = -o * vis * v_alpha;
                            const float gx = v_sigma * (a * dx + b * dy), gy = v_sigma * (b * dx + c * dy);
                            L[0] += gx;
                            L[1] += gy;
                            L[2] += 0.5f * v_sigma * dx * dx;
                            L[3] += v_sigma * dx * dy;
                            L[4] += 0.5f * v_sigma * dy * dy;
                            L[5] += vis * v_alpha;
                            L[6 + D] += fabsf(gx);
                            L[7 + D] += fabsf(gy);
                        }
                        for (int k = 0; k < D; ++k) S_buf[k] += (double)(col[k] * fac);
                    }
                }
            }
            for (int i = 0; i < n; ++i) {
                const double *L = local + (size_t)i * S;
                double *A = acc + (size_t)flatten_ids[lo + i] * S;
                for (int s = 0; s < S; ++s)
                    if (L[s] != 0.0) atomic_add_d(A + s, L[s]);
            }
            if (backgrounds)
                for (int k = 0; k < D; ++k) atomic_add_d(bgacc + (size_t)cam * D + k, bgl[k]);
        }
        free(local);
        free(S_buf);
        free(vout);
        free(bgl);
    }
#pragma omp parallel for
    for (int64_t g = 0; g < P; ++g) {
        const double *A = acc + (size_t)g * S;
        v_means2d[2 * g] = (float)A[0];
        v_means2d[2 * g + 1] = (float)A[1];
        v_conics[3 * g] = (float)A[2];
        v_conics[3 * g + 1] = (float)A[3];
        v_conics[3 * g + 2] = (float)A[4];
        v_opacities[g] = (float)A[5];
        for (int k = 0; k < D; ++k) v_colors[(int64_t)D * g + k] = (float)A[6 + k];
        if (v_means2d_abs) {
            v_means2d_abs[2 * g] = (float)A[6 + D];
            v_means2d_abs[2 * g + 1] = (float)A[7 + D];
        }
    }
    if (v_backgrounds && backgrounds)
        for (int64_t i = 0; i < (int64_t)C * D; ++i) v_backgrounds[i] = (float)bgacc[i];
    free(acc);
    free(bgacc);
    return 0;
}

int gso_abi_version(void) { return 1; }
