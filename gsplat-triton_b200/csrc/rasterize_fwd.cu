// rasterize_to_pixels forward: front-to-back alpha compositing, one CTA per 16x16 tile.
//
// Replaces (behaviour, not code): gsplat/triton_impl/rasterize_to_pixels_fwd.py:21-196 and
// gsplat/cuda/csrc/RasterizeToPixels3DGSFwd.cu:18-188. Oracle: oracle/gs_oracle.py::rasterize_to_pixels.
//
// Per pixel p = (x+.5, y+.5), over the tile's Gaussians in sorted order:
//   sigma = .5(a dx^2 + c dy^2) + b dx dy ; alpha = min(.999, o e^-sigma)
//   skip if sigma < 0 or alpha < 1/255 ; stop (exclusive) when T(1-alpha) <= 1e-4
//   colour += alpha T col ; T *= 1-alpha ; last_id = sorted index of the last contributor
//
// B200 design: each batch of 256 sorted intersections is gathered once per tile through the packed
// 2 x float4 splat records and *compacted* into shared memory, dropping splats whose
// {alpha >= 1/255} bounding box misses the tile; the staging thread also records which of the 8
// warps (8x4 pixel blocks) the box reaches. Each warp turns those bits into a 32-wide ballot and
// walks only its own splats (find-first-set), so skipped splats cost no issue slots, and stops as
// soon as its 32 pixels are saturated. Issue / FP32-pipe bound; the records live in L2.
#include "rasterize_common.cuh"

namespace raster {

__global__ void __launch_bounds__(256)
splat_pack_kernel(int64_t P, const float *__restrict__ means2d, const float *__restrict__ conics,
                  const float *__restrict__ opacities, float4 *__restrict__ recA, float4 *__restrict__ recB) {
    gsb_chain_enter();
    int64_t g = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (g >= P) return;
    float2 m = *reinterpret_cast<const float2 *>(means2d + 2 * g);
    const float *cp = conics + 3 * g;
    float a = cp[0], b = cp[1], c = cp[2];
    float o = opacities[g];
    float hx, hy;
    splat_extent(a, b, c, o, hx, hy);
    recA[g] = make_float4(m.x, m.y, hx, hy);
    recB[g] = make_float4(a, b, c, o);
}

#ifndef GSB_FWD_MIN_BLOCKS
#define GSB_FWD_MIN_BLOCKS 5
#endif

// COUNT (diagnostics, D = 1): every contribution weighs 1 instead of alpha T, so with unit colours the
// "image" is the number of Gaussians composited into each pixel (bench.py's useful-work count).
template <int D, bool COUNT = false>
__global__ void __launch_bounds__(kThreads, (D <= 4 ? GSB_FWD_MIN_BLOCKS : 1))
rasterize_fwd_kernel(int C, int64_t n_isects, const float4 *__restrict__ recA, const float4 *__restrict__ recB,
                     const float *__restrict__ colors, const float *__restrict__ backgrounds,
                     const uint8_t *__restrict__ masks, int W, int H, int tw, int th,
                     const int32_t *__restrict__ offsets, const int32_t *__restrict__ flatten_ids,
                     float *__restrict__ out_colors, float *__restrict__ out_alphas,
                     int32_t *__restrict__ last_ids, const int64_t *__restrict__ n_isects_dev,
                     uint8_t *__restrict__ block_masks, unsigned long long *__restrict__ warp_pairs) {
    // speculatively sized intersection buffers: n_isects is their capacity, the count lives on the device
    if (n_isects_dev != nullptr) n_isects = min(n_isects, (int64_t)*n_isects_dev);
    // staged batch: three float4 arrays with the same 16-byte stride, so one address computation serves all
    //   P = (x, y, sorted index, warp mask)   B = (conic xx, xy, yy, opacity)   Cq = colours when D <= 4
    constexpr bool kPacked = (D <= 4);
    struct __align__(16) Stage {
        float4 P[kThreads];
        float4 B[kThreads];
        float4 Cq[kPacked ? kThreads : 1];
        float col[kPacked ? 1 : kThreads * D];
        int wcnt[kWarps];
    };
    __shared__ Stage st;
    int *s_wcnt = st.wcnt;

    const TileGeom tg = tile_geometry(n_isects, W, H, tw, th, C, offsets);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    float pix[D];
#pragma unroll
    for (int k = 0; k < D; ++k) pix[k] = 0.f;
    // |T| is the transmittance; T < 0 marks a finished pixel (saturated, or outside the image)
    float T = tg.inside ? 1.0f : -1.0f;
    int cur = 0;
    unsigned n_pairs = 0;  // COUNT only: (warp, splat) pairs this warp walked

    const bool render = (masks == nullptr) || masks[tg.tile_id];
    if (render) {
        const float pxf = (float)tg.px + 0.5f, pyf = (float)tg.py + 0.5f;
        bool warp_done = __all_sync(0xffffffffu, T < 0.f);

        for (int base = tg.range_start; base < tg.range_end; base += kThreads) {
            // all 256 pixels saturated (or outside the image): nothing left to composite
            if (__syncthreads_count(T < 0.f) == kThreads) break;
            const int idx = base + tid;
            unsigned mask = 0u;
            int g = 0;
            float4 A, B;
            if (idx < tg.range_end) {
                g = flatten_ids[idx];
                A = __ldg(recA + g);
                B = __ldg(recB + g);
                mask = splat_block_mask(A, B, tg.tile_x * kTile, tg.tile_y * kTile);
                // the backward pass re-stages the same intersections: hand it the culling result (one byte each)
                if (block_masks != nullptr) block_masks[idx] = (uint8_t)mask;
            }
            int total;
            const int slot = stage_slot(mask != 0u, s_wcnt, total);
            if (mask != 0u) {
                st.P[slot] = make_float4(A.x, A.y, __int_as_float(idx), __uint_as_float(mask));
                st.B[slot] = stage_conic(B);
                const float *cp = colors + (size_t)g * D;
                if (kPacked) {
                    float4 c4 = make_float4(0.f, 0.f, 0.f, 0.f);
                    c4.x = cp[0];
                    if (D > 1) c4.y = cp[1];
                    if (D > 2) c4.z = cp[2];
                    if (D > 3) c4.w = cp[3];
                    st.Cq[slot] = c4;
                } else {
#pragma unroll
                    for (int k = 0; k < D; ++k) st.col[slot * D + k] = cp[k];
                }
            }
            __syncthreads();
            if (warp_done) continue;
            for (int c0 = 0; c0 < total; c0 += 32) {
                const int jl = c0 + lane;
                unsigned todo =
                    __ballot_sync(0xffffffffu, jl < total && ((__float_as_uint(st.P[jl].w) >> warp) & 1u));
                while (todo) {
                    const int j = c0 + __ffs(todo) - 1;
                    todo &= todo - 1u;
                    if (COUNT) ++n_pairs;
                    const float4 p4 = st.P[j];
                    const float4 b4 = st.B[j];
                    const float dx = p4.x - pxf, dy = p4.y - pyf;
                    const float ex = gauss_log2(b4, dx, dy);  // log2(e^-sigma)
                    const float alpha = fminf(kAlphaMax, b4.w * exp2f(ex));
                    const float next_T = fmaf(-T, alpha, T);  // T (1 - alpha); < 0 on finished pixels
                    if (!(ex > 0.f) && !(alpha < kAlphaMin)) {
                        if (next_T <= kTmin) {
                            T = -fabsf(T);  // stop *before* this splat; no-op when already finished
                        } else {
                            const float w = COUNT ? 1.0f : alpha * T;
                            if (kPacked) {
                                const float4 c4 = st.Cq[j];
                                pix[0] = fmaf(c4.x, w, pix[0]);
                                if (D > 1) pix[1] = fmaf(c4.y, w, pix[1]);
                                if (D > 2) pix[2] = fmaf(c4.z, w, pix[2]);
                                if (D > 3) pix[3] = fmaf(c4.w, w, pix[3]);
                            } else {
#pragma unroll
                                for (int k = 0; k < D; ++k) pix[k] = fmaf(st.col[j * D + k], w, pix[k]);
                            }
                            cur = __float_as_int(p4.z);
                            T = next_T;
                        }
                    }
                }
                // all 32 pixels finished? Checked once per 32 staged splats (about every 4th of this warp's pairs), not
                // per pair: a finished pixel ignores further splats (next_T < 0 keeps T), so the few extra pairs are
                // harmless and cost less than three instructions on every pair.
                if (__all_sync(0xffffffffu, T < 0.f)) {
                    warp_done = true;
                    break;
                }
            }
        }
    }
    T = fabsf(T);
    if (COUNT && warp_pairs != nullptr && lane == 0 && n_pairs) atomicAdd(warp_pairs, (unsigned long long)n_pairs);

    if (tg.inside) {
        float *oc = out_colors + tg.pix_id * D;
        const float *bg = backgrounds != nullptr ? backgrounds + (size_t)tg.cam * D : nullptr;
#pragma unroll
        for (int k = 0; k < D; ++k) oc[k] = bg != nullptr ? fmaf(T, bg[k], pix[k]) : pix[k];
        out_alphas[tg.pix_id] = 1.0f - T;
        last_ids[tg.pix_id] = cur;
    }
}

template <int D>
int launch_fwd(int C, int64_t n_isects, const float *recA, const float *recB, const float *colors,
               const float *backgrounds, const uint8_t *masks, int W, int H, int tw, int th,
               const int32_t *offsets, const int32_t *flatten_ids, float *out_colors, float *out_alphas,
               int32_t *last_ids, const int64_t *n_isects_dev, uint8_t *block_masks, cudaStream_t stream) {
    dim3 grid(tw, th, C);
    rasterize_fwd_kernel<D><<<grid, kThreads, 0, stream>>>(
        C, n_isects, reinterpret_cast<const float4 *>(recA), reinterpret_cast<const float4 *>(recB), colors,
        backgrounds, masks, W, H, tw, th, offsets, flatten_ids, out_colors, out_alphas, last_ids, n_isects_dev,
        block_masks, nullptr);
    return gsb_check_launch("rasterize_fwd");
}

}  // namespace raster

// rasterize_fwd2.cu
int gsb_rasterize_fwd2_dispatch(int C, int64_t n_isects, const int64_t *n_isects_dev, int D, const float *recA,
                                const float *recB, const float *colors, const float *backgrounds,
                                const uint8_t *masks, int width, int height, int tile_width, int tile_height,
                                const int32_t *isect_offsets, const int32_t *flatten_ids, float *render_colors,
                                float *render_alphas, int32_t *last_ids, uint8_t *block_masks, cudaStream_t stream);

extern "C" int gsb_rasterize_pack(int64_t P, const float *means2d, const float *conics, const float *opacities,
                                  float *recA, float *recB, cudaStream_t stream) {
    GSB_REQUIRE(P >= 0, "negative size");
    if (P == 0) return GSB_OK;
    gsb_launch_chained(GSB_CHAIN_PACK, raster::splat_pack_kernel, gsb_cdiv(P, 256), 256, 0, stream, P, means2d, conics, opacities,
                       reinterpret_cast<float4 *>(recA), reinterpret_cast<float4 *>(recB));
    return gsb_check_launch("rasterize_pack");
}

extern "C" int gsb_rasterize_fwd(int C, int64_t n_isects, int D, const float *recA, const float *recB,
                                 const float *colors, const float *backgrounds, const uint8_t *masks, int width,
                                 int height, int tile_size, int tile_width, int tile_height,
                                 const int32_t *isect_offsets, const int32_t *flatten_ids, float *render_colors,
                                 float *render_alphas, int32_t *last_ids, cudaStream_t stream) {
    return gsb_rasterize_fwd_cap(C, n_isects, nullptr, D, recA, recB, colors, backgrounds, masks, width, height,
                                 tile_size, tile_width, tile_height, isect_offsets, flatten_ids, render_colors,
                                 render_alphas, last_ids, nullptr, stream);
}

extern "C" int gsb_rasterize_fwd_cap(int C, int64_t n_isects, const int64_t *n_isects_dev, int D, const float *recA,
                                     const float *recB, const float *colors, const float *backgrounds,
                                     const uint8_t *masks, int width, int height, int tile_size, int tile_width,
                                     int tile_height, const int32_t *isect_offsets, const int32_t *flatten_ids,
                                     float *render_colors, float *render_alphas, int32_t *last_ids,
                                     uint8_t *block_masks, cudaStream_t stream) {
    GSB_REQUIRE(C >= 0 && n_isects >= 0 && width > 0 && height > 0, "bad sizes");
    GSB_REQUIRE(n_isects < ((int64_t)1 << 31), "more than 2^31-1 intersections");
    if (tile_size != raster::kTile) {
        gsb_set_error("rasterize_fwd: tile_size %d unsupported (the b200 compositor is built for 16)", tile_size);
        return GSB_ERR_UNSUPPORTED;
    }
    GSB_REQUIRE(tile_width * tile_size >= width && tile_height * tile_size >= height, "tile grid smaller than image");
    GSB_REQUIRE(tile_height <= 65535 && C <= 65535, "grid too large");
    if (C == 0) return GSB_OK;
    if (D <= 4 && gsb_option(GSB_OPT_FWD_VARIANT) >= 2)
        return gsb_rasterize_fwd2_dispatch(C, n_isects, n_isects_dev, D, recA, recB, colors, backgrounds, masks, width,
                                           height, tile_width, tile_height, isect_offsets, flatten_ids,
                                           render_colors, render_alphas, last_ids, block_masks, stream);
#define GSB_FWD_CASE(DD)                                                                                          \
    case DD:                                                                                                      \
        return raster::launch_fwd<DD>(C, n_isects, recA, recB, colors, backgrounds, masks, width, height,         \
                                      tile_width, tile_height, isect_offsets, flatten_ids, render_colors,         \
                                      render_alphas, last_ids, n_isects_dev, block_masks, stream);
    switch (D) {
        GSB_FWD_CASE(1)
        GSB_FWD_CASE(2)
        GSB_FWD_CASE(3)
        GSB_FWD_CASE(4)
        GSB_FWD_CASE(8)
        GSB_FWD_CASE(16)
        GSB_FWD_CASE(32)
        default:
            gsb_set_error("rasterize_fwd: channel count %d unsupported (1,2,3,4,8,16,32)", D);
            return GSB_ERR_UNSUPPORTED;
    }
#undef GSB_FWD_CASE
}

extern "C" int gsb_rasterize_count(int C, int64_t n_isects, const float *recA, const float *recB, const float *ones,
                                   int width, int height, int tile_size, int tile_width, int tile_height,
                                   const int32_t *isect_offsets, const int32_t *flatten_ids, float *counts,
                                   float *render_alphas, int32_t *last_ids, int64_t *warp_pairs,
                                   cudaStream_t stream) {
    GSB_REQUIRE(C >= 0 && n_isects >= 0 && width > 0 && height > 0, "bad sizes");
    GSB_REQUIRE(tile_size == raster::kTile, "tile_size must be 16");
    GSB_REQUIRE(tile_width * tile_size >= width && tile_height * tile_size >= height, "tile grid smaller than image");
    if (C == 0) return GSB_OK;
    dim3 grid(tile_width, tile_height, C);
    raster::rasterize_fwd_kernel<1, true><<<grid, raster::kThreads, 0, stream>>>(
        C, n_isects, reinterpret_cast<const float4 *>(recA), reinterpret_cast<const float4 *>(recB), ones, nullptr,
        nullptr, width, height, tile_width, tile_height, isect_offsets, flatten_ids, counts, render_alphas, last_ids,
        nullptr, nullptr, reinterpret_cast<unsigned long long *>(warp_pairs));
    return gsb_check_launch("rasterize_count");
}
