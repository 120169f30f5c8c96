// rasterize_to_pixels backward: back-to-front VJP of the tile compositor.
//
// Replaces (behaviour, not code): gsplat/triton_impl/rasterize_to_pixels_bwd.py:21-337 and
// gsplat/cuda/csrc/RasterizeToPixels3DGSBwd.cu:16-278. Oracle: torch autograd through
// oracle/gs_oracle.py::rasterize_to_pixels.
//
// Per pixel, walking its contributors from last_id back to the tile's first intersection:
//   ra = 1/(1-alpha); T *= ra (T before this splat); fac = alpha T
//   v_col   = fac v_out
//   v_alpha = sum_k (col_k T - S_k ra) v_out_k + T_final ra (v_alpha_out - sum_k bg_k v_out_k)
//   S      += col fac
//   if o e^-sigma <= .999:  v_sigma = -o e^-sigma v_alpha ; v_conic = (.5 v_sigma dx^2, v_sigma dx dy, .5 v_sigma dy^2)
//                           v_xy = v_sigma (a dx + b dy, b dx + c dy) ; v_o = e^-sigma v_alpha
//
// B200 design (what differs from the reference's "warpSum of every value, lane-0 atomics"):
//   * same compacted staging, exact per-warp block culling and ballot-compacted splat lists as the
//     forward kernel;
//   * two-phase reduction without a cross-lane reduction tree. Phase 1 (lane = pixel, sequential over
//     the warp's splats because of the transmittance recurrence) only produces two scalars per
//     (pixel, splat): w1 = -d loss / d sigma and w2 = alpha T, parked in a per-warp shared table.
//     Phase 2 (every 16 splats; lane = splat x half block) walks the 32 pixels of the block and
//     accumulates in registers the moments of w1 about the block centre (sum w1 {1, x, y, x^2, xy, y^2})
//     and sum w2 v_out; the conic / mean / opacity gradients are closed forms of those moments.
//     ~11 issue slots per (warp, splat) instead of ~70 for a shuffle tree over 9 values;
//   * one set of global reductions per (warp, splat), issued by the lane that owns the splat.
#include "rasterize_common.cuh"

#include <atomic>

namespace raster {

#ifndef GSB_BWD_MIN_BLOCKS
#define GSB_BWD_MIN_BLOCKS 4  // resident CTAs per SM the D <= 4 instantiations are register-limited to
#endif
constexpr int kChunk = 16;       // splats per splat-parallel reduction pass
constexpr int kChunkStride = 33; // float2 row stride of the per-warp weight table (bank-conflict free)

template <int D>
struct BwdSmem {
    static constexpr int DP = (D + 3) & ~3;  // per-pixel cotangent row, padded for 16-byte loads
    static constexpr int CS = D <= 4 ? 4 : D;  // colour row of a staged splat (one float4 when D <= 4)
    static constexpr size_t bytes = sizeof(float4) * 2 * kThreads                       // s_P, s_B
                                    + sizeof(float) * kThreads * CS                     // s_col
                                    + sizeof(float2) * kWarps * kChunk * kChunkStride   // s_w
                                    + sizeof(float) * kWarps * (32 * DP + 4)            // s_vout
                                    + sizeof(int) * (2 * kWarps);                       // counters
};

// Second half of the backward pass for up to kChunk splats of ONE warp (8x4 pixel block):
// lane (s = lane & 15, h = lane >> 4) owns splat s and sums over the 16 pixels of block rows 2h, 2h+1.
// Inputs per (splat, pixel): w1 = -d loss / d sigma (sign folded into the epilogue), w2 = alpha * T (weight of the colour cotangent).
// Geometry gradients come from the moments of w1 over the block, taken about the block centre:
//   sum w1 dx = ux M0 - Mx, sum w1 dx^2 = ux^2 M0 - 2 ux Mx + Mxx, ...   (dx = ux - cx, ux = mu_x - centre_x)
template <int D, bool ABS>
__device__ __forceinline__ void reduce_chunk(int count, int lane, const float2 *s_w, const float *s_vout,
                                             int chunk_j, const float4 *s_P, const float4 *s_B,
                                             const int32_t *__restrict__ flatten_ids,
                                             float bcx, float bcy, const GradTargets &out) {
    constexpr int DP = BwdSmem<D>::DP;
    __syncwarp();
    const int s = lane & (kChunk - 1), h = lane >> 4;
    const float2 *wrow = s_w + s * kChunkStride + h * 16;
    const float *vrow = s_vout + h * (16 * DP + 4);
    float R0[2] = {0.f, 0.f}, Rx[2] = {0.f, 0.f}, Rxx[2] = {0.f, 0.f};
    float col[D];
#pragma unroll
    for (int k = 0; k < D; ++k) col[k] = 0.f;
    float abs_x = 0.f, abs_y = 0.f;
    // lane n holds the staged slot of the chunk's n-th splat (slots past `count` repeat the first: masked below)
    const int j = __shfl_sync(0xffffffffu, chunk_j, s < count ? s : 0);
    const float4 mu = s_P[j];  // (x, y, sorted index, -)
    float4 b4 = s_B[j];        // staged (pre-scaled) conic -> conic
    b4.x *= kUnscaleDiag;
    b4.y *= kUnscaleOff;
    b4.z *= kUnscaleDiag;
    const float ux = mu.x - bcx, uy = mu.y - bcy;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const float2 w = wrow[i];
        const float cx = (float)(i & 7) - 3.5f;
        const int r = i >> 3;
        R0[r] += w.x;
        Rx[r] = fmaf(w.x, cx, Rx[r]);
        Rxx[r] = fmaf(w.x, cx * cx, Rxx[r]);
#pragma unroll
        for (int k4 = 0; k4 < DP / 4; ++k4) {
            const float4 v = *reinterpret_cast<const float4 *>(vrow + i * DP + 4 * k4);
            if (4 * k4 + 0 < D) col[4 * k4 + 0] = fmaf(w.y, v.x, col[4 * k4 + 0]);
            if (4 * k4 + 1 < D) col[4 * k4 + 1] = fmaf(w.y, v.y, col[4 * k4 + 1]);
            if (4 * k4 + 2 < D) col[4 * k4 + 2] = fmaf(w.y, v.z, col[4 * k4 + 2]);
            if (4 * k4 + 3 < D) col[4 * k4 + 3] = fmaf(w.y, v.w, col[4 * k4 + 3]);
        }
        if (ABS) {
            const float dx = ux - cx, dy = uy - ((float)(2 * h + r) - 1.5f);
            abs_x += fabsf(w.x * (b4.x * dx + b4.y * dy));
            abs_y += fabsf(w.x * (b4.y * dx + b4.z * dy));
        }
    }
    // rows 2h, 2h+1 have centred y = 2h - 1.5, 2h - 0.5
    const float y0 = (float)(2 * h) - 1.5f, y1 = y0 + 1.0f;
    float M0 = R0[0] + R0[1], Mx = Rx[0] + Rx[1], Mxx = Rxx[0] + Rxx[1];
    float My = y0 * R0[0] + y1 * R0[1], Mxy = y0 * Rx[0] + y1 * Rx[1], Myy = y0 * y0 * R0[0] + y1 * y1 * R0[1];
    // other half of the block
    M0 += __shfl_xor_sync(0xffffffffu, M0, 16);
    Mx += __shfl_xor_sync(0xffffffffu, Mx, 16);
    My += __shfl_xor_sync(0xffffffffu, My, 16);
    Mxx += __shfl_xor_sync(0xffffffffu, Mxx, 16);
    Mxy += __shfl_xor_sync(0xffffffffu, Mxy, 16);
    Myy += __shfl_xor_sync(0xffffffffu, Myy, 16);
#pragma unroll
    for (int k = 0; k < D; ++k) col[k] += __shfl_xor_sync(0xffffffffu, col[k], 16);
    if (ABS) {
        abs_x += __shfl_xor_sync(0xffffffffu, abs_x, 16);
        abs_y += __shfl_xor_sync(0xffffffffu, abs_y, 16);
    }
    if (h == 0 && s < count) {
        const int64_t g = flatten_ids[__float_as_int(mu.z)];
        const float Sx = ux * M0 - Mx, Sy = uy * M0 - My;
        const float Sxx = fmaf(ux, fmaf(ux, M0, -2.f * Mx), Mxx);
        const float Syy = fmaf(uy, fmaf(uy, M0, -2.f * My), Myy);
        const float Sxy = fmaf(ux, fmaf(uy, M0, -My), fmaf(-uy, Mx, Mxy));
        float v[6];
        // the moments are of -w1 (phase 1 stores -d loss / d sigma): the signs are folded in here
        v[0] = -(b4.x * Sx + b4.y * Sy);
        v[1] = -(b4.y * Sx + b4.z * Sy);
        v[2] = -0.5f * Sxx;
        v[3] = -Sxy;
        v[4] = -0.5f * Syy;
        v[5] = __fdividef(M0, b4.w);  // v_opacity = sum e^-sigma v_alpha = -sum w1 / opacity
        if (v[0] != 0.f) atomicAdd(out.v_means2d + 2 * g, v[0]);
        if (v[1] != 0.f) atomicAdd(out.v_means2d + 2 * g + 1, v[1]);
        if (v[2] != 0.f) atomicAdd(out.v_conics + 3 * g, v[2]);
        if (v[3] != 0.f) atomicAdd(out.v_conics + 3 * g + 1, v[3]);
        if (v[4] != 0.f) atomicAdd(out.v_conics + 3 * g + 2, v[4]);
        if (v[5] != 0.f) atomicAdd(out.v_opacities + g, v[5]);
#pragma unroll
        for (int k = 0; k < D; ++k)
            if (col[k] != 0.f) atomicAdd(out.v_colors + (int64_t)D * g + k, col[k]);
        if (ABS) {
            if (abs_x != 0.f) atomicAdd(out.v_abs + 2 * g, abs_x);
            if (abs_y != 0.f) atomicAdd(out.v_abs + 2 * g + 1, abs_y);
        }
    }
    __syncwarp();
}

template <int D, bool ABS>
__global__ void __launch_bounds__(kThreads, (D <= 4 ? GSB_BWD_MIN_BLOCKS : (D <= 16 ? 2 : 1)))
rasterize_bwd_kernel(int C, int64_t n_isects, const float4 *__restrict__ recA, const float4 *__restrict__ recB,
                     const float *__restrict__ colors, const float *__restrict__ backgrounds,
                     const uint8_t *__restrict__ masks, int W, int H, int tw, int th,
                     const int32_t *__restrict__ offsets, const int32_t *__restrict__ flatten_ids,
                     const float *__restrict__ render_alphas, const int32_t *__restrict__ last_ids,
                     const uint8_t *__restrict__ block_masks, const float *__restrict__ v_render_colors,
                     const float *__restrict__ v_render_alphas, GradTargets out) {
    constexpr int DP = BwdSmem<D>::DP;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int CS = BwdSmem<D>::CS;
    // staged batch: P = (x, y, sorted index, warp mask), B = (conic xx, xy, yy, opacity), colours; same stride
    float4 *s_P = reinterpret_cast<float4 *>(smem_raw);
    float4 *s_B = s_P + kThreads;
    float *s_col = reinterpret_cast<float *>(s_B + kThreads);
    float2 *s_w_all = reinterpret_cast<float2 *>(s_col + kThreads * CS);
    float *s_vout_all = reinterpret_cast<float *>(s_w_all + kWarps * kChunk * kChunkStride);
    int *s_wcnt = reinterpret_cast<int *>(s_vout_all + kWarps * (32 * DP + 4));
    int *s_red = s_wcnt + kWarps;

    const TileGeom tg = tile_geometry(n_isects, W, H, tw, th, C, offsets);
    if (masks != nullptr && !masks[tg.tile_id]) return;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const float pxf = (float)tg.px + 0.5f, pyf = (float)tg.py + 0.5f;
    float2 *s_w = s_w_all + warp * kChunk * kChunkStride;
    float *s_vout = s_vout_all + warp * (32 * DP + 4);
    // centre of this warp's 8x4 block, in pixel-centre coordinates
    const float bcx = (float)(tg.tile_x * kTile + (warp & 1) * 8) + 4.0f;
    const float bcy = (float)(tg.tile_y * kTile + (warp >> 1) * 4) + 2.0f;

    // per-pixel state
    const float T_final = tg.inside ? 1.0f - render_alphas[tg.pix_id] : 1.0f;
    float T = T_final;
    const int bin_final = tg.inside ? last_ids[tg.pix_id] : -1;
    float v_out[D];
    float v_out_a = 0.f, bg_dot = 0.f;
#pragma unroll
    for (int k = 0; k < D; ++k) v_out[k] = tg.inside ? v_render_colors[tg.pix_id * D + k] : 0.f;
    if (tg.inside) v_out_a = v_render_alphas[tg.pix_id];
    if (backgrounds != nullptr) {
#pragma unroll
        for (int k = 0; k < D; ++k) bg_dot = fmaf(backgrounds[(size_t)tg.cam * D + k], v_out[k], bg_dot);
    }
    // d(out)/d(alpha_i) carries T_final/(1-alpha_i) * (v_alpha_out - bg . v_out)
    // Everything behind splat i reaches v_alpha_i only through the scalar
    //   u_i = T_final (v_alpha_out - bg . v_out) - sum_{j behind i} w2_j (c_j . v_out):
    // v_alpha_i = (c_i . v_out) T_i + u_i / (1 - alpha_i). One recurrence instead of one per channel.
    float u = T_final * (v_out_a - bg_dot);
    {   // colour cotangents of the block's pixels, for the splat-parallel pass
        float *dst = s_vout + lane * DP + (lane >> 4) * 4;
#pragma unroll
        for (int k = 0; k < DP; ++k) dst[k] = k < D ? v_out[k] : 0.f;
    }

    const int warp_last = warp_max_i(bin_final);
    if (lane == 0) s_red[warp] = warp_last;
    __syncthreads();
    int tile_last = s_red[0];
#pragma unroll
    for (int w = 1; w < kWarps; ++w) tile_last = max(tile_last, s_red[w]);

    const int hi_start = min(tg.range_end - 1, tile_last);
    for (int hi = hi_start; hi >= tg.range_start; hi -= kThreads) {
        const int idx = hi - tid;
        unsigned mask = 0u;
        int g = 0;
        float4 A, B;
        if (idx >= tg.range_start) {
            if (block_masks != nullptr) {
                // culling result of the forward pass (every intersection up to the tile's last contributor was
                // staged there): culled intersections cost one byte, no record gathers and no arithmetic
                mask = block_masks[idx];
                if (mask != 0u) {
                    g = flatten_ids[idx];
                    A = __ldg(recA + g);
                    B = __ldg(recB + g);
                }
            } else {
                g = flatten_ids[idx];
                A = __ldg(recA + g);
                B = __ldg(recB + g);
                mask = splat_block_mask(A, B, tg.tile_x * kTile, tg.tile_y * kTile);
            }
        }
        int total;
        const int slot = stage_slot(mask != 0u, s_wcnt, total);  // syncs: previous batch fully consumed
        if (mask != 0u) {
            s_P[slot] = make_float4(A.x, A.y, __int_as_float(idx), __uint_as_float(mask));
            s_B[slot] = stage_conic(B);
            const float *cp = colors + (size_t)g * D;
#pragma unroll
            for (int k = 0; k < CS; ++k) s_col[slot * CS + k] = k < D ? cp[k] : 0.f;
        }
        __syncthreads();

        int n_chunk = 0, chunk_j = 0;
        float2 *w_slot = s_w + lane;  // this lane's entry of the chunk row being filled
        for (int c0 = 0; c0 < total; c0 += 32) {
            const int jl = c0 + lane;
            // this warp's splats of the chunk: may touch its 8x4 block and not behind every pixel's last id
            bool mine = false;
            if (jl < total) {
                const float4 p4 = s_P[jl];
                mine = ((__float_as_uint(p4.w) >> warp) & 1u) && __float_as_int(p4.z) <= warp_last;
            }
            unsigned todo = __ballot_sync(0xffffffffu, mine);
            while (todo) {
                const int j = c0 + __ffs(todo) - 1;
                todo &= todo - 1u;
                const float4 p4 = s_P[j];
                const float4 b4 = s_B[j];
                const float dx = p4.x - pxf, dy = p4.y - pyf;
                const float ex = gauss_log2(b4, dx, dy);  // log2(e^-sigma), same expression as the forward pass
                const float ov = b4.w * exp2f(ex);
                const float alpha = fminf(kAlphaMax, ov);
                const bool hit = (__float_as_int(p4.z) <= bin_final) && !(ex > 0.f) && !(alpha < kAlphaMin);
                if (!__any_sync(0xffffffffu, hit)) continue;

                // branch-free over the lanes: a lane without a hit runs the same arithmetic with alpha = 0
                // (1 / (1 - 0) and T * 1 are exact, w2 = 0 leaves u alone) and a zero weight on v_alpha
                const float alpha_h = hit ? alpha : 0.f;
                const float ov_h = ov <= kAlphaMax ? alpha_h : 0.f;  // = ov on a hit; no gradient through the clamp
                const float ra = __fdividef(1.0f, 1.0f - alpha_h);
                T *= ra;
                const float w2 = alpha_h * T;
                float cv = 0.f;
#pragma unroll
                for (int k = 0; k < D; ++k) cv = fmaf(s_col[j * CS + k], v_out[k], cv);
                const float w1 = ov_h * fmaf(cv, T, u * ra);  // = -d loss / d sigma: reduce_chunk flips the sign once
                u = fmaf(-cv, w2, u);
                *w_slot = make_float2(w1, w2);
                w_slot += kChunkStride;
                chunk_j = lane == n_chunk ? j : chunk_j;  // the chunk's splat slots live in lanes 0..15, not in memory
                if (++n_chunk == kChunk) {
                    reduce_chunk<D, ABS>(n_chunk, lane, s_w, s_vout, chunk_j, s_P, s_B, flatten_ids, bcx, bcy, out);
                    n_chunk = 0;
                    w_slot = s_w + lane;
                }
            }
        }
        if (n_chunk > 0)
            reduce_chunk<D, ABS>(n_chunk, lane, s_w, s_vout, chunk_j, s_P, s_B, flatten_ids, bcx, bcy, out);
    }
}

template <int D, bool ABS>
int launch_bwd(int C, int64_t n_isects, const float *recA, const float *recB, const float *colors,
               const float *backgrounds, const uint8_t *masks, int W, int H, int tw, int th,
               const int32_t *offsets, const int32_t *flatten_ids, const float *render_alphas,
               const int32_t *last_ids, const uint8_t *block_masks, const float *v_render_colors,
               const float *v_render_alphas, GradTargets out, cudaStream_t stream) {
    constexpr size_t smem = BwdSmem<D>::bytes;
    auto kernel = rasterize_bwd_kernel<D, ABS>;
    // one instantiation per (D, ABS); the attribute is per DEVICE: bit d = configured on device d
    static std::atomic<unsigned long long> configured{0ull};
    int device = 0;
    cudaGetDevice(&device);
    if (!((configured.load(std::memory_order_acquire) >> (device & 63)) & 1ull)) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) {
            gsb_set_error("rasterize_bwd: cannot reserve %zu B of shared memory: %s", smem, cudaGetErrorString(e));
            return GSB_ERR_CUDA;
        }
        configured.fetch_or(1ull << (device & 63), std::memory_order_release);
    }
    dim3 grid(tw, th, C);
    kernel<<<grid, kThreads, smem, stream>>>(C, n_isects, reinterpret_cast<const float4 *>(recA),
                                             reinterpret_cast<const float4 *>(recB), colors, backgrounds, masks, W,
                                             H, tw, th, offsets, flatten_ids, render_alphas, last_ids, block_masks,
                                             v_render_colors, v_render_alphas, out);
    return gsb_check_launch("rasterize_bwd");
}

template <int D>
int dispatch_abs(bool abs, int C, int64_t n_isects, const float *recA, const float *recB, const float *colors,
                 const float *backgrounds, const uint8_t *masks, int W, int H, int tw, int th,
                 const int32_t *offsets, const int32_t *flatten_ids, const float *render_alphas,
                 const int32_t *last_ids, const uint8_t *block_masks, const float *v_render_colors,
                 const float *v_render_alphas, GradTargets out, cudaStream_t stream) {
    if (abs)
        return launch_bwd<D, true>(C, n_isects, recA, recB, colors, backgrounds, masks, W, H, tw, th, offsets,
                                   flatten_ids, render_alphas, last_ids, block_masks, v_render_colors,
                                   v_render_alphas, out, stream);
    return launch_bwd<D, false>(C, n_isects, recA, recB, colors, backgrounds, masks, W, H, tw, th, offsets,
                                flatten_ids, render_alphas, last_ids, block_masks, v_render_colors, v_render_alphas,
                                out, stream);
}

}  // namespace raster

// rasterize_bwd2.cu
int gsb_rasterize_bwd2_dispatch(int C, int64_t n_isects, int D, const float *recA, const float *recB,
                                const float *colors, const float *backgrounds, const uint8_t *masks, int width,
                                int height, int tile_width, int tile_height, const int32_t *isect_offsets,
                                const int32_t *flatten_ids, const float *render_alphas, const int32_t *last_ids,
                                const uint8_t *block_masks, const float *v_render_colors,
                                const float *v_render_alphas, float *v_means2d_abs, float *v_means2d, float *v_conics,
                                float *v_colors, float *v_opacities, cudaStream_t stream);

extern "C" int gsb_rasterize_bwd(int C, int64_t n_isects, int D, const float *recA, const float *recB,
                                 const float *colors, const float *backgrounds, const uint8_t *masks, int width,
                                 int height, int tile_size, int tile_width, int tile_height,
                                 const int32_t *isect_offsets, const int32_t *flatten_ids,
                                 const float *render_alphas, const int32_t *last_ids, const uint8_t *block_masks,
                                 const float *v_render_colors, const float *v_render_alphas, float *v_means2d_abs, float *v_means2d,
                                 float *v_conics, float *v_colors, float *v_opacities, cudaStream_t stream) {
    GSB_REQUIRE(C >= 0 && n_isects >= 0 && width > 0 && height > 0, "bad sizes");
    GSB_REQUIRE(n_isects < ((int64_t)1 << 31), "more than 2^31-1 intersections");
    if (tile_size != raster::kTile) {
        gsb_set_error("rasterize_bwd: tile_size %d unsupported (the b200 compositor is built for 16)", tile_size);
        return GSB_ERR_UNSUPPORTED;
    }
    GSB_REQUIRE(tile_height <= 65535 && C <= 65535, "grid too large");
    if (C == 0 || n_isects == 0) return GSB_OK;
    if (D <= 4 && gsb_option(GSB_OPT_BWD_VARIANT) == 2)
        return gsb_rasterize_bwd2_dispatch(C, n_isects, D, recA, recB, colors, backgrounds, masks, width, height,
                                           tile_width, tile_height, isect_offsets, flatten_ids, render_alphas,
                                           last_ids, block_masks, v_render_colors, v_render_alphas, v_means2d_abs,
                                           v_means2d, v_conics, v_colors, v_opacities, stream);
    raster::GradTargets out = {v_means2d, v_conics, v_opacities, v_colors, v_means2d_abs};
    const bool abs = v_means2d_abs != nullptr;
#define GSB_BWD_CASE(DD)                                                                                          \
    case DD:                                                                                                      \
        return raster::dispatch_abs<DD>(abs, C, n_isects, recA, recB, colors, backgrounds, masks, width,     \
                                             height, tile_width, tile_height, isect_offsets, flatten_ids,         \
                                             render_alphas, last_ids, block_masks, v_render_colors,               \
                                             v_render_alphas, out, stream);
    switch (D) {
        GSB_BWD_CASE(1)
        GSB_BWD_CASE(2)
        GSB_BWD_CASE(3)
        GSB_BWD_CASE(4)
        GSB_BWD_CASE(8)
        GSB_BWD_CASE(16)
        GSB_BWD_CASE(32)
        default:
            gsb_set_error("rasterize_bwd: channel count %d unsupported (1,2,3,4,8,16,32)", D);
            return GSB_ERR_UNSUPPORTED;
    }
#undef GSB_BWD_CASE
}
