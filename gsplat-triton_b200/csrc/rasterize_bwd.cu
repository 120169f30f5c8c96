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
//   * same compacted staging, per-warp block masks and ballot-compacted splat lists as the forward kernel;
//   * the 6+D(+2) per-pixel gradient values are reduced over the warp with a *transposing*
//     butterfly: every xor step halves the number of live values per lane, so 9 values cost
//     12 shuffles instead of 45; the sums end up one per lane;
//   * those lanes add into one row of sums per staged splat in shared memory (9 KB at D=3, so 4
//     CTAs stay resident per SM); after the batch one thread per staged splat issues a single set
//     of global reductions per (tile, splat) instead of one per (warp, splat).
#include "rasterize_common.cuh"

namespace raster {

// ---- transposing warp reduction ---------------------------------------------------------------
// After the call, v[i] for i < ceil(N/32) holds the warp-wide sum of the original value
// reduce_origin<N>(lane, i) (or garbage when that returns -1).
template <int N>
__device__ __forceinline__ void warp_transpose_sum(float (&v)[N], int lane) {
    int n = N;
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1) {
        if (n > 1) {
            const int h = (n + 1) >> 1;
            const bool up = (lane & m) != 0;
#pragma unroll
            for (int i = 0; i < (N + 1) / 2; ++i) {
                if (i < h) {
                    const float lo = v[i];
                    const float hi = (i + h < n) ? v[i + h] : 0.f;
                    const float send = up ? lo : hi;
                    const float keep = up ? hi : lo;
                    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, m);
                }
            }
            n = h;
        } else {
            v[0] += __shfl_xor_sync(0xffffffffu, v[0], m);
        }
    }
}

// which original value ends up in v[i] of this lane; -1 = none (padding or a duplicate copy)
template <int N>
__device__ __forceinline__ int reduce_origin(int lane, int i) {
    int cnt[6];
    cnt[0] = N;
#pragma unroll
    for (int k = 0; k < 5; ++k) cnt[k + 1] = cnt[k] > 1 ? (cnt[k] + 1) >> 1 : 1;
    int p = i;
    bool ok = i < cnt[5];
#pragma unroll
    for (int k = 4; k >= 0; --k) {
        const bool up = (lane & (16 >> k)) != 0;
        if (cnt[k] > 1) {
            p += up ? cnt[k + 1] : 0;
            ok = ok && (p < cnt[k]);
        } else {
            ok = ok && !up;  // plain butterfly step: every lane holds a copy, the low lane owns it
        }
    }
    return ok ? p : -1;
}

// Layout of the reduced values: [0,1] v_xy, [2..4] v_conic, [5] v_opacity, [6..6+D) v_color, then [abs_x, abs_y]
template <int D, bool ABS>
struct GradLayout {
    static constexpr int kN = 6 + D + (ABS ? 2 : 0);
    static constexpr int kPerLane = (kN + 31) / 32;
};

struct GradTargets {
    float *v_means2d, *v_conics, *v_opacities, *v_colors, *v_abs;
};

template <int D, bool ABS>
__device__ __forceinline__ float *grad_address(const GradTargets &t, int value, int64_t g) {
    if (value < 2) return t.v_means2d + 2 * g + value;
    if (value < 5) return t.v_conics + 3 * g + (value - 2);
    if (value == 5) return t.v_opacities + g;
    if (value < 6 + D) return t.v_colors + (int64_t)D * g + (value - 6);
    return t.v_abs + 2 * g + (value - 6 - D);
}

// Shared-memory layout (dynamic): the staged batch + one row of NV gradient sums per staged splat.
template <int D, bool ABS>
struct BwdSmem {
    static constexpr int NV = GradLayout<D, ABS>::kN;
    static constexpr size_t bytes = sizeof(float4) * kThreads + sizeof(float2) * kThreads +
                                    sizeof(int) * (3 * kThreads + 2 * kWarps) +
                                    sizeof(float) * kThreads * (D + NV);
};

template <int D, bool ABS>
__global__ void __launch_bounds__(kThreads, (D <= 4 ? 4 : (D <= 16 ? 3 : 2)))
rasterize_bwd_kernel(int C, int64_t n_isects, const float4 *__restrict__ recA, const float4 *__restrict__ recB,
                     const float *__restrict__ colors, const float *__restrict__ backgrounds,
                     const uint8_t *__restrict__ masks, int W, int H, int tw, int th,
                     const int32_t *__restrict__ offsets, const int32_t *__restrict__ flatten_ids,
                     const float *__restrict__ render_alphas, const int32_t *__restrict__ last_ids,
                     const float *__restrict__ v_render_colors, const float *__restrict__ v_render_alphas,
                     GradTargets out) {
    using L = GradLayout<D, ABS>;
    constexpr int NV = L::kN;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4 *s_B = reinterpret_cast<float4 *>(smem_raw);
    float2 *s_xy = reinterpret_cast<float2 *>(s_B + kThreads);
    int *s_idx = reinterpret_cast<int *>(s_xy + kThreads);
    int *s_gid = s_idx + kThreads;
    unsigned *s_mask = reinterpret_cast<unsigned *>(s_gid + kThreads);
    int *s_wcnt = reinterpret_cast<int *>(s_mask + kThreads);  // kWarps
    int *s_red = s_wcnt + kWarps;                              // kWarps
    float *s_col = reinterpret_cast<float *>(s_red + kWarps);  // kThreads * D
    float *s_acc = s_col + kThreads * D;                       // kThreads * NV, zero between batches

    const TileGeom tg = tile_geometry(n_isects, W, H, tw, th, C, offsets);
    if (masks != nullptr && !masks[tg.tile_id]) return;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const float pxf = (float)tg.px + 0.5f, pyf = (float)tg.py + 0.5f;

    // per-pixel state
    const float T_final = tg.inside ? 1.0f - render_alphas[tg.pix_id] : 1.0f;
    float T = T_final;
    const int bin_final = tg.inside ? last_ids[tg.pix_id] : -1;
    float S[D], v_out[D];
    float v_out_a = 0.f, bg_dot = 0.f;
#pragma unroll
    for (int k = 0; k < D; ++k) {
        S[k] = 0.f;
        v_out[k] = tg.inside ? v_render_colors[tg.pix_id * D + k] : 0.f;
    }
    if (tg.inside) v_out_a = v_render_alphas[tg.pix_id];
    if (backgrounds != nullptr) {
#pragma unroll
        for (int k = 0; k < D; ++k) bg_dot = fmaf(backgrounds[(size_t)tg.cam * D + k], v_out[k], bg_dot);
    }
    // d(out)/d(alpha_i) carries T_final/(1-alpha_i) * (v_alpha_out - bg . v_out)
    const float tail = T_final * (v_out_a - bg_dot);

    const int warp_last = warp_max_i(bin_final);
    if (lane == 0) s_red[warp] = warp_last;
    for (int i = tid; i < kThreads * NV; i += kThreads) s_acc[i] = 0.f;
    __syncthreads();
    int tile_last = s_red[0];
#pragma unroll
    for (int w = 1; w < kWarps; ++w) tile_last = max(tile_last, s_red[w]);

    // lane-constant ownership of the reduced values
    int own[L::kPerLane];
#pragma unroll
    for (int i = 0; i < L::kPerLane; ++i) own[i] = reduce_origin<NV>(lane, i);

    const int hi_start = min(tg.range_end - 1, tile_last);
    for (int hi = hi_start; hi >= tg.range_start; hi -= kThreads) {
        const int idx = hi - tid;
        unsigned mask = 0u;
        int g = 0;
        float4 A;
        if (idx >= tg.range_start) {
            g = flatten_ids[idx];
            A = __ldg(recA + g);
            mask = splat_block_mask(A, tg.tile_x * kTile, tg.tile_y * kTile);
        }
        int total;
        const int slot = stage_slot(mask != 0u, s_wcnt, total);  // syncs: previous batch fully flushed
        if (mask != 0u) {
            s_xy[slot] = make_float2(A.x, A.y);
            s_B[slot] = __ldg(recB + g);
            s_idx[slot] = idx;
            s_gid[slot] = g;
            s_mask[slot] = mask;
            const float *cp = colors + (size_t)g * D;
#pragma unroll
            for (int k = 0; k < D; ++k) s_col[slot * D + k] = cp[k];
        }
        __syncthreads();

        for (int c0 = 0; c0 < total; c0 += 32) {
            const int jl = c0 + lane;
            // this warp's splats of the chunk: box reaches its 8x4 block and not behind every pixel's last id
            unsigned todo = __ballot_sync(0xffffffffu, jl < total && ((s_mask[jl] >> warp) & 1u) &&
                                                           s_idx[jl] <= warp_last);
            while (todo) {
                const int j = c0 + __ffs(todo) - 1;
                todo &= todo - 1u;
                const float2 xy = s_xy[j];
                const float4 b4 = s_B[j];
                const float dx = xy.x - pxf, dy = xy.y - pyf;
                const float sigma = 0.5f * (b4.x * dx * dx + b4.z * dy * dy) + b4.y * dx * dy;
                const float vis = __expf(-sigma);
                const float ov = b4.w * vis;
                const float alpha = fminf(kAlphaMax, ov);
                const bool hit = (s_idx[j] <= bin_final) && !(sigma < 0.f) && !(alpha < kAlphaMin);
                if (!__any_sync(0xffffffffu, hit)) continue;

                float v[NV];
#pragma unroll
                for (int i = 0; i < NV; ++i) v[i] = 0.f;
                if (hit) {
                    const float ra = __fdividef(1.0f, 1.0f - alpha);
                    T *= ra;
                    const float fac = alpha * T;
                    float v_alpha = 0.f;
#pragma unroll
                    for (int k = 0; k < D; ++k) {
                        const float c = s_col[j * D + k];
                        v[6 + k] = fac * v_out[k];
                        v_alpha = fmaf(c * T - S[k] * ra, v_out[k], v_alpha);
                        S[k] = fmaf(c, fac, S[k]);
                    }
                    v_alpha = fmaf(tail, ra, v_alpha);
                    if (ov <= kAlphaMax) {
                        const float v_sigma = -ov * v_alpha;
                        v[0] = v_sigma * (b4.x * dx + b4.y * dy);
                        v[1] = v_sigma * (b4.y * dx + b4.z * dy);
                        v[2] = 0.5f * v_sigma * dx * dx;
                        v[3] = v_sigma * dx * dy;
                        v[4] = 0.5f * v_sigma * dy * dy;
                        v[5] = vis * v_alpha;
                        if (ABS) {
                            v[6 + D] = fabsf(v[0]);
                            v[7 + D] = fabsf(v[1]);
                        }
                    }
                }
                warp_transpose_sum<NV>(v, lane);
                float *dst = s_acc + j * NV;
#pragma unroll
                for (int i = 0; i < L::kPerLane; ++i)
                    if (own[i] >= 0) atomicAdd(dst + own[i], v[i]);
            }
        }

        __syncthreads();
        // one flush per (tile, splat): the row owner adds the tile's sums to global memory and re-zeroes the row
        if (tid < total) {
            const int64_t gid = s_gid[tid];
            float *row = s_acc + tid * NV;
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                const float a = row[i];
                if (a != 0.f) {
                    atomicAdd(grad_address<D, ABS>(out, i, gid), a);
                    row[i] = 0.f;
                }
            }
        }
    }
}

template <int D, bool ABS>
int launch_bwd(int C, int64_t n_isects, const float *recA, const float *recB, const float *colors,
               const float *backgrounds, const uint8_t *masks, int W, int H, int tw, int th,
               const int32_t *offsets, const int32_t *flatten_ids, const float *render_alphas,
               const int32_t *last_ids, const float *v_render_colors, const float *v_render_alphas,
               GradTargets out, cudaStream_t stream) {
    constexpr size_t smem = BwdSmem<D, ABS>::bytes;
    auto kernel = rasterize_bwd_kernel<D, ABS>;
    static bool configured = false;  // one instantiation per (D, ABS)
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) {
            gsb_set_error("rasterize_bwd: cannot reserve %zu B of shared memory: %s", smem, cudaGetErrorString(e));
            return GSB_ERR_CUDA;
        }
        configured = true;
    }
    dim3 grid(tw, th, C);
    kernel<<<grid, kThreads, smem, stream>>>(C, n_isects, reinterpret_cast<const float4 *>(recA),
                                             reinterpret_cast<const float4 *>(recB), colors, backgrounds, masks, W,
                                             H, tw, th, offsets, flatten_ids, render_alphas, last_ids,
                                             v_render_colors, v_render_alphas, out);
    return gsb_check_launch("rasterize_bwd");
}

template <int D>
int dispatch_abs(bool abs, int C, int64_t n_isects, const float *recA, const float *recB, const float *colors,
                 const float *backgrounds, const uint8_t *masks, int W, int H, int tw, int th,
                 const int32_t *offsets, const int32_t *flatten_ids, const float *render_alphas,
                 const int32_t *last_ids, const float *v_render_colors, const float *v_render_alphas,
                 GradTargets out, cudaStream_t stream) {
    if (abs)
        return launch_bwd<D, true>(C, n_isects, recA, recB, colors, backgrounds, masks, W, H, tw, th, offsets,
                                   flatten_ids, render_alphas, last_ids, v_render_colors, v_render_alphas, out,
                                   stream);
    return launch_bwd<D, false>(C, n_isects, recA, recB, colors, backgrounds, masks, W, H, tw, th, offsets,
                                flatten_ids, render_alphas, last_ids, v_render_colors, v_render_alphas, out, stream);
}

}  // namespace raster

extern "C" int gsb_rasterize_bwd(int C, int64_t n_isects, int D, const float *recA, const float *recB,
                                 const float *colors, const float *backgrounds, const uint8_t *masks, int width,
                                 int height, int tile_size, int tile_width, int tile_height,
                                 const int32_t *isect_offsets, const int32_t *flatten_ids,
                                 const float *render_alphas, const int32_t *last_ids, const float *v_render_colors,
                                 const float *v_render_alphas, float *v_means2d_abs, float *v_means2d,
                                 float *v_conics, float *v_colors, float *v_opacities, cudaStream_t stream) {
    GSB_REQUIRE(C >= 0 && n_isects >= 0 && width > 0 && height > 0, "bad sizes");
    GSB_REQUIRE(n_isects < ((int64_t)1 << 31), "more than 2^31-1 intersections");
    if (tile_size != raster::kTile) {
        gsb_set_error("rasterize_bwd: tile_size %d unsupported (the b200 compositor is built for 16)", tile_size);
        return GSB_ERR_UNSUPPORTED;
    }
    GSB_REQUIRE(tile_height <= 65535 && C <= 65535, "grid too large");
    if (C == 0 || n_isects == 0) return GSB_OK;
    raster::GradTargets out = {v_means2d, v_conics, v_opacities, v_colors, v_means2d_abs};
    const bool abs = v_means2d_abs != nullptr;
#define GSB_BWD_CASE(DD)                                                                                          \
    case DD:                                                                                                      \
        return raster::dispatch_abs<DD>(abs, C, n_isects, recA, recB, colors, backgrounds, masks, width,     \
                                             height, tile_width, tile_height, isect_offsets, flatten_ids,         \
                                             render_alphas, last_ids, v_render_colors, v_render_alphas, out,      \
                                             stream);
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
