// rasterize_to_pixels for tile sizes other than 16 (any tile_size with tile_size^2 <= 1024): the plain algorithm.
//
// `rasterization(tile_size=...)` is a public argument of the reference (gsplat/rendering.py:57), honoured by both of
// its backends (RasterizeToPixels3DGSFwd.cu:31,221; triton_impl/rasterize_to_pixels_fwd.py:37,61-62) and used by its
// own tests with tile_size = 4 (triton_tests/test_isect.py:72, test_ras2pix.py:88). The tuned b200 compositor
// (rasterize_fwd2.cu / rasterize_bwd.cu) is built around 16x16 tiles; this file keeps every other tile size
// correct with the textbook structure: one CTA per tile, one thread per pixel, batches of the tile's sorted
// intersections staged in shared memory, front-to-back forward, back-to-front backward with a warp-shuffle
// reduction before the global atomics. Same per-pair arithmetic as the tuned kernels (gauss_log2 / exp2).
#include "rasterize_common.cuh"

namespace raster {

struct GenericStage {
    float4 *P;  // x, y, sorted index, -
    float4 *B;  // staged conic, opacity
    int *G;     // flattened Gaussian id
};

__device__ __forceinline__ GenericStage carve(unsigned char *smem, int n) {
    GenericStage s;
    s.P = reinterpret_cast<float4 *>(smem);
    s.B = s.P + n;
    s.G = reinterpret_cast<int *>(s.B + n);
    return s;
}

template <int D>
__global__ void rasterize_fwd_generic_kernel(int C, int64_t n_isects, const float *__restrict__ means2d,
                                             const float *__restrict__ conics, const float *__restrict__ opacities,
                                             const float *__restrict__ colors, const float *__restrict__ backgrounds,
                                             const uint8_t *__restrict__ masks, int W, int H, int ts, int tw, int th,
                                             const int32_t *__restrict__ offsets,
                                             const int32_t *__restrict__ flatten_ids, float *__restrict__ out_colors,
                                             float *__restrict__ out_alphas, int32_t *__restrict__ last_ids) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int nthreads = blockDim.x, tid = threadIdx.x;
    GenericStage st = carve(smem_raw, nthreads);
    const int cam = blockIdx.z, tile_y = blockIdx.y, tile_x = blockIdx.x;
    const int64_t tile_id = ((int64_t)cam * th + tile_y) * tw + tile_x;
    const int range_start = offsets[tile_id];
    const int range_end = (tile_id + 1 < (int64_t)C * th * tw) ? offsets[tile_id + 1] : (int)n_isects;
    const bool owner = tid < ts * ts;
    const int px = tile_x * ts + (owner ? tid % ts : 0), py = tile_y * ts + (owner ? tid / ts : 0);
    const bool inside = owner && px < W && py < H;
    const float pxf = (float)px + 0.5f, pyf = (float)py + 0.5f;
    float pix[D];
#pragma unroll
    for (int k = 0; k < D; ++k) pix[k] = 0.f;
    float T = 1.0f;
    int cur = 0;
    bool done = !inside;
    const bool render = (masks == nullptr) || masks[tile_id];
    for (int base = range_start; render && base < range_end; base += nthreads) {
        if (__syncthreads_count(done) == nthreads) break;
        const int idx = base + tid;
        if (idx < range_end) {
            const int g = flatten_ids[idx];
            st.G[tid] = g;
            st.P[tid] = make_float4(means2d[2 * (int64_t)g], means2d[2 * (int64_t)g + 1], __int_as_float(idx), 0.f);
            st.B[tid] = stage_conic(make_float4(conics[3 * (int64_t)g], conics[3 * (int64_t)g + 1],
                                                conics[3 * (int64_t)g + 2], opacities[g]));
        }
        __syncthreads();
        const int n = min(nthreads, range_end - base);
        for (int j = 0; j < n && !done; ++j) {
            const float4 p4 = st.P[j], b4 = st.B[j];
            const float ex = gauss_log2(b4, p4.x - pxf, p4.y - pyf);
            const float alpha = fminf(kAlphaMax, b4.w * exp2f(ex));
            if (ex > 0.f || alpha < kAlphaMin) continue;
            const float next_T = T * (1.0f - alpha);
            if (next_T <= kTmin) {
                done = true;
                break;
            }
            const float w = alpha * T;
            const float *cp = colors + (int64_t)st.G[j] * D;
#pragma unroll
            for (int k = 0; k < D; ++k) pix[k] = fmaf(cp[k], w, pix[k]);
            cur = base + j;
            T = next_T;
        }
    }
    if (inside) {
        const int64_t p = ((int64_t)cam * H + py) * W + px;
#pragma unroll
        for (int k = 0; k < D; ++k)
            out_colors[p * D + k] = backgrounds ? fmaf(T, backgrounds[(int64_t)cam * D + k], pix[k]) : pix[k];
        out_alphas[p] = 1.0f - T;
        last_ids[p] = cur;
    }
}

template <int D>
__global__ void rasterize_bwd_generic_kernel(int C, int64_t n_isects, const float *__restrict__ means2d,
                                             const float *__restrict__ conics, const float *__restrict__ opacities,
                                             const float *__restrict__ colors, const float *__restrict__ backgrounds,
                                             const uint8_t *__restrict__ masks, int W, int H, int ts, int tw, int th,
                                             const int32_t *__restrict__ offsets,
                                             const int32_t *__restrict__ flatten_ids,
                                             const float *__restrict__ render_alphas,
                                             const int32_t *__restrict__ last_ids,
                                             const float *__restrict__ v_render_colors,
                                             const float *__restrict__ v_render_alphas, float *__restrict__ v_abs,
                                             float *__restrict__ v_means2d, float *__restrict__ v_conics,
                                             float *__restrict__ v_colors, float *__restrict__ v_opacities) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int nthreads = blockDim.x, tid = threadIdx.x, lane = tid & 31;
    GenericStage st = carve(smem_raw, nthreads);
    const int cam = blockIdx.z, tile_y = blockIdx.y, tile_x = blockIdx.x;
    const int64_t tile_id = ((int64_t)cam * th + tile_y) * tw + tile_x;
    if (masks != nullptr && !masks[tile_id]) return;
    const int range_start = offsets[tile_id];
    const int range_end = (tile_id + 1 < (int64_t)C * th * tw) ? offsets[tile_id + 1] : (int)n_isects;
    const bool owner = tid < ts * ts;
    const int px = tile_x * ts + (owner ? tid % ts : 0), py = tile_y * ts + (owner ? tid / ts : 0);
    const bool inside = owner && px < W && py < H;
    const float pxf = (float)px + 0.5f, pyf = (float)py + 0.5f;
    const int64_t p = ((int64_t)cam * H + min(py, H - 1)) * W + min(px, W - 1);
    const float T_final = inside ? 1.0f - render_alphas[p] : 1.0f;
    const int bin_final = inside ? last_ids[p] : -1;
    float T = T_final;
    float v_out[D], S[D];
    float bg_dot = 0.f;
#pragma unroll
    for (int k = 0; k < D; ++k) {
        v_out[k] = inside ? v_render_colors[p * D + k] : 0.f;
        S[k] = 0.f;
        if (backgrounds) bg_dot = fmaf(backgrounds[(int64_t)cam * D + k], v_out[k], bg_dot);
    }
    const float v_out_a = inside ? v_render_alphas[p] : 0.f;
    // the tile's deepest contributor: nothing behind it matters to any pixel
    __shared__ int s_last;
    if (tid == 0) s_last = -1;
    __syncthreads();
    atomicMax(&s_last, bin_final);
    __syncthreads();
    const int hi_start = min(range_end - 1, s_last);
    for (int hi = hi_start; hi >= range_start; hi -= nthreads) {
        __syncthreads();  // previous batch consumed
        const int idx = hi - tid;
        if (idx >= range_start) {
            const int g = flatten_ids[idx];
            st.G[tid] = g;
            st.P[tid] = make_float4(means2d[2 * (int64_t)g], means2d[2 * (int64_t)g + 1], __int_as_float(idx), 0.f);
            st.B[tid] = stage_conic(make_float4(conics[3 * (int64_t)g], conics[3 * (int64_t)g + 1],
                                                conics[3 * (int64_t)g + 2], opacities[g]));
        }
        __syncthreads();
        const int n = min(nthreads, hi - range_start + 1);
        for (int j = 0; j < n; ++j) {
            const float4 p4 = st.P[j], b4 = st.B[j];
            const float dx = p4.x - pxf, dy = p4.y - pyf;
            const float ex = gauss_log2(b4, dx, dy);
            const float ov = b4.w * exp2f(ex);
            const float alpha = fminf(kAlphaMax, ov);
            const bool hit = inside && (hi - j) <= bin_final && !(ex > 0.f) && !(alpha < kAlphaMin);
            if (!__any_sync(0xffffffffu, hit)) continue;
            float g_xy0 = 0.f, g_xy1 = 0.f, g_c0 = 0.f, g_c1 = 0.f, g_c2 = 0.f, g_o = 0.f, a_x = 0.f, a_y = 0.f;
            float g_col[D];
#pragma unroll
            for (int k = 0; k < D; ++k) g_col[k] = 0.f;
            const int g = st.G[j];
            if (hit) {
                const float *cp = colors + (int64_t)g * D;
                const float ra = 1.0f / (1.0f - alpha);
                T *= ra;
                const float fac = alpha * T;
                float v_alpha = 0.f;
#pragma unroll
                for (int k = 0; k < D; ++k) {
                    const float c = cp[k];
                    g_col[k] = fac * v_out[k];
                    v_alpha += (c * T - S[k] * ra) * v_out[k];
                    S[k] = fmaf(c, fac, S[k]);
                }
                v_alpha += T_final * ra * (v_out_a - bg_dot);
                if (ov <= kAlphaMax) {  // no gradient through the clamp
                    const float a = b4.x * kUnscaleDiag, b = b4.y * kUnscaleOff, c = b4.z * kUnscaleDiag;
                    const float v_sigma = -ov * v_alpha;
                    g_xy0 = v_sigma * (a * dx + b * dy);
                    g_xy1 = v_sigma * (b * dx + c * dy);
                    g_c0 = 0.5f * v_sigma * dx * dx;
                    g_c1 = v_sigma * dx * dy;
                    g_c2 = 0.5f * v_sigma * dy * dy;
                    g_o = __fdividef(ov, b4.w) * v_alpha;  // e^-sigma v_alpha
                    a_x = fabsf(g_xy0);
                    a_y = fabsf(g_xy1);
                }
            }
            g_xy0 = warp_sum(g_xy0);
            g_xy1 = warp_sum(g_xy1);
            g_c0 = warp_sum(g_c0);
            g_c1 = warp_sum(g_c1);
            g_c2 = warp_sum(g_c2);
            g_o = warp_sum(g_o);
#pragma unroll
            for (int k = 0; k < D; ++k) g_col[k] = warp_sum(g_col[k]);
            if (v_abs != nullptr) {
                a_x = warp_sum(a_x);
                a_y = warp_sum(a_y);
            }
            if (lane == 0) {
                atomicAdd(v_means2d + 2 * (int64_t)g, g_xy0);
                atomicAdd(v_means2d + 2 * (int64_t)g + 1, g_xy1);
                atomicAdd(v_conics + 3 * (int64_t)g, g_c0);
                atomicAdd(v_conics + 3 * (int64_t)g + 1, g_c1);
                atomicAdd(v_conics + 3 * (int64_t)g + 2, g_c2);
                atomicAdd(v_opacities + g, g_o);
#pragma unroll
                for (int k = 0; k < D; ++k) atomicAdd(v_colors + (int64_t)D * g + k, g_col[k]);
                if (v_abs != nullptr) {
                    atomicAdd(v_abs + 2 * (int64_t)g, a_x);
                    atomicAdd(v_abs + 2 * (int64_t)g + 1, a_y);
                }
            }
        }
    }
}

}  // namespace raster

namespace {
int generic_threads(int tile_size) { return ((tile_size * tile_size + 31) / 32) * 32; }
size_t generic_smem(int threads) { return (size_t)threads * (2 * sizeof(float4) + sizeof(int)); }
}  // namespace

extern "C" int gsb_rasterize_fwd_generic(int C, int64_t n_isects, int D, const float *means2d, const float *conics,
                                         const float *opacities, const float *colors, const float *backgrounds,
                                         const uint8_t *masks, int width, int height, int tile_size, int tile_width,
                                         int tile_height, const int32_t *isect_offsets, const int32_t *flatten_ids,
                                         float *render_colors, float *render_alphas, int32_t *last_ids,
                                         cudaStream_t stream) {
    GSB_REQUIRE(C >= 0 && n_isects >= 0 && width > 0 && height > 0, "bad sizes");
    GSB_REQUIRE(tile_size >= 1 && tile_size * tile_size <= 1024, "tile_size must satisfy 1 <= tile_size^2 <= 1024");
    GSB_REQUIRE(tile_width * tile_size >= width && tile_height * tile_size >= height, "tile grid smaller than image");
    GSB_REQUIRE(tile_height <= 65535 && C <= 65535, "grid too large");
    if (C == 0) return GSB_OK;
    const int threads = generic_threads(tile_size);
    dim3 grid(tile_width, tile_height, C);
#define GSB_CASE(DD)                                                                                               \
    case DD:                                                                                                       \
        raster::rasterize_fwd_generic_kernel<DD><<<grid, threads, generic_smem(threads), stream>>>(                \
            C, n_isects, means2d, conics, opacities, colors, backgrounds, masks, width, height, tile_size,         \
            tile_width, tile_height, isect_offsets, flatten_ids, render_colors, render_alphas, last_ids);          \
        break;
    switch (D) {
        GSB_CASE(1) GSB_CASE(2) GSB_CASE(3) GSB_CASE(4) GSB_CASE(8) GSB_CASE(16) GSB_CASE(32)
        default:
            gsb_set_error("rasterize_fwd_generic: channel count %d unsupported (1,2,3,4,8,16,32)", D);
            return GSB_ERR_UNSUPPORTED;
    }
#undef GSB_CASE
    return gsb_check_launch("rasterize_fwd_generic");
}

extern "C" int gsb_rasterize_bwd_generic(int C, int64_t n_isects, int D, const float *means2d, const float *conics,
                                         const float *opacities, const float *colors, const float *backgrounds,
                                         const uint8_t *masks, int width, int height, int tile_size, int tile_width,
                                         int tile_height, const int32_t *isect_offsets, const int32_t *flatten_ids,
                                         const float *render_alphas, const int32_t *last_ids,
                                         const float *v_render_colors, const float *v_render_alphas,
                                         float *v_means2d_abs, float *v_means2d, float *v_conics, float *v_colors,
                                         float *v_opacities, cudaStream_t stream) {
    GSB_REQUIRE(C >= 0 && n_isects >= 0 && width > 0 && height > 0, "bad sizes");
    GSB_REQUIRE(tile_size >= 1 && tile_size * tile_size <= 1024, "tile_size must satisfy 1 <= tile_size^2 <= 1024");
    GSB_REQUIRE(tile_height <= 65535 && C <= 65535, "grid too large");
    if (C == 0 || n_isects == 0) return GSB_OK;
    const int threads = generic_threads(tile_size);
    dim3 grid(tile_width, tile_height, C);
#define GSB_CASE(DD)                                                                                               \
    case DD:                                                                                                       \
        raster::rasterize_bwd_generic_kernel<DD><<<grid, threads, generic_smem(threads), stream>>>(                \
            C, n_isects, means2d, conics, opacities, colors, backgrounds, masks, width, height, tile_size,         \
            tile_width, tile_height, isect_offsets, flatten_ids, render_alphas, last_ids, v_render_colors,         \
            v_render_alphas, v_means2d_abs, v_means2d, v_conics, v_colors, v_opacities);                           \
        break;
    switch (D) {
        GSB_CASE(1) GSB_CASE(2) GSB_CASE(3) GSB_CASE(4) GSB_CASE(8) GSB_CASE(16) GSB_CASE(32)
        default:
            gsb_set_error("rasterize_bwd_generic: channel count %d unsupported (1,2,3,4,8,16,32)", D);
            return GSB_ERR_UNSUPPORTED;
    }
#undef GSB_CASE
    return gsb_check_launch("rasterize_bwd_generic");
}
