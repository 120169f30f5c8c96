// rasterize_to_indices_in_range: which (Gaussian, pixel) pairs the compositor blends, as index lists.
//
// Replaces (behaviour, not code): gsplat/cuda/csrc/RasterizeToIndices3DGS.cu:14-170 behind
// gsplat/cuda/_wrapper.py:577-648 (the op the reference's pure-torch compositor, _torch_impl.py:522-598, is
// built on). Oracle: oracle/gs_oracle.py::rasterize_to_indices_in_range.
//
// A "batch" is tile_size^2 consecutive sorted intersections of a tile; a call walks batches
// [range_start, range_end) of every tile, starting from the given per-pixel transmittance, with the compositor's
// rules (skip sigma < 0 or alpha < 1/255; stop before the splat that makes T (1 - alpha) <= 1e-4). Two passes over
// the same traversal: count per pixel, then (after the caller's exclusive scan) write. Output order: pixel-major
// (camera, row, column), sorted order inside a pixel; gaussian_ids are within-camera ids (flat id mod N).
// Not on the training hot path: one thread per pixel, plain loads (every pixel of a tile reads the same splat:
// broadcast).
#include "common.cuh"

namespace {

template <bool WRITE>
__global__ void __launch_bounds__(256)
indices_kernel(int C, int N, int64_t n_isects, int range_start, int range_end, const float *__restrict__ transmittances,
               const float *__restrict__ means2d, const float *__restrict__ conics,
               const float *__restrict__ opacities, int W, int H, int tile_size, int tw, int th,
               const int32_t *__restrict__ offsets, const int32_t *__restrict__ flatten_ids,
               int32_t *__restrict__ counts, const int64_t *__restrict__ starts, int64_t *__restrict__ gaussian_ids,
               int64_t *__restrict__ indices) {
    const int cam = blockIdx.z;
    const int px = blockIdx.x * blockDim.x + threadIdx.x, py = blockIdx.y * blockDim.y + threadIdx.y;
    if (px >= W || py >= H) return;
    const int64_t tile_id = ((int64_t)cam * th + py / tile_size) * tw + px / tile_size;
    const int64_t lo = offsets[tile_id];
    const int64_t hi = (tile_id + 1 < (int64_t)C * th * tw) ? offsets[tile_id + 1] : n_isects;
    const int64_t block = (int64_t)tile_size * tile_size;
    const int64_t first = lo + block * range_start;
    const int64_t last = min(hi, lo + block * (int64_t)range_end);
    const int64_t pix = ((int64_t)cam * H + py) * W + px;
    float T = transmittances[pix];
    const float fx = (float)px + 0.5f, fy = (float)py + 0.5f;
    int64_t out = WRITE ? starts[pix] : 0;
    int cnt = 0;
    for (int64_t i = first; i < last; ++i) {
        const int g = flatten_ids[i];
        const float dx = means2d[2 * (size_t)g] - fx, dy = means2d[2 * (size_t)g + 1] - fy;
        const float *cp = conics + 3 * (size_t)g;
        const float sigma = 0.5f * (cp[0] * dx * dx + cp[2] * dy * dy) + cp[1] * dx * dy;
        const float alpha = fminf(0.999f, opacities[g] * __expf(-sigma));
        if (sigma < 0.f || alpha < 1.0f / 255.0f) continue;
        const float next_T = T * (1.0f - alpha);
        if (next_T <= 1e-4f) break;
        if (WRITE) {
            gaussian_ids[out] = (int64_t)(g % N);
            indices[out] = pix;
            ++out;
        }
        ++cnt;
        T = next_T;
    }
    if (!WRITE) counts[pix] = cnt;
}

template <bool WRITE>
int launch(int C, int N, int64_t n_isects, int range_start, int range_end, const float *transmittances,
           const float *means2d, const float *conics, const float *opacities, int width, int height, int tile_size,
           int tile_width, int tile_height, const int32_t *isect_offsets, const int32_t *flatten_ids, int32_t *counts,
           const int64_t *starts, int64_t *gaussian_ids, int64_t *indices, cudaStream_t stream) {
    GSB_REQUIRE(C >= 0 && N >= 0 && n_isects >= 0 && width > 0 && height > 0 && tile_size > 0, "bad sizes");
    GSB_REQUIRE(range_start >= 0 && range_end >= range_start, "bad batch range");
    GSB_REQUIRE(tile_width * tile_size >= width && tile_height * tile_size >= height, "tile grid smaller than image");
    GSB_REQUIRE(C <= 65535, "too many cameras");
    if (C == 0) return GSB_OK;
    dim3 block(16, 16), grid(gsb_cdiv(width, 16), gsb_cdiv(height, 16), C);
    indices_kernel<WRITE><<<grid, block, 0, stream>>>(C, N, n_isects, range_start, range_end, transmittances, means2d,
                                                      conics, opacities, width, height, tile_size, tile_width,
                                                      tile_height, isect_offsets, flatten_ids, counts, starts,
                                                      gaussian_ids, indices);
    return gsb_check_launch(WRITE ? "rasterize_indices_write" : "rasterize_indices_count");
}

}  // namespace

extern "C" int gsb_rasterize_indices_count(int C, int N, int64_t n_isects, int range_start, int range_end,
                                           const float *transmittances, const float *means2d, const float *conics,
                                           const float *opacities, int width, int height, int tile_size,
                                           int tile_width, int tile_height, const int32_t *isect_offsets,
                                           const int32_t *flatten_ids, int32_t *counts, cudaStream_t stream) {
    return launch<false>(C, N, n_isects, range_start, range_end, transmittances, means2d, conics, opacities, width,
                         height, tile_size, tile_width, tile_height, isect_offsets, flatten_ids, counts, nullptr,
                         nullptr, nullptr, stream);
}

extern "C" int gsb_rasterize_indices_write(int C, int N, int64_t n_isects, int range_start, int range_end,
                                           const float *transmittances, const float *means2d, const float *conics,
                                           const float *opacities, int width, int height, int tile_size,
                                           int tile_width, int tile_height, const int32_t *isect_offsets,
                                           const int32_t *flatten_ids, const int64_t *pixel_starts,
                                           int64_t *gaussian_ids, int64_t *indices, cudaStream_t stream) {
    return launch<true>(C, N, n_isects, range_start, range_end, transmittances, means2d, conics, opacities, width,
                        height, tile_size, tile_width, tile_height, isect_offsets, flatten_ids, nullptr, pixel_starts,
                        gaussian_ids, indices, stream);
}
