// Row pack / unpack for the distributed exchange of projected Gaussians.
//
// Replaces (behaviour, not code) the torch.cat / split / view packing around the reference's
// all-to-all calls: gsplat/distributed.py:229-257 as used by gsplat/rendering.py:413-494.
// The reference exchanges radii (int32) in a separate collective because its dtype differs;
// here one row carries means2d(2) depths(1) conics(3) opacities(1) colors(D) and the radius bit
// pattern, so a single all-to-all per direction moves everything. HBM-bound streaming kernels.
#include "common.cuh"

namespace {

constexpr int kThreads = 256;

__global__ void __launch_bounds__(kThreads)
a2a_pack_kernel(int64_t rows, int D, const float *__restrict__ means2d, const float *__restrict__ depths,
                const float *__restrict__ conics, const float *__restrict__ opacities,
                const float *__restrict__ colors, const int32_t *__restrict__ radii, float *__restrict__ packed) {
    const int width = 8 + D;
    int64_t total = rows * width;
    for (int64_t e = (int64_t)blockIdx.x * kThreads + threadIdx.x; e < total; e += (int64_t)gridDim.x * kThreads) {
        int64_t r = e / width;
        int k = (int)(e - r * width);
        float v;
        if (k < 2) v = means2d[2 * r + k];
        else if (k == 2) v = depths[r];
        else if (k < 6) v = conics[3 * r + (k - 3)];
        else if (k == 6) v = opacities[r];
        else if (k < 7 + D) v = colors[(int64_t)D * r + (k - 7)];
        else v = __int_as_float(radii[r]);
        packed[e] = v;
    }
}

__global__ void __launch_bounds__(kThreads)
a2a_unpack_kernel(int64_t rows, int D, const float *__restrict__ packed, float *__restrict__ means2d,
                  float *__restrict__ depths, float *__restrict__ conics, float *__restrict__ opacities,
                  float *__restrict__ colors, int32_t *__restrict__ radii) {
    const int width = 8 + D;
    int64_t total = rows * width;
    for (int64_t e = (int64_t)blockIdx.x * kThreads + threadIdx.x; e < total; e += (int64_t)gridDim.x * kThreads) {
        int64_t r = e / width;
        int k = (int)(e - r * width);
        float v = packed[e];
        if (k < 2) { if (means2d) means2d[2 * r + k] = v; }
        else if (k == 2) { if (depths) depths[r] = v; }
        else if (k < 6) { if (conics) conics[3 * r + (k - 3)] = v; }
        else if (k == 6) { if (opacities) opacities[r] = v; }
        else if (k < 7 + D) { if (colors) colors[(int64_t)D * r + (k - 7)] = v; }
        else { if (radii) radii[r] = __float_as_int(v); }
    }
}

unsigned grid_for(int64_t total) {
    int64_t b = (total + kThreads - 1) / kThreads;
    int64_t cap = 148 * 16;
    return (unsigned)(b < cap ? (b > 0 ? b : 1) : cap);
}

}  // namespace

extern "C" int gsb_a2a_pack(int64_t rows, int D, const float *means2d, const float *depths, const float *conics,
                            const float *opacities, const float *colors, const int32_t *radii, float *packed,
                            cudaStream_t stream) {
    GSB_REQUIRE(rows >= 0 && D >= 0, "negative size");
    if (rows == 0) return GSB_OK;
    a2a_pack_kernel<<<grid_for(rows * (8 + D)), kThreads, 0, stream>>>(rows, D, means2d, depths, conics, opacities,
                                                                      colors, radii, packed);
    return gsb_check_launch("a2a_pack");
}

extern "C" int gsb_a2a_unpack(int64_t rows, int D, const float *packed, float *means2d, float *depths,
                              float *conics, float *opacities, float *colors, int32_t *radii,
                              cudaStream_t stream) {
    GSB_REQUIRE(rows >= 0 && D >= 0, "negative size");
    if (rows == 0) return GSB_OK;
    a2a_unpack_kernel<<<grid_for(rows * (8 + D)), kThreads, 0, stream>>>(rows, D, packed, means2d, depths, conics,
                                                                        opacities, colors, radii);
    return gsb_check_launch("a2a_unpack");
}
