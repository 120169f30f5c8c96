// Layout of a mailbox region (csrc/peer.cu) shared with the kernels that store their rows straight into the camera
// owners' mailboxes (projection.cu, sh.cu): a structure of arrays, field after field (means2d 2, depths 1, conics 3,
// opacities 1, colours D, radii 1 [, camera ids 1, Gaussian ids 1: packed rows]), each starting on a 16-byte boundary.
#pragma once
#include <stdint.h>

namespace peer {

constexpr int kMaxPeers = 16;

__host__ __device__ __forceinline__ int64_t round_up4(int64_t v) { return (v + 3) & ~(int64_t)3; }
__host__ __device__ __forceinline__ int64_t field_offset(const int *width, int field, int64_t rows) {
    int64_t off = 0;
    for (int g = 0; g < field; ++g) off += round_up4(rows * width[g]);
    return off;
}

// "Push" destination of the unpacked exchange: row (camera c, local Gaussian n) of this rank belongs to rank
// c / C_local, at row (c % C_local) * N_total + n_base + n of its [C_local, N_total] table. world == 0: no push.
struct PushRows {
    float *region[kMaxPeers];  // peer mailbox base + region offset, as floats
    int world, C_local;
    int64_t N_total, n_base;
    int64_t off_means2d, off_depths, off_conics, off_opacities, off_colors, off_radii;  // float offsets of the fields
    const float *opacities;  // [N_local]: the mailbox holds opacity x compensation (what rasterization() composites)
};

__device__ __forceinline__ float *push_row(const PushRows &p, int c, int64_t n, int64_t &row) {
    const int dest = c / p.C_local;
    row = (int64_t)(c - dest * p.C_local) * p.N_total + p.n_base + n;
    return p.region[dest];
}

}  // namespace peer
