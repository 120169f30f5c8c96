// Shared pieces of the tile compositor (rasterize_to_pixels forward / backward).
//
// Layout: one CTA of 256 threads per 16x16 tile; warp w owns the 8x4 pixel block
// (bx = w & 1, by = w >> 1), lane l the pixel (l & 7, l >> 3) inside it.
//
// Per-Gaussian "splat records" are two float4 arrays built once per call by splat_pack_kernel:
//   recA[g] = (x, y, hx, hy)     hx/hy: half extents of the bbox of {alpha >= 1/255}
//   recB[g] = (conic_xx, conic_xy, conic_yy, opacity)
// so that a tile gathers a Gaussian with two aligned 16-byte loads. While staging a batch, the
// loading thread turns the box into an 8-bit mask of the warps (8x4 pixel blocks) it can touch;
// every warp then walks only its own splats through ballot-compacted bit lists, so Gaussians whose
// alpha would be < 1/255 on all pixels of a block cost that warp nothing. The reject test is conservative (see splat_extent), so the composited result
// is the one the reference produces (RasterizeToPixels3DGSFwd.cu:140-169): rejected Gaussians
// are exactly ones the reference `continue`s over.
#pragma once
#include "common.cuh"

namespace raster {

constexpr int kTile = 16;
constexpr int kThreads = kTile * kTile;
constexpr int kWarps = kThreads / 32;
constexpr float kAlphaMin = 1.0f / 255.0f;
constexpr float kAlphaMax = 0.999f;
constexpr float kTmin = 1e-4f;

// The staged conic is pre-multiplied so that the per-pixel exponent needs no extra scaling:
//   log2(e^-sigma) = dx (A' dx + B' dy) + C' dy^2,  A' = -log2(e)/2 a,  B' = -log2(e) b,  C' = -log2(e)/2 c
// (5 FP32 instructions + one MUFU.EX2 instead of 8 + a scaling multiply). sigma < 0  <=>  exponent > 0.
// Forward and backward evaluate the same expression, so they agree on every alpha >= 1/255 decision.
constexpr float kLog2e = 1.4426950408889634f;
__device__ __forceinline__ float4 stage_conic(const float4 &B) {
    return make_float4(-0.5f * kLog2e * B.x, -kLog2e * B.y, -0.5f * kLog2e * B.z, B.w);
}
__device__ __forceinline__ float gauss_log2(const float4 &b4, float dx, float dy) {
    const float t = fmaf(b4.y, dy, b4.x * dx);
    return fmaf(b4.z * dy, dy, dx * t);
}
// staged (A', B', C') -> conic (a, b, c)
constexpr float kUnscaleDiag = -2.0f / kLog2e, kUnscaleOff = -1.0f / kLog2e;

__device__ __forceinline__ void splat_extent(float a, float b, float c, float opac, float &hx, float &hy) {
    const float inf = __int_as_float(0x7f800000);
    hx = inf;
    hy = inf;
    if (opac < kAlphaMin) {  // alpha = opac * exp(-sigma) <= opac for every counted (sigma >= 0) pixel
        hx = -inf;
        hy = -inf;
        return;
    }
    float det = a * c - b * b;
    if (!(det > 0.f) || !(a > 0.f) || !(c > 0.f) || !(opac <= 1e30f)) return;  // not an ellipse: never reject
    // alpha >= 1/255  <=>  sigma <= ln(255 opac) =: tau; inflate tau for float rounding of sigma
    float tau = __logf(255.0f * opac) * 1.001f + 0.05f;
    float ex = sqrtf(2.0f * tau * c / det) * 1.001f + 0.05f;
    float ey = sqrtf(2.0f * tau * a / det) * 1.001f + 0.05f;
    if (ex == ex) hx = ex;
    if (ey == ey) hy = ey;
}

// compacting append of this thread's item into the staged batch; returns the slot or -1.
// Must be called by all 256 threads. s_wcnt: kWarps ints. total = number kept in the batch.
__device__ __forceinline__ int stage_slot(bool keep, int *s_wcnt, int &total) {
    unsigned ballot = __ballot_sync(0xffffffffu, keep);
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) s_wcnt[warp] = __popc(ballot);
    __syncthreads();
    int base = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
        int c = s_wcnt[w];
        base += (w < warp) ? c : 0;
        tot += c;
    }
    total = tot;
    return keep ? base + __popc(ballot & ((1u << lane) - 1u)) : -1;
}

// pixel-centre bounds of a rectangle of pixels [x0, x0+w) x [y0, y0+h)
struct Rect {
    float xlo, xhi, ylo, yhi;
};
__device__ __forceinline__ Rect pixel_rect(int x0, int y0, int w, int h) {
    Rect r;
    r.xlo = (float)x0 + 0.5f;
    r.xhi = (float)(x0 + w) - 0.5f;
    r.ylo = (float)y0 + 0.5f;
    r.yhi = (float)(y0 + h) - 0.5f;
    return r;
}
// false when no pixel centre of the rectangle can see alpha >= 1/255 from this splat (A = x, y, hx, hy)
__device__ __forceinline__ bool splat_touches(const float4 &A, const Rect &r) {
    return !((A.x + A.z < r.xlo) | (A.x - A.z > r.xhi) | (A.y + A.w < r.ylo) | (A.y - A.w > r.yhi));
}

// Smallest sigma(p) = .5 (a ex^2 + c ey^2) + b ex ey, e = p - mu, over the rectangle of pixel centres
// [X0, X1] x [Y0, Y1], minus a bound on the float error of evaluating it. sigma is convex, so unless
// mu lies inside the rectangle the minimum sits on one of the (at most two) edges facing mu.
__device__ __forceinline__ float min_sigma_over_rect(float mx, float my, float a, float b, float c, float b_over_a,
                                                     float b_over_c, float X0, float X1, float Y0, float Y1) {
    const float px = fminf(fmaxf(mx, X0), X1), py = fminf(fmaxf(my, Y0), Y1);
    const float ex = px - mx, ey = py - my;
    if (ex == 0.f && ey == 0.f) return 0.f;
    // edge x = px: free coordinate y, unconstrained optimum at my - (b/c) ex
    const float ey1 = fminf(fmaxf(my - b_over_c * ex, Y0), Y1) - my;
    const float s1 = 0.5f * (a * ex * ex + c * ey1 * ey1), t1 = b * ex * ey1;
    // edge y = py
    const float ex2 = fminf(fmaxf(mx - b_over_a * ey, X0), X1) - mx;
    const float s2 = 0.5f * (a * ex2 * ex2 + c * ey * ey), t2 = b * ex2 * ey;
    const float q1 = (s1 + t1) - 1e-5f * (s1 + fabsf(t1));
    const float q2 = (s2 + t2) - 1e-5f * (s2 + fabsf(t2));
    // an edge that does not face mu (ex == 0 resp. ey == 0) degenerates to a line through the
    // rectangle that contains the other candidate's optimum or a larger value: min() stays valid
    return fminf(ex != 0.f ? q1 : q2, ey != 0.f ? q2 : q1);
}

// bit w set <=> warp w's 8x4 pixel block (bx = w & 1, by = w >> 1) of the tile whose first pixel is
// (x0, y0) may contain a pixel with alpha >= 1/255 for this splat. 0 = the whole tile can skip it.
// Two levels: the cheap bounding box of the {alpha >= 1/255} ellipse, then for the blocks that
// survive the exact minimum of sigma over the block against tau = ln(255 opacity) (+ margins).
__device__ __forceinline__ unsigned splat_block_mask(const float4 &A, const float4 &B, int x0, int y0) {
    const float xlo = A.x - A.z, xhi = A.x + A.z, ylo = A.y - A.w, yhi = A.y + A.w;
    const float fx = (float)x0, fy = (float)y0;
    unsigned xb = 0u, yb = 0u;
#pragma unroll
    for (int bx = 0; bx < 2; ++bx)  // pixel centres fx + 8 bx + .5 .. fx + 8 bx + 7.5
        xb |= (!((xhi < fx + 8.f * bx + 0.5f) | (xlo > fx + 8.f * bx + 7.5f))) ? (1u << bx) : 0u;
#pragma unroll
    for (int by = 0; by < 4; ++by)
        yb |= (!((yhi < fy + 4.f * by + 0.5f) | (ylo > fy + 4.f * by + 3.5f))) ? (1u << by) : 0u;
    unsigned m = 0u;
#pragma unroll
    for (int by = 0; by < 4; ++by) m |= ((yb >> by) & 1u) ? (xb << (2 * by)) : 0u;
    if (m == 0u || !(A.z < 3.0e38f)) return m;  // box misses the tile / not an ellipse (never culled)
    const float tau = __logf(255.0f * B.w) * 1.0001f + 2e-3f;
    const float b_over_a = __fdividef(B.y, B.x), b_over_c = __fdividef(B.y, B.z);
    unsigned keep = 0u;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
        if ((m >> w) & 1u) {
            const float X0 = fx + 8.f * (w & 1) + 0.5f, Y0 = fy + 4.f * (w >> 1) + 0.5f;
            const float q = min_sigma_over_rect(A.x, A.y, B.x, B.y, B.z, b_over_a, b_over_c, X0, X0 + 7.f, Y0,
                                                Y0 + 3.f);
            keep |= (q <= tau) ? (1u << w) : 0u;
        }
    }
    return keep;
}

// ---- Blackwell packed FP32 (PTX *.f32x2 -> SASS FFMA2 / FMUL2 / FADD2): one issued instruction, two lanes of work. A
// scalar enters as a broadcast operand at no cost (ptxas folds `mov.b64 {a, a}` into the operand modifier R.F32).
typedef unsigned long long f32x2;

__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float &lo, float &hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
    f32x2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
    f32x2 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 bcast2(float a) { return pack2(a, a); }
// acc += a * b (accumulate form: no register moves around the asm)
__device__ __forceinline__ void fma2_acc(f32x2 &acc, f32x2 a, f32x2 b) {
    asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(a), "l"(b));
}
__device__ __forceinline__ float4 lds128(unsigned addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}

struct GradTargets {
    float *v_means2d, *v_conics, *v_opacities, *v_colors, *v_abs;
};

struct TileGeom {
    int cam, tile_x, tile_y;
    int64_t tile_id;      // flat (camera, tile) index
    int range_start, range_end;
    int px, py;           // this thread's pixel
    bool inside;
    int64_t pix_id;       // flat (camera, y, x) index
};

__device__ __forceinline__ TileGeom tile_geometry(int64_t n_isects, int W, int H, int tw, int th, int C,
                                                  const int32_t *__restrict__ offsets) {
    TileGeom g;
    g.cam = blockIdx.z;
    g.tile_y = blockIdx.y;
    g.tile_x = blockIdx.x;
    g.tile_id = ((int64_t)g.cam * th + g.tile_y) * tw + g.tile_x;
    g.range_start = offsets[g.tile_id];
    g.range_end = (g.tile_id + 1 < (int64_t)C * th * tw) ? offsets[g.tile_id + 1] : (int)n_isects;
    int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    g.px = g.tile_x * kTile + (warp & 1) * 8 + (lane & 7);
    g.py = g.tile_y * kTile + (warp >> 1) * 4 + (lane >> 3);
    g.inside = (g.px < W) & (g.py < H);
    g.pix_id = ((int64_t)g.cam * H + min(g.py, H - 1)) * W + min(g.px, W - 1);
    return g;
}

}  // namespace raster
