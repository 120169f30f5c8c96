// fully_fused_projection forward / backward, unpacked and packed (COO) variants.
//
// Replaces (behaviour, not code):
//   gsplat/triton_impl/fused_projection_fwd.py:18-228, fused_projection_bwd.py:26-363
//   gsplat/cuda/csrc/ProjectionEWA3DGSFused.cu / ProjectionEWA3DGSPacked.cu (packed + covars + ortho)
// Math contract: SURVEY.md Appendix A; oracle: oracle/gs_oracle.py::fully_fused_projection.
//
// B200 design: one thread per Gaussian, cameras looped *inside* the thread, so the
// quat/scale -> covariance work is done once per Gaussian and, in the backward pass,
// v_means / v_covar accumulate in registers across cameras and are written with plain
// stores (no atomics; the reference backends use one atomic per (camera, Gaussian, float)).
// All kernels are HBM-bound element-wise passes.
#include "common.cuh"
#include "gsplat_b200.h"
#include "peer_layout.cuh"

namespace {

constexpr int kThreads = 256;
enum { MODEL_PINHOLE = 0, MODEL_ORTHO = 1, MODEL_FISHEYE = 2 };

struct Sym3 {
    float xx, xy, xz, yy, yz, zz;
};

struct Camera {
    float r00, r01, r02, r10, r11, r12, r20, r21, r22;
    float t0, t1, t2;
    float fx, fy, cx, cy;
};

__device__ __forceinline__ Camera load_camera(const float *__restrict__ viewmats, const float *__restrict__ Ks, int c) {
    const float *v = viewmats + 16 * (size_t)c;
    const float *k = Ks + 9 * (size_t)c;
    Camera cam;
    cam.r00 = __ldg(v + 0); cam.r01 = __ldg(v + 1); cam.r02 = __ldg(v + 2);  cam.t0 = __ldg(v + 3);
    cam.r10 = __ldg(v + 4); cam.r11 = __ldg(v + 5); cam.r12 = __ldg(v + 6);  cam.t1 = __ldg(v + 7);
    cam.r20 = __ldg(v + 8); cam.r21 = __ldg(v + 9); cam.r22 = __ldg(v + 10); cam.t2 = __ldg(v + 11);
    cam.fx = __ldg(k + 0); cam.cx = __ldg(k + 2); cam.fy = __ldg(k + 4); cam.cy = __ldg(k + 5);
    return cam;
}

// rotation matrix of the normalised quaternion (w,x,y,z); also returns 1/|q|
__device__ __forceinline__ void quat_to_rotmat(float4 q, float R[9], float &inv_norm) {
    inv_norm = rsqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
    float w = q.x * inv_norm, x = q.y * inv_norm, y = q.z * inv_norm, z = q.w * inv_norm;
    float x2 = x * x, y2 = y * y, z2 = z * z;
    float xy = x * y, xz = x * z, yz = y * z, wx = w * x, wy = w * y, wz = w * z;
    R[0] = 1.f - 2.f * (y2 + z2); R[1] = 2.f * (xy - wz);       R[2] = 2.f * (xz + wy);
    R[3] = 2.f * (xy + wz);       R[4] = 1.f - 2.f * (x2 + z2); R[5] = 2.f * (yz - wx);
    R[6] = 2.f * (xz - wy);       R[7] = 2.f * (yz + wx);       R[8] = 1.f - 2.f * (x2 + y2);
}

// Sigma = (R diag(s)) (R diag(s))^T
__device__ __forceinline__ Sym3 covar_from_rot_scale(const float R[9], float s0, float s1, float s2) {
    float m00 = R[0] * s0, m01 = R[1] * s1, m02 = R[2] * s2;
    float m10 = R[3] * s0, m11 = R[4] * s1, m12 = R[5] * s2;
    float m20 = R[6] * s0, m21 = R[7] * s1, m22 = R[8] * s2;
    Sym3 S;
    S.xx = m00 * m00 + m01 * m01 + m02 * m02;
    S.xy = m00 * m10 + m01 * m11 + m02 * m12;
    S.xz = m00 * m20 + m01 * m21 + m02 * m22;
    S.yy = m10 * m10 + m11 * m11 + m12 * m12;
    S.yz = m10 * m20 + m11 * m21 + m12 * m22;
    S.zz = m20 * m20 + m21 * m21 + m22 * m22;
    return S;
}

__device__ __forceinline__ Sym3 load_covar(const float *__restrict__ covars, const float *__restrict__ quats,
                                           const float *__restrict__ scales, int n) {
    if (covars != nullptr) {
        const float *p = covars + 6 * (size_t)n;
        Sym3 S;
        S.xx = p[0]; S.xy = p[1]; S.xz = p[2]; S.yy = p[3]; S.yz = p[4]; S.zz = p[5];
        return S;
    }
    float4 q = *reinterpret_cast<const float4 *>(quats + 4 * (size_t)n);
    const float *sp = scales + 3 * (size_t)n;
    float R[9], inv_norm;
    quat_to_rotmat(q, R, inv_norm);
    return covar_from_rot_scale(R, sp[0], sp[1], sp[2]);
}

// Sigma_c = Rc Sigma Rc^T
__device__ __forceinline__ Sym3 rotate_covar(const Camera &c, const Sym3 &S) {
    // A = Rc * Sigma
    float a00 = c.r00 * S.xx + c.r01 * S.xy + c.r02 * S.xz;
    float a01 = c.r00 * S.xy + c.r01 * S.yy + c.r02 * S.yz;
    float a02 = c.r00 * S.xz + c.r01 * S.yz + c.r02 * S.zz;
    float a10 = c.r10 * S.xx + c.r11 * S.xy + c.r12 * S.xz;
    float a11 = c.r10 * S.xy + c.r11 * S.yy + c.r12 * S.yz;
    float a12 = c.r10 * S.xz + c.r11 * S.yz + c.r12 * S.zz;
    float a20 = c.r20 * S.xx + c.r21 * S.xy + c.r22 * S.xz;
    float a21 = c.r20 * S.xy + c.r21 * S.yy + c.r22 * S.yz;
    float a22 = c.r20 * S.xz + c.r21 * S.yz + c.r22 * S.zz;
    Sym3 o;
    o.xx = a00 * c.r00 + a01 * c.r01 + a02 * c.r02;
    o.xy = a00 * c.r10 + a01 * c.r11 + a02 * c.r12;
    o.xz = a00 * c.r20 + a01 * c.r21 + a02 * c.r22;
    o.yy = a10 * c.r10 + a11 * c.r11 + a12 * c.r12;
    o.yz = a10 * c.r20 + a11 * c.r21 + a12 * c.r22;
    o.zz = a20 * c.r20 + a21 * c.r21 + a22 * c.r22;
    return o;
}

struct ProjParams {
    int width, height;
    float eps2d, near_plane, far_plane, radius_clip;
};

struct Proj2D {
    float depth, mx, my;           // camera-space z and projected mean
    float cxx, cxy, cyy;           // conic (inverse of blurred 2D covariance)
    float comp;                    // sqrt(max(0, det/det_blur))
    int radius;                    // 0 when culled
};

// Jacobian entries of the (clamped) perspective projection, shared by fwd and bwd.
struct PinholeJ {
    float iz, jxx, jxz, jyy, jyz;
    bool clamp_x, clamp_y;
};

__device__ __forceinline__ PinholeJ pinhole_jacobian(const Camera &c, float tx, float ty, float tz, int W, int H) {
    PinholeJ j;
    j.iz = 1.0f / tz;
    float margin_x = 0.15f * (float)W / c.fx;
    float lo_x = -margin_x - c.cx / c.fx, hi_x = margin_x + ((float)W - c.cx) / c.fx;
    float sx = tx * j.iz;
    j.clamp_x = (sx < lo_x) | (sx > hi_x);
    sx = fminf(fmaxf(sx, lo_x), hi_x);
    float margin_y = 0.15f * (float)H / c.fy;
    float lo_y = -margin_y - c.cy / c.fy, hi_y = margin_y + ((float)H - c.cy) / c.fy;
    float sy = ty * j.iz;
    j.clamp_y = (sy < lo_y) | (sy > hi_y);
    sy = fminf(fmaxf(sy, lo_y), hi_y);
    j.jxx = c.fx * j.iz;
    j.jxz = -c.fx * sx * j.iz;
    j.jyy = c.fy * j.iz;
    j.jyz = -c.fy * sy * j.iz;
    return j;
}

// Equidistant fisheye (reference: _torch_impl._fisheye_proj :123-182, Utils.cuh:554-592): with r = |(x, y)|,
// theta = atan2(r, z), the image point is f * (x, y) * g + c where g = theta / r. Its 2x3 Jacobian is
//   [ fx (g + x^2 h)   fx x y h        -fx x k ]        h = (dg/dr) / r = z k / r^2 - theta / r^3
//   [ fy x y h         fy (g + y^2 h)  -fy y k ]        k = 1 / (r^2 + z^2)
// and every derivative of J (backward) needs only one more radial function m = (dh/dr) / r plus dh/dz = 2 k^2.
// Near the optical axis (r < 0.3 z) g, h, m are evaluated by their series in t = r/z, where the closed forms cancel
// catastrophically (the reference guards that region with eps = 1e-7 instead; differences stay below 1e-5).
struct FisheyeJ {
    float g, h, k, m;
    float j00, j01, j02, j10, j11, j12;
};

__device__ __forceinline__ FisheyeJ fisheye_jacobian(const Camera &c, float x, float y, float z) {
    FisheyeJ j;
    const float r2 = x * x + y * y, r = sqrtf(r2);
    j.k = 1.0f / (r2 + z * z);
    if (z > 0.f && r < 0.3f * z) {
        // atan(t)/t = sum_n a_n t^2n, a_n = (-1)^n / (2n+1):  g = sum a_n t^2n / z,
        // h = sum_{n>=1} 2n a_n t^(2n-2) / z^3,  m = sum_{n>=2} 2n (2n-2) a_n t^(2n-4) / z^5   (10 terms: < 1e-6 at t = .3)
        const float iz = 1.0f / z, t2 = r2 * iz * iz, iz2 = iz * iz;
        float sg = 0.f, sh = 0.f, sm = 0.f;
#pragma unroll
        for (int n = 9; n >= 0; --n) {
            const float a = ((n & 1) ? -1.0f : 1.0f) / (float)(2 * n + 1);
            sg = fmaf(sg, t2, a);
            if (n >= 1) sh = fmaf(sh, t2, (float)(2 * n) * a);
            if (n >= 2) sm = fmaf(sm, t2, (float)(2 * n * (2 * n - 2)) * a);
        }
        j.g = iz * sg;
        j.h = iz * iz2 * sh;
        j.m = iz * iz2 * iz2 * sm;
    } else {
        const float theta = atan2f(r, z), ir = 1.0f / r, ir2 = ir * ir;
        j.g = theta * ir;
        j.h = z * j.k * ir2 - theta * ir2 * ir;
        j.m = -2.0f * z * j.k * j.k * ir2 - 3.0f * z * j.k * ir2 * ir2 + 3.0f * theta * ir2 * ir2 * ir;
    }
    j.j00 = c.fx * (j.g + x * x * j.h);
    j.j01 = c.fx * x * y * j.h;
    j.j02 = -c.fx * x * j.k;
    j.j10 = c.fy * x * y * j.h;
    j.j11 = c.fy * (j.g + y * y * j.h);
    j.j12 = -c.fy * y * j.k;
    return j;
}

template <int MODEL>
__device__ __forceinline__ Proj2D project_one(const Camera &c, float px, float py, float pz, const Sym3 &S,
                                              const ProjParams &p) {
    Proj2D o;
    float tx = c.r00 * px + c.r01 * py + c.r02 * pz + c.t0;
    float ty = c.r10 * px + c.r11 * py + c.r12 * pz + c.t1;
    float tz = c.r20 * px + c.r21 * py + c.r22 * pz + c.t2;
    o.depth = tz;
    bool keep = (tz > p.near_plane) & (tz < p.far_plane);
    Sym3 Sc = rotate_covar(c, S);
    float v_xx, v_xy, v_yy;  // 2D covariance
    if (MODEL == MODEL_PINHOLE) {
        PinholeJ j = pinhole_jacobian(c, tx, ty, tz, p.width, p.height);
        v_xx = j.jxx * Sc.xx * j.jxx + 2.0f * j.jxx * Sc.xz * j.jxz + j.jxz * Sc.zz * j.jxz;
        v_xy = j.jxx * (Sc.xy * j.jyy + Sc.xz * j.jyz) + j.jxz * (Sc.yz * j.jyy + Sc.zz * j.jyz);
        v_yy = j.jyy * Sc.yy * j.jyy + 2.0f * j.jyy * Sc.yz * j.jyz + j.jyz * Sc.zz * j.jyz;
        o.mx = c.fx * tx * j.iz + c.cx;
        o.my = c.fy * ty * j.iz + c.cy;
    } else if (MODEL == MODEL_FISHEYE) {
        const FisheyeJ j = fisheye_jacobian(c, tx, ty, tz);
        // rows of J * Sigma_c
        const float a0 = j.j00 * Sc.xx + j.j01 * Sc.xy + j.j02 * Sc.xz;
        const float a1 = j.j00 * Sc.xy + j.j01 * Sc.yy + j.j02 * Sc.yz;
        const float a2 = j.j00 * Sc.xz + j.j01 * Sc.yz + j.j02 * Sc.zz;
        const float b0 = j.j10 * Sc.xx + j.j11 * Sc.xy + j.j12 * Sc.xz;
        const float b1 = j.j10 * Sc.xy + j.j11 * Sc.yy + j.j12 * Sc.yz;
        const float b2 = j.j10 * Sc.xz + j.j11 * Sc.yz + j.j12 * Sc.zz;
        v_xx = a0 * j.j00 + a1 * j.j01 + a2 * j.j02;
        v_xy = a0 * j.j10 + a1 * j.j11 + a2 * j.j12;
        v_yy = b0 * j.j10 + b1 * j.j11 + b2 * j.j12;
        o.mx = c.fx * tx * j.g + c.cx;
        o.my = c.fy * ty * j.g + c.cy;
    } else {
        v_xx = c.fx * c.fx * Sc.xx;
        v_xy = c.fx * c.fy * Sc.xy;
        v_yy = c.fy * c.fy * Sc.yy;
        o.mx = c.fx * tx + c.cx;
        o.my = c.fy * ty + c.cy;
    }
    float det0 = v_xx * v_yy - v_xy * v_xy;
    v_xx += p.eps2d;
    v_yy += p.eps2d;
    float det = v_xx * v_yy - v_xy * v_xy;
    o.comp = sqrtf(fmaxf(det0 / det, 0.0f));
    keep &= (det > 0.0f);
    float inv_det = 1.0f / det;
    o.cxx = inv_det * v_yy;
    o.cxy = -inv_det * v_xy;
    o.cyy = inv_det * v_xx;
    float b = 0.5f * (v_xx + v_yy);
    float v1 = b + sqrtf(fmaxf(0.01f, b * b - det));
    float radius = ceilf(3.0f * sqrtf(v1));
    keep &= (radius > p.radius_clip) & (o.mx + radius > 0.0f) & (o.mx - radius < (float)p.width) &
            (o.my + radius > 0.0f) & (o.my - radius < (float)p.height);
    o.radius = keep ? (int)radius : 0;
    return o;
}

// ---------------------------------------------------------------- forward, unpacked
template <int MODEL, bool PUSH>
__global__ void __launch_bounds__(kThreads)
projection_fwd_kernel(int C, int N, const float *__restrict__ means, const float *__restrict__ covars,
                      const float *__restrict__ quats, const float *__restrict__ scales,
                      const float *__restrict__ viewmats, const float *__restrict__ Ks, ProjParams p,
                      int32_t *__restrict__ radii, float *__restrict__ means2d, float *__restrict__ depths,
                      float *__restrict__ conics, float *__restrict__ comps, peer::PushRows push) {
    gsb_chain_enter();
    int n = blockIdx.x * kThreads + threadIdx.x;
    if (n >= N) return;
    const float *mp = means + 3 * (size_t)n;
    float px = mp[0], py = mp[1], pz = mp[2];
    Sym3 S = load_covar(covars, quats, scales, n);
    for (int c = blockIdx.y; c < C; c += gridDim.y) {
        Camera cam = load_camera(viewmats, Ks, c);
        Proj2D o = project_one<MODEL>(cam, px, py, pz, S, p);
        size_t row = (size_t)c * N + n;
        bool v = o.radius > 0;
        radii[row] = o.radius;
        depths[row] = o.depth;
        *reinterpret_cast<float2 *>(means2d + 2 * row) = v ? make_float2(o.mx, o.my) : make_float2(0.f, 0.f);
        float *cp = conics + 3 * row;
        cp[0] = v ? o.cxx : 0.f;
        cp[1] = v ? o.cxy : 0.f;
        cp[2] = v ? o.cyy : 0.f;
        if (comps != nullptr) comps[row] = v ? o.comp : 0.f;
        if (PUSH) {
            // distributed, unpacked: the same row goes straight into the mailbox of the rank that owns camera c
            // (NVLink peer stores; consecutive lanes = consecutive rows: coalesced), no separate put kernel
            int64_t drow;
            float *dst = peer::push_row(push, c, n, drow);
            reinterpret_cast<int32_t *>(dst + push.off_radii)[drow] = o.radius;
            dst[push.off_depths + drow] = o.depth;
            *reinterpret_cast<float2 *>(dst + push.off_means2d + 2 * drow) =
                v ? make_float2(o.mx, o.my) : make_float2(0.f, 0.f);
            float *dc = dst + push.off_conics + 3 * drow;
            dc[0] = v ? o.cxx : 0.f;
            dc[1] = v ? o.cxy : 0.f;
            dc[2] = v ? o.cyy : 0.f;
            const float opac = push.opacities[n];
            dst[push.off_opacities + drow] = comps != nullptr ? opac * (v ? o.comp : 0.f) : opac;
        }
    }
    if (PUSH) __threadfence_system();  // this thread's peer stores are performed before the grid can be seen complete
}

// ---------------------------------------------------------------- backward pieces
struct GradIn {
    float v_mx, v_my, v_depth, v_cxx, v_cxy, v_cyy, v_comp;
    float cxx, cxy, cyy, comp;
};

struct GradCam {  // gradient w.r.t. camera-space quantities and viewmat
    float v_px, v_py, v_pz;  // world mean
    Sym3 vS;                 // world covariance, per-entry values of the full symmetric matrix
    float vR[9], vT[3];      // viewmat rotation / translation
};

template <int MODEL, bool NEED_VIEWMAT>
__device__ __forceinline__ void project_one_vjp(const Camera &c, float px, float py, float pz, const Sym3 &S,
                                                const ProjParams &p, bool calc_comp, const GradIn &g, GradCam &o) {
    // 1) conic -> blurred 2D covariance (d inverse), halves the off-diagonal: the external
    //    gradient of conic_xy is the sum of the xy and yx entries
    float vxx = g.v_cxx, vxy = 0.5f * g.v_cxy, vyy = g.v_cyy;
    float A = g.cxx, B = g.cxy, Cc = g.cyy;
    float d_xx = -(A * vxx * A + B * vyy * B + 2.0f * B * vxy * A);
    float d_xy = -(A * vxy * Cc + B * vxy * B + B * vyy * Cc + A * vxx * B);
    float d_yy = -(B * vxx * B + Cc * vyy * Cc + 2.0f * Cc * vxy * B);
    if (calc_comp) {
        float det_q = A * Cc - B * B;
        float d_a = 0.5f * g.v_comp / (g.comp + 1e-6f);
        float oma = 1.0f - g.comp * g.comp;
        d_xx += d_a * (oma * A - p.eps2d * det_q);
        d_xy += d_a * (oma * B);
        d_yy += d_a * (oma * Cc - p.eps2d * det_q);
    }
    // 2) recompute camera-space mean / covariance
    float tx = c.r00 * px + c.r01 * py + c.r02 * pz + c.t0;
    float ty = c.r10 * px + c.r11 * py + c.r12 * pz + c.t1;
    float tz = c.r20 * px + c.r21 * py + c.r22 * pz + c.t2;
    Sym3 Sc = rotate_covar(c, S);
    float v_tx, v_ty, v_tz;
    Sym3 vSc;
    if (MODEL == MODEL_PINHOLE) {
        PinholeJ j = pinhole_jacobian(c, tx, ty, tz, p.width, p.height);
        vSc.xx = j.jxx * d_xx * j.jxx;
        vSc.yy = j.jyy * d_yy * j.jyy;
        vSc.zz = j.jxz * d_xx * j.jxz + j.jyz * d_yy * j.jyz + 2.0f * j.jyz * d_xy * j.jxz;
        vSc.xy = j.jxx * d_xy * j.jyy;
        vSc.xz = j.jxx * d_xx * j.jxz + j.jxx * d_xy * j.jyz;
        vSc.yz = j.jyy * d_xy * j.jxz + j.jyy * d_yy * j.jyz;
        // mean path (exact derivative of fx*tx/tz + cx)
        float iz2 = j.iz * j.iz;
        v_tx = c.fx * j.iz * g.v_mx;
        v_ty = c.fy * j.iz * g.v_my;
        v_tz = -(c.fx * tx * g.v_mx + c.fy * ty * g.v_my) * iz2;
        // Jacobian path: v_J = 2 * d_cov2d * J * Sigma_c
        float jc_xx = j.jxx * Sc.xx + j.jxz * Sc.xz;
        float jc_xy = j.jxx * Sc.xy + j.jxz * Sc.yz;
        float jc_xz = j.jxx * Sc.xz + j.jxz * Sc.zz;
        float jc_yx = j.jyy * Sc.xy + j.jyz * Sc.xz;
        float jc_yy = j.jyy * Sc.yy + j.jyz * Sc.yz;
        float jc_yz = j.jyy * Sc.yz + j.jyz * Sc.zz;
        float v_jxx = 2.0f * (d_xx * jc_xx + d_xy * jc_yx);
        float v_jxz = 2.0f * (d_xx * jc_xz + d_xy * jc_yz);
        float v_jyy = 2.0f * (d_xy * jc_xy + d_yy * jc_yy);
        float v_jyz = 2.0f * (d_xy * jc_xz + d_yy * jc_yz);
        v_tx += j.clamp_x ? 0.0f : -v_jxz * c.fx * iz2;
        v_ty += j.clamp_y ? 0.0f : -v_jyz * c.fy * iz2;
        float tmp = v_jxx * j.jxx + v_jyy * j.jyy + 2.0f * (v_jxz * j.jxz + v_jyz * j.jyz);
        tmp -= j.clamp_x ? v_jxz * j.jxz : 0.0f;
        tmp -= j.clamp_y ? v_jyz * j.jyz : 0.0f;
        v_tz -= j.iz * tmp;
    } else if (MODEL == MODEL_FISHEYE) {
        const FisheyeJ j = fisheye_jacobian(c, tx, ty, tz);
        // v_Sigma_c = J^T d J (d symmetric)
        const float e00 = d_xx * j.j00 + d_xy * j.j10, e01 = d_xx * j.j01 + d_xy * j.j11, e02 = d_xx * j.j02 + d_xy * j.j12;
        const float e10 = d_xy * j.j00 + d_yy * j.j10, e11 = d_xy * j.j01 + d_yy * j.j11, e12 = d_xy * j.j02 + d_yy * j.j12;
        vSc.xx = j.j00 * e00 + j.j10 * e10;
        vSc.xy = j.j00 * e01 + j.j10 * e11;
        vSc.xz = j.j00 * e02 + j.j10 * e12;
        vSc.yy = j.j01 * e01 + j.j11 * e11;
        vSc.yz = j.j01 * e02 + j.j11 * e12;
        vSc.zz = j.j02 * e02 + j.j12 * e12;
        // mean path: v_t = J^T v_mean2d
        v_tx = j.j00 * g.v_mx + j.j10 * g.v_my;
        v_ty = j.j01 * g.v_mx + j.j11 * g.v_my;
        v_tz = j.j02 * g.v_mx + j.j12 * g.v_my;
        // Jacobian path: v_J = 2 d (J Sigma_c), then v_t += sum_ij v_J[i][j] dJ[i][j]/dt
        const float a0 = j.j00 * Sc.xx + j.j01 * Sc.xy + j.j02 * Sc.xz;
        const float a1 = j.j00 * Sc.xy + j.j01 * Sc.yy + j.j02 * Sc.yz;
        const float a2 = j.j00 * Sc.xz + j.j01 * Sc.yz + j.j02 * Sc.zz;
        const float b0 = j.j10 * Sc.xx + j.j11 * Sc.xy + j.j12 * Sc.xz;
        const float b1 = j.j10 * Sc.xy + j.j11 * Sc.yy + j.j12 * Sc.yz;
        const float b2 = j.j10 * Sc.xz + j.j11 * Sc.yz + j.j12 * Sc.zz;
        const float q00 = 2.0f * c.fx * (d_xx * a0 + d_xy * b0), q01 = 2.0f * c.fx * (d_xx * a1 + d_xy * b1);
        const float q02 = 2.0f * c.fx * (d_xx * a2 + d_xy * b2);  // v_J row 0, pre-multiplied by fx
        const float q10 = 2.0f * c.fy * (d_xy * a0 + d_yy * b0), q11 = 2.0f * c.fy * (d_xy * a1 + d_yy * b1);
        const float q12 = 2.0f * c.fy * (d_xy * a2 + d_yy * b2);  // v_J row 1, pre-multiplied by fy
        const float x = tx, y = ty, z = tz, h = j.h, m = j.m, k = j.k, k2 = 2.0f * j.k * j.k;
        const float xx = x * x, yy = y * y, xy = x * y;
        // d/dx, d/dy, d/dz of the unit-focal entries (g + x^2 h, x y h, -x k ; x y h, g + y^2 h, -y k)
        v_tx += q00 * (3.0f * x * h + xx * x * m) + (q01 + q10) * (y * h + xx * y * m) + q02 * (xx * k2 - k) +
                q11 * (x * h + x * yy * m) + q12 * (xy * k2);
        v_ty += q00 * (y * h + xx * y * m) + (q01 + q10) * (x * h + x * yy * m) + q02 * (xy * k2) +
                q11 * (3.0f * y * h + yy * y * m) + q12 * (yy * k2 - k);
        v_tz += q00 * (xx * k2 - k) + (q01 + q10) * (xy * k2) + q02 * (x * z * k2) + q11 * (yy * k2 - k) +
                q12 * (y * z * k2);
    } else {
        vSc.xx = c.fx * c.fx * d_xx;
        vSc.yy = c.fy * c.fy * d_yy;
        vSc.xy = c.fx * c.fy * d_xy;
        vSc.xz = vSc.yz = vSc.zz = 0.0f;
        v_tx = c.fx * g.v_mx;
        v_ty = c.fy * g.v_my;
        v_tz = 0.0f;
    }
    v_tz += g.v_depth;
    // 3) camera -> world
    o.v_px = c.r00 * v_tx + c.r10 * v_ty + c.r20 * v_tz;
    o.v_py = c.r01 * v_tx + c.r11 * v_ty + c.r21 * v_tz;
    o.v_pz = c.r02 * v_tx + c.r12 * v_ty + c.r22 * v_tz;
    // B = vSc * Rc (vSc as full symmetric matrix)
    float b00 = vSc.xx * c.r00 + vSc.xy * c.r10 + vSc.xz * c.r20;
    float b01 = vSc.xx * c.r01 + vSc.xy * c.r11 + vSc.xz * c.r21;
    float b02 = vSc.xx * c.r02 + vSc.xy * c.r12 + vSc.xz * c.r22;
    float b10 = vSc.xy * c.r00 + vSc.yy * c.r10 + vSc.yz * c.r20;
    float b11 = vSc.xy * c.r01 + vSc.yy * c.r11 + vSc.yz * c.r21;
    float b12 = vSc.xy * c.r02 + vSc.yy * c.r12 + vSc.yz * c.r22;
    float b20 = vSc.xz * c.r00 + vSc.yz * c.r10 + vSc.zz * c.r20;
    float b21 = vSc.xz * c.r01 + vSc.yz * c.r11 + vSc.zz * c.r21;
    float b22 = vSc.xz * c.r02 + vSc.yz * c.r12 + vSc.zz * c.r22;
    // vS = Rc^T * B
    o.vS.xx = c.r00 * b00 + c.r10 * b10 + c.r20 * b20;
    o.vS.xy = c.r00 * b01 + c.r10 * b11 + c.r20 * b21;
    o.vS.xz = c.r00 * b02 + c.r10 * b12 + c.r20 * b22;
    o.vS.yy = c.r01 * b01 + c.r11 * b11 + c.r21 * b21;
    o.vS.yz = c.r01 * b02 + c.r11 * b12 + c.r21 * b22;
    o.vS.zz = c.r02 * b02 + c.r12 * b12 + c.r22 * b22;
    if (NEED_VIEWMAT) {
        // v_Rc = v_t mu^T + 2 * (vSc Rc) Sigma ; v_T = v_t
        o.vR[0] = v_tx * px + 2.0f * (b00 * S.xx + b01 * S.xy + b02 * S.xz);
        o.vR[1] = v_tx * py + 2.0f * (b00 * S.xy + b01 * S.yy + b02 * S.yz);
        o.vR[2] = v_tx * pz + 2.0f * (b00 * S.xz + b01 * S.yz + b02 * S.zz);
        o.vR[3] = v_ty * px + 2.0f * (b10 * S.xx + b11 * S.xy + b12 * S.xz);
        o.vR[4] = v_ty * py + 2.0f * (b10 * S.xy + b11 * S.yy + b12 * S.yz);
        o.vR[5] = v_ty * pz + 2.0f * (b10 * S.xz + b11 * S.yz + b12 * S.zz);
        o.vR[6] = v_tz * px + 2.0f * (b20 * S.xx + b21 * S.xy + b22 * S.xz);
        o.vR[7] = v_tz * py + 2.0f * (b20 * S.xy + b21 * S.yy + b22 * S.yz);
        o.vR[8] = v_tz * pz + 2.0f * (b20 * S.xz + b21 * S.yz + b22 * S.zz);
        o.vT[0] = v_tx; o.vT[1] = v_ty; o.vT[2] = v_tz;
    }
}

// gradient of Sigma = (R(q) diag(s)) (..)^T w.r.t. q (unnormalised) and s
__device__ __forceinline__ void covar_vjp_quat_scale(float4 q, float s0, float s1, float s2, const Sym3 &G,
                                                     float vq[4], float vs[3]) {
    float R[9], inv_norm;
    quat_to_rotmat(q, R, inv_norm);
    float m00 = R[0] * s0, m01 = R[1] * s1, m02 = R[2] * s2;
    float m10 = R[3] * s0, m11 = R[4] * s1, m12 = R[5] * s2;
    float m20 = R[6] * s0, m21 = R[7] * s1, m22 = R[8] * s2;
    // v_M = 2 * G * M
    float vm00 = 2.f * (G.xx * m00 + G.xy * m10 + G.xz * m20);
    float vm01 = 2.f * (G.xx * m01 + G.xy * m11 + G.xz * m21);
    float vm02 = 2.f * (G.xx * m02 + G.xy * m12 + G.xz * m22);
    float vm10 = 2.f * (G.xy * m00 + G.yy * m10 + G.yz * m20);
    float vm11 = 2.f * (G.xy * m01 + G.yy * m11 + G.yz * m21);
    float vm12 = 2.f * (G.xy * m02 + G.yy * m12 + G.yz * m22);
    float vm20 = 2.f * (G.xz * m00 + G.yz * m10 + G.zz * m20);
    float vm21 = 2.f * (G.xz * m01 + G.yz * m11 + G.zz * m21);
    float vm22 = 2.f * (G.xz * m02 + G.yz * m12 + G.zz * m22);
    vs[0] = R[0] * vm00 + R[3] * vm10 + R[6] * vm20;
    vs[1] = R[1] * vm01 + R[4] * vm11 + R[7] * vm21;
    vs[2] = R[2] * vm02 + R[5] * vm12 + R[8] * vm22;
    float d00 = vm00 * s0, d01 = vm01 * s1, d02 = vm02 * s2;
    float d10 = vm10 * s0, d11 = vm11 * s1, d12 = vm12 * s2;
    float d20 = vm20 * s0, d21 = vm21 * s1, d22 = vm22 * s2;
    float w = q.x * inv_norm, x = q.y * inv_norm, y = q.z * inv_norm, z = q.w * inv_norm;
    float a = d21 - d12, b = d02 - d20, c = d10 - d01;  // antisymmetric parts
    float e = d01 + d10, f = d02 + d20, h = d12 + d21;  // symmetric parts
    float dw = 2.f * (x * a + y * b + z * c);
    float dx = 2.f * (-2.f * x * (d11 + d22) + y * e + z * f + w * a);
    float dy = 2.f * (x * e - 2.f * y * (d00 + d22) + z * h + w * b);
    float dz = 2.f * (x * f + y * h - 2.f * z * (d00 + d11) + w * c);
    float dot = w * dw + x * dx + y * dy + z * dz;
    vq[0] = (dw - w * dot) * inv_norm;
    vq[1] = (dx - x * dot) * inv_norm;
    vq[2] = (dy - y * dot) * inv_norm;
    vq[3] = (dz - z * dot) * inv_norm;
}

__device__ __forceinline__ void sym_add(Sym3 &a, const Sym3 &b) {
    a.xx += b.xx; a.xy += b.xy; a.xz += b.xz; a.yy += b.yy; a.yz += b.yz; a.zz += b.zz;
}

// block-level reduction of the 12 viewmat gradient values -> one atomic per value per block
__device__ __forceinline__ void reduce_viewmat_grad(float vals[12], float *smem12, float *__restrict__ v_viewmat) {
    if (threadIdx.x < 12) smem12[threadIdx.x] = 0.f;
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 12; ++i) {
        float s = warp_sum(vals[i]);
        if ((threadIdx.x & 31) == 0 && s != 0.f) atomicAdd(smem12 + i, s);
    }
    __syncthreads();
    if (threadIdx.x < 12) {
        // rows 0..2 of the 4x4: [R | T]
        int r = threadIdx.x < 9 ? threadIdx.x / 3 : threadIdx.x - 9;
        int col = threadIdx.x < 9 ? threadIdx.x % 3 : 3;
        float s = smem12[threadIdx.x];
        if (s != 0.f) atomicAdd(v_viewmat + 4 * r + col, s);
    }
    __syncthreads();
}

// writes the per-Gaussian outputs from the accumulated world-space gradients
__device__ __forceinline__ void finalize_gaussian_grads(int n, float vpx, float vpy, float vpz, const Sym3 &G,
                                                        const float *__restrict__ quats,
                                                        const float *__restrict__ scales, float *__restrict__ v_means,
                                                        float *__restrict__ v_covars, float *__restrict__ v_quats,
                                                        float *__restrict__ v_scales) {
    if (v_means != nullptr) {
        float *o = v_means + 3 * (size_t)n;
        o[0] = vpx; o[1] = vpy; o[2] = vpz;
    }
    if (v_covars != nullptr) {
        float *o = v_covars + 6 * (size_t)n;
        o[0] = G.xx; o[1] = 2.f * G.xy; o[2] = 2.f * G.xz; o[3] = G.yy; o[4] = 2.f * G.yz; o[5] = G.zz;
    }
    if (v_quats != nullptr || v_scales != nullptr) {
        float4 q = *reinterpret_cast<const float4 *>(quats + 4 * (size_t)n);
        const float *sp = scales + 3 * (size_t)n;
        float vq[4], vs[3];
        covar_vjp_quat_scale(q, sp[0], sp[1], sp[2], G, vq, vs);
        if (v_quats != nullptr) *reinterpret_cast<float4 *>(v_quats + 4 * (size_t)n) = make_float4(vq[0], vq[1], vq[2], vq[3]);
        if (v_scales != nullptr) {
            float *o = v_scales + 3 * (size_t)n;
            o[0] = vs[0]; o[1] = vs[1]; o[2] = vs[2];
        }
    }
}

// ---------------------------------------------------------------- backward, unpacked
template <int MODEL, bool NEED_VIEWMAT>
__global__ void __launch_bounds__(kThreads)
projection_bwd_kernel(int C, int N, const float *__restrict__ means, const float *__restrict__ covars,
                      const float *__restrict__ quats, const float *__restrict__ scales,
                      const float *__restrict__ viewmats, const float *__restrict__ Ks, ProjParams p,
                      const int32_t *__restrict__ radii, const float *__restrict__ conics,
                      const float *__restrict__ comps, const float *__restrict__ v_means2d,
                      const float *__restrict__ v_depths, const float *__restrict__ v_conics,
                      const float *__restrict__ v_comps, float *__restrict__ v_means, float *__restrict__ v_covars,
                      float *__restrict__ v_quats, float *__restrict__ v_scales, float *__restrict__ v_viewmats) {
    gsb_chain_enter();
    __shared__ float s_vm[12];
    int n = blockIdx.x * kThreads + threadIdx.x;
    bool in_range = n < N;
    float px = 0.f, py = 0.f, pz = 0.f;
    Sym3 S = {1.f, 0.f, 0.f, 1.f, 0.f, 1.f};
    if (in_range) {
        const float *mp = means + 3 * (size_t)n;
        px = mp[0]; py = mp[1]; pz = mp[2];
        S = load_covar(covars, quats, scales, n);
    }
    float vpx = 0.f, vpy = 0.f, vpz = 0.f;
    Sym3 G = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    const bool calc_comp = (comps != nullptr) && (v_comps != nullptr);
    for (int c = 0; c < C; ++c) {
        size_t row = (size_t)c * N + n;
        bool valid = in_range && radii[row] > 0;
        float vm[12];
#pragma unroll
        for (int i = 0; i < 12; ++i) vm[i] = 0.f;
        if (valid) {
            Camera cam = load_camera(viewmats, Ks, c);
            GradIn g;
            float2 vm2 = *reinterpret_cast<const float2 *>(v_means2d + 2 * row);
            g.v_mx = vm2.x; g.v_my = vm2.y;
            g.v_depth = v_depths != nullptr ? v_depths[row] : 0.f;
            const float *vc = v_conics + 3 * row;
            g.v_cxx = vc[0]; g.v_cxy = vc[1]; g.v_cyy = vc[2];
            const float *cc = conics + 3 * row;
            g.cxx = cc[0]; g.cxy = cc[1]; g.cyy = cc[2];
            g.comp = calc_comp ? comps[row] : 1.f;
            g.v_comp = calc_comp ? v_comps[row] : 0.f;
            GradCam o;
            project_one_vjp<MODEL, NEED_VIEWMAT>(cam, px, py, pz, S, p, calc_comp, g, o);
            vpx += o.v_px; vpy += o.v_py; vpz += o.v_pz;
            sym_add(G, o.vS);
            if (NEED_VIEWMAT) {
#pragma unroll
                for (int i = 0; i < 9; ++i) vm[i] = o.vR[i];
                vm[9] = o.vT[0]; vm[10] = o.vT[1]; vm[11] = o.vT[2];
            }
        }
        if (NEED_VIEWMAT) reduce_viewmat_grad(vm, s_vm, v_viewmats + 16 * (size_t)c);
    }
    if (in_range) finalize_gaussian_grads(n, vpx, vpy, vpz, G, quats, scales, v_means, v_covars, v_quats, v_scales);
}

// ---------------------------------------------------------------- packed forward
// pass 1: number of visible (camera, Gaussian) pairs per block; blocks are camera-major so the
// compacted order is row-major (camera, then Gaussian) like the reference (Packed.cu:195-233)
template <int MODEL>
__global__ void __launch_bounds__(kThreads)
projection_packed_count_kernel(int C, int N, const float *__restrict__ means, const float *__restrict__ covars,
                               const float *__restrict__ quats, const float *__restrict__ scales,
                               const float *__restrict__ viewmats, const float *__restrict__ Ks, ProjParams p,
                               int32_t *__restrict__ block_counts) {
    int n = blockIdx.x * kThreads + threadIdx.x;
    int c = blockIdx.y;
    bool valid = false;
    if (n < N) {
        const float *mp = means + 3 * (size_t)n;
        Sym3 S = load_covar(covars, quats, scales, n);
        Camera cam = load_camera(viewmats, Ks, c);
        Proj2D o = project_one<MODEL>(cam, mp[0], mp[1], mp[2], S, p);
        valid = o.radius > 0;
    }
    int cnt = __syncthreads_count(valid);
    if (threadIdx.x == 0) block_counts[(size_t)c * gridDim.x + blockIdx.x] = cnt;
}

template <int MODEL>
__global__ void __launch_bounds__(kThreads)
projection_packed_write_kernel(int C, int N, const float *__restrict__ means, const float *__restrict__ covars,
                               const float *__restrict__ quats, const float *__restrict__ scales,
                               const float *__restrict__ viewmats, const float *__restrict__ Ks, ProjParams p,
                               const int64_t *__restrict__ block_offsets, int64_t *__restrict__ camera_ids,
                               int64_t *__restrict__ gaussian_ids, int32_t *__restrict__ radii,
                               float *__restrict__ means2d, float *__restrict__ depths, float *__restrict__ conics,
                               float *__restrict__ comps) {
    __shared__ int s_warp[kThreads / 32];
    int n = blockIdx.x * kThreads + threadIdx.x;
    int c = blockIdx.y;
    Proj2D o;
    o.radius = 0;
    if (n < N) {
        const float *mp = means + 3 * (size_t)n;
        Sym3 S = load_covar(covars, quats, scales, n);
        Camera cam = load_camera(viewmats, Ks, c);
        o = project_one<MODEL>(cam, mp[0], mp[1], mp[2], S, p);
    }
    bool valid = o.radius > 0;
    unsigned ballot = __ballot_sync(0xffffffffu, valid);
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) s_warp[warp] = __popc(ballot);
    __syncthreads();
    int base = 0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) base += (w < warp) ? s_warp[w] : 0;
    if (!valid) return;
    int64_t row = block_offsets[(size_t)c * gridDim.x + blockIdx.x] + base + __popc(ballot & ((1u << lane) - 1u));
    camera_ids[row] = c;
    gaussian_ids[row] = n;
    radii[row] = o.radius;
    depths[row] = o.depth;
    *reinterpret_cast<float2 *>(means2d + 2 * row) = make_float2(o.mx, o.my);
    float *cp = conics + 3 * row;
    cp[0] = o.cxx; cp[1] = o.cxy; cp[2] = o.cyy;
    if (comps != nullptr) comps[row] = o.comp;
}

// ---------------------------------------------------------------- packed backward
// One thread per packed row. dense mode: atomics into acc[N,9] = (v_mean 3, v_Sigma 6) followed by
// projection_finalize_kernel; sparse mode: per-row gradients [nnz, 3|4|3|6] with plain stores.
template <int MODEL, bool NEED_VIEWMAT, bool SPARSE>
__global__ void __launch_bounds__(kThreads)
projection_packed_bwd_kernel(int64_t nnz, const float *__restrict__ means, const float *__restrict__ covars,
                             const float *__restrict__ quats, const float *__restrict__ scales,
                             const float *__restrict__ viewmats, const float *__restrict__ Ks, ProjParams p,
                             const int64_t *__restrict__ camera_ids, const int64_t *__restrict__ gaussian_ids,
                             const float *__restrict__ conics, const float *__restrict__ comps,
                             const float *__restrict__ v_means2d, const float *__restrict__ v_depths,
                             const float *__restrict__ v_conics, const float *__restrict__ v_comps,
                             float *__restrict__ acc, float *__restrict__ v_means, float *__restrict__ v_covars,
                             float *__restrict__ v_quats, float *__restrict__ v_scales,
                             float *__restrict__ v_viewmats) {
    int64_t row = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    bool valid = row < nnz;
    const bool calc_comp = (comps != nullptr) && (v_comps != nullptr);
    GradCam o;
    int c = 0;
    if (valid) {
        c = (int)camera_ids[row];
        int n = (int)gaussian_ids[row];
        const float *mp = means + 3 * (size_t)n;
        float px = mp[0], py = mp[1], pz = mp[2];
        Sym3 S = load_covar(covars, quats, scales, n);
        Camera cam = load_camera(viewmats, Ks, c);
        GradIn g;
        float2 vm2 = *reinterpret_cast<const float2 *>(v_means2d + 2 * row);
        g.v_mx = vm2.x; g.v_my = vm2.y;
        g.v_depth = v_depths != nullptr ? v_depths[row] : 0.f;
        const float *vc = v_conics + 3 * row;
        g.v_cxx = vc[0]; g.v_cxy = vc[1]; g.v_cyy = vc[2];
        const float *cc = conics + 3 * row;
        g.cxx = cc[0]; g.cxy = cc[1]; g.cyy = cc[2];
        g.comp = calc_comp ? comps[row] : 1.f;
        g.v_comp = calc_comp ? v_comps[row] : 0.f;
        project_one_vjp<MODEL, NEED_VIEWMAT>(cam, px, py, pz, S, p, calc_comp, g, o);
        if (SPARSE) {
            finalize_gaussian_grads((int)0, 0.f, 0.f, 0.f, o.vS, quats + 4 * (size_t)n - 0, scales + 3 * (size_t)n - 0,
                                    nullptr, v_covars ? v_covars + 6 * row : nullptr,
                                    v_quats ? v_quats + 4 * row : nullptr, v_scales ? v_scales + 3 * row : nullptr);
            if (v_means != nullptr) {
                float *om = v_means + 3 * row;
                om[0] = o.v_px; om[1] = o.v_py; om[2] = o.v_pz;
            }
        } else {
            float *a = acc + 9 * (size_t)n;
            atomicAdd(a + 0, o.v_px); atomicAdd(a + 1, o.v_py); atomicAdd(a + 2, o.v_pz);
            atomicAdd(a + 3, o.vS.xx); atomicAdd(a + 4, o.vS.xy); atomicAdd(a + 5, o.vS.xz);
            atomicAdd(a + 6, o.vS.yy); atomicAdd(a + 7, o.vS.yz); atomicAdd(a + 8, o.vS.zz);
        }
    }
    if (NEED_VIEWMAT) {
        // rows are camera-major, so a warp almost always holds a single camera: reduce over the
        // lanes that share lane 0's camera, the stragglers fall back to per-lane atomics
        int c0 = __shfl_sync(0xffffffffu, c, 0);
        bool same = valid && (c == c0);
#pragma unroll
        for (int i = 0; i < 12; ++i) {
            float v = i < 9 ? o.vR[i < 9 ? i : 0] : o.vT[i >= 9 ? i - 9 : 0];
            int r = i < 9 ? i / 3 : i - 9, col = i < 9 ? i % 3 : 3;
            float s = warp_sum(same ? v : 0.f);
            if ((threadIdx.x & 31) == 0 && s != 0.f) atomicAdd(v_viewmats + 16 * (size_t)c0 + 4 * r + col, s);
            if (valid && !same) atomicAdd(v_viewmats + 16 * (size_t)c + 4 * r + col, v);
        }
    }
}

__global__ void __launch_bounds__(kThreads)
projection_finalize_kernel(int N, const float *__restrict__ acc, const float *__restrict__ quats,
                           const float *__restrict__ scales, float *__restrict__ v_means,
                           float *__restrict__ v_covars, float *__restrict__ v_quats, float *__restrict__ v_scales) {
    int n = blockIdx.x * kThreads + threadIdx.x;
    if (n >= N) return;
    const float *a = acc + 9 * (size_t)n;
    Sym3 G = {a[3], a[4], a[5], a[6], a[7], a[8]};
    finalize_gaussian_grads(n, a[0], a[1], a[2], G, quats, scales, v_means, v_covars, v_quats, v_scales);
}

int check_model(int camera_model) {
    if (camera_model != MODEL_PINHOLE && camera_model != MODEL_ORTHO && camera_model != MODEL_FISHEYE) {
        gsb_set_error("camera_model %d is not supported by the b200 backend (0=pinhole, 1=ortho, 2=fisheye)",
                      camera_model);
        return GSB_ERR_UNSUPPORTED;
    }
    return GSB_OK;
}

}  // namespace

#define DISPATCH_MODEL(model, ...)                         \
    if ((model) == MODEL_PINHOLE) {                        \
        constexpr int M = MODEL_PINHOLE;                   \
        __VA_ARGS__;                                       \
    } else if ((model) == MODEL_ORTHO) {                   \
        constexpr int M = MODEL_ORTHO;                     \
        __VA_ARGS__;                                       \
    } else {                                               \
        constexpr int M = MODEL_FISHEYE;                   \
        __VA_ARGS__;                                       \
    }

namespace {
int projection_fwd_launch(int C, int N, const float *means, const float *covars, const float *quats,
                          const float *scales, const float *viewmats, const float *Ks, int width, int height,
                          float eps2d, float near_plane, float far_plane, float radius_clip, int camera_model,
                          int32_t *radii, float *means2d, float *depths, float *conics, float *compensations,
                          const peer::PushRows *push, cudaStream_t stream) {
    GSB_REQUIRE(C >= 0 && N >= 0, "negative size");
    GSB_REQUIRE(covars != nullptr || (quats != nullptr && scales != nullptr), "covars or (quats, scales) required");
    if (int e = check_model(camera_model)) return e;
    if (C == 0 || N == 0) return GSB_OK;
    ProjParams p = {width, height, eps2d, near_plane, far_plane, radius_clip};
    unsigned nblk = gsb_cdiv(N, kThreads);
    // split cameras over grid.y only when there are too few Gaussian blocks to fill the 148 SMs
    unsigned gy = 1;
    if (nblk < 148u * 4u) gy = (unsigned)min((long long)C, (long long)gsb_cdiv(148 * 4, nblk));
    dim3 grid(nblk, gy);
    if (push != nullptr) {
        DISPATCH_MODEL(camera_model, (gsb_launch_chained(GSB_CHAIN_PROJECTION, projection_fwd_kernel<M, true>, grid, kThreads, 0, stream, C, N,
                                                         means, covars, quats, scales, viewmats, Ks, p, radii, means2d,
                                                         depths, conics, compensations, *push)));
    } else {
        peer::PushRows none = {};
        DISPATCH_MODEL(camera_model, (gsb_launch_chained(GSB_CHAIN_PROJECTION, projection_fwd_kernel<M, false>, grid, kThreads, 0, stream, C, N,
                                                         means, covars, quats, scales, viewmats, Ks, p, radii, means2d,
                                                         depths, conics, compensations, none)));
    }
    return gsb_check_launch("projection_fwd");
}
}  // namespace

extern "C" int gsb_projection_fwd(int C, int N, const float *means, const float *covars, const float *quats,
                                  const float *scales, const float *viewmats, const float *Ks, int width, int height,
                                  float eps2d, float near_plane, float far_plane, float radius_clip,
                                  int camera_model, int32_t *radii, float *means2d, float *depths, float *conics,
                                  float *compensations, cudaStream_t stream) {
    return projection_fwd_launch(C, N, means, covars, quats, scales, viewmats, Ks, width, height, eps2d, near_plane,
                                 far_plane, radius_clip, camera_model, radii, means2d, depths, conics, compensations,
                                 nullptr, stream);
}

// peer.cu
int gsb_fill_push_rows(peer::PushRows &push, int world, int C_local, int64_t N_total, int64_t n_base, int D,
                       const void *const *peer_bases_host, int64_t region_offset, const float *opacities);

extern "C" int gsb_projection_fwd_push(int C, int N, const float *means, const float *covars, const float *quats,
                                       const float *scales, const float *viewmats, const float *Ks, int width,
                                       int height, float eps2d, float near_plane, float far_plane, float radius_clip,
                                       int camera_model, int32_t *radii, float *means2d, float *depths, float *conics,
                                       float *compensations, const float *opacities, int world, int C_local,
                                       int64_t N_total, int64_t n_base, int D, const void *const *peer_bases_host,
                                       int64_t region_offset, cudaStream_t stream) {
    GSB_REQUIRE(opacities != nullptr && world >= 1 && C_local >= 1 && C == world * C_local, "bad push geometry");
    GSB_REQUIRE(n_base >= 0 && n_base + N <= N_total, "rows do not fit the destination table");
    peer::PushRows push;
    if (int e = gsb_fill_push_rows(push, world, C_local, N_total, n_base, D, peer_bases_host, region_offset, opacities))
        return e;
    return projection_fwd_launch(C, N, means, covars, quats, scales, viewmats, Ks, width, height, eps2d, near_plane,
                                 far_plane, radius_clip, camera_model, radii, means2d, depths, conics, compensations,
                                 &push, stream);
}

extern "C" int gsb_projection_bwd(int C, int N, const float *means, const float *covars, const float *quats,
                                  const float *scales, const float *viewmats, const float *Ks, int width, int height,
                                  float eps2d, int camera_model, const int32_t *radii, const float *conics,
                                  const float *compensations, const float *v_means2d, const float *v_depths,
                                  const float *v_conics, const float *v_compensations, float *v_means,
                                  float *v_covars, float *v_quats, float *v_scales, float *v_viewmats,
                                  cudaStream_t stream) {
    GSB_REQUIRE(C >= 0 && N >= 0, "negative size");
    if (int e = check_model(camera_model)) return e;
    if (v_viewmats != nullptr && C > 0) {
        cudaError_t err = cudaMemsetAsync(v_viewmats, 0, sizeof(float) * 16 * (size_t)C, stream);
        if (err != cudaSuccess) { gsb_set_error("memset v_viewmats: %s", cudaGetErrorString(err)); return GSB_ERR_CUDA; }
    }
    if (N == 0) return GSB_OK;
    ProjParams p = {width, height, eps2d, 0.f, 0.f, 0.f};
    unsigned nblk = gsb_cdiv(N, kThreads);
    if (v_viewmats != nullptr) {
        DISPATCH_MODEL(camera_model, (gsb_launch_chained(GSB_CHAIN_PROJECTION, projection_bwd_kernel<M, true>, nblk, kThreads, 0, stream, C, N,
                                                         means, covars, quats, scales, viewmats, Ks, p, radii, conics,
                                                         compensations, v_means2d, v_depths, v_conics, v_compensations,
                                                         v_means, v_covars, v_quats, v_scales, v_viewmats)));
    } else {
        DISPATCH_MODEL(camera_model, (gsb_launch_chained(GSB_CHAIN_PROJECTION, projection_bwd_kernel<M, false>, nblk, kThreads, 0, stream, C, N,
                                                         means, covars, quats, scales, viewmats, Ks, p, radii, conics,
                                                         compensations, v_means2d, v_depths, v_conics, v_compensations,
                                                         v_means, v_covars, v_quats, v_scales, v_viewmats)));
    }
    return gsb_check_launch("projection_bwd");
}

extern "C" int gsb_projection_packed_count(int C, int N, const float *means, const float *covars, const float *quats,
                                           const float *scales, const float *viewmats, const float *Ks, int width,
                                           int height, float eps2d, float near_plane, float far_plane,
                                           float radius_clip, int camera_model, int32_t *block_counts,
                                           cudaStream_t stream) {
    GSB_REQUIRE(C >= 0 && N >= 0, "negative size");
    if (int e = check_model(camera_model)) return e;
    if (C == 0 || N == 0) return GSB_OK;
    ProjParams p = {width, height, eps2d, near_plane, far_plane, radius_clip};
    dim3 grid(gsb_cdiv(N, kThreads), C);
    DISPATCH_MODEL(camera_model, (projection_packed_count_kernel<M><<<grid, kThreads, 0, stream>>>(
                                     C, N, means, covars, quats, scales, viewmats, Ks, p, block_counts)));
    return gsb_check_launch("projection_packed_count");
}

extern "C" int gsb_projection_packed_write(int C, int N, const float *means, const float *covars, const float *quats,
                                           const float *scales, const float *viewmats, const float *Ks, int width,
                                           int height, float eps2d, float near_plane, float far_plane,
                                           float radius_clip, int camera_model, const int64_t *block_offsets,
                                           int64_t *camera_ids, int64_t *gaussian_ids, int32_t *radii,
                                           float *means2d, float *depths, float *conics, float *compensations,
                                           cudaStream_t stream) {
    GSB_REQUIRE(C >= 0 && N >= 0, "negative size");
    if (int e = check_model(camera_model)) return e;
    if (C == 0 || N == 0) return GSB_OK;
    ProjParams p = {width, height, eps2d, near_plane, far_plane, radius_clip};
    dim3 grid(gsb_cdiv(N, kThreads), C);
    DISPATCH_MODEL(camera_model, (projection_packed_write_kernel<M><<<grid, kThreads, 0, stream>>>(
                                     C, N, means, covars, quats, scales, viewmats, Ks, p, block_offsets, camera_ids,
                                     gaussian_ids, radii, means2d, depths, conics, compensations)));
    return gsb_check_launch("projection_packed_write");
}

extern "C" int gsb_projection_packed_bwd(int C, int N, int64_t nnz, const float *means, const float *covars,
                                         const float *quats, const float *scales, const float *viewmats,
                                         const float *Ks, int width, int height, float eps2d, int camera_model,
                                         const int64_t *camera_ids, const int64_t *gaussian_ids, const float *conics,
                                         const float *compensations, const float *v_means2d, const float *v_depths,
                                         const float *v_conics, const float *v_compensations, int sparse_grad,
                                         float *acc, float *v_means, float *v_covars, float *v_quats, float *v_scales,
                                         float *v_viewmats, cudaStream_t stream) {
    GSB_REQUIRE(C >= 0 && N >= 0 && nnz >= 0, "negative size");
    GSB_REQUIRE(sparse_grad || acc != nullptr, "dense packed backward needs the acc[N,9] scratch buffer");
    if (int e = check_model(camera_model)) return e;
    cudaError_t err = cudaSuccess;
    if (v_viewmats != nullptr && C > 0) err = cudaMemsetAsync(v_viewmats, 0, sizeof(float) * 16 * (size_t)C, stream);
    if (err == cudaSuccess && !sparse_grad && N > 0) err = cudaMemsetAsync(acc, 0, sizeof(float) * 9 * (size_t)N, stream);
    if (err != cudaSuccess) { gsb_set_error("memset: %s", cudaGetErrorString(err)); return GSB_ERR_CUDA; }
    ProjParams p = {width, height, eps2d, 0.f, 0.f, 0.f};
    if (nnz > 0) {
        unsigned nblk = gsb_cdiv(nnz, kThreads);
#define LAUNCH_PACKED_BWD(VM, SP)                                                                              \
    DISPATCH_MODEL(camera_model, (projection_packed_bwd_kernel<M, VM, SP><<<nblk, kThreads, 0, stream>>>(       \
                                     nnz, means, covars, quats, scales, viewmats, Ks, p, camera_ids,            \
                                     gaussian_ids, conics, compensations, v_means2d, v_depths, v_conics,        \
                                     v_compensations, acc, v_means, v_covars, v_quats, v_scales, v_viewmats)))
        if (v_viewmats != nullptr) {
            if (sparse_grad) { LAUNCH_PACKED_BWD(true, true); } else { LAUNCH_PACKED_BWD(true, false); }
        } else {
            if (sparse_grad) { LAUNCH_PACKED_BWD(false, true); } else { LAUNCH_PACKED_BWD(false, false); }
        }
#undef LAUNCH_PACKED_BWD
        if (int e = gsb_check_launch("projection_packed_bwd")) return e;
    }
    if (!sparse_grad && N > 0) {
        projection_finalize_kernel<<<gsb_cdiv(N, kThreads), kThreads, 0, stream>>>(N, acc, quats, scales, v_means,
                                                                                  v_covars, v_quats, v_scales);
        if (int e = gsb_check_launch("projection_finalize")) return e;
    }
    return GSB_OK;
}
