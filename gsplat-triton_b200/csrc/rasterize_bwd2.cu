// rasterize_to_pixels backward, second generation ("v2"): half-warp per 8x4 pixel block, two pixels per lane, packed
// FP32 (FFMA2 / FMUL2 / FADD2). Same semantics and the same per-pair arithmetic as rasterize_bwd.cu (every value that
// phase 1 produces is bit-identical; only the order of the sums over pixels differs), D <= 4.
//
// Why (ncu on v1, cfg2: 68 % issue-active, short-scoreboard + barrier stalls next; the forward kernel's move to this
// layout cut its instruction count per evaluation by a third and its time by 13 %):
//   * phase 1 (lane = two vertically adjacent pixels, sequential over the half-warp's splats): list bookkeeping, the
//     three 16-byte staged loads, the exponent, T / u recurrences and both weights are issued once per two pixels and
//     per TWO splats (the two halves of a warp walk the lists of two vertically adjacent blocks);
//   * phase 2 (every 8 splats of a half-warp; lane = splat x half of the block): a lane sums 8 table entries of
//     two pixels each with packed instructions, one shuffle stage joins the two halves of the block;
//   * CTA = 128 threads, 27 KB of shared memory: occupancy is set by registers, not by the table. (A 16-splat
//     chunk with one lane per splat needs no shuffle stage and ~7 fewer instructions per splat, but its 42 KB cap
//     the SM at 5 CTAs = 20 warps: ncu showed 54 % issue-active, one eligible warp per scheduler and cycle.)
// Phase-1 table: per half-warp 8 rows (splats) x 16 float4 (lane: v_sigma of its two pixels, -alpha T of its two
// pixels), the float4 column XOR-swizzled by the row so that phase 1 (a row at a time) and phase 2 (a column walk by
// 8 lanes on 8 different rows) are both bank-conflict free without padding.
#include "rasterize_common.cuh"

namespace raster {

constexpr int kB2Threads = 128;
constexpr int kB2Warps = kB2Threads / 32;
constexpr int kB2Halves = 2 * kB2Warps;
constexpr int kB2Chunk = 8;

#ifndef GSB_BWD2_MIN_BLOCKS
#define GSB_BWD2_MIN_BLOCKS 7
#endif

struct __align__(256) Bwd2Smem {
    float4 Wt[kB2Halves][kB2Chunk][16];        // phase-1 table (swizzled columns); rows are 256-byte aligned, so
                                               // the swizzle is an XOR on address bits 4-6
    float4 V[kB2Halves][16][2];                // colour cotangents of the half's pixels: (v0[0],v1[0],v0[1],v1[1]), (v0[2],v1[2],v0[3],v1[3])
    float4 P[kB2Threads + 1];                  // entry 0 = zero pad; x, y, sorted index, block mask
    float4 B[kB2Threads + 1];                  // staged conic (A', B', C'), MINUS opacity
    float4 Cq[kB2Threads + 1];                 // colour
    int wcnt[kB2Warps];
    int red[kB2Halves];
};

// hit / clamp decisions of one (pixel, Gaussian) pair for the backward pass, in predicate form:
//   hit = index <= last_id && !(ex > 0) && !(minus_alpha > -1/255)
//   minus_alpha_h = hit ? minus_alpha : 0 ; minus_ov_h = (hit && minus_ov >= -0.999) ? minus_ov : 0   (no gradient
//   through the 0.999 clamp). Returns hit.
__device__ __forceinline__ bool bwd_decide(int index, int last_id, float ex, float minus_ov, float minus_alpha,
                                           float &minus_alpha_h, float &minus_ov_h) {
    int hit;
    asm("{\n\t"
        ".reg .pred h, g;\n\t"
        "setp.le.s32 h, %3, %4;\n\t"
        "setp.leu.and.ftz.f32 h, %5, 0f00000000, h;\n\t"
        "setp.leu.and.ftz.f32 h, %7, 0fBB808081, h;\n\t"
        "setp.geu.and.ftz.f32 g, %6, 0fBF7FBE77, h;\n\t"
        "selp.f32 %0, %7, 0f00000000, h;\n\t"
        "selp.f32 %1, %6, 0f00000000, g;\n\t"
        "selp.s32 %2, 1, 0, h;\n\t"
        "}"
        : "=f"(minus_alpha_h), "=f"(minus_ov_h), "=r"(hit)
        : "r"(index), "r"(last_id), "f"(ex), "f"(minus_ov), "f"(minus_alpha));
    return hit != 0;
}

// Phase 2 for the (up to 8) splats each half-warp has parked in its table: lane (s = l16 & 7, q = l16 >> 3) sums, for
// splat s, the 8 table entries of block rows 2q and 2q+1; the two q of a splat are joined by one shuffle stage and
// lane q = 0 issues the global reductions.
template <int D, bool ABS>
__device__ __forceinline__ void reduce_chunk2(Bwd2Smem &sm, int hw, int l16, int count, int entry, float bcx, float bcy,
                                              const int32_t *__restrict__ flatten_ids, const GradTargets &out) {
    __syncwarp();
    const int s_ = l16 & 7, q_ = l16 >> 3;
    const bool valid = s_ < count;
    const int e = valid ? entry : 0;
    const float4 mu = sm.P[e];
    const float4 sb = sm.B[e];
    const float ca = sb.x * kUnscaleDiag, cb = sb.y * kUnscaleOff, cc = sb.z * kUnscaleDiag, opac = -sb.w;
    const float ux = mu.x - bcx, uy = mu.y - bcy;
    f32x2 R0 = pack2(0.f, 0.f), Rx = R0, Rxx = R0;  // (row 2q, row 2q + 1)
    f32x2 col[D];
#pragma unroll
    for (int k = 0; k < D; ++k) col[k] = pack2(0.f, 0.f);
    float abs_x = 0.f, abs_y = 0.f;
    const float4 *row = sm.Wt[hw][s_];
    const float4 *vrow = &sm.V[hw][8 * q_][0];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float4 w = row[(8 * q_ + i) ^ s_];  // (v_sigma pixel 0, v_sigma pixel 1, -alpha T pixel 0, -alpha T pixel 1)
        const float4 v0 = vrow[2 * i];
        const f32x2 p = pack2(w.x, w.y), q = pack2(w.z, w.w);
        const float cx = (float)i - 3.5f;
        R0 = add2(R0, p);
        Rx = fma2(p, bcast2(cx), Rx);
        Rxx = fma2(p, bcast2(cx * cx), Rxx);
        col[0] = fma2(q, pack2(v0.x, v0.y), col[0]);
        if (D > 1) col[1] = fma2(q, pack2(v0.z, v0.w), col[1]);
        if (D > 2) {
            const float4 v1 = vrow[2 * i + 1];
            col[2] = fma2(q, pack2(v1.x, v1.y), col[2]);
            if (D > 3) col[3] = fma2(q, pack2(v1.z, v1.w), col[3]);
        }
        if (ABS) {
            const float dx = ux - cx, dy0 = uy - ((float)(2 * q_) - 1.5f), dy1 = dy0 - 1.0f;
            abs_x += fabsf(w.x * (ca * dx + cb * dy0)) + fabsf(w.y * (ca * dx + cb * dy1));
            abs_y += fabsf(w.x * (cb * dx + cc * dy0)) + fabsf(w.y * (cb * dx + cc * dy1));
        }
    }
    // moments of v_sigma about the block centre over this lane's two rows (centres at y0, y0 + 1), then over the block
    float a0, a1, x0, x1, q0, q1;
    unpack2(R0, a0, a1);
    unpack2(Rx, x0, x1);
    unpack2(Rxx, q0, q1);
    const float y0 = (float)(2 * q_) - 1.5f, y1 = y0 + 1.0f;
    float M0 = a0 + a1, Mx = x0 + x1, Mxx = q0 + q1;
    float My = y0 * a0 + y1 * a1, Mxy = y0 * x0 + y1 * x1, Myy = y0 * y0 * a0 + y1 * y1 * a1;
    M0 += __shfl_xor_sync(0xffffffffu, M0, 8);
    Mx += __shfl_xor_sync(0xffffffffu, Mx, 8);
    My += __shfl_xor_sync(0xffffffffu, My, 8);
    Mxx += __shfl_xor_sync(0xffffffffu, Mxx, 8);
    Mxy += __shfl_xor_sync(0xffffffffu, Mxy, 8);
    Myy += __shfl_xor_sync(0xffffffffu, Myy, 8);
    float cs[D];
#pragma unroll
    for (int k = 0; k < D; ++k) {
        float lo, hi;
        unpack2(col[k], lo, hi);
        cs[k] = lo + hi;
        cs[k] += __shfl_xor_sync(0xffffffffu, cs[k], 8);
    }
    if (ABS) {
        abs_x += __shfl_xor_sync(0xffffffffu, abs_x, 8);
        abs_y += __shfl_xor_sync(0xffffffffu, abs_y, 8);
    }
    if (valid && q_ == 0) {
        const float Sx = ux * M0 - Mx, Sy = uy * M0 - My;
        const float Sxx = fmaf(ux, fmaf(ux, M0, -2.f * Mx), Mxx);
        const float Syy = fmaf(uy, fmaf(uy, M0, -2.f * My), Myy);
        const float Sxy = fmaf(ux, fmaf(uy, M0, -My), fmaf(-uy, Mx, Mxy));
        const int64_t g = flatten_ids[__float_as_int(mu.z)];
        float v[6];
        v[0] = ca * Sx + cb * Sy;
        v[1] = cb * Sx + cc * Sy;
        v[2] = 0.5f * Sxx;
        v[3] = Sxy;
        v[4] = 0.5f * Syy;
        v[5] = -__fdividef(M0, opac);  // v_opacity = sum e^-sigma v_alpha = -sum v_sigma / opacity
        // (v_means2d / absgrad as ONE 8-byte reduction, REDG.E.ADD.F32x2: -3 us per step at cfg2, but the five
        // gradient arrays are views of one buffer and only 4-byte aligned in general, e.g. absgrad behind 9 P floats
        // with P odd: misaligned address. Scalar reductions stay.)
        if (v[0] != 0.f) atomicAdd(out.v_means2d + 2 * g, v[0]);
        if (v[1] != 0.f) atomicAdd(out.v_means2d + 2 * g + 1, v[1]);
        if (v[2] != 0.f) atomicAdd(out.v_conics + 3 * g, v[2]);
        if (v[3] != 0.f) atomicAdd(out.v_conics + 3 * g + 1, v[3]);
        if (v[4] != 0.f) atomicAdd(out.v_conics + 3 * g + 2, v[4]);
        if (v[5] != 0.f) atomicAdd(out.v_opacities + g, v[5]);
#pragma unroll
        for (int k = 0; k < D; ++k)
            if (cs[k] != 0.f) atomicAdd(out.v_colors + (int64_t)D * g + k, -cs[k]);  // the table holds MINUS alpha T
        if (ABS) {
            if (abs_x != 0.f) atomicAdd(out.v_abs + 2 * g, abs_x);
            if (abs_y != 0.f) atomicAdd(out.v_abs + 2 * g + 1, abs_y);
        }
    }
    __syncwarp();
}

template <int D, bool ABS>
__global__ void __launch_bounds__(kB2Threads, GSB_BWD2_MIN_BLOCKS)
rasterize_bwd2_kernel(int C, int64_t n_isects, const float4 *__restrict__ recA, const float4 *__restrict__ recB,
                      const float *__restrict__ colors, const float *__restrict__ backgrounds,
                      const uint8_t *__restrict__ masks, int W, int H, int tw, int th,
                      const int32_t *__restrict__ offsets, const int32_t *__restrict__ flatten_ids,
                      const float *__restrict__ render_alphas, const int32_t *__restrict__ last_ids,
                      const uint8_t *__restrict__ block_masks, const float *__restrict__ v_render_colors,
                      const float *__restrict__ v_render_alphas, GradTargets out) {
    static_assert(D >= 1 && D <= 4, "v2 keeps the colour of a staged splat in one float4");
    gsb_chain_enter();
    __shared__ Bwd2Smem sm;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, half = lane >> 4, l16 = lane & 15;
    const int hw = 2 * warp + half;
    const int cam = blockIdx.z, tile_y = blockIdx.y, tile_x = blockIdx.x;
    const int64_t tile_id = ((int64_t)cam * th + tile_y) * tw + tile_x;
    if (masks != nullptr && !masks[tile_id]) return;
    const int range_start = offsets[tile_id];
    const int range_end = (tile_id + 1 < (int64_t)C * th * tw) ? offsets[tile_id + 1] : (int)n_isects;
    // the halves of a warp own vertically adjacent 8x4 blocks: (bx, by) = (warp & 1, 2 (warp >> 1) + half)
    const int bx = warp & 1, by = 2 * (warp >> 1) + half;
    const int px = tile_x * kTile + bx * 8 + (l16 & 7);
    const int py0 = tile_y * kTile + by * 4 + (l16 >> 3) * 2, py1 = py0 + 1;
    const bool in0 = (px < W) & (py0 < H), in1 = (px < W) & (py1 < H);
    const int64_t p0 = ((int64_t)cam * H + min(py0, H - 1)) * W + min(px, W - 1);
    const int64_t p1 = ((int64_t)cam * H + min(py1, H - 1)) * W + min(px, W - 1);
    const float bcx = (float)(tile_x * kTile + bx * 8) + 4.0f, bcy = (float)(tile_y * kTile + by * 4) + 2.0f;
    if (tid == 0) {
        sm.P[0] = make_float4(0.f, 0.f, 0.f, 0.f);
        sm.B[0] = make_float4(0.f, 0.f, 0.f, 0.f);
        sm.Cq[0] = make_float4(0.f, 0.f, 0.f, 0.f);
    }

    // per-pixel state (two pixels)
    const float Tf0 = in0 ? 1.0f - render_alphas[p0] : 1.0f, Tf1 = in1 ? 1.0f - render_alphas[p1] : 1.0f;
    float T0 = Tf0, T1 = Tf1;
    const int bin0 = in0 ? last_ids[p0] : -1, bin1 = in1 ? last_ids[p1] : -1;
    float vo0[4] = {0.f, 0.f, 0.f, 0.f}, vo1[4] = {0.f, 0.f, 0.f, 0.f};
    float bg0 = 0.f, bg1 = 0.f;
#pragma unroll
    for (int k = 0; k < D; ++k) {
        vo0[k] = in0 ? v_render_colors[p0 * D + k] : 0.f;
        vo1[k] = in1 ? v_render_colors[p1 * D + k] : 0.f;
        if (backgrounds != nullptr) {
            const float b = backgrounds[(size_t)cam * D + k];
            bg0 = fmaf(b, vo0[k], bg0);
            bg1 = fmaf(b, vo1[k], bg1);
        }
    }
    // everything behind splat i reaches v_alpha_i through ONE scalar per pixel (see rasterize_bwd.cu):
    //   u_i = T_final (v_alpha_out - bg . v_out) - sum_{j behind i} alpha_j T_j (c_j . v_out)
    float u0 = Tf0 * ((in0 ? v_render_alphas[p0] : 0.f) - bg0), u1 = Tf1 * ((in1 ? v_render_alphas[p1] : 0.f) - bg1);
    sm.V[hw][l16][0] = make_float4(vo0[0], vo1[0], vo0[1], vo1[1]);
    sm.V[hw][l16][1] = make_float4(vo0[2], vo1[2], vo0[3], vo1[3]);
    f32x2 vo[D];
#pragma unroll
    for (int k = 0; k < D; ++k) vo[k] = pack2(vo0[k], vo1[k]);

    // deepest contributor of this half-warp's block and of the tile
    int half_last = max(bin0, bin1);
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) half_last = max(half_last, __shfl_xor_sync(0xffffffffu, half_last, o));
    if (l16 == 0) sm.red[hw] = half_last;
    __syncthreads();
    int tile_last = sm.red[0];
#pragma unroll
    for (int h = 1; h < kB2Halves; ++h) tile_last = max(tile_last, sm.red[h]);
    const int last_a = __shfl_sync(0xffffffffu, half_last, 0), last_b = __shfl_sync(0xffffffffu, half_last, 16);

    const float pxf = (float)px + 0.5f;
    const f32x2 npy = pack2(-((float)py0 + 0.5f), -((float)py1 + 0.5f));
    const int bit0 = (4 * (warp >> 1)) + bx, bit1 = bit0 + 2;
    unsigned half_mask = half ? 0xffffffffu : 0u, half_lanes = half ? 0xffff0000u : 0x0000ffffu;
    asm volatile("mov.u32 %0, %0;" : "+r"(half_mask));
    unsigned sP = (unsigned)__cvta_generic_to_shared(sm.P);
    unsigned sW = (unsigned)__cvta_generic_to_shared(&sm.Wt[hw][0][0]) + 16u * (unsigned)l16;
    asm volatile("mov.u32 %0, %0;" : "+r"(sP));
    asm volatile("mov.u32 %0, %0;" : "+r"(sW));
    constexpr unsigned kOffB = sizeof(float4) * (kB2Threads + 1), kOffC = 2 * kOffB;

    const int hi_start = min(range_end - 1, tile_last);
    for (int hi = hi_start; hi >= range_start; hi -= kB2Threads) {
        const int idx = hi - tid;
        unsigned mask = 0u;
        int g = 0;
        float4 A, B;
        if (idx >= range_start) {
            if (block_masks != nullptr) {
                // the forward pass staged every intersection up to the tile's last contributor and kept its culling
                // result: culled intersections cost one byte here
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
                mask = splat_block_mask(A, B, tile_x * kTile, tile_y * kTile);
            }
        }
        const unsigned ballot = __ballot_sync(0xffffffffu, mask != 0u);
        if (lane == 0) sm.wcnt[warp] = __popc(ballot);
        __syncthreads();  // also: the previous batch (staged entries, tables) is fully consumed
        int slot = 0, total = 0;
#pragma unroll
        for (int w = 0; w < kB2Warps; ++w) {
            const int c = sm.wcnt[w];
            slot += (w < warp) ? c : 0;
            total += c;
        }
        slot += __popc(ballot & ((1u << lane) - 1u)) + 1;  // entry 0 is the pad
        if (mask != 0u) {
            sm.P[slot] = make_float4(A.x, A.y, __int_as_float(idx), __uint_as_float(mask));
            float4 sb = stage_conic(B);
            sb.w = -sb.w;
            sm.B[slot] = sb;
            const float *cp = colors + (size_t)g * D;
            float4 c4 = make_float4(0.f, 0.f, 0.f, 0.f);
            c4.x = cp[0];
            if (D > 1) c4.y = cp[1];
            if (D > 2) c4.z = cp[2];
            if (D > 3) c4.w = cp[3];
            sm.Cq[slot] = c4;
        }
        __syncthreads();

        int n_chunk = 0, entry = 0;  // this half's chunk fill; lanes l16 & 7 == s keep the staged entry of its s-th splat
        for (int c0 = 0; c0 < total; c0 += 32) {
            const int jl = c0 + lane;
            bool mine0 = false, mine1 = false;
            if (jl < total) {
                const float4 p4 = sm.P[jl + 1];
                const unsigned m = __float_as_uint(p4.w);
                const int sidx = __float_as_int(p4.z);
                mine0 = ((m >> bit0) & 1u) && sidx <= last_a;
                mine1 = ((m >> bit1) & 1u) && sidx <= last_b;
            }
            unsigned todo0 = __ballot_sync(0xffffffffu, mine0);
            unsigned todo1 = __ballot_sync(0xffffffffu, mine1);
            // (Measured and dropped: two list entries per trip, with the T/u-independent part of both issued ahead of
            // the recurrences, raised issue-active from 54 % to 67 % but also the instruction count by 18 %: 0.669 ms
            // against 0.661 ms for this loop at cfg2.)
            while (todo0 | todo1) {
                const unsigned todo = todo0 ^ ((todo0 ^ todo1) & half_mask);
                const int e = c0 + __ffs(todo);  // staged entry (pad-shifted); c0 = a harmless finite entry when exhausted
                const unsigned addr = sP + 16u * (unsigned)e;
                todo0 &= todo0 - 1u;
                todo1 &= todo1 - 1u;
                const float4 p4 = lds128(addr);
                const float4 b4 = lds128(addr + kOffB);
                const float4 c4 = lds128(addr + kOffC);
                const float dx = p4.x - pxf;
                const f32x2 dy = add2(bcast2(p4.y), npy);
                const f32x2 t = fma2(bcast2(b4.y), dy, bcast2(b4.x * dx));
                const f32x2 ex = fma2(mul2(bcast2(b4.z), dy), dy, mul2(bcast2(dx), t));  // log2(e^-sigma), as gauss_log2
                float ex0, ex1;
                unpack2(ex, ex0, ex1);
                const float nopac = todo != 0u ? b4.w : 0.f;  // an exhausted half never hits
                float nov0, nov1;  // MINUS opacity e^-sigma
                unpack2(mul2(bcast2(nopac), pack2(exp2f(ex0), exp2f(ex1))), nov0, nov1);
                const int sidx = __float_as_int(p4.z);
                float na0, na1, nvh0, nvh1;
                const bool hit0 = bwd_decide(sidx, bin0, ex0, nov0, fmaxf(-kAlphaMax, nov0), na0, nvh0);
                const bool hit1 = bwd_decide(sidx, bin1, ex1, nov1, fmaxf(-kAlphaMax, nov1), na1, nvh1);
                const unsigned hits = __ballot_sync(0xffffffffu, hit0 | hit1);
                // branch-free over the lanes: a pixel without a hit runs the same arithmetic with alpha = 0
                // (1 / (1 - 0) and T * 1 are exact, a zero weight leaves u alone)
                const f32x2 na = pack2(na0, na1);
                float om0, om1;
                unpack2(add2(bcast2(1.0f), na), om0, om1);
                const f32x2 ra = pack2(__fdividef(1.0f, om0), __fdividef(1.0f, om1));
                f32x2 T2 = mul2(pack2(T0, T1), ra);  // transmittance in front of this splat
                unpack2(T2, T0, T1);
                const f32x2 nw2 = mul2(na, T2);  // MINUS alpha T
                f32x2 cv = mul2(bcast2(c4.x), vo[0]);
                if (D > 1) cv = fma2(bcast2(c4.y), vo[1], cv);
                if (D > 2) cv = fma2(bcast2(c4.z), vo[2], cv);
                if (D > 3) cv = fma2(bcast2(c4.w), vo[3], cv);
                const f32x2 u2 = pack2(u0, u1);
                // v_sigma = -(opacity e^-sigma) v_alpha, v_alpha = (c . v_out) T + u / (1 - alpha)
                const f32x2 vs = mul2(pack2(nvh0, nvh1), fma2(cv, T2, mul2(u2, ra)));
                unpack2(fma2(cv, nw2, u2), u0, u1);
                // park (v_sigma, -alpha T) of the two pixels in row n_chunk of the half's table
                float s0, s1, w0, w1;
                unpack2(vs, s0, s1);
                unpack2(nw2, w0, w1);
                const unsigned waddr = (sW + ((unsigned)n_chunk << 8)) ^ (((unsigned)n_chunk & 7u) << 4);
                asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(waddr), "f"(s0), "f"(s1), "f"(w0), "f"(w1)
                             : "memory");
                if (hits & half_lanes) {  // some pixel of this half's block was hit: the row is kept
                    entry = ((l16 & 7) == n_chunk) ? e : entry;  // both lanes (q = 0, 1) of the chunk's splat keep it
                    ++n_chunk;
                }
                if (__any_sync(0xffffffffu, n_chunk == kB2Chunk)) {
                    reduce_chunk2<D, ABS>(sm, hw, l16, n_chunk, entry, bcx, bcy, flatten_ids, out);
                    n_chunk = 0;
                }
            }
        }
        if (__any_sync(0xffffffffu, n_chunk > 0))
            reduce_chunk2<D, ABS>(sm, hw, l16, n_chunk, entry, bcx, bcy, flatten_ids, out);
    }
}

template <int D, bool ABS>
int launch_bwd2(int C, int64_t n_isects, const float *recA, const float *recB, const float *colors,
                const float *backgrounds, const uint8_t *masks, int W, int H, int tw, int th, const int32_t *offsets,
                const int32_t *flatten_ids, const float *render_alphas, const int32_t *last_ids,
                const uint8_t *block_masks, const float *v_render_colors, const float *v_render_alphas,
                GradTargets out, cudaStream_t stream) {
    dim3 grid(tw, th, C);
    gsb_launch_chained(GSB_CHAIN_BWD, rasterize_bwd2_kernel<D, ABS>, grid, kB2Threads, 0, stream, C, n_isects,
                       reinterpret_cast<const float4 *>(recA), reinterpret_cast<const float4 *>(recB), colors, backgrounds,
                       masks, W, H, tw, th, offsets, flatten_ids, render_alphas, last_ids, block_masks, v_render_colors,
                       v_render_alphas, out);
    return gsb_check_launch("rasterize_bwd2");
}

}  // namespace raster

// D <= 4 only; GSB_ERR_UNSUPPORTED otherwise (the caller falls back to v1's instantiations)
int gsb_rasterize_bwd2_dispatch(int C, int64_t n_isects, int D, const float *recA, const float *recB,
                                const float *colors, const float *backgrounds, const uint8_t *masks, int width,
                                int height, int tile_width, int tile_height, const int32_t *isect_offsets,
                                const int32_t *flatten_ids, const float *render_alphas, const int32_t *last_ids,
                                const uint8_t *block_masks, const float *v_render_colors,
                                const float *v_render_alphas, float *v_means2d_abs, float *v_means2d, float *v_conics,
                                float *v_colors, float *v_opacities, cudaStream_t stream) {
    raster::GradTargets out = {v_means2d, v_conics, v_opacities, v_colors, v_means2d_abs};
#define GSB_BWD2_CASE(DD)                                                                                          \
    case DD:                                                                                                       \
        if (v_means2d_abs != nullptr)                                                                              \
            return raster::launch_bwd2<DD, true>(C, n_isects, recA, recB, colors, backgrounds, masks, width, height, \
                                                 tile_width, tile_height, isect_offsets, flatten_ids, render_alphas, \
                                                 last_ids, block_masks, v_render_colors, v_render_alphas, out,     \
                                                 stream);                                                          \
        return raster::launch_bwd2<DD, false>(C, n_isects, recA, recB, colors, backgrounds, masks, width, height,  \
                                              tile_width, tile_height, isect_offsets, flatten_ids, render_alphas,  \
                                              last_ids, block_masks, v_render_colors, v_render_alphas, out, stream);
    switch (D) {
        GSB_BWD2_CASE(1)
        GSB_BWD2_CASE(2)
        GSB_BWD2_CASE(3)
        GSB_BWD2_CASE(4)
        default:
            return GSB_ERR_UNSUPPORTED;
    }
#undef GSB_BWD2_CASE
}
