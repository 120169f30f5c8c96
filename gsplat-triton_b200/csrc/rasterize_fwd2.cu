// rasterize_to_pixels forward, second generation ("v2"): half-warp per 8x4 pixel block, two pixels per lane,
// Blackwell packed FP32 (FFMA2 / FMUL2 / FADD2, PTX `*.f32x2`).
//
// Same semantics, same staging, same exact block culling and the SAME arithmetic per (pixel, Gaussian) pair as
// rasterize_fwd.cu (every rounding identical: the two kernels produce bit-identical images; tests assert it). What
// changes is the issue-slot economy, the measured limiter of v1 (ncu: 79 % issue-active, FMA pipe 34 %):
//   * a warp owns TWO vertically adjacent 8x4 blocks of the tile (lanes 0-15 / 16-31); each half walks the ballot
//     list of its own block, so one loop iteration composites two different Gaussians onto 2 x 32 pixels;
//   * a lane owns two vertically adjacent pixels and evaluates them with packed instructions (one FFMA2 = two
//     FMAs, per-Gaussian scalars enter as broadcast operands at no cost), so the list bookkeeping, the three 16-byte
//     shared-memory loads and most of the arithmetic are issued once per two pixels;
//   * the blend is branch-free (select, zero weight) — no divergence between the two halves or the two pixels.
// ~47 issued instructions per iteration (64 pixel-Gaussian evaluations) against 37 per 32 in v1.
#include "rasterize_common.cuh"

namespace raster {

// The four decisions of one (pixel, Gaussian) pair in predicate form (FSETP with a predicate operand, predicated
// moves) — written out because the compiler's version of the same logic spends twice the instructions on PLOP3 / SEL:
//   hit  = !(ex > 0) && !(minus_alpha > -1/255)        go = hit && !(next_T <= 1e-4)
//   w    = go ? minus_w : 0 ;  if (go) { cur = index; T = next_T; }  else if (hit) T = -|T| (sign bit: stopped)
__device__ __forceinline__ float blend_decide(float ex, float minus_alpha, float next_T, float minus_w, int index,
                                              int &cur, float &T) {
    float w;
    unsigned t = __float_as_uint(T);
    asm("{\n\t"
        ".reg .pred h, g;\n\t"
        "setp.leu.ftz.f32 h, %3, 0f00000000;\n\t"
        "setp.leu.and.ftz.f32 h, %4, 0fBB808081, h;\n\t"
        "setp.gtu.and.ftz.f32 g, %5, 0f38D1B717, h;\n\t"
        "selp.f32 %0, %6, 0f00000000, g;\n\t"
        "@g mov.b32 %1, %7;\n\t"
        "@h or.b32 %2, %2, 0x80000000;\n\t"  // every hit marks the pixel stopped ...
        "@g mov.b32 %2, %8;\n\t"              // ... and a composited one gets its new transmittance back
        "}"
        : "=f"(w), "+r"(cur), "+r"(t)
        : "f"(ex), "f"(minus_alpha), "f"(next_T), "f"(minus_w), "r"(index), "r"(__float_as_uint(next_T)));
    T = __uint_as_float(t);
    return w;
}

constexpr int kThreads2 = 128;  // 4 warps x 2 blocks
constexpr int kWarps2 = kThreads2 / 32;

#ifndef GSB_FWD2_MIN_BLOCKS
#define GSB_FWD2_MIN_BLOCKS 8
#endif

// COUNT (diagnostics): `warp_iters` receives the number of loop iterations (each = 64 lane evaluations)
template <int D, bool COUNT = false>
__global__ void __launch_bounds__(kThreads2, GSB_FWD2_MIN_BLOCKS)
rasterize_fwd2_kernel(int C, int64_t n_isects, const float4 *__restrict__ recA, const float4 *__restrict__ recB,
                      const float *__restrict__ colors, const float *__restrict__ backgrounds,
                      const uint8_t *__restrict__ masks, int W, int H, int tw, int th,
                      const int32_t *__restrict__ offsets, const int32_t *__restrict__ flatten_ids,
                      float *__restrict__ out_colors, float *__restrict__ out_alphas,
                      int32_t *__restrict__ last_ids, const int64_t *__restrict__ n_isects_dev,
                      uint8_t *__restrict__ block_masks, unsigned long long *__restrict__ warp_iters) {
    static_assert(D >= 1 && D <= 4, "v2 keeps the colour of a staged splat in one float4");
    gsb_chain_enter();
    if (n_isects_dev != nullptr) n_isects = min(n_isects, (int64_t)*n_isects_dev);
    // (Measured and dropped: a double-buffered, uncompacted batch needs one CTA barrier per batch instead of three, but
    // the warps then scan every entry's mask instead of the survivors': 0.420 ms against 0.408 ms at cfg2. The
    // barriers top the stall samples, yet they are not what bounds the kernel: issue slots and the ALU pipe are.)
    // Staged batch. Entry 0 of every array is a zeroed pad: a half-warp whose list is exhausted reads it (index
    // c0 + ffs(0) = c0, harmless finite values, discarded through a zero opacity) while the other half goes on.
    struct __align__(16) Stage {
        float4 P[kThreads2 + 32];   // x, y, sorted index, block mask
        float4 B[kThreads2 + 32];   // staged conic (A', B', C'), MINUS opacity
        float4 Cq[kThreads2 + 32];  // colour
        int wcnt[kWarps2];
    };
    __shared__ Stage st;
    for (int i = threadIdx.x; i < kThreads2 + 32; i += kThreads2) {
        st.P[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        st.B[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        st.Cq[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    unsigned sP = (unsigned)__cvta_generic_to_shared(st.P);
    asm volatile("mov.u32 %0, %0;" : "+r"(sP));  // opaque: kept in a register instead of being recomputed per iteration
    constexpr unsigned kOffB = sizeof(float4) * (kThreads2 + 32), kOffC = 2 * kOffB;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, half = lane >> 4, l16 = lane & 15;
    const int cam = blockIdx.z, tile_y = blockIdx.y, tile_x = blockIdx.x;
    const int64_t tile_id = ((int64_t)cam * th + tile_y) * tw + tile_x;
    const int range_start = offsets[tile_id];
    const int range_end = (tile_id + 1 < (int64_t)C * th * tw) ? offsets[tile_id + 1] : (int)n_isects;
    // The two halves of a warp own VERTICALLY adjacent 8x4 blocks (centres 4 pixels apart; horizontal neighbours are
    // 8 apart): their Gaussian lists overlap most, so the two list lengths, whose maximum the loop runs for, differ
    // least. Block (bx, by) = (warp & 1, 2 (warp >> 1) + half); its bit in the staged mask is by * 2 + bx as in v1.
    const int bx = warp & 1, by = 2 * (warp >> 1) + half;
    const int px = tile_x * kTile + bx * 8 + (l16 & 7);
    const int py0 = tile_y * kTile + by * 4 + (l16 >> 3) * 2, py1 = py0 + 1;
    const bool in0 = (px < W) & (py0 < H), in1 = (px < W) & (py1 < H);

    // MINUS the colour sums of the two pixels. Carried as scalars and packed at the use: ptxas then keeps the FFMA2
    // accumulation in place (loop-carried 64-bit pairs cost it two register moves per channel and iteration)
    float pa[D], pb[D];
#pragma unroll
    for (int k = 0; k < D; ++k) pa[k] = pb[k] = 0.f;
    // |T| is the transmittance; T < 0 marks a finished pixel (saturated, or outside the image)
    float T0 = in0 ? 1.0f : -1.0f, T1 = in1 ? 1.0f : -1.0f;
    int cur0 = 0, cur1 = 0;
    unsigned n_iters = 0;

    const bool render = (masks == nullptr) || masks[tile_id];
    if (render) {
        const float pxf = (float)px + 0.5f;
        const f32x2 npy = pack2(-((float)py0 + 0.5f), -((float)py1 + 0.5f));
        bool warp_done = __all_sync(0xffffffffu, (T0 < 0.f) & (T1 < 0.f));
        const int bit0 = (4 * (warp >> 1)) + bx, bit1 = bit0 + 2;  // by * 2 + bx for half 0 / half 1
        unsigned half_mask = half ? 0xffffffffu : 0u;
        asm volatile("mov.u32 %0, %0;" : "+r"(half_mask));  // opaque, so the list select stays one LOP3

        for (int base = range_start; base < range_end; base += kThreads2) {
            if (__syncthreads_count((T0 < 0.f) & (T1 < 0.f)) == kThreads2) break;
            const int idx = base + tid;
            unsigned mask = 0u;
            int g = 0;
            float4 A, B;
            if (idx < range_end) {
                g = flatten_ids[idx];
                A = __ldg(recA + g);
                B = __ldg(recB + g);
                mask = splat_block_mask(A, B, tile_x * kTile, tile_y * kTile);
                if (block_masks != nullptr) block_masks[idx] = (uint8_t)mask;
            }
            // compacting append (4 warps)
            const unsigned ballot = __ballot_sync(0xffffffffu, mask != 0u);
            if (lane == 0) st.wcnt[warp] = __popc(ballot);
            __syncthreads();
            int slot = 0, total = 0;
#pragma unroll
            for (int w = 0; w < kWarps2; ++w) {
                const int c = st.wcnt[w];
                slot += (w < warp) ? c : 0;
                total += c;
            }
            slot += __popc(ballot & ((1u << lane) - 1u)) + 1;  // entry 0 is the pad
            if (mask != 0u) {
                st.P[slot] = make_float4(A.x, A.y, __int_as_float(idx), __uint_as_float(mask));
                float4 sb = stage_conic(B);
                sb.w = -sb.w;
                st.B[slot] = sb;
                const float *cp = colors + (size_t)g * D;
                float4 c4 = make_float4(0.f, 0.f, 0.f, 0.f);
                c4.x = cp[0];
                if (D > 1) c4.y = cp[1];
                if (D > 2) c4.z = cp[2];
                if (D > 3) c4.w = cp[3];
                st.Cq[slot] = c4;
            }
            __syncthreads();
            if (warp_done) continue;
            for (int c0 = 0; c0 < total; c0 += 32) {
                const int jl = c0 + lane;
                const unsigned m = jl < total ? __float_as_uint(st.P[jl + 1].w) : 0u;
                unsigned todo0 = __ballot_sync(0xffffffffu, (m >> bit0) & 1u);
                unsigned todo1 = __ballot_sync(0xffffffffu, (m >> bit1) & 1u);
                while (todo0 | todo1) {
                    const unsigned todo = todo0 ^ ((todo0 ^ todo1) & half_mask);  // this half's list (one LOP3)
                    const unsigned addr = sP + 16u * (unsigned)(c0 + __ffs(todo));  // entry c0 + ffs (pad-shifted)
                    todo0 &= todo0 - 1u;
                    todo1 &= todo1 - 1u;
                    if (COUNT) ++n_iters;
                    const float4 p4 = lds128(addr);
                    const float4 b4 = lds128(addr + kOffB);
                    const float4 c4 = lds128(addr + kOffC);
                    const float dx = p4.x - pxf;
                    const f32x2 dy = add2(bcast2(p4.y), npy);
                    // log2(e^-sigma) = dx (A' dx + B' dy) + C' dy^2: the expression (and every rounding) of gauss_log2
                    const f32x2 t = fma2(bcast2(b4.y), dy, bcast2(b4.x * dx));
                    const f32x2 ex = fma2(mul2(bcast2(b4.z), dy), dy, mul2(bcast2(dx), t));
                    float ex0, ex1;
                    unpack2(ex, ex0, ex1);
                    // MINUS alpha; an exhausted half composites the pad with opacity 0: never a hit
                    const float nopac = todo != 0u ? b4.w : 0.f;
                    float na0, na1;
                    unpack2(mul2(bcast2(nopac), pack2(exp2f(ex0), exp2f(ex1))), na0, na1);
                    na0 = fmaxf(-kAlphaMax, na0);
                    na1 = fmaxf(-kAlphaMax, na1);
                    const f32x2 na = pack2(na0, na1), T2 = pack2(T0, T1);
                    float nT0, nT1;  // T (1 - alpha); < 0 on finished pixels
                    unpack2(fma2(T2, na, T2), nT0, nT1);
                    float w0, w1;  // MINUS alpha T
                    unpack2(mul2(na, T2), w0, w1);
                    const int sidx = __float_as_int(p4.z);
                    w0 = blend_decide(ex0, na0, nT0, w0, sidx, cur0, T0);
                    w1 = blend_decide(ex1, na1, nT1, w1, sidx, cur1, T1);
                    const f32x2 wn = pack2(w0, w1);
                    unpack2(fma2(bcast2(c4.x), wn, pack2(pa[0], pb[0])), pa[0], pb[0]);
                    if (D > 1) unpack2(fma2(bcast2(c4.y), wn, pack2(pa[1], pb[1])), pa[1], pb[1]);
                    if (D > 2) unpack2(fma2(bcast2(c4.z), wn, pack2(pa[2], pb[2])), pa[2], pb[2]);
                    if (D > 3) unpack2(fma2(bcast2(c4.w), wn, pack2(pa[3], pb[3])), pa[3], pb[3]);
                }
                if (__all_sync(0xffffffffu, (T0 < 0.f) & (T1 < 0.f))) {
                    warp_done = true;
                    break;
                }
            }
        }
    }
    T0 = fabsf(T0);
    T1 = fabsf(T1);
    if (COUNT && warp_iters != nullptr && lane == 0 && n_iters) atomicAdd(warp_iters, (unsigned long long)n_iters);

    const float *bg = backgrounds != nullptr ? backgrounds + (size_t)cam * D : nullptr;
    float c0v[D], c1v[D];
#pragma unroll
    for (int k = 0; k < D; ++k) {
        c0v[k] = pa[k]; c1v[k] = pb[k];
        c0v[k] = -c0v[k];
        c1v[k] = -c1v[k];
    }
    if (in0) {
        const int64_t p = ((int64_t)cam * H + py0) * W + px;
#pragma unroll
        for (int k = 0; k < D; ++k) out_colors[p * D + k] = bg != nullptr ? fmaf(T0, bg[k], c0v[k]) : c0v[k];
        out_alphas[p] = 1.0f - T0;
        last_ids[p] = cur0;
    }
    if (in1) {
        const int64_t p = ((int64_t)cam * H + py1) * W + px;
#pragma unroll
        for (int k = 0; k < D; ++k) out_colors[p * D + k] = bg != nullptr ? fmaf(T1, bg[k], c1v[k]) : c1v[k];
        out_alphas[p] = 1.0f - T1;
        last_ids[p] = cur1;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// v3 = v2 + a producer warp. ncu on v2 (cfg2): the 50-instruction compositing loop draws only 44 % of the stall
// samples; 55 % sit in the per-batch staging code (two dependent global gathers, ~200 instructions of exact culling
// per splat, two CTA barriers) that every warp executed in lock step BEFORE it could composite. Here a fifth warp
// stages batch k+1 (gathers in flight four deep, culling, warp-local ballot compaction: no CTA barrier at all) into
// the second half of a double buffer while the four consumer warps composite batch k; the hand-over is two pairs of
// named barriers (bar.sync / bar.arrive, "full" and "empty" per buffer).
constexpr int kThreads3 = kThreads2 + 32;
constexpr int kBarFull = 1, kBarEmpty = 3;  // + buffer index; barrier 0 is __syncthreads

__device__ __forceinline__ void named_bar_sync(int id) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(kThreads3) : "memory");
}
__device__ __forceinline__ void named_bar_arrive(int id) {
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(kThreads3) : "memory");
}

struct __align__(16) Stage3 {
    float4 P[kThreads2 + 1];   // entry 0 = zero pad; x, y, sorted index, block mask
    float4 B[kThreads2 + 1];   // staged conic (A', B', C'), MINUS opacity
    float4 Cq[kThreads2 + 1];  // colour
};

#ifndef GSB_FWD3_MIN_BLOCKS
#define GSB_FWD3_MIN_BLOCKS 6
#endif

template <int D>
__global__ void __launch_bounds__(kThreads3, GSB_FWD3_MIN_BLOCKS)
rasterize_fwd3_kernel(int C, int64_t n_isects, const float4 *__restrict__ recA, const float4 *__restrict__ recB,
                      const float *__restrict__ colors, const float *__restrict__ backgrounds,
                      const uint8_t *__restrict__ masks, int W, int H, int tw, int th,
                      const int32_t *__restrict__ offsets, const int32_t *__restrict__ flatten_ids,
                      float *__restrict__ out_colors, float *__restrict__ out_alphas,
                      int32_t *__restrict__ last_ids, const int64_t *__restrict__ n_isects_dev,
                      uint8_t *__restrict__ block_masks) {
    static_assert(D >= 1 && D <= 4, "one float4 of colour per staged splat");
    if (n_isects_dev != nullptr) n_isects = min(n_isects, (int64_t)*n_isects_dev);
    __shared__ Stage3 st[2];
    __shared__ int s_total[2];
    __shared__ int s_done_warps;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int cam = blockIdx.z, tile_y = blockIdx.y, tile_x = blockIdx.x;
    const int64_t tile_id = ((int64_t)cam * th + tile_y) * tw + tile_x;
    const int range_start = offsets[tile_id];
    const int range_end = (tile_id + 1 < (int64_t)C * th * tw) ? offsets[tile_id + 1] : (int)n_isects;
    const bool render = (masks == nullptr) || masks[tile_id];
    if (tid < 2) {
        st[tid].P[0] = make_float4(0.f, 0.f, 0.f, 0.f);
        st[tid].B[0] = make_float4(0.f, 0.f, 0.f, 0.f);
        st[tid].Cq[0] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (tid == 0) s_done_warps = 0;
    __syncthreads();

    if (warp == kWarps2) {
        // ---------------- producer warp ----------------
        if (!render) return;
        int k = 0;
        for (int base = range_start;; base += kThreads2, ++k) {
            const int b = k & 1;
            if (k >= 2) named_bar_sync(kBarEmpty + b);  // the consumers are done with this buffer
            const bool stop = *reinterpret_cast<volatile int *>(&s_done_warps) == kWarps2;  // every pixel finished
            if (base >= range_end || stop) {
                if (lane == 0) s_total[b] = -1;  // end marker
                __threadfence_block();
                named_bar_arrive(kBarFull + b);
                return;
            }
            // four rounds of 32 intersections; all ids, then all record gathers, are in flight together
            int g[4];
            float4 A[4], Bq[4];
            float col[4][D];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int idx = base + 32 * r + lane;
                g[r] = idx < range_end ? flatten_ids[idx] : -1;
            }
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                if (g[r] >= 0) {
                    A[r] = __ldg(recA + g[r]);
                    Bq[r] = __ldg(recB + g[r]);
#pragma unroll
                    for (int c = 0; c < D; ++c) col[r][c] = __ldg(colors + (size_t)g[r] * D + c);
                }
            }
            int cnt = 0;
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int idx = base + 32 * r + lane;
                unsigned mask = 0u;
                if (g[r] >= 0) {
                    mask = splat_block_mask(A[r], Bq[r], tile_x * kTile, tile_y * kTile);
                    if (block_masks != nullptr) block_masks[idx] = (uint8_t)mask;
                }
                const unsigned ballot = __ballot_sync(0xffffffffu, mask != 0u);
                if (mask != 0u) {
                    const int slot = 1 + cnt + __popc(ballot & ((1u << lane) - 1u));  // entry 0 is the pad
                    st[b].P[slot] = make_float4(A[r].x, A[r].y, __int_as_float(idx), __uint_as_float(mask));
                    float4 sb = stage_conic(Bq[r]);
                    sb.w = -sb.w;
                    st[b].B[slot] = sb;
                    float4 c4 = make_float4(0.f, 0.f, 0.f, 0.f);
                    c4.x = col[r][0];
                    if (D > 1) c4.y = col[r][1];
                    if (D > 2) c4.z = col[r][2];
                    if (D > 3) c4.w = col[r][3];
                    st[b].Cq[slot] = c4;
                }
                cnt += __popc(ballot);
            }
            if (lane == 0) s_total[b] = cnt;
            __threadfence_block();
            named_bar_arrive(kBarFull + b);
        }
    }

    // ---------------- consumer warps (the v2 compositor) ----------------
    const int half = lane >> 4, l16 = lane & 15;
    const int bx = warp & 1, by = 2 * (warp >> 1) + half;
    const int px = tile_x * kTile + bx * 8 + (l16 & 7);
    const int py0 = tile_y * kTile + by * 4 + (l16 >> 3) * 2, py1 = py0 + 1;
    const bool in0 = (px < W) & (py0 < H), in1 = (px < W) & (py1 < H);
    float pa[D], pb[D];
#pragma unroll
    for (int k = 0; k < D; ++k) pa[k] = pb[k] = 0.f;
    float T0 = in0 ? 1.0f : -1.0f, T1 = in1 ? 1.0f : -1.0f;
    int cur0 = 0, cur1 = 0;
    if (render) {
        const float pxf = (float)px + 0.5f;
        const f32x2 npy = pack2(-((float)py0 + 0.5f), -((float)py1 + 0.5f));
        bool warp_done = __all_sync(0xffffffffu, (T0 < 0.f) & (T1 < 0.f));
        if (warp_done && lane == 0) atomicAdd(&s_done_warps, 1);
        const int bit0 = (4 * (warp >> 1)) + bx, bit1 = bit0 + 2;
        unsigned half_mask = half ? 0xffffffffu : 0u;
        asm volatile("mov.u32 %0, %0;" : "+r"(half_mask));
        unsigned sP0 = (unsigned)__cvta_generic_to_shared(st[0].P);
        asm volatile("mov.u32 %0, %0;" : "+r"(sP0));
        constexpr unsigned kOffB = sizeof(float4) * (kThreads2 + 1), kOffC = 2 * kOffB;
        for (int k = 0;; ++k) {
            const int b = k & 1;
            named_bar_sync(kBarFull + b);
            const int total = *reinterpret_cast<volatile int *>(&s_total[b]);
            if (total < 0) break;
            if (!warp_done) {
                const unsigned sP = sP0 + (unsigned)b * (unsigned)sizeof(Stage3);
                for (int c0 = 0; c0 < total; c0 += 32) {
                    const int jl = c0 + lane;
                    const unsigned m = jl < total ? __float_as_uint(st[b].P[jl + 1].w) : 0u;
                    unsigned todo0 = __ballot_sync(0xffffffffu, (m >> bit0) & 1u);
                    unsigned todo1 = __ballot_sync(0xffffffffu, (m >> bit1) & 1u);
                    while (todo0 | todo1) {
                        const unsigned todo = todo0 ^ ((todo0 ^ todo1) & half_mask);
                        const unsigned addr = sP + 16u * (unsigned)(c0 + __ffs(todo));
                        todo0 &= todo0 - 1u;
                        todo1 &= todo1 - 1u;
                        const float4 p4 = lds128(addr);
                        const float4 b4 = lds128(addr + kOffB);
                        const float4 c4 = lds128(addr + kOffC);
                        const float dx = p4.x - pxf;
                        const f32x2 dy = add2(bcast2(p4.y), npy);
                        const f32x2 t = fma2(bcast2(b4.y), dy, bcast2(b4.x * dx));
                        const f32x2 ex = fma2(mul2(bcast2(b4.z), dy), dy, mul2(bcast2(dx), t));
                        float ex0, ex1;
                        unpack2(ex, ex0, ex1);
                        const float nopac = todo != 0u ? b4.w : 0.f;
                        float na0, na1;
                        unpack2(mul2(bcast2(nopac), pack2(exp2f(ex0), exp2f(ex1))), na0, na1);
                        na0 = fmaxf(-kAlphaMax, na0);
                        na1 = fmaxf(-kAlphaMax, na1);
                        const f32x2 na = pack2(na0, na1), T2 = pack2(T0, T1);
                        float nT0, nT1;
                        unpack2(fma2(T2, na, T2), nT0, nT1);
                        float w0, w1;
                        unpack2(mul2(na, T2), w0, w1);
                        const int sidx = __float_as_int(p4.z);
                        w0 = blend_decide(ex0, na0, nT0, w0, sidx, cur0, T0);
                        w1 = blend_decide(ex1, na1, nT1, w1, sidx, cur1, T1);
                        const f32x2 wn = pack2(w0, w1);
                        unpack2(fma2(bcast2(c4.x), wn, pack2(pa[0], pb[0])), pa[0], pb[0]);
                        if (D > 1) unpack2(fma2(bcast2(c4.y), wn, pack2(pa[1], pb[1])), pa[1], pb[1]);
                        if (D > 2) unpack2(fma2(bcast2(c4.z), wn, pack2(pa[2], pb[2])), pa[2], pb[2]);
                        if (D > 3) unpack2(fma2(bcast2(c4.w), wn, pack2(pa[3], pb[3])), pa[3], pb[3]);
                    }
                    if (__all_sync(0xffffffffu, (T0 < 0.f) & (T1 < 0.f))) {
                        warp_done = true;
                        if (lane == 0) atomicAdd(&s_done_warps, 1);  // the producer stops once all four are
                        break;
                    }
                }
            }
            named_bar_arrive(kBarEmpty + b);
        }
    }
    T0 = fabsf(T0);
    T1 = fabsf(T1);
    const float *bg = backgrounds != nullptr ? backgrounds + (size_t)cam * D : nullptr;
    if (in0) {
        const int64_t p = ((int64_t)cam * H + py0) * W + px;
#pragma unroll
        for (int k = 0; k < D; ++k) out_colors[p * D + k] = bg != nullptr ? fmaf(T0, bg[k], -pa[k]) : -pa[k];
        out_alphas[p] = 1.0f - T0;
        last_ids[p] = cur0;
    }
    if (in1) {
        const int64_t p = ((int64_t)cam * H + py1) * W + px;
#pragma unroll
        for (int k = 0; k < D; ++k) out_colors[p * D + k] = bg != nullptr ? fmaf(T1, bg[k], -pb[k]) : -pb[k];
        out_alphas[p] = 1.0f - T1;
        last_ids[p] = cur1;
    }
}

template <int D>
int launch_fwd3(int C, int64_t n_isects, const float *recA, const float *recB, const float *colors,
                const float *backgrounds, const uint8_t *masks, int W, int H, int tw, int th,
                const int32_t *offsets, const int32_t *flatten_ids, float *out_colors, float *out_alphas,
                int32_t *last_ids, const int64_t *n_isects_dev, uint8_t *block_masks, cudaStream_t stream) {
    dim3 grid(tw, th, C);
    rasterize_fwd3_kernel<D><<<grid, kThreads3, 0, stream>>>(
        C, n_isects, reinterpret_cast<const float4 *>(recA), reinterpret_cast<const float4 *>(recB), colors,
        backgrounds, masks, W, H, tw, th, offsets, flatten_ids, out_colors, out_alphas, last_ids, n_isects_dev,
        block_masks);
    return gsb_check_launch("rasterize_fwd3");
}

template <int D>
int launch_fwd2(int C, int64_t n_isects, const float *recA, const float *recB, const float *colors,
                const float *backgrounds, const uint8_t *masks, int W, int H, int tw, int th,
                const int32_t *offsets, const int32_t *flatten_ids, float *out_colors, float *out_alphas,
                int32_t *last_ids, const int64_t *n_isects_dev, uint8_t *block_masks, cudaStream_t stream) {
    dim3 grid(tw, th, C);
    gsb_launch_chained(GSB_CHAIN_FWD, rasterize_fwd2_kernel<D, false>, grid, kThreads2, 0, stream, C, n_isects,
                       reinterpret_cast<const float4 *>(recA), reinterpret_cast<const float4 *>(recB), colors, backgrounds,
                       masks, W, H, tw, th, offsets, flatten_ids, out_colors, out_alphas, last_ids, n_isects_dev,
                       block_masks, (unsigned long long *)nullptr);
    return gsb_check_launch("rasterize_fwd2");
}

}  // namespace raster

// D <= 4 only; returns GSB_ERR_UNSUPPORTED otherwise (the caller falls back to v1's instantiations)
int gsb_rasterize_fwd2_dispatch(int C, int64_t n_isects, const int64_t *n_isects_dev, int D, const float *recA,
                                const float *recB, const float *colors, const float *backgrounds,
                                const uint8_t *masks, int width, int height, int tile_width, int tile_height,
                                const int32_t *isect_offsets, const int32_t *flatten_ids, float *render_colors,
                                float *render_alphas, int32_t *last_ids, uint8_t *block_masks, cudaStream_t stream) {
#define GSB_FWD2_CASE(DD)                                                                                         \
    case DD:                                                                                                      \
        if (gsb_option(GSB_OPT_FWD_VARIANT) == 3)                                                                 \
            return raster::launch_fwd3<DD>(C, n_isects, recA, recB, colors, backgrounds, masks, width, height,    \
                                           tile_width, tile_height, isect_offsets, flatten_ids, render_colors,    \
                                           render_alphas, last_ids, n_isects_dev, block_masks, stream);           \
        return raster::launch_fwd2<DD>(C, n_isects, recA, recB, colors, backgrounds, masks, width, height,        \
                                       tile_width, tile_height, isect_offsets, flatten_ids, render_colors,        \
                                       render_alphas, last_ids, n_isects_dev, block_masks, stream);
    switch (D) {
        GSB_FWD2_CASE(1)
        GSB_FWD2_CASE(2)
        GSB_FWD2_CASE(3)
        GSB_FWD2_CASE(4)
        default:
            return GSB_ERR_UNSUPPORTED;
    }
#undef GSB_FWD2_CASE
}

// diagnostics for bench.py: iterations of the v2 inner loop (x 64 = executed lane evaluations)
int gsb_rasterize_fwd2_count(int C, int64_t n_isects, const float *recA, const float *recB, const float *ones,
                             int width, int height, int tile_width, int tile_height, const int32_t *isect_offsets,
                             const int32_t *flatten_ids, float *scratch_colors, float *scratch_alphas,
                             int32_t *scratch_last, int64_t *warp_iters, cudaStream_t stream) {
    dim3 grid(tile_width, tile_height, C);
    raster::rasterize_fwd2_kernel<1, true><<<grid, raster::kThreads2, 0, stream>>>(
        C, n_isects, reinterpret_cast<const float4 *>(recA), reinterpret_cast<const float4 *>(recB), ones, nullptr,
        nullptr, width, height, tile_width, tile_height, isect_offsets, flatten_ids, scratch_colors, scratch_alphas,
        scratch_last, nullptr, nullptr, reinterpret_cast<unsigned long long *>(warp_iters));
    return gsb_check_launch("rasterize_fwd2_count");
}
