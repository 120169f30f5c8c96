// Shared helpers for the gsplat b200 backend kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "gsplat_b200.h"

// records the last error string for gsb_last_error()
void gsb_set_error(const char *fmt, ...);
// bumps the counter behind gsb_launch_count()
void gsb_count_launch(int n);
// current value of a GSB_OPT_* switch (gsb_set_option)
int gsb_option(int which);

static inline int gsb_check_launch(const char *what) {
    gsb_count_launch(1);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        gsb_set_error("%s: %s", what, cudaGetErrorString(e));
        return GSB_ERR_CUDA;
    }
    return GSB_OK;
}

static inline unsigned gsb_cdiv(long long a, long long b) { return (unsigned)((a + b - 1) / b); }

// ---- chained launches (programmatic dependent launch, sm_90+) ---------------------------------------------------
// The intersection pipeline is a dozen short, dependent kernels on one stream; between two of them the GPU drains,
// the next grid is launched and ramps up: a few microseconds each time in which nothing runs. A kernel launched with
// gsb_launch_chained may become resident while its predecessor in the stream is still finishing: its CTAs take the SM
// slots the predecessor's last CTAs free and stop at gsb_chain_enter() — `griddepcontrol.wait`: until the predecessor
// grid has completed and its writes are visible — which therefore has to be the FIRST statement of every kernel
// launched this way, executed by every thread, before any early return (a grid whose CTAs skipped it could complete
// before its predecessor, and its own successor would then overtake that one too). `griddepcontrol.launch_dependents`
// right behind it lets the successor in, in turn, once every CTA of this grid has started.
// GSB_OPT_CHAINED_LAUNCH is a mask of kernel groups; a group whose bit is clear is launched the ordinary way (the
// instructions are then no-ops).
#ifdef __CUDACC__
__device__ __forceinline__ void gsb_chain_enter() {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// kernel groups, one bit each in GSB_OPT_CHAINED_LAUNCH
enum { GSB_CHAIN_ISECT = 1, GSB_CHAIN_PROJECTION = 2, GSB_CHAIN_COLOURS = 4, GSB_CHAIN_PACK = 8, GSB_CHAIN_FWD = 16,
       GSB_CHAIN_BWD = 32 };

template <typename... P, typename... A>
inline void gsb_launch_chained(int group, void (*kernel)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                               A &&...args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr;
    attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr.val.programmaticStreamSerializationAllowed = 1;
    if ((gsb_option(GSB_OPT_CHAINED_LAUNCH) & group) != 0) {
        cfg.attrs = &attr;
        cfg.numAttrs = 1;
    }
    cudaLaunchKernelEx(&cfg, kernel, static_cast<P>(args)...);  // errors surface in gsb_check_launch (cudaGetLastError)
}
#endif

#define GSB_REQUIRE(cond, msg)                  \
    do {                                        \
        if (!(cond)) {                          \
            gsb_set_error("%s: %s", __func__, msg); \
            return GSB_ERR_INVALID;             \
        }                                       \
    } while (0)

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ int warp_max_i(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Shared-memory digit histograms (`passes` x 256 counters) for the onesweep sort that follows a producer of 32-bit
// keys (gsb_radix_prepare): one count per live key and pass. (Merging the lanes of a warp that hold the same high digit
// with match.any before the atomic was measured: +10 us per step at cfg2, the match costs more than the serialised
// shared-memory increments it saves. Aggregating only the TOP digit, which the lanes of a warp mostly share, with
// one shuffle and two votes: 1.617-1.626 ms per step against 1.602-1.605, profiles/r04_staging/r04l_*.)
__device__ __forceinline__ void gsb_hist_add(uint32_t *s_hist, uint32_t key, bool live, int passes, int end_bit) {
    if (!live) return;
    for (int p = 0; p < passes; ++p) {
        const int bits = min(8, end_bit - 8 * p);
        atomicAdd(&s_hist[p * 256 + ((key >> (8 * p)) & ((1u << bits) - 1u))], 1u);
    }
}
