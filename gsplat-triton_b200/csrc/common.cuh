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
