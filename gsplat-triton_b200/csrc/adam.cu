// Selective (visibility-masked) fused Adam step: the update either side of the render in a training iteration.
//
// Replaces (behaviour, not code): gsplat/cuda/csrc/AdamCUDA.cu:13-52 behind gsplat/optimizers/selective_adam.py.
//   m = b1 m + (1-b1) g ; v = b2 v + (1-b2) g^2 ; p -= lr m / (sqrt(v) + eps)      (no bias correction, as there)
// for the M = numel / N values of every Gaussian n with visibility[n] != 0; invisible Gaussians keep parameter
// AND moments untouched (that is the point: their zero gradients must not decay the moments).
// The reference indexes the mask with element / size(-1), which runs past the mask for [N, K, 3] tensors; here
// the Gaussian of element e is e / M with M = numel / N, what the Python side computes (selective_adam.py:79).
// HBM-bound streaming kernel: 16 B read + 12 B written per visible value, 1 B per invisible Gaussian.
#include "common.cuh"

namespace {

__global__ void __launch_bounds__(256)
selective_adam_kernel(int64_t total, int64_t M, float *__restrict__ param, const float *__restrict__ grad,
                      float *__restrict__ exp_avg, float *__restrict__ exp_avg_sq,
                      const uint8_t *__restrict__ visibility, float lr, float b1, float b2, float eps) {
    for (int64_t e = (int64_t)blockIdx.x * 256 + threadIdx.x; e < total; e += (int64_t)gridDim.x * 256) {
        if (visibility != nullptr && !visibility[e / M]) continue;
        const float g = grad[e];
        const float m = b1 * exp_avg[e] + (1.0f - b1) * g;
        const float v = b2 * exp_avg_sq[e] + (1.0f - b2) * g * g;
        param[e] += -lr * m / (sqrtf(v) + eps);
        exp_avg[e] = m;
        exp_avg_sq[e] = v;
    }
}

}  // namespace

extern "C" int gsb_selective_adam(int64_t N, int64_t M, float *param, const float *grad, float *exp_avg,
                                  float *exp_avg_sq, const uint8_t *visibility, float lr, float b1, float b2,
                                  float eps, cudaStream_t stream) {
    GSB_REQUIRE(N >= 0 && M >= 1, "bad sizes");
    const int64_t total = N * M;
    if (total == 0) return GSB_OK;
    int64_t blocks = (total + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    selective_adam_kernel<<<(unsigned)blocks, 256, 0, stream>>>(total, M, param, grad, exp_avg, exp_avg_sq, visibility,
                                                               lr, b1, b2, eps);
    return gsb_check_launch("selective_adam");
}

// ---- MCMC relocation (gsplat/relocation.py:10-50, RelocationCUDA.cu:13-45; eq. 9 of arXiv 2404.09591) -----------
// A Gaussian split into n copies keeps its rendered contribution when each copy gets
//   o' = 1 - (1 - o)^(1/n),   s' = s * o / sum_{i=1..n} sum_{k=0..i-1} B[i-1][k] (-1)^k o'^(k+1) / sqrt(k+1)
// with B the caller's binomial table [n_max, n_max]. The double sum is taken column-wise here
// (sum_k term_k * sum_{i>=k} B[i][k]) so each power and square root is formed once per k, by recurrence.
namespace {

__global__ void __launch_bounds__(256)
relocation_kernel(int64_t N, const float *__restrict__ opacities, const float *__restrict__ scales,
                  const int32_t *__restrict__ ratios, const float *__restrict__ binoms, int n_max,
                  float *__restrict__ new_opacities, float *__restrict__ new_scales) {
    const int64_t n = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (n >= N) return;
    const int copies = ratios[n];
    const float o = opacities[n];
    const float o_new = 1.0f - powf(1.0f - o, 1.0f / (float)copies);
    float denom = 0.0f, power = o_new;  // power = (-1)^k o_new^(k+1)
    for (int k = 0; k < copies; ++k) {
        float column = 0.0f;
        for (int i = k; i < copies; ++i) column += binoms[(int64_t)i * n_max + k];
        denom = fmaf(column * power, rsqrtf((float)(k + 1)), denom);
        power *= -o_new;
    }
    const float coeff = o / denom;
    new_opacities[n] = o_new;
    for (int a = 0; a < 3; ++a) new_scales[n * 3 + a] = coeff * scales[n * 3 + a];
}

}  // namespace

extern "C" int gsb_relocation(int64_t N, const float *opacities, const float *scales, const int32_t *ratios,
                              const float *binoms, int n_max, float *new_opacities, float *new_scales,
                              cudaStream_t stream) {
    GSB_REQUIRE(N >= 0 && n_max >= 1, "bad sizes");
    if (N == 0) return GSB_OK;
    relocation_kernel<<<(unsigned)((N + 255) / 256), 256, 0, stream>>>(N, opacities, scales, ratios, binoms, n_max,
                                                                      new_opacities, new_scales);
    return gsb_check_launch("relocation");
}
