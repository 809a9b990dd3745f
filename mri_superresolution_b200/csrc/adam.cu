// Fused Adam step over the flat parameter arena: torch.optim.Adam(lr, betas, eps, weight_decay) semantics
// (L2 decay folded into the gradient, bias-corrected), as constructed in scripts/train.py:186 and stepped in :318.
#include "kernels.cuh"
#include <math.h>

namespace mrisr {
namespace {
__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                   float* __restrict__ m, float* __restrict__ v, int64_t n, float lr,
                                                   float b1, float b2, float eps, float wd, float step_size,
                                                   float inv_sqrt_bc2, float grad_scale) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float pi = p[i];
        const float gi = fmaf(wd, pi, g[i] * grad_scale);
        const float mi = m[i] + (gi - m[i]) * (1.f - b1);           // torch: exp_avg.lerp_(grad, 1-beta1)
        const float vi = b2 * v[i] + (1.f - b2) * gi * gi;          // exp_avg_sq.mul_(b2).addcmul_(g, g, 1-b2)
        const float denom = sqrtf(vi) * inv_sqrt_bc2 + eps;
        p[i] = pi - step_size * (mi / denom);
        m[i] = mi;
        v[i] = vi;
    }
}
// CUDA-graph form: the step number and the learning rate live on the device (state: int64 step, float step_size,
// float 1/sqrt(bc2), float lr), so a captured launch advances the step on every replay and follows the scheduler's lr instead
// of freezing the capture-time values.
__global__ void adam_advance_kernel(long long* __restrict__ step, float* __restrict__ sc, float b1, float b2) {
    const long long t = *step + 1;
    *step = t;
    const float lr = sc[2];
    const double bc1 = 1.0 - pow((double)b1, (double)t);
    const double bc2 = 1.0 - pow((double)b2, (double)t);
    sc[0] = (float)((double)lr / bc1);
    sc[1] = (float)(1.0 / sqrt(bc2));
}
__global__ void __launch_bounds__(256) adam_dev_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                       float* __restrict__ m, float* __restrict__ v, int64_t n, float b1,
                                                       float b2, float eps, float wd, const float* __restrict__ sc,
                                                       float grad_scale) {
    const float step_size = sc[0], inv_sqrt_bc2 = sc[1];
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float pi = p[i];
        const float gi = fmaf(wd, pi, g[i] * grad_scale);
        const float mi = m[i] + (gi - m[i]) * (1.f - b1);
        const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
        const float denom = sqrtf(vi) * inv_sqrt_bc2 + eps;
        p[i] = pi - step_size * (mi / denom);
        m[i] = mi;
        v[i] = vi;
    }
}
}  // namespace

int launch_adam_dev(float* p, const float* g, float* m, float* v, int64_t n, float b1, float b2, float eps, float wd,
                    void* state32, float grad_scale, cudaStream_t s) {
    ProfScope prof_("adam", s);
    long long* step = (long long*)state32;
    float* sc = (float*)((char*)state32 + 8);
    adam_advance_kernel<<<1, 1, 0, s>>>(step, sc, b1, b2);
    MRISR_LAUNCH_CHECK();
    int blocks = (int)ceil_div64(n, 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    adam_dev_kernel<<<blocks, 256, 0, s>>>(p, g, m, v, n, b1, b2, eps, wd, sc, grad_scale);
    MRISR_LAUNCH_CHECK();
    return 0;
}

int launch_adam(float* p, const float* g, float* m, float* v, int64_t n, float lr, float b1, float b2, float eps,
                float wd, int step, float grad_scale, cudaStream_t s) {
    ProfScope prof_("adam", s);
    const double bc1 = 1.0 - pow((double)b1, (double)step);
    const double bc2 = 1.0 - pow((double)b2, (double)step);
    const float step_size = (float)((double)lr / bc1);
    const float inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
    int blocks = (int)ceil_div64(n, 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    adam_kernel<<<blocks, 256, 0, s>>>(p, g, m, v, n, lr, b1, b2, eps, wd, step_size, inv_sqrt_bc2, grad_scale);
    MRISR_LAUNCH_CHECK();
    return 0;
}
}  // namespace mrisr
