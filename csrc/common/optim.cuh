// b200ml — K15/K16 update rules as device functions, shared by the single-GPU optimizer kernels and by the
// fused all-reduce epilogues (csrc/comm/allreduce.cu) so both paths produce bit-identical weights.
// Matches torch.optim.SGD (no momentum; distributed_cnn.py:138, distributed_multilayer_perceptron.py:111) and
// torch.optim.Adam defaults (distributed_lstm.py:141, pytorch_machine_translator.py:129).
#pragma once
#include <cuda_bf16.h>
#include <stdint.h>

namespace b200 {

struct OptimSpec {
    int kind;  // 0 = none (reduce only), 1 = SGD, 2 = Adam
    float lr, beta1, beta2, eps, weight_decay;
    const int* step_ptr;  // device step counter (1-based at the time of the update); Adam bias correction
};

struct AdamCorr {
    float step_size;   // lr / (1 - beta1^t)
    float inv_sqrt_bc2;  // 1 / sqrt(1 - beta2^t)
};

__device__ __forceinline__ AdamCorr adam_corr(const OptimSpec& o) {
    const float t = (float)(o.step_ptr ? *o.step_ptr : 1);
    AdamCorr c;
    c.step_size = o.lr / (1.f - powf(o.beta1, t));
    c.inv_sqrt_bc2 = rsqrtf(1.f - powf(o.beta2, t));
    return c;
}

__device__ __forceinline__ float sgd_update(float p, float g, const OptimSpec& o) {
    if (o.weight_decay != 0.f) g += o.weight_decay * p;
    return p - o.lr * g;
}

__device__ __forceinline__ float adam_update(float p, float g, float& m, float& v, const OptimSpec& o,
                                             const AdamCorr& c) {
    if (o.weight_decay != 0.f) g += o.weight_decay * p;
    m = o.beta1 * m + (1.f - o.beta1) * g;
    v = o.beta2 * v + (1.f - o.beta2) * g * g;
    const float denom = sqrtf(v) * c.inv_sqrt_bc2 + o.eps;
    return p - c.step_size * (m / denom);
}

}  // namespace b200
