// b200ml — counter-based dropout RNG shared by every kernel that draws or re-draws a mask.
// A mask bit depends only on (seed, step, site, element index), so the backward pass regenerates it
// instead of storing it, and a CUDA-graph replay gets fresh masks because `step` lives in device memory.
#pragma once
#include <stdint.h>

namespace b200 {

struct RngState {       // device-resident, 16 bytes
    uint64_t seed;
    uint64_t step;      // bumped once per training step by b200ml_rng_advance
};

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// key for one dropout site in one step
__host__ __device__ __forceinline__ uint64_t rng_key(uint64_t seed, uint64_t step, uint32_t site) {
    return mix64(seed ^ mix64(step * 0x9E3779B97F4A7C15ull + site));
}

// 64 random bits for the group of four consecutive elements starting at element index 4*group
__host__ __device__ __forceinline__ uint64_t rng_bits4(uint64_t key, uint64_t group) {
    return mix64(key + group * 0x9E3779B97F4A7C15ull);
}

// keep decision for element `idx` (16-bit resolution): true = keep
__host__ __device__ __forceinline__ bool rng_keep(uint64_t key, uint64_t idx, uint32_t thr16) {
    uint64_t b = rng_bits4(key, idx >> 2);
    uint32_t r = static_cast<uint32_t>(b >> ((idx & 3) * 16)) & 0xFFFFu;
    return r >= thr16;
}

__host__ __device__ __forceinline__ uint32_t rng_thr16(float p) {
    float t = p * 65536.0f;
    return t <= 0.f ? 0u : (t >= 65535.f ? 65535u : static_cast<uint32_t>(t + 0.5f));
}

}  // namespace b200
