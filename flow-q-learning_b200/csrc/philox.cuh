// Counter-based Philox4x32-10 (Salmon et al., SC'11) used for every random draw of the population step.
// One call = one (seed, step, row, slot) tuple, so draws are order-independent and the numpy restatement in
// oracle/philox_oracle.py reproduces the integer stream bit for bit.
//
// Replaces: np.random.randint in Dataset.sample (fql submodule; call site
// /root/reference/task/offline_task_simulated.py:53-58) and jax.random.{normal,uniform} inside
// FQLAgent.update / sample_actions.
#pragma once
#include <cstdint>

namespace fqlpop {

struct Philox4 {
  uint32_t x, y, z, w;
};

__host__ __device__ __forceinline__ void philox_mulhilo(uint32_t a, uint32_t b, uint32_t& hi, uint32_t& lo) {
  const unsigned long long p = (unsigned long long)a * (unsigned long long)b;
  hi = (uint32_t)(p >> 32);
  lo = (uint32_t)p;
}

// key = seed (lo, hi); counter = (step lo, step hi, row, slot)
__host__ __device__ __forceinline__ Philox4 philox4x32_10(unsigned long long seed, unsigned long long step,
                                                          uint32_t row, uint32_t slot) {
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  uint32_t c0 = (uint32_t)step, c1 = (uint32_t)(step >> 32), c2 = row, c3 = slot;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0, lo0, hi1, lo1;
    philox_mulhilo(0xD2511F53u, c0, hi0, lo0);
    philox_mulhilo(0xCD9E8D57u, c2, hi1, lo1);
    const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return Philox4{c0, c1, c2, c3};
}

// uniform in [0, 1): top 24 bits of word 0
__host__ __device__ __forceinline__ float philox_uniform(unsigned long long seed, unsigned long long step,
                                                         uint32_t row, uint32_t slot) {
  const Philox4 r = philox4x32_10(seed, step, row, slot);
  return (float)(r.x >> 8) * (1.0f / 16777216.0f);
}

// uniform integer in [0, n): multiply-shift of word 0 (n < 2^32)
__host__ __device__ __forceinline__ uint32_t philox_index(unsigned long long seed, unsigned long long step,
                                                          uint32_t row, uint32_t slot, uint32_t n) {
  const Philox4 r = philox4x32_10(seed, step, row, slot);
  return (uint32_t)(((unsigned long long)r.x * (unsigned long long)n) >> 32);
}

// standard normal: Box-Muller on words 0 and 1
__device__ __forceinline__ float philox_normal(unsigned long long seed, unsigned long long step, uint32_t row,
                                               uint32_t slot) {
  const Philox4 r = philox4x32_10(seed, step, row, slot);
  const float u1 = (float)((r.x >> 8) + 1u) * (1.0f / 16777216.0f);  // (0, 1]
  const float u2 = (float)(r.y >> 8) * (1.0f / 16777216.0f);         // [0, 1)
  return sqrtf(-2.0f * logf(u1)) * cospif(2.0f * u2);
}

}  // namespace fqlpop
