// Counter-based Gaussian noise for the DDPM step: Philox4x32-10 keyed by the run seed,
// counter = (element-quad index, global sample index, step), then Box-Muller.  The value
// for (sample, step, element) does not depend on how samples are sharded over GPUs, so an
// R-rank run reproduces the 1-rank run element for element.
// (The reference draws from torch's global generator, diff.py:158/327; bit-matching that
// stream is neither possible nor promised -- parity runs inject noise tensors instead.)
#pragma once
#include <stdint.h>

namespace dmb {

struct Philox4 {
  uint32_t v[4];
};

__host__ __device__ inline uint32_t mulhilo32(uint32_t a, uint32_t b, uint32_t* hi) {
#ifdef __CUDA_ARCH__
  *hi = __umulhi(a, b);
  return a * b;
#else
  uint64_t p = static_cast<uint64_t>(a) * b;
  *hi = static_cast<uint32_t>(p >> 32);
  return static_cast<uint32_t>(p);
#endif
}

__host__ __device__ inline Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                 uint32_t k0, uint32_t k1) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    uint32_t hi0, hi1;
    uint32_t lo0 = mulhilo32(M0, c0, &hi0);
    uint32_t lo1 = mulhilo32(M1, c2, &hi1);
    uint32_t n0 = hi1 ^ c1 ^ k0;
    uint32_t n1 = lo1;
    uint32_t n2 = hi0 ^ c3 ^ k1;
    uint32_t n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += W0; k1 += W1;
  }
  Philox4 r;
  r.v[0] = c0; r.v[1] = c1; r.v[2] = c2; r.v[3] = c3;
  return r;
}

// Four N(0,1) values for element quad `quad` of sample `sample` at step `step`.
__device__ inline void philox_normal4(uint64_t seed, uint64_t sample, uint32_t step, uint32_t quad, float* out) {
  Philox4 r = philox4x32_10(quad, static_cast<uint32_t>(sample), static_cast<uint32_t>(sample >> 32), step,
                            static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
  // (0,1] uniforms from 32 bits, then Box-Muller on two pairs.
  const float k = 2.3283064365386963e-10f;  // 2^-32
  float u0 = (static_cast<float>(r.v[0]) + 1.0f) * k;
  float u1 = static_cast<float>(r.v[1]) * k;
  float u2 = (static_cast<float>(r.v[2]) + 1.0f) * k;
  float u3 = static_cast<float>(r.v[3]) * k;
  u0 = fminf(u0, 1.0f); u2 = fminf(u2, 1.0f);
  float r0 = sqrtf(-2.0f * logf(u0));
  float r1 = sqrtf(-2.0f * logf(u2));
  float s0, c0, s1, c1;
  sincospif(2.0f * u1, &s0, &c0);
  sincospif(2.0f * u3, &s1, &c1);
  out[0] = r0 * c0; out[1] = r0 * s0; out[2] = r1 * c1; out[3] = r1 * s1;
}

}  // namespace dmb
