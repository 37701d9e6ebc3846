// Fused attention core on tcgen05: softmax(Q K^T / sqrt(D)) V for one (sample, head, 128-query
// block) per CTA.  models/unet_cond.py:36, 49 (nn.MultiheadAttention, 4 heads).
//
// The head's Q block, all K and V^T live in shared memory in the UMMA "no-swizzle, K-major"
// canonical layout (8-row x 16-byte core matrices; LBO = 128 B between core matrices along K,
// SBO between 8-row groups), written by the CTA's own threads.  Scores S = Q K^T go to TMEM
// (fp32, 128 lanes = 128 queries); the 128 threads read their own row with tcgen05.ld, so the
// softmax needs no shuffles.  Two passes over the key blocks: pass 1 finds the row maximum
// (scores are recomputed in pass 2 -- with d_head 16..64 the MMAs are nearly free and this
// avoids rescaling O), pass 2 forms P = exp2(..) in bf16 in shared memory and accumulates
// O += P V in TMEM.  The work is bound by the exp throughput, not by the tensor pipe.
// 128 threads, 256 TMEM columns, <= 92 KB shared memory: two CTAs per SM overlap each other.
#pragma once

#include "common.cuh"

namespace dmb {

constexpr int kAttnBlock = 128;  // queries per CTA and keys per block

// byte offset of element (r, k) in a K-major no-swizzle operand whose K extent is kx elements
__device__ __forceinline__ uint32_t cm_off(int r, int k, int kx) {
  return static_cast<uint32_t>((r >> 3) * (kx >> 3) * 128 + (k >> 3) * 128 + (r & 7) * 16 + (k & 7) * 2);
}
__device__ __forceinline__ uint64_t umma_desc_noswz(uint32_t smem_addr, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr >> 4) & 0x3FFF);
  d |= static_cast<uint64_t>(128 >> 4) << 16;                 // LBO: next core matrix along K
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFF) << 32;  // SBO: next 8-row group
  d |= static_cast<uint64_t>(1) << 46;
  return d;
}

inline size_t attention_tc_smem_bytes(int L, int D) {
  const int lpad = (L + kAttnBlock - 1) / kAttnBlock * kAttnBlock;
  return static_cast<size_t>(kAttnBlock) * D * 2 + 2ull * lpad * D * 2 + kAttnBlock * kAttnBlock * 2 + 128;
}

template <int D>
__global__ void __launch_bounds__(128)
attention_tc_kernel(const __nv_bfloat16* __restrict__ qkv, Geo g, int E, __nv_bfloat16* __restrict__ out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar[2];
  __shared__ uint32_t tmem_base_smem;

  const int tid = threadIdx.x, warp = tid >> 5;
  const int L = g.H * g.W;
  const int nkb = (L + kAttnBlock - 1) / kAttnBlock;
  const int lpad = nkb * kAttnBlock;
  const int b = blockIdx.z, h = blockIdx.y, qb = blockIdx.x;
  const int ld = 3 * E;

  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((128u - (raw & 127u)) & 127u);
  uint8_t* Qs = smem;
  uint8_t* Ks = Qs + kAttnBlock * D * 2;
  uint8_t* Vt = Ks + static_cast<size_t>(lpad) * D * 2;
  uint8_t* Ps = Vt + static_cast<size_t>(lpad) * D * 2;

  if (tid == 0) {
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc(&tmem_base_smem, 256);

  // ---- stage Q (this block's queries), K and V^T of the head
  const int qi = qb * kAttnBlock + tid;
  const bool q_valid = qi < L;
  const int qrow = q_valid ? geo_row(g, b, qi / g.W, qi % g.W) : 0;
  {
    const uint4 zero = make_uint4(0, 0, 0, 0);
    const __nv_bfloat16* qp = qkv + static_cast<size_t>(qrow) * ld + h * D;
#pragma unroll
    for (int c = 0; c < D / 8; ++c) {
      const uint4 v = q_valid ? *reinterpret_cast<const uint4*>(qp + 8 * c) : zero;
      *reinterpret_cast<uint4*>(Qs + cm_off(tid, 8 * c, D)) = v;
    }
    for (int idx = tid; idx < lpad * (D / 8); idx += 128) {
      const int c = idx / lpad, key = idx - c * lpad;
      uint4 kv = zero, vv = zero;
      if (key < L) {
        const __nv_bfloat16* rp = qkv + static_cast<size_t>(geo_row(g, b, key / g.W, key % g.W)) * ld + h * D + 8 * c;
        kv = *reinterpret_cast<const uint4*>(rp + E);
        vv = *reinterpret_cast<const uint4*>(rp + 2 * E);
      }
      *reinterpret_cast<uint4*>(Ks + cm_off(key, 8 * c, D)) = kv;
      const __nv_bfloat16* ve = reinterpret_cast<const __nv_bfloat16*>(&vv);
#pragma unroll
      for (int i = 0; i < 8; ++i)
        *reinterpret_cast<__nv_bfloat16*>(Vt + cm_off(8 * c + i, key, lpad)) = ve[i];
    }
  }
  fence_async_proxy();  // generic-proxy smem writes -> visible to the tensor core (async proxy)
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_smem;
  const uint32_t t_lane = tmem + (static_cast<uint32_t>(warp * 32) << 16);

  const uint32_t q_addr = smem_u32(Qs), k_addr = smem_u32(Ks), v_addr = smem_u32(Vt), p_addr = smem_u32(Ps);
  const uint32_t sbo_qk = (D / 8) * 128, sbo_v = (lpad / 8) * 128, sbo_p = (kAttnBlock / 8) * 128;
  const uint32_t idesc_s = umma_idesc(1, kAttnBlock, kAttnBlock);
  const uint32_t idesc_o = umma_idesc(1, kAttnBlock, D);

  auto issue_scores = [&](int j, uint32_t tcol) {  // S_j = Q K_j^T  -> TMEM columns [tcol, tcol+128)
#pragma unroll
    for (int kk = 0; kk < D / 16; ++kk) {
      const uint64_t ad = umma_desc_noswz(q_addr + kk * 256, sbo_qk);
      const uint64_t bd = umma_desc_noswz(k_addr + j * (kAttnBlock / 8) * sbo_qk + kk * 256, sbo_qk);
      umma_ss<false>(tmem + tcol, ad, bd, idesc_s, kk > 0 ? 1u : 0u);
    }
  };

  const float c2 = rsqrtf(static_cast<float>(D)) * 1.4426950408889634f;  // 1/sqrt(D) * log2(e)
  uint32_t ph[2] = {0, 0};

  // ---- pass 1: row maximum, S ping-pongs between TMEM columns [0,128) and [128,256)
  float m = -INFINITY;
  if (tid == 0) {
    issue_scores(0, 0);
    umma_commit(&bar[0]);
  }
  for (int j = 0; j < nkb; ++j) {
    const int buf = j & 1;
    if (tid == 0 && j + 1 < nkb) {
      issue_scores(j + 1, (buf ^ 1) * kAttnBlock);
      umma_commit(&bar[buf ^ 1]);
    }
    mbar_wait(&bar[buf], ph[buf]);
    ph[buf] ^= 1;
    __syncwarp();
    tc_fence_after();
    const int kbase = j * kAttnBlock;
#pragma unroll
    for (int c0 = 0; c0 < kAttnBlock; c0 += 32) {
      float s[32];
      tmem_ld32(t_lane + buf * kAttnBlock + c0, s);
      tmem_ld_wait();
      if (kbase + c0 + 32 <= L) {
#pragma unroll
        for (int i = 0; i < 32; ++i) m = fmaxf(m, s[i]);
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (kbase + c0 + i < L) m = fmaxf(m, s[i]);
      }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
  }
  const float mc = m * c2;

  // ---- pass 2: P = exp2(S c - m c) -> smem (bf16), O += P V_j in TMEM columns [128, 128+D)
  float rowsum = 0.f;
  if (tid == 0) {
    issue_scores(0, 0);
    umma_commit(&bar[0]);
  }
  for (int j = 0; j < nkb; ++j) {
    // bar[0] completes when S_j (and, issued before it, O += P V_{j-1}) has retired
    mbar_wait(&bar[0], ph[0]);
    ph[0] ^= 1;
    __syncwarp();
    tc_fence_after();
    const int kbase = j * kAttnBlock;
#pragma unroll
    for (int c0 = 0; c0 < kAttnBlock; c0 += 32) {
      float s[32];
      tmem_ld32(t_lane + c0, s);
      tmem_ld_wait();
      const bool full = kbase + c0 + 32 <= L;
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        float p = exp2f(fmaf(s[i], c2, -mc));
        if (!full && kbase + c0 + i >= L) p = 0.f;
        s[i] = p;
        rowsum += p;
      }
#pragma unroll
      for (int gq = 0; gq < 4; ++gq) {
        uint4 u;
        __nv_bfloat162* hp = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
        for (int i = 0; i < 4; ++i) hp[i] = __floats2bfloat162_rn(s[gq * 8 + 2 * i], s[gq * 8 + 2 * i + 1]);
        *reinterpret_cast<uint4*>(Ps + cm_off(tid, c0 + gq * 8, kAttnBlock)) = u;
      }
    }
    fence_async_proxy();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
#pragma unroll
      for (int kk = 0; kk < kAttnBlock / 16; ++kk) {
        const uint64_t ad = umma_desc_noswz(p_addr + kk * 256, sbo_p);
        const uint64_t bd = umma_desc_noswz(v_addr + (j * kAttnBlock / 8 + kk * 2) * 128, sbo_v);
        umma_ss<false>(tmem + kAttnBlock, ad, bd, idesc_o, (j > 0 || kk > 0) ? 1u : 0u);
      }
      if (j + 1 < nkb) issue_scores(j + 1, 0);
      umma_commit(&bar[0]);
    }
  }
  mbar_wait(&bar[0], ph[0]);
  __syncwarp();
  tc_fence_after();

  // ---- epilogue: O / rowsum -> out
  {
    float o[D];
#pragma unroll
    for (int c0 = 0; c0 < D; c0 += 16) tmem_ld16(t_lane + kAttnBlock + c0, o + c0);
    tmem_ld_wait();
    if (q_valid) {
      const float inv = 1.0f / rowsum;
      __nv_bfloat16* op = out + static_cast<size_t>(qrow) * E + h * D;
#pragma unroll
      for (int c0 = 0; c0 < D; c0 += 8) {
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = o[c0 + i] * inv;
        store8(op + c0, v);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}

}  // namespace dmb
