// Fused attention core on tcgen05: softmax(Q K^T / sqrt(D)) V.
// models/unet_cond.py:36, 49 (nn.MultiheadAttention, 4 heads, batch_first).
//
// One CTA (128 threads) owns one (sample, head) and a range of 128-query blocks.  The head's
// K and V^T are staged ONCE in shared memory in the UMMA "no-swizzle, K-major" canonical
// layout (8-row x 16-byte core matrices; LBO = 128 B between core matrices along K, SBO
// between 8-row groups), written by the CTA's own threads; each query block then stages Q.
// Scores S = Q K_j^T for 64-key blocks go to TMEM (fp32, lane = query); the 128 threads read
// their own row with tcgen05.ld, so the softmax needs no shuffles.  ONE pass with an online
// softmax: a running row maximum that is only raised when a block exceeds it by more than
// 2^8 (then O, which lives in TMEM, is rescaled by the warp with tcgen05.ld / .st); P =
// exp2(..) goes to shared memory as bf16 and O += P V_j accumulates in TMEM.
// Pipeline: S is double-buffered in TMEM (S_{j+2} is issued as soon as S_j has been read) and
// P is double-buffered in shared memory (P V_j runs while P_{j+1} is produced), so the
// tensor-pipe latency is hidden behind the exp work, which is what bounds this kernel
// (MUFU / issue bound, not tensor bound: d_head is 16..64).
// 192 TMEM columns (256 allocated) and <= 90 KB shared memory: two CTAs per SM.
#pragma once

#include "common.cuh"

namespace dmb {

constexpr int kAttnQ = 128;   // queries per block (MMA M)
constexpr int kAttnK = 64;    // keys per block (MMA N of S, K of P V)

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
  const int lpad = (L + kAttnK - 1) / kAttnK * kAttnK;
  return static_cast<size_t>(kAttnQ) * D * 2 + 2ull * lpad * D * 2 + 2ull * kAttnQ * kAttnK * 2 + 128;
}

template <int D>
__global__ void __launch_bounds__(128, 2)
attention_tc_kernel(const __nv_bfloat16* __restrict__ qkv, Geo g, int E, __nv_bfloat16* __restrict__ out,
                    int qblocks_per_cta) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t s_bar[2];    // S buffer b holds fresh scores
  __shared__ __align__(8) uint64_t pv_bar[2];   // the P V MMA that read P buffer b has retired
  __shared__ uint32_t tmem_base_smem;

  const int tid = threadIdx.x, warp = tid >> 5;
  const int L = g.H * g.W;
  const int nkb = (L + kAttnK - 1) / kAttnK;
  const int lpad = nkb * kAttnK;
  const int nqb = (L + kAttnQ - 1) / kAttnQ;
  const int b = blockIdx.z, h = blockIdx.y;
  const int qb_begin = blockIdx.x * qblocks_per_cta;
  const int qb_end = min(nqb, qb_begin + qblocks_per_cta);
  const int ld = 3 * E;

  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((128u - (raw & 127u)) & 127u);
  uint8_t* Qs = smem;
  uint8_t* Ks = Qs + kAttnQ * D * 2;
  uint8_t* Vt = Ks + static_cast<size_t>(lpad) * D * 2;
  uint8_t* Ps = Vt + static_cast<size_t>(lpad) * D * 2;   // two buffers of kAttnQ x kAttnK bf16

  if (tid == 0) {
    mbar_init(&s_bar[0], 1);
    mbar_init(&s_bar[1], 1);
    mbar_init(&pv_bar[0], 1);
    mbar_init(&pv_bar[1], 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc(&tmem_base_smem, 256);

  // ---- stage K and V^T of the head (once per CTA); 4 keys in flight per thread
  {
    const uint4 zero = make_uint4(0, 0, 0, 0);
    constexpr int VPK = D / 8;  // 16-byte vectors per key
    const int total = lpad * VPK;
    for (int i0 = tid; i0 < total; i0 += 128 * 4) {
      uint4 kv[4], vv[4];
      int key[4], c[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int idx = i0 + u * 128;
        c[u] = idx / lpad;
        key[u] = idx - c[u] * lpad;
        kv[u] = zero; vv[u] = zero;
        if (idx < total && key[u] < L) {
          const __nv_bfloat16* rp =
              qkv + static_cast<size_t>(geo_row_of_pixel(g, b, key[u])) * ld + h * D + 8 * c[u];
          kv[u] = *reinterpret_cast<const uint4*>(rp + E);
          vv[u] = *reinterpret_cast<const uint4*>(rp + 2 * E);
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (i0 + u * 128 < total) {
          *reinterpret_cast<uint4*>(Ks + cm_off(key[u], 8 * c[u], D)) = kv[u];
          const __nv_bfloat16* ve = reinterpret_cast<const __nv_bfloat16*>(&vv[u]);
#pragma unroll
          for (int i = 0; i < 8; ++i)
            *reinterpret_cast<__nv_bfloat16*>(Vt + cm_off(8 * c[u] + i, key[u], lpad)) = ve[i];
        }
      }
    }
  }

  const uint32_t q_addr = smem_u32(Qs), k_addr = smem_u32(Ks), v_addr = smem_u32(Vt), p_addr = smem_u32(Ps);
  const uint32_t sbo_qk = (D / 8) * 128, sbo_v = (lpad / 8) * 128, sbo_p = (kAttnK / 8) * 128;
  const uint32_t idesc_s = umma_idesc(1, kAttnQ, kAttnK);
  const uint32_t idesc_o = umma_idesc(1, kAttnQ, D);
  const float c2 = rsqrtf(static_cast<float>(D)) * 1.4426950408889634f;  // 1/sqrt(D) * log2(e)
  constexpr uint32_t kColO = 2 * kAttnK;                                   // TMEM: S0 | S1 | O
  constexpr uint32_t kPBytes = kAttnQ * kAttnK * 2;

  uint32_t gi = 0;       // key blocks processed so far by this CTA (drives buffer / phase choice)
  uint32_t tmem = 0, t_lane = 0;

  for (int qb = qb_begin; qb < qb_end; ++qb) {
    // ---- stage this block's queries
    const int qi = qb * kAttnQ + tid;
    const bool q_valid = qi < L;
    const int qrow = q_valid ? geo_row_of_pixel(g, b, qi) : 0;
    {
      const uint4 zero = make_uint4(0, 0, 0, 0);
      const __nv_bfloat16* qp = qkv + static_cast<size_t>(qrow) * ld + h * D;
#pragma unroll
      for (int c = 0; c < D / 8; ++c) {
        const uint4 v = q_valid ? *reinterpret_cast<const uint4*>(qp + 8 * c) : zero;
        *reinterpret_cast<uint4*>(Qs + cm_off(tid, 8 * c, D)) = v;
      }
    }
    fence_async_proxy();  // generic-proxy smem writes -> visible to the tensor core (async proxy)
    tc_fence_before();
    __syncthreads();      // also: every thread has finished reading O of the previous block
    tc_fence_after();
    tmem = tmem_base_smem;
    t_lane = tmem + (static_cast<uint32_t>(warp * 32) << 16);

    auto issue_scores = [&](int j, uint32_t buf) {  // S_j = Q K_j^T -> TMEM columns [buf*64, buf*64+64)
#pragma unroll
      for (int kk = 0; kk < D / 16; ++kk) {
        const uint64_t ad = umma_desc_noswz(q_addr + kk * 256, sbo_qk);
        const uint64_t bd = umma_desc_noswz(k_addr + j * (kAttnK / 8) * sbo_qk + kk * 256, sbo_qk);
        umma_ss<false>(tmem + buf * kAttnK, ad, bd, idesc_s, kk > 0 ? 1u : 0u);
      }
      umma_commit(&s_bar[buf]);
    };
    if (tid == 0) {
      issue_scores(0, gi & 1);
      if (nkb > 1) issue_scores(1, (gi + 1) & 1);
    }

    float m_run = -INFINITY, rowsum = 0.f;
    for (int j = 0; j < nkb; ++j) {
      const uint32_t G = gi + j;
      const uint32_t buf = G & 1, ph = (G >> 1) & 1;
      mbar_wait(&s_bar[buf], ph);
      __syncwarp();
      tc_fence_after();
      float s[kAttnK];
      tmem_ld32(t_lane + buf * kAttnK, s);
      tmem_ld32(t_lane + buf * kAttnK + 32, s + 32);
      tmem_ld_wait();
      const int kbase = j * kAttnK;
      if (kbase + kAttnK > L) {
#pragma unroll
        for (int i = 0; i < kAttnK; ++i)
          if (kbase + i >= L) s[i] = -INFINITY;
      }
      float bm;
      {  // four independent chains: a serial max over 64 values would expose 64 x 4 cycles
        float m0 = s[0], m1 = s[1], m2 = s[2], m3 = s[3];
#pragma unroll
        for (int i = 4; i < kAttnK; i += 4) {
          m0 = fmaxf(m0, s[i]); m1 = fmaxf(m1, s[i + 1]); m2 = fmaxf(m2, s[i + 2]); m3 = fmaxf(m3, s[i + 3]);
        }
        bm = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
      }
      if (j == 0) {
        m_run = bm;
      } else {
        const bool need = (bm - m_run) * c2 > 8.0f;
        if (__any_sync(0xffffffffu, need)) {
          // raise the running maximum: O (TMEM) and the row sum are rescaled.  All P V MMAs
          // issued so far must have retired before O is touched.
          const uint32_t Gp = G - 1;
          mbar_wait(&pv_bar[Gp & 1], (Gp >> 1) & 1);
          __syncwarp();
          tc_fence_after();
          const float m_new = need ? bm : m_run;
          const float f = ex2_approx((m_run - m_new) * c2);
          rowsum *= f;
          m_run = m_new;
#pragma unroll
          for (int c0 = 0; c0 < D; c0 += 16) {
            float o[16];
            tmem_ld16(t_lane + kColO + c0, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) o[i] *= f;
            tmem_st16(t_lane + kColO + c0, o);
          }
          tmem_st_wait();
        }
      }
      const float mc = m_run * c2;
      {
        float r0 = 0.f, r1 = 0.f, r2 = 0.f, r3 = 0.f;
#pragma unroll
        for (int i = 0; i < kAttnK; i += 4) {
          s[i] = ex2_approx(fmaf(s[i], c2, -mc));   // exp2(-inf) = 0 for masked keys
          s[i + 1] = ex2_approx(fmaf(s[i + 1], c2, -mc));
          s[i + 2] = ex2_approx(fmaf(s[i + 2], c2, -mc));
          s[i + 3] = ex2_approx(fmaf(s[i + 3], c2, -mc));
          r0 += s[i]; r1 += s[i + 1]; r2 += s[i + 2]; r3 += s[i + 3];
        }
        rowsum += (r0 + r1) + (r2 + r3);
      }
      // P buffer `buf` was last read by the P V MMA of block G-2
      if (G >= 2) {
        mbar_wait(&pv_bar[buf], ph ^ 1);
        __syncwarp();
      }
      uint8_t* Pb = Ps + buf * kPBytes;
#pragma unroll
      for (int gq = 0; gq < kAttnK / 8; ++gq) {
        uint4 u;
        __nv_bfloat162* hp = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
        for (int i = 0; i < 4; ++i) hp[i] = __floats2bfloat162_rn(s[gq * 8 + 2 * i], s[gq * 8 + 2 * i + 1]);
        *reinterpret_cast<uint4*>(Pb + cm_off(tid, gq * 8, kAttnK)) = u;
      }
      fence_async_proxy();
      tc_fence_before();
      __syncthreads();
      if (tid == 0) {
        tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < kAttnK / 16; ++kk) {
          const uint64_t ad = umma_desc_noswz(p_addr + buf * kPBytes + kk * 256, sbo_p);
          const uint64_t bd = umma_desc_noswz(v_addr + (j * (kAttnK / 8) + kk * 2) * 128, sbo_v);
          umma_ss<false>(tmem + kColO, ad, bd, idesc_o, (j > 0 || kk > 0) ? 1u : 0u);
        }
        umma_commit(&pv_bar[buf]);
        if (j + 2 < nkb) issue_scores(j + 2, buf);
      }
    }
    // ---- epilogue of the query block: O / rowsum -> out
    {
      const uint32_t Gl = gi + nkb - 1;
      mbar_wait(&pv_bar[Gl & 1], (Gl >> 1) & 1);
      __syncwarp();
      tc_fence_after();
      float o[D];
#pragma unroll
      for (int c0 = 0; c0 < D; c0 += 16) tmem_ld16(t_lane + kColO + c0, o + c0);
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
    gi += nkb;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base_smem, 256);
  }
}

}  // namespace dmb
