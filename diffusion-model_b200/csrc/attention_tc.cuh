// Fused attention core on tcgen05: softmax(Q K^T / sqrt(D)) V.
// models/unet_cond.py:36, 49 (nn.MultiheadAttention, 4 heads, batch_first).
//
// One CTA (320 threads) owns one (sample, head) and a range of 128-query blocks.  The head's
// K and V^T are staged ONCE in shared memory in the UMMA "no-swizzle, K-major" canonical
// layout (8-row x 16-byte core matrices; LBO = 128 B between core matrices along K, SBO
// between 8-row groups), written by the CTA's own threads.
// Two independent CHAINS run in the CTA; each is one softmax warpgroup (4 warps, thread =
// query row) plus its own MMA-issuing warp, and works through every second query block:
//   MMA warp : S_j = Q K_j^T (64 keys) -> TMEM;  O += P_j V_j and R += P_j 1 -> TMEM
//   softmax  : tcgen05.ld of the own row of S_j (no shuffles) -> release S -> online max with
//              lazy rescale of O/R in TMEM (only when a block exceeds the running maximum by
//              more than 2^8) -> P = exp2(..) as bf16 into shared memory -> signal the MMA warp
// All hand-offs are mbarriers (s_bar, sfree_bar, p_bar, pv_bar); there is no CTA-wide barrier
// in the key loop and no softmax thread ever issues an MMA.  S_{j+1} is issued as soon as S_j
// has been read, so one S buffer per chain is enough.  The row sum comes from the tensor core
// too (R = P x ones), from exactly the bf16 P that enters P V; P is rounded by truncation (a
// byte permute per pair instead of an F2FP per pair, which shares the MUFU pipe) and the
// truncation bias cancels in O / R.
// The kernel is bound by MUFU.EX2 (one per score): two CTAs per SM give 4 softmax warps per
// scheduler to keep that pipe fed.  TMEM: 2 x (64 + D + 16) columns.
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

// exp2 on the FMA pipe for a share of the scores (the MUFU pipe bounds this kernel): round to
// the nearest integer with the 1.5 * 2^23 trick, cubic minimax polynomial for 2^f on
// [-0.5, 0.5] (relative error 7.5e-5, below the bf16 rounding of P), exponent added as an integer.
#ifndef DMB_ATTN_POLY_MASK
#define DMB_ATTN_POLY_MASK 3   // scores with (i & 3) == MASK take the FMA-pipe exp2; -1 disables
#endif
constexpr bool kPolyExp = DMB_ATTN_POLY_MASK >= 0;
__device__ __forceinline__ float ex2_poly(float x) {
  x = fmaxf(x, -120.0f);
  const float t = x + 12582912.0f;
  const float f = x - (t - 12582912.0f);
  float p = fmaf(f, 0.05517145f, 0.24261062f);
  p = fmaf(p, f, 0.69326097f);
  p = fmaf(p, f, 0.9999281f);
  return __uint_as_float(__float_as_uint(p) + (__float_as_uint(t) << 23));
}


inline size_t attention_tc_smem_bytes(int L, int D) {
  const int lpad = (L + kAttnK - 1) / kAttnK * kAttnK;
  return 2ull * kAttnQ * D * 2 + 2ull * lpad * D * 2 + 2ull * kAttnQ * kAttnK * 2 + 16 * kAttnK * 2 + 128;
}

constexpr int kAttnThreads = 320;   // 2 chains x (4 softmax warps + 1 MMA warp)

template <int D>
__global__ void __launch_bounds__(kAttnThreads, 2)
attention_tc_kernel(const __nv_bfloat16* __restrict__ qkv, Geo g, int E, __nv_bfloat16* __restrict__ out,
                    int qblocks_per_cta) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t s_bar[2];      // chain c: S holds fresh scores        (MMA -> softmax)
  __shared__ __align__(8) uint64_t sfree_bar[2];  // chain c: S has been read             (softmax -> MMA)
  __shared__ __align__(8) uint64_t p_bar[2];      // chain c: P is in shared memory       (softmax -> MMA)
  __shared__ __align__(8) uint64_t pv_bar[2];     // chain c: the P V MMA has retired     (MMA -> softmax)
  __shared__ uint32_t tmem_base_smem;

  pdl_trigger();
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int L = g.H * g.W;
  const int nkb = (L + kAttnK - 1) / kAttnK;
  const int lpad = nkb * kAttnK;
  const int nqb = (L + kAttnQ - 1) / kAttnQ;
  const int b = blockIdx.z, h = blockIdx.y;
  const int qb_begin = blockIdx.x * qblocks_per_cta;
  const int qb_end = min(nqb, qb_begin + qblocks_per_cta);
  const int ld = 3 * E;
  constexpr uint32_t kQBytes = kAttnQ * D * 2;
  constexpr uint32_t kPBytes = kAttnQ * kAttnK * 2;
  constexpr uint32_t kChainCols = kAttnK + D + 16;   // TMEM per chain: S | O | R
  // d_head 64 (7x7 maps: a single query block) runs one chain, so that 256 columns and two
  // CTAs per SM are enough there as well
  constexpr int kChains = 2 * kChainCols <= 256 ? 2 : 1;
  constexpr uint32_t kTmemCols = 256;
  static_assert(kChains * kChainCols <= kTmemCols, "TMEM budget");

  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((128u - (raw & 127u)) & 127u);
  uint8_t* Qs = smem;                                        // one Q tile per chain
  uint8_t* Ks = Qs + 2 * kQBytes;
  uint8_t* Vt = Ks + static_cast<size_t>(lpad) * D * 2;
  uint8_t* Ps = Vt + static_cast<size_t>(lpad) * D * 2;      // one P tile per chain
  uint8_t* Ones = Ps + 2 * kPBytes;                          // B operand [16 x kAttnK]: row 0 = 1, rest 0

  if (tid == 0) {
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      mbar_init(&s_bar[c], 1);
      mbar_init(&sfree_bar[c], 4);
      mbar_init(&p_bar[c], 4);
      mbar_init(&pv_bar[c], 1);
    }
    fence_mbar_init();
  }
  if (warp == 8) tmem_alloc(&tmem_base_smem, kTmemCols);
  for (int i = tid; i < 16 * kAttnK; i += kAttnThreads) {
    const int n = i / kAttnK, k = i % kAttnK;
    *reinterpret_cast<__nv_bfloat16*>(Ones + cm_off(n, k, kAttnK)) = __float2bfloat16_rn(n == 0 ? 1.0f : 0.0f);
  }

  pdl_wait();   // (barrier init, TMEM allocation and the ones tile overlap the previous kernel's tail)
  // ---- stage K and V^T of the head (once per CTA); 4 keys in flight per thread
  {
    const uint4 zero = make_uint4(0, 0, 0, 0);
    constexpr int VPK = D / 8;  // 16-byte vectors per key
    const int total = lpad * VPK;
    for (int i0 = tid; i0 < total; i0 += kAttnThreads * 4) {
      uint4 kv[4], vv[4];
      int key[4], c[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int idx = i0 + u * kAttnThreads;
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
        if (i0 + u * kAttnThreads < total) {
          *reinterpret_cast<uint4*>(Ks + cm_off(key[u], 8 * c[u], D)) = kv[u];
          const __nv_bfloat16* ve = reinterpret_cast<const __nv_bfloat16*>(&vv[u]);
#pragma unroll
          for (int i = 0; i < 8; ++i)
            *reinterpret_cast<__nv_bfloat16*>(Vt + cm_off(8 * c[u] + i, key[u], lpad)) = ve[i];
        }
      }
    }
  }
  fence_async_proxy();  // generic-proxy smem writes -> visible to the tensor core (async proxy)
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_smem;

  const uint32_t sbo_qk = (D / 8) * 128, sbo_v = (lpad / 8) * 128, sbo_p = (kAttnK / 8) * 128;
  const float c2 = rsqrtf(static_cast<float>(D)) * 1.4426950408889634f;  // 1/sqrt(D) * log2(e)

  if (warp >= 8) {
    // =================== MMA warp of chain c (warp-uniform; one elected lane issues) ===================
    const int c = warp - 8;
    const uint32_t q_addr = smem_u32(Qs) + c * kQBytes, k_addr = smem_u32(Ks), v_addr = smem_u32(Vt);
    const uint32_t p_addr = smem_u32(Ps) + c * kPBytes, ones_addr = smem_u32(Ones);
    const uint32_t idesc_s = umma_idesc(1, kAttnQ, kAttnK);
    const uint32_t idesc_o = umma_idesc(1, kAttnQ, D);
    const uint32_t idesc_r = umma_idesc(1, kAttnQ, 16);
    const uint32_t col_s = tmem + c * kChainCols, col_o = col_s + kAttnK, col_r = col_o + D;
    uint32_t G = 0;
    for (int qb = qb_begin + c; qb < qb_end && c < kChains; qb += kChains) {
      named_bar_sync(1 + c, 160);   // Q of this block is staged; O / S of the previous one are read
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk)
          umma_ss<false>(col_s, umma_desc_noswz(q_addr + kk * 256, sbo_qk),
                         umma_desc_noswz(k_addr + kk * 256, sbo_qk), idesc_s, kk > 0 ? 1u : 0u);
        umma_commit(&s_bar[c]);
      }
      __syncwarp();
      for (int j = 0; j < nkb; ++j, ++G) {
        mbar_wait(&sfree_bar[c], G & 1);
        tc_fence_after();
        if (j + 1 < nkb && elect_one()) {
#pragma unroll
          for (int kk = 0; kk < D / 16; ++kk)
            umma_ss<false>(col_s, umma_desc_noswz(q_addr + kk * 256, sbo_qk),
                           umma_desc_noswz(k_addr + (j + 1) * (kAttnK / 8) * sbo_qk + kk * 256, sbo_qk), idesc_s,
                           kk > 0 ? 1u : 0u);
          umma_commit(&s_bar[c]);
        }
        __syncwarp();
        mbar_wait(&p_bar[c], G & 1);
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < kAttnK / 16; ++kk)
            umma_ss<false>(col_o, umma_desc_noswz(p_addr + kk * 256, sbo_p),
                           umma_desc_noswz(v_addr + (j * (kAttnK / 8) + kk * 2) * 128, sbo_v), idesc_o,
                           (j > 0 || kk > 0) ? 1u : 0u);
#pragma unroll
          for (int kk = 0; kk < kAttnK / 16; ++kk)
            umma_ss<false>(col_r, umma_desc_noswz(p_addr + kk * 256, sbo_p),
                           umma_desc_noswz(ones_addr + kk * 256, sbo_p), idesc_r, (j > 0 || kk > 0) ? 1u : 0u);
          umma_commit(&pv_bar[c]);
        }
        __syncwarp();
      }
    }
  } else {
    // =================== softmax warpgroup of chain c: thread = query row ===================
    const int c = warp >> 2;
    const int row = tid & 127;
    uint8_t* Qc = Qs + c * kQBytes;
    uint8_t* Pc = Ps + c * kPBytes;
    const uint32_t t_lane = tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16) + c * kChainCols;
    constexpr uint32_t kColO = kAttnK;
    uint32_t G = 0;
    for (int qb = qb_begin + c; qb < qb_end && c < kChains; qb += kChains) {
      const int qi = qb * kAttnQ + row;
      const bool q_valid = qi < L;
      const int qrow = q_valid ? geo_row_of_pixel(g, b, qi) : 0;
      {
        const uint4 zero = make_uint4(0, 0, 0, 0);
        const __nv_bfloat16* qp = qkv + static_cast<size_t>(qrow) * ld + h * D;
#pragma unroll
        for (int cc = 0; cc < D / 8; ++cc) {
          const uint4 v = q_valid ? *reinterpret_cast<const uint4*>(qp + 8 * cc) : zero;
          *reinterpret_cast<uint4*>(Qc + cm_off(row, 8 * cc, D)) = v;
        }
      }
      fence_async_proxy();
      tc_fence_before();
      named_bar_sync(1 + c, 160);

      if (qb * kAttnQ + (row & ~31) >= L) {
        // every query row of this warp is padding (tail block): keep the hand-shakes in step
        // and skip the arithmetic.  The P rows stay whatever they were; their O rows are never read.
        for (int j = 0; j < nkb; ++j, ++G) {
          mbar_wait(&s_bar[c], G & 1);
          if (lane == 0) mbar_arrive(&sfree_bar[c]);
          // S_G can be ready before the other warps have delivered P_{G-1}: an arrival now would
          // be counted in that earlier phase
          mbar_wait(&p_bar[c], (G - 1) & 1);
          if (lane == 0) mbar_arrive(&p_bar[c]);
          __syncwarp();
        }
        continue;
      }

      float m_run = -INFINITY;
      for (int j = 0; j < nkb; ++j, ++G) {
        mbar_wait(&s_bar[c], G & 1);
        tc_fence_after();
        float s[kAttnK];
        tmem_ld32(t_lane, s);
        tmem_ld32(t_lane + 32, s + 32);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sfree_bar[c]);   // the MMA warp may overwrite S with the next block
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
        bool pv_done = (j == 0);   // O / R / P of this chain are not in use by the tensor core
        if (j == 0) {
          m_run = bm;
        } else {
          const bool need = (bm - m_run) * c2 > 8.0f;
          if (__any_sync(0xffffffffu, need)) {
            // raise the running maximum: O and R (TMEM) are rescaled once the P V MMAs issued
            // so far have retired
            mbar_wait(&pv_bar[c], (G - 1) & 1);
            tc_fence_after();
            pv_done = true;
            const float m_new = need ? bm : m_run;
            const float f = ex2_approx((m_run - m_new) * c2);
            m_run = m_new;
#pragma unroll
            for (int c0 = 0; c0 < D + 16; c0 += 16) {
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
        const int nval = L - kbase;   // valid keys of this block; whole groups of 16 padding keys cost nothing
        if (!pv_done) mbar_wait(&pv_bar[c], (G - 1) & 1);   // P was last read by the P V MMA of block G-1
        // exp2, bf16 truncation and the shared-memory store of P in ONE loop over 8-key chunks:
        // a warp issues in order, so MUFU instructions that queue for the XU pipe should have
        // independent FMA / ALU / LSU work right behind them rather than 63 more MUFUs.
        // fp32 -> bf16 by truncation: one byte-permute per PAIR (the F2FP pack shares the MUFU pipe);
        // the truncation bias cancels because the row sum comes from the same truncated values.
#pragma unroll
        for (int ch = 0; ch < kAttnK / 8; ++ch) {
          uint4 u = make_uint4(0u, 0u, 0u, 0u);
          if ((ch >> 1) * 16 < nval) {
            float e[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float x = fmaf(s[ch * 8 + i], c2, -mc);
              e[i] = (kPolyExp && (DMB_ATTN_POLY_MASK >= 4 ? (i & 1) == 1 : (i & 3) == DMB_ATTN_POLY_MASK)) ? ex2_poly(x)
                                                                                                            : ex2_approx(x);   // exp2(-inf) = 0 for masked keys
            }
            u.x = __byte_perm(__float_as_uint(e[0]), __float_as_uint(e[1]), 0x7632);
            u.y = __byte_perm(__float_as_uint(e[2]), __float_as_uint(e[3]), 0x7632);
            u.z = __byte_perm(__float_as_uint(e[4]), __float_as_uint(e[5]), 0x7632);
            u.w = __byte_perm(__float_as_uint(e[6]), __float_as_uint(e[7]), 0x7632);
          }
          *reinterpret_cast<uint4*>(Pc + cm_off(row, ch * 8, kAttnK)) = u;
        }
        fence_async_proxy();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_bar[c]);
      }
      // ---- epilogue of the query block: O / R -> out
      if (nkb > 0) {
        mbar_wait(&pv_bar[c], (G - 1) & 1);
        tc_fence_after();
        float o[D + 16];
#pragma unroll
        for (int c0 = 0; c0 < D + 16; c0 += 16) tmem_ld16(t_lane + kColO + c0, o + c0);
        tmem_ld_wait();
        tc_fence_before();
        if (q_valid) {
          const float inv = 1.0f / o[D];   // column 0 of R = sum of the row's P
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
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) {
    tc_fence_after();
    tmem_dealloc(tmem, kTmemCols);
  }
}

}  // namespace dmb
