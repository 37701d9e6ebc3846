// Attention core on warp-level tensor-core MMAs: softmax(Q K^T / sqrt(D)) V.
// models/unet_cond.py:36, 49 (nn.MultiheadAttention, 4 heads, batch_first).
//
// Why not tcgen05 here (attention_tc.cuh stays for comparison, DMB200_ATTN_MMA=0): with d_head
// 16 / 32 / 64 and 49 .. 784 tokens the tensor work of this block is tiny (7 % tensor-pipe
// activity in the tcgen05 kernel); what bounds it is the softmax -- one MUFU.EX2 and ~4 issue
// slots per score.  The tcgen05 kernel couples the four softmax warps of a 128-query tile to one
// another and to an MMA warp through four mbarriers per 64 keys, stages V transposed with scalar
// stores, and runs 7 query blocks on 2 chains (4 + 3): ncu (profiles/r2_ncu_attention_tc_sa6.md)
// shows 53 % issue utilisation, 37 % XU utilisation, 21 % of the softmax warps' samples waiting
// for S / P V hand-offs and 20 % idle behind the CTA's last query block.  Here every warp is
// independent: it owns 16 query rows, keeps scores, P and the O accumulator in registers
// (mma.sync.m16n8k16, bf16 in, fp32 accumulate), reads K and V fragments with ldmatrix from a
// swizzled shared-memory copy staged once per CTA with 16-byte loads, and never waits for
// another warp after the one barrier behind the staging.  784 tokens = 49 tiles of 16 rows = 7
// warps x 7 tiles: no tail.
//
// Arithmetic (same rounding points as the tcgen05 kernel): S = Q K^T in fp32 from bf16
// operands; online softmax in base 2 with the scale folded into one FFMA per score; P truncated
// to bf16 (one byte-permute per pair: the F2FP pack would share the MUFU pipe); the row sum is
// accumulated by the tensor core from exactly that truncated P (an all-ones B tile), so the
// truncation bias cancels in O / R; O = P V in fp32, divided and rounded to bf16 once.
// The result of a query row does not depend on how tiles are spread over warps / CTAs.
#pragma once

#include <type_traits>

#include "common.cuh"

namespace dmb {

constexpr int kAmKeys = 64;        // keys per softmax step
constexpr int kAmMaxWarps = 7;    // 224 threads: three CTAs per SM may use 96 registers per thread

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void am_mma_16816(float* d, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// exp2 on the FMA pipe (see attention_tc.cuh: cubic minimax on [-0.5, 0.5], rel. error 7.5e-5)
__device__ __forceinline__ float am_ex2_poly(float x) {
  x = fmaxf(x, -120.0f);
  const float t = x + 12582912.0f;
  const float f = x - (t - 12582912.0f);
  float p = fmaf(f, 0.05517145f, 0.24261062f);
  p = fmaf(p, f, 0.69326097f);
  p = fmaf(p, f, 0.9999281f);
  return __uint_as_float(__float_as_uint(p) + (__float_as_uint(t) << 23));
}

// the same for two scores at once on packed fp32x2 instructions (FADD2 / FFMA2, sm_100): 5 issue
// slots per score instead of 8
__device__ __forceinline__ float2 am_ex2_poly2(float2 x) {
  x.x = fmaxf(x.x, -120.0f); x.y = fmaxf(x.y, -120.0f);
  const float2 magic = make_float2(12582912.0f, 12582912.0f);
  const float2 t = __fadd2_rn(x, magic);
  const float2 r = __fadd2_rn(t, make_float2(-12582912.0f, -12582912.0f));
  const float2 f = __ffma2_rn(r, make_float2(-1.0f, -1.0f), x);
  float2 p = __ffma2_rn(f, make_float2(0.05517145f, 0.05517145f), make_float2(0.24261062f, 0.24261062f));
  p = __ffma2_rn(p, f, make_float2(0.69326097f, 0.69326097f));
  p = __ffma2_rn(p, f, make_float2(0.9999281f, 0.9999281f));
  return make_float2(__uint_as_float(__float_as_uint(p.x) + (__float_as_uint(t.x) << 23)),
                     __uint_as_float(__float_as_uint(p.y) + (__float_as_uint(t.y) << 23)));
}

inline size_t attention_mma_smem_bytes(int L, int D) {
  const int lpad = (L + kAmKeys - 1) / kAmKeys * kAmKeys;
  return 2ull * lpad * D * 2 + 16;
}

// 16-byte chunk c of key row r sits at chunk (c ^ swz(r)): the eight rows of one ldmatrix 8x8
// block then fall into eight different 16-byte bank groups for every row length (32 / 64 / 128 B)
template <int D>
__device__ __forceinline__ uint32_t am_off(int r, int c) {
  constexpr int CPR = D / 8;
  return static_cast<uint32_t>((r * CPR + (c ^ ((r / (8 / CPR)) & (CPR - 1)))) * 16);
}

// kPolyMask: scores whose index i within a thread's chunk has (i & 3) == kPolyMask take the
// FMA-pipe exp2 (a quarter of them); -1: all on MUFU.
// kVar (experiment bits): 1 = raise the running maximum only when a step exceeds it by more than 2^8
// (P <= 2^8 stays exact in bf16 / fp32; the rescale of O and its two MUFU then almost never run);
// 2 = row sum by FADD of the fp32 exponentials instead of an all-ones MMA tile, with the truncation
// of P de-biased (see kDebias); 4 = packed fp32x2 arithmetic for the exponent (and for the FMA-pipe
// exp2 when kPolyMask >= 0).
template <int D, int kPolyMask, int kVar>
__global__ void __launch_bounds__(kAmMaxWarps * 32, (D == 16) ? 3 : 2)
attention_mma_kernel(const __nv_bfloat16* __restrict__ qkv, Geo g, int E, __nv_bfloat16* __restrict__ out,
                     int tiles_per_cta) {
  extern __shared__ uint8_t am_smem_raw[];
  constexpr int CPR = D / 8;      // 16-byte chunks per key row
  constexpr bool kLazy = (kVar & 1) != 0, kFaddSum = (kVar & 2) != 0, kPacked = (kVar & 4) != 0;
  constexpr int NO = D / 8 + (kFaddSum ? 0 : 1);   // output n-tiles: D / 8 of V (+ the all-ones tile = row sum)
  // kFaddSum: P is truncated to bf16, which loses 2^-9 of its value on average.  The exponent gets
  // log2(1 + 2^-9) added (folded into the per-row constant, free), so the truncated P is an unbiased
  // estimate of exp2(x) with the same error width as round-to-nearest; the fp32 row sum carries the
  // factor (1 + 2^-9), which the final division takes out again.
  constexpr float kDebias = kFaddSum ? 0.00281502f : 0.f;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, nw = blockDim.x >> 5;
  const int L = g.H * g.W;
  const int nch = (L + kAmKeys - 1) / kAmKeys;
  const int lpad = nch * kAmKeys;
  const int ntiles = (L + 15) >> 4;
  const int b = blockIdx.z, h = blockIdx.y;
  const int ld = 3 * E;
  const uint32_t sraw = smem_u32(am_smem_raw);
  uint8_t* Ks = am_smem_raw + ((16u - (sraw & 15u)) & 15u);
  uint8_t* Vs = Ks + static_cast<size_t>(lpad) * D * 2;

  pdl_trigger();
  pdl_wait();
  // ---- stage K and V of the (sample, head): 16-byte loads, kSU key chunks of each in flight per thread.  d_head
  // 32 / 64: the whole head in one round trip (L = 196, d 32: 82 -> 78 us at batch 256); d_head 16 keeps two per
  // round -- measured with eight: L = 784 419 -> 425 us, L = 196 49 -> 51 us (three CTAs per SM already overlap
  // one CTA's staging with the others' MMAs, and the deeper batch costs registers)
  {
    constexpr int kSU = (D == 16) ? 2 : 8;
    const int total = lpad * CPR;
    const uint4 zero = make_uint4(0, 0, 0, 0);
    for (int i0 = tid; i0 < total; i0 += kSU * blockDim.x) {
      uint4 kv[kSU], vv[kSU];
#pragma unroll
      for (int u = 0; u < kSU; ++u) {
        const int idx = i0 + u * blockDim.x;
        const int key = idx / CPR, c = idx - key * CPR;
        kv[u] = zero; vv[u] = zero;
        if (idx < total && key < L) {
          const __nv_bfloat16* rp = qkv + static_cast<size_t>(geo_row_of_pixel(g, b, key)) * ld + h * D + 8 * c;
          kv[u] = *reinterpret_cast<const uint4*>(rp + E);
          vv[u] = *reinterpret_cast<const uint4*>(rp + 2 * E);
        }
      }
#pragma unroll
      for (int u = 0; u < kSU; ++u) {
        const int idx = i0 + u * blockDim.x;
        if (idx < total) {
          const int key = idx / CPR, c = idx - key * CPR;
          *reinterpret_cast<uint4*>(Ks + am_off<D>(key, c)) = kv[u];
          *reinterpret_cast<uint4*>(Vs + am_off<D>(key, c)) = vv[u];
        }
      }
    }
  }
  __syncthreads();

  const uint32_t ks_addr = smem_u32(Ks), vs_addr = smem_u32(Vs);
  const float c2 = rsqrtf(static_cast<float>(D)) * 1.4426950408889634f;  // 1/sqrt(D) * log2(e)
  const int qr = lane >> 2, qc = (lane & 3) * 2;   // fragment row / first column of this lane
  const int lm = lane >> 3, lr = lane & 7;         // ldmatrix: matrix index / row this lane addresses
  const int tile_end = min(ntiles, (static_cast<int>(blockIdx.x) + 1) * tiles_per_cta);

  for (int tile = blockIdx.x * tiles_per_cta + warp; tile < tile_end; tile += nw) {
    const int rowA = tile * 16 + qr, rowB = rowA + 8;
    const bool vA = rowA < L, vB = rowB < L;
    const int prA = geo_row_of_pixel(g, b, vA ? rowA : 0), prB = geo_row_of_pixel(g, b, vB ? rowB : 0);
    // Q fragments (A operand, row-major 16 x 16 per K step)
    uint32_t qa[D / 16][4];
    {
      const __nv_bfloat16* qA = qkv + static_cast<size_t>(prA) * ld + h * D + qc;
      const __nv_bfloat16* qB = qkv + static_cast<size_t>(prB) * ld + h * D + qc;
#pragma unroll
      for (int ks = 0; ks < D / 16; ++ks) {
        qa[ks][0] = vA ? *reinterpret_cast<const uint32_t*>(qA + ks * 16) : 0u;
        qa[ks][1] = vB ? *reinterpret_cast<const uint32_t*>(qB + ks * 16) : 0u;
        qa[ks][2] = vA ? *reinterpret_cast<const uint32_t*>(qA + ks * 16 + 8) : 0u;
        qa[ks][3] = vB ? *reinterpret_cast<const uint32_t*>(qB + ks * 16 + 8) : 0u;
      }
    }
    float o[NO][4];
#pragma unroll
    for (int n = 0; n < NO; ++n) { o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f; }
    float mA = -INFINITY, mB = -INFINITY;
    float rsA = 0.f, rsB = 0.f;   // kFaddSum: this lane's share of the two row sums

    // K fragments of one 64-key step: matrices [n-tile 2np: d lo, d hi], [n-tile 2np + 1: d lo, d hi]
    // (prefetched one step ahead where the registers allow it: d_head 16)
    constexpr bool kPrefetchK = D == 16;
    uint32_t kf[D / 16][4][4];
    auto load_k = [&](int kb) {
#pragma unroll
      for (int ks = 0; ks < D / 16; ++ks)
#pragma unroll
        for (int np = 0; np < 4; ++np) {
          const int key = kb + (2 * np + (lm >> 1)) * 8 + lr;
          ldsm_x4(ks_addr + am_off<D>(key, ks * 2 + (lm & 1)), kf[ks][np][0], kf[ks][np][1], kf[ks][np][2], kf[ks][np][3]);
        }
    };
    // One step = 64 keys.  `tail` (compile-time) marks the last step of a sequence whose length is not
    // a multiple of 64: only its 16-key groups that hold real keys are computed; the full steps carry
    // no checks at all.
    auto step = [&](auto tail, int j) {
      constexpr bool kTail = decltype(tail)::value;
      const int kb = j * kAmKeys;
      const int ng = kTail ? (L - kb + 15) >> 4 : 4;   // 16-key groups with at least one real key
      if (!kPrefetchK) load_k(kb);
      // ---- S = Q K^T: 8 n-tiles of 8 keys
      float s[8][4];
#pragma unroll
      for (int n = 0; n < 8; ++n) { s[n][0] = s[n][1] = s[n][2] = s[n][3] = kTail ? -INFINITY : 0.f; }
#pragma unroll
      for (int np = 0; np < 4; ++np) {
        if (!kTail || np < ng) {
          if (kTail) { s[2 * np][0] = s[2 * np][1] = s[2 * np][2] = s[2 * np][3] = 0.f;
                       s[2 * np + 1][0] = s[2 * np + 1][1] = s[2 * np + 1][2] = s[2 * np + 1][3] = 0.f; }
#pragma unroll
          for (int ks = 0; ks < D / 16; ++ks) {
            am_mma_16816(s[2 * np], qa[ks], kf[ks][np][0], kf[ks][np][1]);
            am_mma_16816(s[2 * np + 1], qa[ks], kf[ks][np][2], kf[ks][np][3]);
          }
        }
      }
      if (kPrefetchK && j + 1 < nch) load_k(kb + kAmKeys);   // the next step's K fragments land while this step's softmax runs
      if (kTail) {   // keys >= L do not exist
#pragma unroll
        for (int n = 0; n < 8; ++n) {
          const int k0 = kb + n * 8 + qc;
          if (k0 >= L) { s[n][0] = -INFINITY; s[n][2] = -INFINITY; }
          if (k0 + 1 >= L) { s[n][1] = -INFINITY; s[n][3] = -INFINITY; }
        }
      }
      // ---- running maximum of the two rows (the four lanes of a quad hold one row)
      float xA = fmaxf(s[0][0], s[0][1]), xB = fmaxf(s[0][2], s[0][3]);
#pragma unroll
      for (int n = 1; n < 8; ++n) {
        xA = fmaxf(xA, fmaxf(s[n][0], s[n][1]));
        xB = fmaxf(xB, fmaxf(s[n][2], s[n][3]));
      }
      xA = fmaxf(xA, __shfl_xor_sync(0xffffffffu, xA, 1));
      xB = fmaxf(xB, __shfl_xor_sync(0xffffffffu, xB, 1));
      xA = fmaxf(xA, __shfl_xor_sync(0xffffffffu, xA, 2));
      xB = fmaxf(xB, __shfl_xor_sync(0xffffffffu, xB, 2));
      // rescale O / R only when some row of the warp has to raise its maximum (otherwise the factor
      // is exactly 1, so skipping changes no bit).  kLazy: a row keeps its maximum until a step exceeds
      // it by more than 2^8 in the exponent's units.
      const bool upA = kLazy ? (xA - mA) * c2 > 8.0f : xA > mA;
      const bool upB = kLazy ? (xB - mB) * c2 > 8.0f : xB > mB;
      if (__any_sync(0xffffffffu, upA || upB)) {
        const float nA = upA ? xA : mA, nB = upB ? xB : mB;
        const float fA = ex2_approx((mA - nA) * c2), fB = ex2_approx((mB - nB) * c2);   // exp2(-inf) = 0 at j = 0
        mA = nA; mB = nB;
#pragma unroll
        for (int n = 0; n < NO; ++n) { o[n][0] *= fA; o[n][1] *= fA; o[n][2] *= fB; o[n][3] *= fB; }
        if (kFaddSum) { rsA *= fA; rsB *= fB; }
      }
      const float cA = kDebias - mA * c2, cB = kDebias - mB * c2;
      // ---- P = exp2(..), truncated to bf16, straight into A fragments; O += P V, R += P 1
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {   // 16 keys per K step
        if (!kTail || kk < ng) {
          float e[8];
          if constexpr (kPacked) {
            // two scores per FFMA2 (the accumulator pairs of a row sit in adjacent registers); with
            // kPolyMask >= 0 the pair kPolyMask of each group of four takes the packed FMA-pipe exp2
#pragma unroll
            for (int pr = 0; pr < 4; ++pr) {
              const int n = 2 * kk + (pr >> 1), q = (pr & 1) * 2;
              const float cc = (pr & 1) ? cB : cA;
              const float2 x = __ffma2_rn(make_float2(s[n][q], s[n][q + 1]), make_float2(c2, c2), make_float2(cc, cc));
              if (kPolyMask >= 0 && pr == (kPolyMask & 3)) {
                const float2 r = am_ex2_poly2(x);
                e[2 * pr] = r.x; e[2 * pr + 1] = r.y;
              } else {
                e[2 * pr] = ex2_approx(x.x); e[2 * pr + 1] = ex2_approx(x.y);
              }
            }
          } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int n = 2 * kk + (i >> 2), q = i & 3;
              const float x = fmaf(s[n][q], c2, (q < 2) ? cA : cB);
              e[i] = (kPolyMask >= 0 && ((kk * 8 + i) & 3) == kPolyMask) ? am_ex2_poly(x) : ex2_approx(x);
            }
          }
          if (kFaddSum) { rsA += (e[0] + e[1]) + (e[4] + e[5]); rsB += (e[2] + e[3]) + (e[6] + e[7]); }
          uint32_t pa[4];
          pa[0] = __byte_perm(__float_as_uint(e[0]), __float_as_uint(e[1]), 0x7632);
          pa[1] = __byte_perm(__float_as_uint(e[2]), __float_as_uint(e[3]), 0x7632);
          pa[2] = __byte_perm(__float_as_uint(e[4]), __float_as_uint(e[5]), 0x7632);
          pa[3] = __byte_perm(__float_as_uint(e[6]), __float_as_uint(e[7]), 0x7632);
          const int key = kb + kk * 16 + (lm & 1) * 8 + lr;
#pragma unroll
          for (int np = 0; np < D / 16; ++np) {
            // transposed 8x8 blocks of V: [d tile 2np: keys lo, keys hi], [d tile 2np + 1: keys lo, keys hi]
            uint32_t b0, b1, b2, b3;
            ldsm_x4_t(vs_addr + am_off<D>(key, 2 * np + (lm >> 1)), b0, b1, b2, b3);
            am_mma_16816(o[2 * np], pa, b0, b1);
            am_mma_16816(o[2 * np + 1], pa, b2, b3);
          }
          if (!kFaddSum) am_mma_16816(o[NO - 1], pa, 0x3f803f80u, 0x3f803f80u);   // B = ones: every column = the row sum
        }
      }
    };
    if (kPrefetchK) load_k(0);
    const int nfull = L / kAmKeys;   // steps whose 64 keys are all real
    for (int j = 0; j < nfull; ++j) step(std::false_type{}, j);
    if (nfull < nch) step(std::true_type{}, nfull);
    // ---- O / R -> out (bf16)
    float iA, iB;
    if (kFaddSum) {
      rsA += __shfl_xor_sync(0xffffffffu, rsA, 1); rsB += __shfl_xor_sync(0xffffffffu, rsB, 1);
      rsA += __shfl_xor_sync(0xffffffffu, rsA, 2); rsB += __shfl_xor_sync(0xffffffffu, rsB, 2);
      iA = 1.001953125f / rsA; iB = 1.001953125f / rsB;   // (1 + 2^-9): see kDebias
    } else {
      iA = 1.0f / o[NO - 1][0]; iB = 1.0f / o[NO - 1][2];
    }
    __nv_bfloat16* oA = out + static_cast<size_t>(prA) * E + h * D + qc;
    __nv_bfloat16* oB = out + static_cast<size_t>(prB) * E + h * D + qc;
#pragma unroll
    for (int n = 0; n < D / 8; ++n) {
      if (vA) *reinterpret_cast<__nv_bfloat162*>(oA + n * 8) = __floats2bfloat162_rn(o[n][0] * iA, o[n][1] * iA);
      if (vB) *reinterpret_cast<__nv_bfloat162*>(oB + n * 8) = __floats2bfloat162_rn(o[n][2] * iB, o[n][3] * iB);
    }
  }
}

// ---------------------------------------------------------------------------
// The same kernel for the fp32-storage / TF32-operand mode (DMB200_PRECISION=tf32), where the
// CUDA-core attention took 5.8 of the step's 10.4 ms: mma.sync.m16n8k8 with TF32 operands (the
// tensor core reads the upper 19 bits of the fp32 words), fp32 accumulate, fp32 softmax.  K and V sit
// in shared memory as fp32 rows of D + 4 words (the padding makes the per-lane fragment loads
// conflict-free: lane (g, t) reads word key * (D + 4) + d with key = 8 n + g or 2 t).  For P V the
// key order inside an 8-key step is permuted (A column t <-> key 2 t, column t + 4 <-> key 2 t + 1) so
// that the score fragments ARE the A fragments; V rows are loaded in the matching order.  The row
// sum again comes from the tensor core (B = ones) from exactly the truncated P that enters P V.
// ---------------------------------------------------------------------------
constexpr int kAfMaxWarps = 14;

__device__ __forceinline__ void am_mma_tf32(float* d, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

inline size_t attention_mma_f32_smem_bytes(int L, int D) {
  const int lpad = (L + kAmKeys - 1) / kAmKeys * kAmKeys;
  return 2ull * lpad * (D + 4) * 4 + 16;
}

template <int D>
__global__ void __launch_bounds__(kAfMaxWarps * 32, 1)
attention_mma_f32_kernel(const float* __restrict__ qkv, Geo g, int E, float* __restrict__ out, int tiles_per_cta) {
  extern __shared__ uint8_t am_smem_raw[];
  constexpr int LD = D + 4;       // words per key row in shared memory
  constexpr int NO = D / 8 + 1;   // output n-tiles: D / 8 of V plus the all-ones tile (row sum)
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, nw = blockDim.x >> 5;
  const int L = g.H * g.W;
  const int nch = (L + kAmKeys - 1) / kAmKeys;
  const int lpad = nch * kAmKeys;
  const int ntiles = (L + 15) >> 4;
  const int b = blockIdx.z, h = blockIdx.y;
  const int ld = 3 * E;
  const uint32_t sraw = smem_u32(am_smem_raw);
  float* Ks = reinterpret_cast<float*>(am_smem_raw + ((16u - (sraw & 15u)) & 15u));
  float* Vs = Ks + static_cast<size_t>(lpad) * LD;

  // ---- stage K and V of the (sample, head): 16-byte loads
  {
    constexpr int VPK = D / 4;   // float4 per key
    const int total = lpad * VPK;
    for (int idx = tid; idx < total; idx += blockDim.x) {
      const int key = idx / VPK, c = idx - key * VPK;
      float4 kv = make_float4(0.f, 0.f, 0.f, 0.f), vv = kv;
      if (key < L) {
        const float* rp = qkv + static_cast<size_t>(geo_row_of_pixel(g, b, key)) * ld + h * D + 4 * c;
        kv = *reinterpret_cast<const float4*>(rp + E);
        vv = *reinterpret_cast<const float4*>(rp + 2 * E);
      }
      *reinterpret_cast<float4*>(Ks + key * LD + 4 * c) = kv;
      *reinterpret_cast<float4*>(Vs + key * LD + 4 * c) = vv;
    }
  }
  __syncthreads();

  const float c2 = rsqrtf(static_cast<float>(D)) * 1.4426950408889634f;
  const int gq = lane >> 2, tq = lane & 3;
  const int tile_end = min(ntiles, (static_cast<int>(blockIdx.x) + 1) * tiles_per_cta);
  const uint32_t one = __float_as_uint(1.0f);

  for (int tile = blockIdx.x * tiles_per_cta + warp; tile < tile_end; tile += nw) {
    const int rowA = tile * 16 + gq, rowB = rowA + 8;
    const bool vA = rowA < L, vB = rowB < L;
    const int prA = geo_row_of_pixel(g, b, vA ? rowA : 0), prB = geo_row_of_pixel(g, b, vB ? rowB : 0);
    uint32_t qa[D / 8][4];
    {
      const float* qA = qkv + static_cast<size_t>(prA) * ld + h * D + tq;
      const float* qB = qkv + static_cast<size_t>(prB) * ld + h * D + tq;
#pragma unroll
      for (int ks = 0; ks < D / 8; ++ks) {
        qa[ks][0] = vA ? __float_as_uint(qA[ks * 8]) : 0u;
        qa[ks][1] = vB ? __float_as_uint(qB[ks * 8]) : 0u;
        qa[ks][2] = vA ? __float_as_uint(qA[ks * 8 + 4]) : 0u;
        qa[ks][3] = vB ? __float_as_uint(qB[ks * 8 + 4]) : 0u;
      }
    }
    float o[NO][4];
#pragma unroll
    for (int n = 0; n < NO; ++n) { o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f; }
    float mA = -INFINITY, mB = -INFINITY;

    for (int j = 0; j < nch; ++j) {
      const int kb = j * kAmKeys;
      const int ng = min(8, (L - kb + 7) >> 3);   // 8-key groups with at least one real key
      float s[8][4];
#pragma unroll
      for (int n = 0; n < 8; ++n) {
        s[n][0] = s[n][1] = s[n][2] = s[n][3] = -INFINITY;
        if (n < ng) {
          s[n][0] = s[n][1] = s[n][2] = s[n][3] = 0.f;
          const float* kp = Ks + (kb + n * 8 + gq) * LD + tq;
#pragma unroll
          for (int ks = 0; ks < D / 8; ++ks)
            am_mma_tf32(s[n], qa[ks], __float_as_uint(kp[ks * 8]), __float_as_uint(kp[ks * 8 + 4]));
        }
      }
      if (kb + kAmKeys > L) {
#pragma unroll
        for (int n = 0; n < 8; ++n) {
          const int k0 = kb + n * 8 + 2 * tq;
          if (k0 >= L) { s[n][0] = -INFINITY; s[n][2] = -INFINITY; }
          if (k0 + 1 >= L) { s[n][1] = -INFINITY; s[n][3] = -INFINITY; }
        }
      }
      float xA = fmaxf(s[0][0], s[0][1]), xB = fmaxf(s[0][2], s[0][3]);
#pragma unroll
      for (int n = 1; n < 8; ++n) {
        xA = fmaxf(xA, fmaxf(s[n][0], s[n][1]));
        xB = fmaxf(xB, fmaxf(s[n][2], s[n][3]));
      }
      xA = fmaxf(xA, __shfl_xor_sync(0xffffffffu, xA, 1));
      xB = fmaxf(xB, __shfl_xor_sync(0xffffffffu, xB, 1));
      xA = fmaxf(xA, __shfl_xor_sync(0xffffffffu, xA, 2));
      xB = fmaxf(xB, __shfl_xor_sync(0xffffffffu, xB, 2));
      const bool upA = (xA - mA) * c2 > 8.0f, upB = (xB - mB) * c2 > 8.0f;   // lazy maximum, as in the bf16 kernel
      if (__any_sync(0xffffffffu, upA || upB)) {
        const float nA = upA ? xA : mA, nB = upB ? xB : mB;
        const float fA = ex2_approx((mA - nA) * c2), fB = ex2_approx((mB - nB) * c2);
        mA = nA; mB = nB;
#pragma unroll
        for (int n = 0; n < NO; ++n) { o[n][0] *= fA; o[n][1] *= fA; o[n][2] *= fB; o[n][3] *= fB; }
      }
      const float cA = -mA * c2, cB = -mB * c2;
#pragma unroll
      for (int n = 0; n < 8; ++n) {   // one 8-key step per score n-tile
        if (n < ng) {
          uint32_t pa[4];   // A columns t / t + 4 <-> keys 2 t / 2 t + 1: the score fragment as it is
          pa[0] = __float_as_uint(ex2_approx(fmaf(s[n][0], c2, cA)));
          pa[2] = __float_as_uint(ex2_approx(fmaf(s[n][1], c2, cA)));
          pa[1] = __float_as_uint(ex2_approx(fmaf(s[n][2], c2, cB)));
          pa[3] = __float_as_uint(ex2_approx(fmaf(s[n][3], c2, cB)));
          const float* vp = Vs + (kb + n * 8 + 2 * tq) * LD + gq;
#pragma unroll
          for (int nt = 0; nt < D / 8; ++nt)
            am_mma_tf32(o[nt], pa, __float_as_uint(vp[nt * 8]), __float_as_uint(vp[LD + nt * 8]));
          am_mma_tf32(o[NO - 1], pa, one, one);
        }
      }
    }
    const float iA = 1.0f / o[NO - 1][0], iB = 1.0f / o[NO - 1][2];
    float* oA = out + static_cast<size_t>(prA) * E + h * D + 2 * tq;
    float* oB = out + static_cast<size_t>(prB) * E + h * D + 2 * tq;
#pragma unroll
    for (int n = 0; n < D / 8; ++n) {
      if (vA) *reinterpret_cast<float2*>(oA + n * 8) = make_float2(o[n][0] * iA, o[n][1] * iA);
      if (vB) *reinterpret_cast<float2*>(oB + n * 8) = make_float2(o[n][2] * iB, o[n][3] * iB);
    }
  }
}

}  // namespace dmb
