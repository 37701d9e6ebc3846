// Tail of the attention block as ONE kernel (models/unet_cond.py:36-50, after the MHA core):
//
//   o   = av Wp^T + bp + x_ln                (out_proj + the residual of the block)
//   h   = LayerNorm(o)                       (ff_self.0)
//   f1  = GELU(h W1^T + b1)                  (ff_self.1, ff_self.2)
//   out = f1 W2^T + b2 + o                   (ff_self.3 + residual)
//
// Every step is local to a token (a PM row) and K = N = C <= 128, so a 128-row tile runs the
// whole chain on chip: the three weight matrices stay resident in shared memory for the
// lifetime of the (persistent) CTA, av / x_ln arrive by TMA, the three GEMMs are tcgen05 MMAs
// into one TMEM accumulator, and the intermediate activations (o, h, f1) only ever exist as
// bf16 tiles in shared memory.  HBM traffic per token: 2 C reads + C writes (bf16) instead of
// the 11 C of the four separate launches (three Linear GEMMs + LayerNorm).
//
// The chain of one tile is strictly sequential (MMA -> epilogue -> MMA ...), so a CTA runs
// kChains independent chains.  A chain is C / 32 warpgroups: a thread owns 32 columns of one row
// for the epilogues (TMEM lane = row; warps with the same warp % 4 may read the same lanes, so
// the columns of a row are split over warpgroups and the row's LayerNorm sums are exchanged
// through shared memory); the chain's first warp issues its TMA loads and MMAs (warp-uniform
// code, elect_one() around the asm only).  32 warps per SM keep the short dependent epilogue
// sequences of several tiles in flight.  Buffers per chain: A (av -> h -> f1 -> out, each the
// K-major 128B-swizzled operand of the next MMA / source of the TMA store) and R (x_ln -> o).
// Rounding points are those of the separate launches (o, h, f1 and out are rounded to bf16);
// the LayerNorm variance is the one-pass E[o^2] - mean^2 in fp32.
#pragma once

#include "common.cuh"

namespace dmb {

struct TailParams {
  Geo geo;
  int m_tiles;
  const float* bp;     // out_proj bias
  const float* ln_w;   // ff_self.0 weight / bias
  const float* ln_b;
  const float* b1;     // ff_self.1 bias
  const float* b2;     // ff_self.3 bias
};

// Epilogue arithmetic on register PAIRS (fp32x2 instructions, one bf16x2 conversion per pair): the
// scalar form spent ~20 issue slots per element on bf16 pack / unpack shuffles (SASS), and these
// kernels are bound by issue slots and the latency of their dependent phases, not by HBM.
__device__ __forceinline__ float2 bf2_to_f2(uint32_t w) {      // exact
  return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
}
__device__ __forceinline__ uint32_t f2_to_bf2(float2 v) {      // round to nearest even, .x in the low half
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(v.y), "f"(v.x));
  return r;
}
// eight consecutive floats of a parameter vector as four pairs (kSmem: the vector was staged in
// shared memory; otherwise read-only global loads)
template <bool kSmem>
__device__ __forceinline__ void ld_pairs(const float* src, float2* dst) {
  float4 a, b;
  if constexpr (kSmem) {
    a = *reinterpret_cast<const float4*>(src); b = *reinterpret_cast<const float4*>(src + 4);
  } else {
    a = __ldg(reinterpret_cast<const float4*>(src)); b = __ldg(reinterpret_cast<const float4*>(src + 4));
  }
  dst[0] = make_float2(a.x, a.y); dst[1] = make_float2(a.z, a.w);
  dst[2] = make_float2(b.x, b.y); dst[3] = make_float2(b.z, b.w);
}
// the five per-channel vectors of the tail (bp, ln_w, ln_b, b1, b2) -> prm[5][C]; they are model
// parameters (no kernel of the stream writes them), so this may run before pdl_wait()
template <int C, int kThreads>
__device__ __forceinline__ void stage_tail_params(float* prm, const float* bp, const float* ln_w, const float* ln_b,
                                                  const float* b1, const float* b2) {
  for (int i = threadIdx.x; i < 5 * C; i += kThreads) {
    const int k = i / C, c = i - k * C;
    const float* src = k == 0 ? bp : k == 1 ? ln_w : k == 2 ? ln_b : k == 3 ? b1 : b2;
    prm[i] = __ldg(src + c);
  }
}

constexpr int kTailCols = 32;        // accumulator columns per thread
constexpr int kTailThreads = 1024;   // kChains * (C / kTailCols) * 128 for both instantiations

template <int C, int kChains>
constexpr int attn_tail_smem_bytes() {
  return 3 * C * C * 2 + kChains * 2 * (C / 64) * (128 * 128) + 1024;
}

template <int C, int kChains>
__global__ void __launch_bounds__(kTailThreads, 1)
attn_tail_kernel(const __grid_constant__ CUtensorMap tmap_av, const __grid_constant__ CUtensorMap tmap_xl,
                 const __grid_constant__ CUtensorMap tmap_wp, const __grid_constant__ CUtensorMap tmap_w1,
                 const __grid_constant__ CUtensorMap tmap_w2, const __grid_constant__ CUtensorMap tmap_out,
                 const TailParams p) {
  static_assert(C == 64 || C == 128, "weights of wider blocks do not fit in shared memory");
  constexpr int kSplit = C / kTailCols;       // warpgroups sharing the columns of one tile
  constexpr int kChainThreads = 128 * kSplit;
  static_assert(kChains * kChainThreads == kTailThreads, "thread layout");
  constexpr int kKC = C / 64;                 // 128-byte K chunks (= 64-column output blocks)
  constexpr uint32_t kChunkBytes = 128 * 128; // one 128-row x 128-byte sub-tile
  constexpr uint32_t kWChunkBytes = C * 128;  // one K chunk of a weight: C rows x 128 bytes
  constexpr uint32_t kTmemCols = kChains * C; // 256
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t w_bar;
  __shared__ __align__(8) uint64_t ld_bar[kChains];
  __shared__ __align__(8) uint64_t mma_bar[kChains];
  __shared__ uint32_t tmem_base_smem;
  // the per-channel vectors bp | ln_w | ln_b | b1 | b2 in shared memory where there is room (C = 64;
  // at C = 128 the weights and two chains leave 2 KB of the 227 KB)
  constexpr bool kPrmSmem = C == 64;
  __shared__ __align__(16) float prm_smem[kPrmSmem ? 5 * C : 4];

  pdl_trigger();
  if constexpr (kPrmSmem) stage_tail_params<C, kTailThreads>(prm_smem, p.bp, p.ln_w, p.ln_b, p.b1, p.b2);
  const int warp = threadIdx.x >> 5;
  const int chain = warp / (4 * kSplit);
  const int split = (warp >> 2) % kSplit;       // which 32 columns of the row this thread owns
  const int q = warp & 3;                       // TMEM lane quadrant of this warp
  const bool lead = (warp % (4 * kSplit)) == 0; // the chain's first warp issues its TMA and MMAs
  const int t = q * 32 + (threadIdx.x & 31);    // row of the tile
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  uint8_t* wsm = smem;                                         // [3][kKC][C rows][128 B]
  uint8_t* bufA = smem + 3 * kKC * kWChunkBytes + chain * (2 * kKC * kChunkBytes);
  uint8_t* bufR = bufA + kKC * kChunkBytes;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmap_av); tma_prefetch_desc(&tmap_xl); tma_prefetch_desc(&tmap_out);
    tma_prefetch_desc(&tmap_wp); tma_prefetch_desc(&tmap_w1); tma_prefetch_desc(&tmap_w2);
    mbar_init(&w_bar, 1);
    for (int i = 0; i < kChains; ++i) { mbar_init(&ld_bar[i], 1); mbar_init(&mma_bar[i], 1); }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(&tmem_base_smem, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  pdl_wait();

  if (threadIdx.x == 0) {   // the three weight matrices, once per CTA
    mbar_expect_tx(&w_bar, 3 * kKC * kWChunkBytes);
    for (int kc = 0; kc < kKC; ++kc) {
      tma_load_2d(wsm + (0 * kKC + kc) * kWChunkBytes, &tmap_wp, &w_bar, kc * 64, 0);
      tma_load_2d(wsm + (1 * kKC + kc) * kWChunkBytes, &tmap_w1, &w_bar, kc * 64, 0);
      tma_load_2d(wsm + (2 * kKC + kc) * kWChunkBytes, &tmap_w2, &w_bar, kc * 64, 0);
    }
  }

  const uint32_t idesc = umma_idesc(1u, 128, C);
  const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(chain * C);
  const int col0 = split * kTailCols;           // first column of this thread
  const float* prm[5];                          // this thread's 32 columns of the five vectors
  if constexpr (kPrmSmem) {
#pragma unroll
    for (int i = 0; i < 5; ++i) prm[i] = prm_smem + i * C + col0;
  } else {
    prm[0] = p.bp + col0; prm[1] = p.ln_w + col0; prm[2] = p.ln_b + col0; prm[3] = p.b1 + col0; prm[4] = p.b2 + col0;
  }
  const uint32_t t_row = d_tmem + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(col0);
  const uint32_t a_addr = smem_u32(bufA);
  const uint32_t w_addr = smem_u32(wsm);
  const int bar_id = 1 + chain;
  const int sw = t & 7;
  // this thread's four 16-byte vectors of a row: chunk (col0 / 64), vectors j0 .. j0 + 3 of its 128 bytes
  const uint32_t row_off = (col0 >> 6) * kChunkBytes + t * 128;
  const int j0 = (col0 & 63) >> 3;
  uint8_t* rowA = bufA + row_off;
  uint8_t* rowR = bufR + row_off;
  // LayerNorm partials of the row's column groups are exchanged through the (then idle) A buffer
  float2* scratch = reinterpret_cast<float2*>(bufA);
  uint32_t ph_ld = 0, ph_mma = 0;
  bool w_ready = false;

  // one GEMM of the chain: D = A(bufA) * W[which]^T, issued by the chain's first warp
  auto issue_gemm = [&](int which) {
    if (lead) {
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int kc = 0; kc < kKC; ++kc) {
          const uint64_t ad = umma_desc_sw128(a_addr + kc * kChunkBytes);
          const uint64_t bd = umma_desc_sw128(w_addr + (which * kKC + kc) * kWChunkBytes);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_ss<false>(d_tmem, ad + (k * 2), bd + (k * 2), idesc, (kc > 0 || k > 0) ? 1u : 0u);
        }
        umma_commit(&mma_bar[chain]);
      }
      __syncwarp();
    }
  };
  auto wait_gemm = [&]() {
    mbar_wait(&mma_bar[chain], ph_mma);
    ph_mma ^= 1;
    tc_fence_after();
  };

  // consecutive tiles go to different CTAs first, so a small launch still spreads over the chip
  for (int tile = blockIdx.x + gridDim.x * chain; tile < p.m_tiles; tile += gridDim.x * kChains) {
    const int m0 = tile * 128;
    int b = 0, y = 0, x = 0;
    const bool valid = geo_decode(p.geo, m0 + t, b, y, x);

    if (lead) {
      if (elect_one()) {
        tma_store_wait_read<0>();   // the previous tile's store has read bufA
        mbar_expect_tx(&ld_bar[chain], 2 * kKC * kChunkBytes);
#pragma unroll
        for (int kc = 0; kc < kKC; ++kc) {
          tma_load_2d(bufA + kc * kChunkBytes, &tmap_av, &ld_bar[chain], kc * 64, m0);
          tma_load_2d(bufR + kc * kChunkBytes, &tmap_xl, &ld_bar[chain], kc * 64, m0);
        }
      }
      __syncwarp();
    }
    if (!w_ready) { mbar_wait(&w_bar, 0); w_ready = true; }
    mbar_wait(&ld_bar[chain], ph_ld);
    ph_ld ^= 1;
    issue_gemm(0);
    wait_gemm();

    // ---- o = acc + bp + x_ln (kept in R as bf16); h = LayerNorm(o) -> A
    {
      float o[kTailCols];
      tmem_ld32(t_row, o);
      tmem_ld_wait();
      float2 s2 = make_float2(0.f, 0.f), ss2 = make_float2(0.f, 0.f);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint4* rp = reinterpret_cast<uint4*>(rowR + (((j0 + j) ^ sw) << 4));
        const uint4 u = *rp;
        const uint32_t uw[4] = {u.x, u.y, u.z, u.w};
        float2 bv[4];
        ld_pairs<kPrmSmem>(prm[0] + 8 * j, bv);
        uint32_t ww[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 acc = make_float2(o[8 * j + 2 * i], o[8 * j + 2 * i + 1]);
          ww[i] = f2_to_bf2(__fadd2_rn(__fadd2_rn(acc, bv[i]), bf2_to_f2(uw[i])));
          const float2 of = bf2_to_f2(ww[i]);   // LayerNorm sees the rounded o, as it would from HBM
          o[8 * j + 2 * i] = of.x; o[8 * j + 2 * i + 1] = of.y;
          s2 = __fadd2_rn(s2, of);
          ss2 = __ffma2_rn(of, of, ss2);
        }
        *rp = make_uint4(ww[0], ww[1], ww[2], ww[3]);
      }
      scratch[split * 128 + t] = make_float2(s2.x + s2.y, ss2.x + ss2.y);
      named_bar_sync(bar_id, kChainThreads);
      float ts = 0.f, tss = 0.f;
#pragma unroll
      for (int g = 0; g < kSplit; ++g) {   // the same order in every thread of the row
        const float2 v = scratch[g * 128 + t];
        ts += v.x; tss += v.y;
      }
      const float mean = ts * (1.0f / C);
      const float rstd = rsqrtf(fmaxf(tss * (1.0f / C) - mean * mean, 0.f) + 1e-5f);
      named_bar_sync(bar_id, kChainThreads);   // the partials are read: A may be overwritten
      const float2 nm = make_float2(-mean, -mean), rs = make_float2(rstd, rstd);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float2 wv[4], bv[4];
        ld_pairs<kPrmSmem>(prm[1] + 8 * j, wv);
        ld_pairs<kPrmSmem>(prm[2] + 8 * j, bv);
        uint32_t ww[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 d = __fmul2_rn(__fadd2_rn(make_float2(o[8 * j + 2 * i], o[8 * j + 2 * i + 1]), nm), rs);
          ww[i] = f2_to_bf2(__ffma2_rn(d, wv[i], bv[i]));
        }
        *reinterpret_cast<uint4*>(rowA + (((j0 + j) ^ sw) << 4)) = make_uint4(ww[0], ww[1], ww[2], ww[3]);
      }
    }
    fence_async_proxy();
    tc_fence_before();
    named_bar_sync(bar_id, kChainThreads);
    issue_gemm(1);
    wait_gemm();

    // ---- f1 = GELU(acc + b1) -> A
    {
      float v[kTailCols];
      tmem_ld32(t_row, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float2 bv[4];
        ld_pairs<kPrmSmem>(prm[3] + 8 * j, bv);
        float h[8];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 a = __fadd2_rn(bv[i], make_float2(v[8 * j + 2 * i], v[8 * j + 2 * i + 1]));
          h[2 * i] = a.x; h[2 * i + 1] = a.y;
        }
        gelu_inplace<__nv_bfloat16, 8>(h);
        uint32_t ww[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) ww[i] = f2_to_bf2(make_float2(h[2 * i], h[2 * i + 1]));
        *reinterpret_cast<uint4*>(rowA + (((j0 + j) ^ sw) << 4)) = make_uint4(ww[0], ww[1], ww[2], ww[3]);
      }
    }
    fence_async_proxy();
    tc_fence_before();
    named_bar_sync(bar_id, kChainThreads);
    issue_gemm(2);
    wait_gemm();

    // ---- out = acc + b2 + o -> A -> TMA store (pad pixels are stored as zeros)
    {
      float v[kTailCols];
      tmem_ld32(t_row, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint32_t off = ((j0 + j) ^ sw) << 4;
        uint4 w4 = make_uint4(0u, 0u, 0u, 0u);
        if (valid) {
          const uint4 u = *reinterpret_cast<const uint4*>(rowR + off);
          const uint32_t uw[4] = {u.x, u.y, u.z, u.w};
          float2 bv[4];
          ld_pairs<kPrmSmem>(prm[4] + 8 * j, bv);
          uint32_t ww[4];
#pragma unroll
          for (int i = 0; i < 4; ++i)
            ww[i] = f2_to_bf2(__fadd2_rn(__fadd2_rn(make_float2(v[8 * j + 2 * i], v[8 * j + 2 * i + 1]), bv[i]), bf2_to_f2(uw[i])));
          w4 = make_uint4(ww[0], ww[1], ww[2], ww[3]);
        }
        *reinterpret_cast<uint4*>(rowA + off) = w4;
      }
    }
    fence_async_proxy();
    tc_fence_before();
    named_bar_sync(bar_id, kChainThreads);
    if (lead) {
      if (elect_one()) {
#pragma unroll
        for (int kc = 0; kc < kKC; ++kc) tma_store_2d(&tmap_out, bufA + kc * kChunkBytes, kc * 64, m0);
        tma_store_commit();
      }
      __syncwarp();
    }
  }
  // a bulk-store group belongs to the thread that committed it; elect_one() picks the same lane
  // every time in converged code, and waiting in every lane of the issuing warps is harmless
  if (lead) tma_store_wait_all<0>();

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

// ---------------------------------------------------------------------------
// Head of the attention block as ONE kernel (models/unet_cond.py:47-49, in front of the MHA core):
//
//   x_ln = LayerNorm(x)                      (self.ln; also the residual of the block, so it is stored)
//   qkv  = x_ln Wqkv^T + bqkv                (nn.MultiheadAttention in_proj)
//
// Same shape as attn_tail_kernel: the in-projection's three [C, C] blocks (q | k | v rows of in_proj_weight)
// stay resident, a 128-row tile of x arrives by TMA, is normalised in place (thread = one row x 32 columns,
// the row sums cross warpgroups through shared memory) and is then both the source of the x_ln TMA store and
// the A operand of three MMAs; each 128 x C accumulator gets its bias and leaves through the second buffer.
// Replaces a LayerNorm launch and a Linear launch where the LayerNorm is not already applied by the producing
// GroupNorm kernel (large activations): x_ln is not re-read from HBM and two passes become one.
// Rounding points are those of the separate launches (x_ln and qkv rounded to bf16); the LayerNorm variance
// is the one-pass E[x^2] - mean^2 in fp32 (the separate kernel subtracts the mean first).
// ---------------------------------------------------------------------------
struct LnQkvParams {
  Geo geo;
  int m_tiles;
  const float* ln_w;
  const float* ln_b;
  const float* bqkv;   // in_proj_bias [3 C]
};

template <int C, int kChains>
__global__ void __launch_bounds__(kTailThreads, 1)
ln_qkv_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
              const __grid_constant__ CUtensorMap tmap_xln, const __grid_constant__ CUtensorMap tmap_qkv,
              const LnQkvParams p) {
  static_assert(C == 64 || C == 128, "weights of wider blocks do not fit in shared memory");
  constexpr int kSplit = C / kTailCols;
  constexpr int kChainThreads = 128 * kSplit;
  static_assert(kChains * kChainThreads == kTailThreads, "thread layout");
  constexpr int kKC = C / 64;
  constexpr uint32_t kChunkBytes = 128 * 128;
  constexpr uint32_t kWChunkBytes = C * 128;
  constexpr uint32_t kTmemCols = kChains * C;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t w_bar;
  __shared__ __align__(8) uint64_t ld_bar[kChains];
  __shared__ __align__(8) uint64_t mma_bar[kChains];
  __shared__ uint32_t tmem_base_smem;
  constexpr bool kPrmSmem = C == 64;   // ln_w | ln_b | bq | bk | bv
  __shared__ __align__(16) float prm_smem[kPrmSmem ? 5 * C : 4];

  pdl_trigger();
  if constexpr (kPrmSmem) stage_tail_params<C, kTailThreads>(prm_smem, p.ln_w, p.ln_b, p.bqkv, p.bqkv + C, p.bqkv + 2 * C);
  const int warp = threadIdx.x >> 5;
  const int chain = warp / (4 * kSplit);
  const int split = (warp >> 2) % kSplit;
  const int q = warp & 3;
  const bool lead = (warp % (4 * kSplit)) == 0;
  const int t = q * 32 + (threadIdx.x & 31);
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  uint8_t* wsm = smem;                                         // [3][kKC][C rows][128 B]
  uint8_t* bufA = smem + 3 * kKC * kWChunkBytes + chain * (2 * kKC * kChunkBytes);   // x -> x_ln
  uint8_t* bufO = bufA + kKC * kChunkBytes;                                          // q / k / v staging

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmap_x); tma_prefetch_desc(&tmap_w); tma_prefetch_desc(&tmap_xln); tma_prefetch_desc(&tmap_qkv);
    mbar_init(&w_bar, 1);
    for (int i = 0; i < kChains; ++i) { mbar_init(&ld_bar[i], 1); mbar_init(&mma_bar[i], 1); }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(&tmem_base_smem, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  pdl_wait();

  if (threadIdx.x == 0) {   // the three [C, C] blocks of in_proj_weight, once per CTA
    mbar_expect_tx(&w_bar, 3 * kKC * kWChunkBytes);
    for (int part = 0; part < 3; ++part)
      for (int kc = 0; kc < kKC; ++kc)
        tma_load_2d(wsm + (part * kKC + kc) * kWChunkBytes, &tmap_w, &w_bar, kc * 64, part * C);
  }

  const uint32_t idesc = umma_idesc(1u, 128, C);
  const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(chain * C);
  const int col0 = split * kTailCols;
  const float* prm[5];
  if constexpr (kPrmSmem) {
#pragma unroll
    for (int i = 0; i < 5; ++i) prm[i] = prm_smem + i * C + col0;
  } else {
    prm[0] = p.ln_w + col0; prm[1] = p.ln_b + col0;
    prm[2] = p.bqkv + col0; prm[3] = p.bqkv + C + col0; prm[4] = p.bqkv + 2 * C + col0;
  }
  const uint32_t t_row = d_tmem + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(col0);
  const uint32_t a_addr = smem_u32(bufA);
  const uint32_t w_addr = smem_u32(wsm);
  const int bar_id = 1 + chain;
  const int sw = t & 7;
  const uint32_t row_off = (col0 >> 6) * kChunkBytes + t * 128;
  const int j0 = (col0 & 63) >> 3;
  uint8_t* rowA = bufA + row_off;
  uint8_t* rowO = bufO + row_off;
  float2* scratch = reinterpret_cast<float2*>(bufO);   // LayerNorm partials (bufO is idle until the first epilogue)
  uint32_t ph_ld = 0, ph_mma = 0;
  bool w_ready = false;

  auto issue_gemm = [&](int which) {
    if (lead) {
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int kc = 0; kc < kKC; ++kc) {
          const uint64_t ad = umma_desc_sw128(a_addr + kc * kChunkBytes);
          const uint64_t bd = umma_desc_sw128(w_addr + (which * kKC + kc) * kWChunkBytes);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_ss<false>(d_tmem, ad + (k * 2), bd + (k * 2), idesc, (kc > 0 || k > 0) ? 1u : 0u);
        }
        umma_commit(&mma_bar[chain]);
      }
      __syncwarp();
    }
  };

  for (int tile = blockIdx.x + gridDim.x * chain; tile < p.m_tiles; tile += gridDim.x * kChains) {
    const int m0 = tile * 128;
    int b = 0, y = 0, x = 0;
    const bool valid = geo_decode(p.geo, m0 + t, b, y, x);

    if (lead) {
      if (elect_one()) {
        tma_store_wait_read<0>();   // the previous tile's stores have read bufA and bufO
        mbar_expect_tx(&ld_bar[chain], kKC * kChunkBytes);
#pragma unroll
        for (int kc = 0; kc < kKC; ++kc) tma_load_2d(bufA + kc * kChunkBytes, &tmap_x, &ld_bar[chain], kc * 64, m0);
      }
      __syncwarp();
    }
    if (!w_ready) { mbar_wait(&w_bar, 0); w_ready = true; }
    mbar_wait(&ld_bar[chain], ph_ld);
    ph_ld ^= 1;

    // ---- x_ln = LayerNorm(x), in place in A (pad pixels stay zero, as the separate kernel leaves them)
    {
      float2 xv[16];
      float2 s2 = make_float2(0.f, 0.f), ss2 = make_float2(0.f, 0.f);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint4 u = *reinterpret_cast<const uint4*>(rowA + (((j0 + j) ^ sw) << 4));
        const uint32_t uw[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          xv[4 * j + i] = bf2_to_f2(uw[i]);
          s2 = __fadd2_rn(s2, xv[4 * j + i]);
          ss2 = __ffma2_rn(xv[4 * j + i], xv[4 * j + i], ss2);
        }
      }
      scratch[split * 128 + t] = make_float2(s2.x + s2.y, ss2.x + ss2.y);
      named_bar_sync(bar_id, kChainThreads);
      float ts = 0.f, tss = 0.f;
#pragma unroll
      for (int g = 0; g < kSplit; ++g) {   // the same order in every thread of the row
        const float2 v = scratch[g * 128 + t];
        ts += v.x; tss += v.y;
      }
      const float mean = ts * (1.0f / C);
      const float rstd = rsqrtf(fmaxf(tss * (1.0f / C) - mean * mean, 0.f) + 1e-5f);
      const float2 nm = make_float2(-mean, -mean), rs = make_float2(rstd, rstd);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint4 w4 = make_uint4(0u, 0u, 0u, 0u);
        if (valid) {
          float2 wv[4], bv[4];
          ld_pairs<kPrmSmem>(prm[0] + 8 * j, wv);
          ld_pairs<kPrmSmem>(prm[1] + 8 * j, bv);
          uint32_t ww[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float2 d = __fmul2_rn(__fadd2_rn(xv[4 * j + i], nm), rs);
            ww[i] = f2_to_bf2(__ffma2_rn(d, wv[i], bv[i]));
          }
          w4 = make_uint4(ww[0], ww[1], ww[2], ww[3]);
        }
        *reinterpret_cast<uint4*>(rowA + (((j0 + j) ^ sw) << 4)) = w4;
      }
    }
    fence_async_proxy();
    tc_fence_before();
    named_bar_sync(bar_id, kChainThreads);
    if (lead) {
      if (elect_one()) {
#pragma unroll
        for (int kc = 0; kc < kKC; ++kc) tma_store_2d(&tmap_xln, bufA + kc * kChunkBytes, kc * 64, m0);
        tma_store_commit();
      }
      __syncwarp();
    }
    issue_gemm(0);

    // ---- q, k, v: acc + bias -> O -> TMA store into columns [part C, part C + C) of qkv
#pragma unroll
    for (int part = 0; part < 3; ++part) {
      mbar_wait(&mma_bar[chain], ph_mma);
      ph_mma ^= 1;
      tc_fence_after();
      float v[kTailCols];
      tmem_ld32(t_row, v);
      tmem_ld_wait();
      tc_fence_before();
      if (lead) {
        if (elect_one()) tma_store_wait_read<0>();   // the previous part's store has read O
        __syncwarp();
      }
      named_bar_sync(bar_id, kChainThreads);   // every thread holds its accumulator columns: TMEM and O are free
      if (part < 2) issue_gemm(part + 1);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint4 w4 = make_uint4(0u, 0u, 0u, 0u);
        if (valid) {
          float2 bv[4];
          ld_pairs<kPrmSmem>(prm[2 + part] + 8 * j, bv);
          uint32_t ww[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) ww[i] = f2_to_bf2(__fadd2_rn(make_float2(v[8 * j + 2 * i], v[8 * j + 2 * i + 1]), bv[i]));
          w4 = make_uint4(ww[0], ww[1], ww[2], ww[3]);
        }
        *reinterpret_cast<uint4*>(rowO + (((j0 + j) ^ sw) << 4)) = w4;
      }
      fence_async_proxy();
      named_bar_sync(bar_id, kChainThreads);
      if (lead) {
        if (elect_one()) {
#pragma unroll
          for (int kc = 0; kc < kKC; ++kc) tma_store_2d(&tmap_qkv, bufO + kc * kChunkBytes, part * C + kc * 64, m0);
          tma_store_commit();
        }
        __syncwarp();
      }
    }
  }
  if (lead) tma_store_wait_all<0>();

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

// ---------------------------------------------------------------------------
// The same tail for C = 256 (sa2 on 7x7 maps, sa3 on 3x3 maps): the three weight matrices (3 x 128
// KB) do not fit next to a tile, so ONE matrix is resident at a time.  Its four 32 KB K chunks are
// loaded by TMA as soon as the previous GEMM's MMAs have retired, i.e. they land while the epilogue
// between two GEMMs runs; o (the residual of the block) stays in registers as packed bf16 and x_ln
// is read straight from global memory, so shared memory holds only the A tile (64 KB) and the
// current weight (128 KB).  One chain per CTA: eight warpgroups share the 256 columns of a 128-row
// tile (32 columns per thread).  Replaces three Linear launches and one LayerNorm launch per block
// (sa2: 62 + 15 us, sa3: 37 + 15 us at batch 256; four ~5 us launches each at batch 1).
// ---------------------------------------------------------------------------
constexpr int kTail256Threads = 1024;
constexpr int attn_tail256_smem_bytes() { return 4 * 128 * 128 + 4 * 256 * 128 + 1024; }

__global__ void __launch_bounds__(kTail256Threads, 1)
attn_tail256_kernel(const __grid_constant__ CUtensorMap tmap_av, const __grid_constant__ CUtensorMap tmap_wp,
                    const __grid_constant__ CUtensorMap tmap_w1, const __grid_constant__ CUtensorMap tmap_w2,
                    const __grid_constant__ CUtensorMap tmap_out, const __nv_bfloat16* __restrict__ xl,
                    const TailParams p) {
  constexpr int C = 256, kSplit = 8, kKC = 4;
  constexpr uint32_t kChunkBytes = 128 * 128;   // one 128-row x 128-byte sub-tile of A
  constexpr uint32_t kWChunkBytes = C * 128;    // one K chunk of a weight: 256 rows x 128 bytes
  constexpr uint32_t kTmemCols = 256;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t w_bar;
  __shared__ __align__(8) uint64_t ld_bar;
  __shared__ __align__(8) uint64_t mma_bar;
  __shared__ uint32_t tmem_base_smem;
  __shared__ __align__(16) float prm_smem[5 * 256];   // bp | ln_w | ln_b | b1 | b2

  pdl_trigger();
  stage_tail_params<256, kTail256Threads>(prm_smem, p.bp, p.ln_w, p.ln_b, p.b1, p.b2);
  const int warp = threadIdx.x >> 5;
  const int split = warp >> 2;                  // which 32 columns of the row this thread owns
  const int q = warp & 3;                       // TMEM lane quadrant of this warp
  const bool lead = warp == 0;                  // issues every TMA and MMA
  const int t = q * 32 + (threadIdx.x & 31);    // row of the tile
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  uint8_t* bufA = smem;                          // [4][128 rows][128 B]: av -> h -> f1 -> out
  uint8_t* wsm = smem + kKC * kChunkBytes;       // [4][256 rows][128 B]: the current weight

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmap_av); tma_prefetch_desc(&tmap_out);
    tma_prefetch_desc(&tmap_wp); tma_prefetch_desc(&tmap_w1); tma_prefetch_desc(&tmap_w2);
    mbar_init(&w_bar, 1); mbar_init(&ld_bar, 1); mbar_init(&mma_bar, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(&tmem_base_smem, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t d_tmem = tmem_base_smem;
  pdl_wait();

  const uint32_t idesc = umma_idesc(1u, 128, C);
  const int col0 = split * kTailCols;
  const float* prm = prm_smem + col0;            // this thread's 32 columns; vector k at prm + k * C
  const uint32_t t_row = d_tmem + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(col0);
  const uint32_t a_addr = smem_u32(bufA), w_addr = smem_u32(wsm);
  const int sw = t & 7;
  const uint32_t row_off = (col0 >> 6) * kChunkBytes + t * 128;
  const int j0 = (col0 & 63) >> 3;
  uint8_t* rowA = bufA + row_off;
  float2* scratch = reinterpret_cast<float2*>(bufA);
  uint32_t ph_ld = 0, ph_mma = 0, ph_w = 0;

  auto load_w = [&](const CUtensorMap* tm) {   // all four K chunks of one weight matrix
    if (lead) {
      if (elect_one()) {
        mbar_expect_tx(&w_bar, kKC * kWChunkBytes);
#pragma unroll
        for (int kc = 0; kc < kKC; ++kc) tma_load_2d(wsm + kc * kWChunkBytes, tm, &w_bar, kc * 64, 0);
      }
      __syncwarp();
    }
  };
  auto issue_gemm = [&]() {   // D = A(bufA) * W(current)^T
    if (lead) {
      mbar_wait(&w_bar, ph_w);
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int kc = 0; kc < kKC; ++kc) {
          const uint64_t ad = umma_desc_sw128(a_addr + kc * kChunkBytes);
          const uint64_t bd = umma_desc_sw128(w_addr + kc * kWChunkBytes);
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_ss<false>(d_tmem, ad + (k * 2), bd + (k * 2), idesc, (kc > 0 || k > 0) ? 1u : 0u);
        }
        umma_commit(&mma_bar);
      }
      __syncwarp();
    }
    ph_w ^= 1;
  };
  auto wait_gemm = [&]() {
    mbar_wait(&mma_bar, ph_mma);
    ph_mma ^= 1;
    tc_fence_after();
  };

  if (static_cast<int>(blockIdx.x) < p.m_tiles) load_w(&tmap_wp);
  for (int tile = blockIdx.x; tile < p.m_tiles; tile += gridDim.x) {
    const int m0 = tile * 128;
    int b = 0, y = 0, x = 0;
    const bool valid = geo_decode(p.geo, m0 + t, b, y, x);
    if (lead) {
      if (elect_one()) {
        tma_store_wait_read<0>();   // the previous tile's store has read bufA
        mbar_expect_tx(&ld_bar, kKC * kChunkBytes);
#pragma unroll
        for (int kc = 0; kc < kKC; ++kc) tma_load_2d(bufA + kc * kChunkBytes, &tmap_av, &ld_bar, kc * 64, m0);
      }
      __syncwarp();
    }
    // this thread's 32 columns of x_ln (the residual of the block), straight from global memory
    uint4 xr[4];
    {
      const uint4* xp = reinterpret_cast<const uint4*>(xl + static_cast<size_t>(m0 + t) * C + col0);
#pragma unroll
      for (int j = 0; j < 4; ++j) xr[j] = __ldg(xp + j);
    }
    mbar_wait(&ld_bar, ph_ld);
    ph_ld ^= 1;
    issue_gemm();
    wait_gemm();
    load_w(&tmap_w1);

    // ---- o = acc + bp + x_ln (kept in registers as packed bf16); h = LayerNorm(o) -> A
    uint32_t ob[16];
    {
      float2 s2 = make_float2(0.f, 0.f), ss2 = make_float2(0.f, 0.f);
#pragma unroll
      for (int half = 0; half < 2; ++half) {   // 16 accumulator columns at a time: 64 registers per thread
        float o[16];
        tmem_ld16(t_row + 16 * half, o);
        tmem_ld_wait();
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) {
          const int j = 2 * half + jj;
          const uint32_t uw[4] = {xr[j].x, xr[j].y, xr[j].z, xr[j].w};
          float2 bv[4];
          ld_pairs<true>(prm + 0 * C + 8 * j, bv);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float2 acc = make_float2(o[8 * jj + 2 * i], o[8 * jj + 2 * i + 1]);
            ob[4 * j + i] = f2_to_bf2(__fadd2_rn(__fadd2_rn(acc, bv[i]), bf2_to_f2(uw[i])));
            const float2 of = bf2_to_f2(ob[4 * j + i]);   // LayerNorm sees the rounded o, as it would from HBM
            s2 = __fadd2_rn(s2, of);
            ss2 = __ffma2_rn(of, of, ss2);
          }
        }
      }
      scratch[split * 128 + t] = make_float2(s2.x + s2.y, ss2.x + ss2.y);   // (GEMM 0 has retired: bufA is idle)
      named_bar_sync(1, kTail256Threads);
      float ts = 0.f, tss = 0.f;
#pragma unroll
      for (int g = 0; g < kSplit; ++g) {   // the same order in every thread of the row
        const float2 v = scratch[g * 128 + t];
        ts += v.x; tss += v.y;
      }
      const float mean = ts * (1.0f / C);
      const float rstd = rsqrtf(fmaxf(tss * (1.0f / C) - mean * mean, 0.f) + 1e-5f);
      named_bar_sync(1, kTail256Threads);   // the partials are read: A may be overwritten
      const float2 nm = make_float2(-mean, -mean), rs = make_float2(rstd, rstd);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float2 wv[4], bv[4];
        ld_pairs<true>(prm + 1 * C + 8 * j, wv);
        ld_pairs<true>(prm + 2 * C + 8 * j, bv);
        uint32_t ww[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 d = __fmul2_rn(__fadd2_rn(bf2_to_f2(ob[4 * j + i]), nm), rs);
          ww[i] = f2_to_bf2(__ffma2_rn(d, wv[i], bv[i]));
        }
        *reinterpret_cast<uint4*>(rowA + (((j0 + j) ^ sw) << 4)) = make_uint4(ww[0], ww[1], ww[2], ww[3]);
      }
    }
    fence_async_proxy();
    tc_fence_before();
    named_bar_sync(1, kTail256Threads);
    issue_gemm();
    wait_gemm();
    load_w(&tmap_w2);

    // ---- f1 = GELU(acc + b1) -> A
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      float v[16];
      tmem_ld16(t_row + 16 * half, v);
      tmem_ld_wait();
#pragma unroll
      for (int jj = 0; jj < 2; ++jj) {
        const int j = 2 * half + jj;
        float2 bv[4];
        ld_pairs<true>(prm + 3 * C + 8 * j, bv);
        float h[8];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 a = __fadd2_rn(bv[i], make_float2(v[8 * jj + 2 * i], v[8 * jj + 2 * i + 1]));
          h[2 * i] = a.x; h[2 * i + 1] = a.y;
        }
        gelu_inplace<__nv_bfloat16, 8>(h);
        uint32_t ww[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) ww[i] = f2_to_bf2(make_float2(h[2 * i], h[2 * i + 1]));
        *reinterpret_cast<uint4*>(rowA + (((j0 + j) ^ sw) << 4)) = make_uint4(ww[0], ww[1], ww[2], ww[3]);
      }
    }
    fence_async_proxy();
    tc_fence_before();
    named_bar_sync(1, kTail256Threads);
    issue_gemm();
    wait_gemm();
    if (tile + static_cast<int>(gridDim.x) < p.m_tiles) load_w(&tmap_wp);   // the next tile's first weight

    // ---- out = acc + b2 + o -> A -> TMA store (pad pixels are stored as zeros)
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      float v[16];
      tmem_ld16(t_row + 16 * half, v);
      tmem_ld_wait();
#pragma unroll
      for (int jj = 0; jj < 2; ++jj) {
        const int j = 2 * half + jj;
        uint4 w4 = make_uint4(0u, 0u, 0u, 0u);
        if (valid) {
          float2 bv[4];
          ld_pairs<true>(prm + 4 * C + 8 * j, bv);
          uint32_t ww[4];
#pragma unroll
          for (int i = 0; i < 4; ++i)
            ww[i] = f2_to_bf2(__fadd2_rn(__fadd2_rn(make_float2(v[8 * jj + 2 * i], v[8 * jj + 2 * i + 1]), bv[i]),
                                         bf2_to_f2(ob[4 * j + i])));
          w4 = make_uint4(ww[0], ww[1], ww[2], ww[3]);
        }
        *reinterpret_cast<uint4*>(rowA + (((j0 + j) ^ sw) << 4)) = w4;
      }
    }
    fence_async_proxy();
    tc_fence_before();
    named_bar_sync(1, kTail256Threads);
    if (lead) {
      if (elect_one()) {
#pragma unroll
        for (int kc = 0; kc < kKC; ++kc) tma_store_2d(&tmap_out, bufA + kc * kChunkBytes, kc * 64, m0);
        tma_store_commit();
      }
      __syncwarp();
    }
  }
  if (lead) tma_store_wait_all<0>();

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(d_tmem, kTmemCols);
  }
}

}  // namespace dmb
