// Non-GEMM kernels of the sampling path: embedding, first/last convolutions of the
// U-Net, GroupNorm application, pooling / upsampling glue, LayerNorm, the attention core,
// the fused CFG-combine + DDPM update, and the VAE decoder's GroupNorm(8) pair.
// All activations are padded pixel matrices (common.cuh); T is the storage type.
#pragma once

#include "common.cuh"
#include "philox.cuh"

namespace dmb {

constexpr int kEmbDim = 256;
constexpr int kCondDim = 12;
constexpr int kProjCols = 896;  // 128 + 256 + 256 + 128 + 64 + 64 emb_layer outputs
constexpr int kLatC = 4;
constexpr int kLatH = 28, kLatW = 28;   // latent map (diff.py:9, img_size 28)

// ---------------------------------------------------------------------------
// Embedding: pos_encoding(t) + class_emb[y] + cond_mlp([vals|mask]) and the six
// Linear(SiLU(emb)) projections of Down/Up.  unet_cond.py:155-167, 62-65, 82-85;
// unet_cond_geom.py:89-95.  One CTA (256 threads) per kEmbSamples samples (both guidance halves).
// ---------------------------------------------------------------------------
struct EmbWeights {
  const float* class_emb;  // [4][256]
  const float* w1t;        // [24][256]  cond_mlp.0 transposed
  const float* b1;         // [256]
  const float* w2t;        // [256][256] cond_mlp.2 transposed
  const float* b2;         // [256]
  const float* inv_freq;   // [128]
  const float* projt;      // [256][896] six emb_layer weights, transposed, concatenated
  const float* projb;      // [896]
};

constexpr int kEmbSamples = 2;   // samples per CTA (x 2 guidance halves): every weight is loaded once for all of them
constexpr int kEmbSplit = 4;     // blockIdx.y: slice of the projection columns (the small embedding part is recomputed)
static_assert(kProjCols % kEmbSplit == 0 && kProjCols / kEmbSplit <= 256, "one projection column per thread");

static __global__ void __launch_bounds__(256)
emb_kernel(EmbWeights w, const long long* __restrict__ t_vec, const long long* __restrict__ y,
           const int* __restrict__ dyn_flags /* {null_label, has_cond} */,
           const float* __restrict__ vals, const float* __restrict__ mask,
           int B, int doubled, float* __restrict__ emb_out, float* __restrict__ proj_out) {
  __shared__ float feat[kEmbSamples][2 * kCondDim];
  __shared__ float h1[kEmbSamples][kEmbDim];
  __shared__ float semb[2 * kEmbSamples][kEmbDim];   // [sample][half]
  const int b0 = blockIdx.x * kEmbSamples;
  const int nb = min(kEmbSamples, B - b0);
  const int tid = threadIdx.x;
  const int null_label = dyn_flags[0];
  const bool has_cond = dyn_flags[1] != 0;

  // cond_mlp([vals | mask]) does not depend on the guidance half: once per sample
  float cond[kEmbSamples];
#pragma unroll
  for (int s = 0; s < kEmbSamples; ++s) cond[s] = 0.f;
  if (has_cond) {
    if (tid < kEmbSamples * 2 * kCondDim) {
      const int s = tid / (2 * kCondDim), i = tid % (2 * kCondDim);
      float v = 0.f;
      if (s < nb) v = (i < kCondDim) ? vals[(b0 + s) * kCondDim + i] : mask[(b0 + s) * kCondDim + i - kCondDim];
      feat[s][i] = v;
    }
    __syncthreads();
    float h[kEmbSamples];
#pragma unroll
    for (int s = 0; s < kEmbSamples; ++s) h[s] = w.b1[tid];
#pragma unroll
    for (int i = 0; i < 2 * kCondDim; ++i) {
      const float wv = w.w1t[i * kEmbDim + tid];
#pragma unroll
      for (int s = 0; s < kEmbSamples; ++s) h[s] = fmaf(feat[s][i], wv, h[s]);
    }
#pragma unroll
    for (int s = 0; s < kEmbSamples; ++s) h1[s][tid] = silu(h[s]);
    __syncthreads();
#pragma unroll
    for (int s = 0; s < kEmbSamples; ++s) cond[s] = w.b2[tid];
#pragma unroll 16
    for (int i = 0; i < kEmbDim; ++i) {
      const float wv = w.w2t[i * kEmbDim + tid];
#pragma unroll
      for (int s = 0; s < kEmbSamples; ++s) cond[s] = fmaf(h1[s][i], wv, cond[s]);
    }
  }
  const int halves = doubled ? 2 : 1;
#pragma unroll
  for (int s = 0; s < kEmbSamples; ++s) {
    if (s < nb) {
      const int b = b0 + s;
      const float tf = static_cast<float>(t_vec[b]);
      const float arg = tf * w.inv_freq[tid & 127];
      const float pe = (tid < 128) ? sinf(arg) : cosf(arg);
      for (int hf = 0; hf < halves; ++hf) {
        // doubled: effective sample hf * B + b; half 0 carries the null label (diff.py:147-150)
        // labels outside class_emb's 4 rows never index memory (nn.Embedding raises for them, and so
        // does the Python layer before the launch): clamped here as a last line of defence
        long long yv = (doubled && hf == 0) ? static_cast<long long>(null_label) : y[b];
        yv = yv < 0 ? 0 : (yv > 3 ? 3 : yv);
        const int be = doubled ? hf * B + b : b;
        float e = pe + w.class_emb[yv * kEmbDim + tid];
        if (has_cond) e = e + cond[s];
        if (emb_out && blockIdx.y == 0) emb_out[be * kEmbDim + tid] = e;
        semb[2 * s + hf][tid] = silu(e);
      }
    }
  }
  __syncthreads();
  if (tid < kProjCols / kEmbSplit) {
    const int o = blockIdx.y * (kProjCols / kEmbSplit) + tid;
    float acc[2 * kEmbSamples];
#pragma unroll
    for (int q = 0; q < 2 * kEmbSamples; ++q) acc[q] = w.projb[o];
#pragma unroll 16
    for (int i = 0; i < kEmbDim; ++i) {
      const float wv = w.projt[i * kProjCols + o];
#pragma unroll
      for (int q = 0; q < 2 * kEmbSamples; ++q) acc[q] = fmaf(semb[q][i], wv, acc[q]);
    }
#pragma unroll
    for (int q = 0; q < 2 * kEmbSamples; ++q) {
      const int s = q >> 1, hf = q & 1;
      if (s < nb && hf < halves) {
        const int be = doubled ? hf * B + b0 + s : b0 + s;
        proj_out[be * kProjCols + o] = acc[q];
      }
    }
  }
}

// ---------------------------------------------------------------------------
// Direct 3x3 / pad 1 convolution from an fp32 NCHW latent (4 channels) into a PM
// activation.  Used for inc.double_conv.0 (4->64, no bias, GroupNorm row statistics)
// and the VAE decoder's dec.0 (4->256, bias, input divided by the VAE scale factor).
// Thread = one pixel x 64 output channels; blockIdx.z selects the 64-channel group.
// ---------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(128)
conv_in_kernel(const float* __restrict__ x, int B_src, const float* __restrict__ w_oihw,
               const float* __restrict__ bias, float in_div, Geo g, int cout,
               T* __restrict__ out, long long* __restrict__ sample_stats) {
  __shared__ float ws[36 * 64];
  const int co0 = blockIdx.z * 64;
  for (int i = threadIdx.x; i < 36 * 64; i += blockDim.x) {
    const int k = i / 64, co = i % 64;  // k = ci*9 + ky*3 + kx
    ws[i] = w_oihw[(co0 + co) * 36 + k];
  }
  __syncthreads();
  const int be = blockIdx.y;
  const int pix_raw = blockIdx.x * blockDim.x + threadIdx.x;
  const bool pix_ok = pix_raw < g.H * g.W;   // (no early return: the statistics are committed by whole warps)
  const int pix = pix_ok ? pix_raw : 0;
  const int y = pix / g.W, xx = pix % g.W;
  const float* xs = x + static_cast<size_t>(be % B_src) * kLatC * g.H * g.W;
  float in[36];
#pragma unroll
  for (int ci = 0; ci < kLatC; ++ci)
#pragma unroll
    for (int ky = 0; ky < 3; ++ky)
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const int iy = y + ky - 1, ix = xx + kx - 1;
        float v = 0.f;
        if (iy >= 0 && iy < g.H && ix >= 0 && ix < g.W) v = xs[(ci * g.H + iy) * g.W + ix] / in_div;
        in[ci * 9 + ky * 3 + kx] = v;
      }
  float acc[64];
#pragma unroll
  for (int co = 0; co < 64; ++co) acc[co] = bias ? bias[co0 + co] : 0.f;
#pragma unroll
  for (int k = 0; k < 36; ++k) {
    const float v = in[k];
#pragma unroll
    for (int co = 0; co < 64; ++co) acc[co] = fmaf(v, ws[k * 64 + co], acc[co]);
  }
  const int r = geo_row(g, be, y, xx);
  T* op = out + static_cast<size_t>(r) * cout + co0;
  if (pix_ok) {
#pragma unroll
    for (int co = 0; co < 64; co += 8) store8(op + co, acc + co);
  }
  if (sample_stats) {
    StatAcc sa;
#pragma unroll
    for (int co = 0; co < 64; co += 16) sa.add(stat_unit16(acc + co));
    stat_commit(sample_stats, be, pix_ok, sa.s1, sa.s2);
  }
}

// ---------------------------------------------------------------------------
// The same convolution for the bf16 mode, on the warp-level tensor-core path (mma.sync
// m16n8k16, bf16 in / fp32 accumulate): K = 4 channels x 9 taps = 36, padded to 48, is far too
// short for a tcgen05 pipeline, and the CUDA-core version above is bound by FFMA issue (15
// TFLOP/s).  One CTA = one sample x 64 output channels: the zero-bordered 4 x 30 x 30 latent and
// the weights sit in shared memory, every thread keeps its B fragments (48 registers) for the
// whole image, and each warp walks 16-pixel tiles, gathering the im2col A fragments straight
// from the staged latent.  Row statistics for the following GroupNorm are reduced in the quad.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ void mma_bf16_16816(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

constexpr int kConvInThreads = 256;
static __global__ void __launch_bounds__(kConvInThreads)
conv_in_mma_kernel(const float* __restrict__ x, int B_src, const float* __restrict__ w_oihw,
                   const float* __restrict__ bias, float in_div, Geo g, int cout,
                   __nv_bfloat16* __restrict__ out, long long* __restrict__ sample_stats) {
  pdl_trigger();
  pdl_wait();
  constexpr int PW = kLatW + 2, PH = kLatH + 2, PLANE = PW * PH;   // zero-bordered latent plane
  __shared__ float img[kLatC * PLANE];
  __shared__ __align__(16) float ws[64 * 36];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int be = blockIdx.x, co0 = blockIdx.y * 64;
  const float inv_div = 1.0f / in_div;
  const float* xs = x + static_cast<size_t>(be % B_src) * kLatC * kLatH * kLatW;
  // Staging: every global load of the CTA is issued before the first one is used (one memory latency in
  // front of the MMA loop; as a load -> store loop the compiler kept them serial: 15 + 9 dependent
  // round trips, about half of this kernel's time).  The latent is 784 float4 (a 28-float image row is
  // seven of them), the 64 x 36 weights 576; both bases are 256-byte aligned (Pool / upload()).
  {
    constexpr int kXV = kLatC * kLatH * kLatW / 4, kWV = 64 * 36 / 4;
    constexpr int kXU = (kXV + kConvInThreads - 1) / kConvInThreads, kWU = (kWV + kConvInThreads - 1) / kConvInThreads;
    static_assert(kLatW % 4 == 0, "image rows are whole float4s");
    float4 xv[kXU], wv[kWU];
    const float4* x4 = reinterpret_cast<const float4*>(xs);
    const bool x_al = (reinterpret_cast<uintptr_t>(xs) & 15) == 0;
    const float4* w4 = reinterpret_cast<const float4*>(w_oihw + static_cast<size_t>(co0) * 36);
#pragma unroll
    for (int u = 0; u < kXU; ++u)
      if (tid + u * kConvInThreads < kXV) {
        if (x_al) {
          xv[u] = __ldg(x4 + tid + u * kConvInThreads);
        } else {   // a caller's latent that is not 16-byte aligned (VAE decode takes the pointer as given)
          const float* q = xs + 4 * (tid + u * kConvInThreads);
          xv[u] = make_float4(__ldg(q), __ldg(q + 1), __ldg(q + 2), __ldg(q + 3));
        }
      }
#pragma unroll
    for (int u = 0; u < kWU; ++u)
      if (tid + u * kConvInThreads < kWV) wv[u] = __ldg(w4 + tid + u * kConvInThreads);
    // the zero border of the four planes (two rows, two columns each)
    for (int i = tid; i < kLatC * PLANE; i += kConvInThreads) {
      const int ci = i / PLANE, rem = i - ci * PLANE, py = rem / PW, px = rem - py * PW;
      if (py == 0 || py == PH - 1 || px == 0 || px == PW - 1) img[i] = 0.f;
    }
#pragma unroll
    for (int u = 0; u < kXU; ++u) {
      const int e = (tid + u * kConvInThreads) * 4;
      if (e < kXV * 4) {
        const int ci = e / (kLatH * kLatW), rem = e - ci * (kLatH * kLatW), py = rem / kLatW, px = rem - py * kLatW;
        float* d = img + ci * PLANE + (py + 1) * PW + px + 1;
        d[0] = xv[u].x * inv_div; d[1] = xv[u].y * inv_div; d[2] = xv[u].z * inv_div; d[3] = xv[u].w * inv_div;
      }
    }
#pragma unroll
    for (int u = 0; u < kWU; ++u)
      if (tid + u * kConvInThreads < kWV) *reinterpret_cast<float4*>(ws + (tid + u * kConvInThreads) * 4) = wv[u];
  }
  __syncthreads();

  const int gq = lane >> 2, tq = lane & 3;
  // this thread's K indices: kstep ks -> k = ks*16 + 2*tq + {0, 1, 8, 9}; k = ci*9 + ky*3 + kx
  int koff[3][4];
  uint32_t bfrag[8][3][2];
#pragma unroll
  for (int ks = 0; ks < 3; ++ks) {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int k = ks * 16 + 2 * tq + (e & 1) + (e >> 1) * 8;
      const int ci = k / 9, ky = (k - ci * 9) / 3, kx = k - ci * 9 - ky * 3;
      koff[ks][e] = k < 36 ? ci * PLANE + ky * PW + kx : 0;
    }
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      float wv[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int k = ks * 16 + 2 * tq + (e & 1) + (e >> 1) * 8;
        wv[e] = k < 36 ? ws[(nt * 8 + gq) * 36 + k] : 0.f;
      }
      bfrag[nt][ks][0] = pack_bf16x2(wv[0], wv[1]);
      bfrag[nt][ks][1] = pack_bf16x2(wv[2], wv[3]);
    }
  }
  float bv[8][2];
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) {
    bv[nt][0] = bias ? bias[co0 + nt * 8 + 2 * tq] : 0.f;
    bv[nt][1] = bias ? bias[co0 + nt * 8 + 2 * tq + 1] : 0.f;
  }

  constexpr int NPIX = kLatH * kLatW;
  StatAcc sacc;   // this lane's share of the sample's GroupNorm statistics
  // gridDim.z CTAs share the 49 pixel tiles of a sample (small batches: one CTA per sample would be the
  // whole latency of this launch)
  for (int mt = warp + (kConvInThreads / 32) * blockIdx.z; mt * 16 < NPIX; mt += (kConvInThreads / 32) * gridDim.z) {
    const int p0 = mt * 16 + gq, p1 = p0 + 8;
    const bool v0 = p0 < NPIX, v1 = p1 < NPIX;
    const int y0 = v0 ? p0 / kLatW : 0, x0 = v0 ? p0 - y0 * kLatW : 0;
    const int y1 = v1 ? p1 / kLatW : 0, x1 = v1 ? p1 - y1 * kLatW : 0;
    const float* i0 = img + y0 * PW + x0;   // top-left of the 3x3 window in the bordered plane
    const float* i1 = img + y1 * PW + x1;
    float acc[8][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) { acc[nt][0] = bv[nt][0]; acc[nt][1] = bv[nt][1]; acc[nt][2] = bv[nt][0]; acc[nt][3] = bv[nt][1]; }
#pragma unroll
    for (int ks = 0; ks < 3; ++ks) {
      uint32_t a[4];
      a[0] = pack_bf16x2(i0[koff[ks][0]], i0[koff[ks][1]]);
      a[1] = pack_bf16x2(i1[koff[ks][0]], i1[koff[ks][1]]);
      a[2] = pack_bf16x2(i0[koff[ks][2]], i0[koff[ks][3]]);
      a[3] = pack_bf16x2(i1[koff[ks][2]], i1[koff[ks][3]]);
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) mma_bf16_16816(acc[nt], a, bfrag[nt][ks][0], bfrag[nt][ks][1]);
    }
    // rows p0 (acc[.][0..1]) and p1 (acc[.][2..3]); channels co0 + nt*8 + 2*tq + {0, 1}
    float s0 = 0.f, q0 = 0.f, s1 = 0.f, q1 = 0.f;
    const int r0 = geo_row(g, be, y0, x0), r1 = geo_row(g, be, y1, x1);
    __nv_bfloat16* o0 = out + static_cast<size_t>(r0) * cout + co0 + 2 * tq;
    __nv_bfloat16* o1 = out + static_cast<size_t>(r1) * cout + co0 + 2 * tq;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      s0 += acc[nt][0] + acc[nt][1]; q0 = fmaf(acc[nt][0], acc[nt][0], fmaf(acc[nt][1], acc[nt][1], q0));
      s1 += acc[nt][2] + acc[nt][3]; q1 = fmaf(acc[nt][2], acc[nt][2], fmaf(acc[nt][3], acc[nt][3], q1));
      if (v0) *reinterpret_cast<uint32_t*>(o0 + nt * 8) = pack_bf16x2(acc[nt][0], acc[nt][1]);
      if (v1) *reinterpret_cast<uint32_t*>(o1 + nt * 8) = pack_bf16x2(acc[nt][2], acc[nt][3]);
    }
    if (sample_stats) {
      // this lane's 16 values of each of its two rows (fp32, fixed order) -> fixed point
      if (v0) sacc.add(make_float2(s0, q0));
      if (v1) sacc.add(make_float2(s1, q1));
    }
  }
  if (sample_stats) stat_commit(sample_stats, be, true, sacc.s1, sacc.s2);   // one CTA = one sample
}

// ---------------------------------------------------------------------------
// GroupNorm(1, C) application.  The producing conv left the sample's (sum, sumsq) as two
// fixed-point integers (common.cuh, StatAcc); each CTA normalises its slice.  mode: 0 affine only, 1 GELU, 2 GELU(residual + .), 3 + emb[b, c].
// models/unet_cond.py:18-28, 67-70, 98-100.
// ---------------------------------------------------------------------------
enum GnMode : int { GN_NONE = 0, GN_GELU = 1, GN_RES_GELU = 2, GN_ADD_EMB = 3 };

// Streams one sample's PM block [S rows x C] contiguously: a thread owns one 8-channel vector
// column (scale / shift live in registers) and walks the rows; pad rows are skipped (they
// stay zero in `out`).  grid = (splits, n).  MODE is a template parameter to keep the
// register count low enough for 4 CTAs per SM (the erf GELU makes this kernel issue-bound
// as much as HBM-bound).
// kLN (with GN_ADD_EMB, C <= 256): the block's output feeds nn.LayerNorm(C) per token and
// nothing else (unet_cond.py:47: x_ln = x = self.ln(x)), and the C/8 lanes that share a row sit
// in one warp, so the LayerNorm is applied here -- on the bf16-rounded value, exactly what the
// separate launch would have read back -- and `out` receives LayerNorm(x); the un-normalised x
// is stored to `out_pre` only when the caller keeps it for inspection (taps).
template <typename T, int MODE, bool kLN = false>
__global__ void __launch_bounds__(256, 3)
gn1_apply_kernel(const T* __restrict__ raw, const long long* __restrict__ sample_stats,
                 const float* __restrict__ gamma, const float* __restrict__ beta,
                 const T* __restrict__ residual, const float* __restrict__ embproj,
                 int emb_off, Geo g, int C, T* __restrict__ out, int n_src, double inv_cnt,
                 const float* __restrict__ ln_w = nullptr, const float* __restrict__ ln_b = nullptr,
                 T* __restrict__ out_pre = nullptr) {
  pdl_trigger();
  pdl_wait();
  // `raw` holds n_src samples; output sample b reads raw sample b mod n_src (n_src < gridDim.y
  // only where the classifier-free-guidance pair still shares its activations, see
  // build_unet_plan).
  // Samples are walked from the LAST to the first: the producing conv wrote its tiles in ascending
  // row order, so the highest samples are the ones still in the 126 MB L2; this pass leaves the
  // lowest samples most recently written, which is where the next conv (ascending) starts.
  const int b = gridDim.y - 1 - blockIdx.y;
  const int b_src = b >= n_src ? b - n_src : b;
  const int tid = threadIdx.x;
  const int vpr = C >> 3;                  // vectors per row: 8, 16, 32 or 64
  const int rows_per_iter = 256 / vpr;
  const int cv = tid % vpr;
  const int c0 = cv * 8;
  // affine parameters (and the embedding row) do not depend on the statistics: their loads are
  // issued first, so they are in flight together with the partials below
  float gm[8], bt[8];
  {
    const float4 ga = *reinterpret_cast<const float4*>(gamma + c0), gb = *reinterpret_cast<const float4*>(gamma + c0 + 4);
    const float4 ba = *reinterpret_cast<const float4*>(beta + c0), bb = *reinterpret_cast<const float4*>(beta + c0 + 4);
    gm[0] = ga.x; gm[1] = ga.y; gm[2] = ga.z; gm[3] = ga.w; gm[4] = gb.x; gm[5] = gb.y; gm[6] = gb.z; gm[7] = gb.w;
    bt[0] = ba.x; bt[1] = ba.y; bt[2] = ba.z; bt[3] = ba.w; bt[4] = bb.x; bt[5] = bb.y; bt[6] = bb.z; bt[7] = bb.w;
    if (MODE == GN_ADD_EMB) {
      const float* e = embproj + static_cast<size_t>(b) * kProjCols + emb_off + c0;
      const float4 ea = *reinterpret_cast<const float4*>(e), eb = *reinterpret_cast<const float4*>(e + 4);
      bt[0] += ea.x; bt[1] += ea.y; bt[2] += ea.z; bt[3] += ea.w; bt[4] += eb.x; bt[5] += eb.y; bt[6] += eb.z; bt[7] += eb.w;
    }
  }
  // ---- statistics: two fixed-point sums per sample, left by the producing conv (StatAcc):
  // every thread reads them itself -- no reduction, no barrier in front of the streaming loop
  float mean, rstd;
  {
    const long long s1 = sample_stats[2 * b_src], s2 = sample_stats[2 * b_src + 1];
    // inv_cnt = 1 / (H W C) from the host: no double-precision division on this path
    const double dmean = static_cast<double>(s1) * kStatInv1 * inv_cnt;
    double var = static_cast<double>(s2) * kStatInv2 * inv_cnt - dmean * dmean;
    if (var < 0.0) var = 0.0;
    mean = static_cast<float>(dmean);
    rstd = rsqrtf(static_cast<float>(var) + 1e-5f);
  }
  float sc[8], sh[8];   // y = x * sc + sh  ==  (x - mean) * rstd * gamma + beta (+ emb)
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    sc[i] = rstd * gm[i];
    sh[i] = fmaf(-mean, sc[i], bt[i]);
  }
  // rows of this CTA: [r_begin, r_end) of the sample's S rows, skipping the leading pad row
  const int rows = g.S - g.Wp;
  const int per = (rows + gridDim.x - 1) / gridDim.x;
  const int r_begin = g.Wp + blockIdx.x * per;
  const int r_end = min(g.S, r_begin + per);
  const size_t base = static_cast<size_t>(b) * g.S;
  // Two register batches of U 16-byte loads per stream: the next batch is issued BEFORE the
  // current one is consumed, so loads stay in flight while the GELU runs (the kernel is bound by
  // its load -> compute -> store latency chain, not by HBM or by issue slots).
  constexpr int U = (MODE == GN_RES_GELU || kLN) ? 2 : 4;
  float lw[8], lb[8];
  uint32_t row_mask = 0xffffffffu;   // the lanes that share this thread's row
  if constexpr (kLN) {
    const float4 wa = *reinterpret_cast<const float4*>(ln_w + c0), wb = *reinterpret_cast<const float4*>(ln_w + c0 + 4);
    const float4 ba = *reinterpret_cast<const float4*>(ln_b + c0), bb = *reinterpret_cast<const float4*>(ln_b + c0 + 4);
    lw[0] = wa.x; lw[1] = wa.y; lw[2] = wa.z; lw[3] = wa.w; lw[4] = wb.x; lw[5] = wb.y; lw[6] = wb.z; lw[7] = wb.w;
    lb[0] = ba.x; lb[1] = ba.y; lb[2] = ba.z; lb[3] = ba.w; lb[4] = bb.x; lb[5] = bb.y; lb[6] = bb.z; lb[7] = bb.w;
    if (vpr < 32) row_mask = ((1u << vpr) - 1u) << ((tid & 31) & ~(vpr - 1));
  }
  T* pp = nullptr;
  // Incremental addressing: the row advances by rows_per_iter per load, its column inside the
  // padded image row (x) is tracked modulo Wp, so the loop has no division and no 64-bit
  // multiply.
  int rr = r_begin + tid / vpr;       // row cursor of the consumer
  int ri = rr;                        // row cursor of the load issue (one batch ahead)
  int x = rr - fdiv(rr, g.mWp, g.Wp) * g.Wp;
  const int xstep = rows_per_iter - fdiv(rows_per_iter, g.mWp, g.Wp) * g.Wp;  // rows_per_iter mod Wp
  const size_t first = (base + rr) * C + c0;
  const T* rp = raw + (static_cast<size_t>(b_src) * g.S + rr) * C + c0;
  const T* sp = (MODE == GN_RES_GELU) ? residual + first : nullptr;
  T* op = out + first;
  if constexpr (kLN) pp = out_pre ? out_pre + first : nullptr;
  const int stride = rows_per_iter * C;
  auto issue = [&](Raw8<T>* xr, Raw8<T>* rs, bool* ok) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      ok[u] = (ri + u * rows_per_iter < r_end) && (x != g.W);
      if (ok[u]) {
        xr[u].load(rp + u * stride);
        if (MODE == GN_RES_GELU) rs[u].load(sp + u * stride);
      }
      x += xstep;
      if (x >= g.Wp) x -= g.Wp;
    }
    ri += U * rows_per_iter;
    rp += U * stride;
    if (MODE == GN_RES_GELU) sp += U * stride;
  };
  auto consume = [&](const Raw8<T>* xr, const Raw8<T>* rs, const bool* ok) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (ok[u]) {
        float xv[8];
        xr[u].unpack(xv);
        if constexpr (sizeof(T) == 2 && (MODE == GN_GELU || MODE == GN_RES_GELU)) {
          // bf16 mode: affine, residual add and GELU on packed fp32x2 instructions, GELU four
          // elements per reciprocal (common.cuh, gelu_fast4)
          float rv[8];
          if (MODE == GN_RES_GELU) rs[u].unpack(rv);
#pragma unroll
          for (int i = 0; i < 8; i += 4) {
            float2 va = __ffma2_rn(make_float2(xv[i], xv[i + 1]), make_float2(sc[i], sc[i + 1]), make_float2(sh[i], sh[i + 1]));
            float2 vb = __ffma2_rn(make_float2(xv[i + 2], xv[i + 3]), make_float2(sc[i + 2], sc[i + 3]), make_float2(sh[i + 2], sh[i + 3]));
            if (MODE == GN_RES_GELU) {
              va = __fadd2_rn(make_float2(rv[i], rv[i + 1]), va);
              vb = __fadd2_rn(make_float2(rv[i + 2], rv[i + 3]), vb);
            }
            if constexpr (kGelu4) {
              gelu_fast4(va, vb);
            } else {
              va = gelu_fast2(va); vb = gelu_fast2(vb);
            }
            xv[i] = va.x; xv[i + 1] = va.y; xv[i + 2] = vb.x; xv[i + 3] = vb.y;
          }
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) xv[i] = fmaf(xv[i], sc[i], sh[i]);
          if (MODE == GN_GELU) {
#pragma unroll
            for (int i = 0; i < 8; ++i) xv[i] = gelu_for<T>(xv[i]);
          } else if (MODE == GN_RES_GELU) {
            float rv[8];
            rs[u].unpack(rv);
#pragma unroll
            for (int i = 0; i < 8; ++i) xv[i] = gelu_for<T>(rv[i] + xv[i]);
          }
        }
        if constexpr (kLN) {
          // (the lanes of a row take this branch together: ok[u] depends on the row only)
#pragma unroll
          for (int i = 0; i < 8; ++i) xv[i] = to_f32(from_f32<T>(xv[i]));   // what the stored tensor holds
          if (pp) store8(pp + u * stride, xv);
          float sum = 0.f;
#pragma unroll
          for (int i = 0; i < 8; ++i) sum += xv[i];
          for (int o = vpr >> 1; o > 0; o >>= 1) sum += __shfl_xor_sync(row_mask, sum, o);
          const float mu = sum / static_cast<float>(C);
          float q = 0.f;
#pragma unroll
          for (int i = 0; i < 8; ++i) { const float d = xv[i] - mu; q += d * d; }
          for (int o = vpr >> 1; o > 0; o >>= 1) q += __shfl_xor_sync(row_mask, q, o);
          const float rs_ln = rsqrtf(q / static_cast<float>(C) + 1e-5f);
#pragma unroll
          for (int i = 0; i < 8; ++i) xv[i] = (xv[i] - mu) * rs_ln * lw[i] + lb[i];
        }
        store8(op + u * stride, xv);
      }
    }
    rr += U * rows_per_iter;
    op += U * stride;
    if constexpr (kLN) { if (pp) pp += U * stride; }
  };
  Raw8<T> xa[U], ra[U], xb[U], rb[U];
  bool oka[U], okb[U];
  issue(xa, ra, oka);
  while (rr < r_end) {
    issue(xb, rb, okb);
    consume(xa, ra, oka);
    if (rr >= r_end) break;
    issue(xa, ra, oka);
    consume(xb, rb, okb);
  }
}

// MaxPool2d(2), floor.  models/unet_cond.py:58.  grid = (chunks of the sample's output vectors, n): the index
// arithmetic is one multiply-high division per thread (the first version decoded a 64-bit linear index with six
// run-time divisions, ~400 instructions around five 16-byte memory operations); the four loads go out together.
template <typename T>
__global__ void __launch_bounds__(256)
maxpool2_kernel(const T* __restrict__ in, Geo gi, Geo go, int C, T* __restrict__ out) {
  pdl_trigger();
  pdl_wait();
  const int vpr = C >> 3;                      // a power of two (C = 64 ... 512)
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= go.H * go.W * vpr) return;
  const int b = blockIdx.y;
  const int c0 = (v & (vpr - 1)) * 8;
  const int pix = v / vpr;
  const int y = fdiv(pix, go.mW, go.W), x = pix - y * go.W;
  const T* src = in + static_cast<size_t>(geo_row(gi, b, 2 * y, 2 * x)) * C + c0;
  Raw8<T> r[4];
  r[0].load(src);
  r[1].load(src + C);
  r[2].load(src + static_cast<size_t>(gi.Wp) * C);
  r[3].load(src + static_cast<size_t>(gi.Wp + 1) * C);
  float m[8], t[8];
  r[0].unpack(m);
#pragma unroll
  for (int k = 1; k < 4; ++k) {
    r[k].unpack(t);
#pragma unroll
    for (int i = 0; i < 8; ++i) m[i] = fmaxf(m[i], t[i]);
  }
  store8(out + static_cast<size_t>(geo_row(go, b, y, x)) * C + c0, m);
}

// Up-stage input: cat([skip, pad(bilinear_x2_align_corners(x))], channel).
// models/unet_cond.py:76, 87-97.  Output geometry = skip geometry.
// grid = (blocks_a + blocks_b, n): the first blocks_a blocks of a sample copy the skip channels, the others
// interpolate the upsampled channels -- a warp is all copy or all interpolation (with both kinds of vector in
// one row of threads every warp ran both paths with half of its lanes idle, and the kernel is bound by issue
// slots).  A thread moves kUpcatU vectors of the same channel group, all loads first (a 16-byte vector per
// thread does not keep HBM busy).
constexpr int kUpcatU = 4;
template <typename T>
__global__ void __launch_bounds__(256, 4)
upcat_kernel(const T* __restrict__ skip, int Cs, int n_skip, const T* __restrict__ xin, Geo gx,
             int Cx, Geo go, T* __restrict__ out, int blocks_a, int quarter_a, int quarter_b
             /* vectors per sample and part / kUpcatU, multiples of the part's vectors per pixel */) {
  pdl_trigger();
  pdl_wait();
  const int C = Cs + Cx;
  const bool copy = static_cast<int>(blockIdx.x) < blocks_a;
  const int vpr = (copy ? Cs : Cx) / 8;            // this part's vectors per pixel
  const int quarter = copy ? quarter_a : quarter_b;
  const int per_sample = go.H * go.W * vpr;
  const int v0 = (static_cast<int>(blockIdx.x) - (copy ? 0 : blocks_a)) * blockDim.x + threadIdx.x;
  if (v0 >= quarter) return;
  const int b = blockIdx.y;
  const int pix0 = v0 / vpr;
  const int cpart = (v0 - pix0 * vpr) * 8;    // channel inside the part; the same for all kUpcatU vectors: quarter % vpr == 0
  const int pstep = quarter / vpr;
  int row[kUpcatU], py[kUpcatU], px[kUpcatU];
  bool ok[kUpcatU];
#pragma unroll
  for (int u = 0; u < kUpcatU; ++u) {
    const int pix = pix0 + u * pstep;
    ok[u] = v0 + u * quarter < per_sample;
    py[u] = fdiv(pix, go.mW, go.W);
    px[u] = pix - py[u] * go.W;
    row[u] = geo_row(go, b, py[u], px[u]);
  }
  if (copy) {
    // the skip tensor may hold only the shared half of a guidance pair (n_skip = go.n / 2)
    const int shift = b >= n_skip ? n_skip * go.S : 0;
    Raw8<T> r[kUpcatU];
#pragma unroll
    for (int u = 0; u < kUpcatU; ++u)
      if (ok[u]) r[u].load(skip + static_cast<size_t>(row[u] - shift) * Cs + cpart);
#pragma unroll
    for (int u = 0; u < kUpcatU; ++u)
      if (ok[u]) r[u].store(out + static_cast<size_t>(row[u]) * C + cpart);
    return;
  }
  const int cx = cpart;
  const int c0 = Cs + cpart;
  const int Hu = 2 * gx.H, Wu = 2 * gx.W;
  // align_corners=True: src = dst * (in - 1) / (out - 1), computed in fp32 like ATen.
  const float sy = (Hu > 1) ? static_cast<float>(gx.H - 1) / static_cast<float>(Hu - 1) : 0.f;
  const float sx = (Wu > 1) ? static_cast<float>(gx.W - 1) / static_cast<float>(Wu - 1) : 0.f;
#pragma unroll
  for (int u = 0; u < kUpcatU; ++u) {
    if (!ok[u]) continue;
    const int y = py[u], x = px[u];
    float o[8];
    if (y < Hu && x < Wu) {
      const float fy = sy * y, fx = sx * x;
      const int y0 = static_cast<int>(fy), x0 = static_cast<int>(fx);
      const int y1 = y0 + ((y0 < gx.H - 1) ? 1 : 0), x1 = x0 + ((x0 < gx.W - 1) ? 1 : 0);
      const float ly1 = fy - y0, ly0 = 1.f - ly1, lx1 = fx - x0, lx0 = 1.f - lx1;
      Raw8<T> q[4];
      q[0].load(xin + static_cast<size_t>(geo_row(gx, b, y0, x0)) * Cx + cx);
      q[1].load(xin + static_cast<size_t>(geo_row(gx, b, y0, x1)) * Cx + cx);
      q[2].load(xin + static_cast<size_t>(geo_row(gx, b, y1, x0)) * Cx + cx);
      q[3].load(xin + static_cast<size_t>(geo_row(gx, b, y1, x1)) * Cx + cx);
      float v00[8], v01[8], v10[8], v11[8];
      q[0].unpack(v00); q[1].unpack(v01); q[2].unpack(v10); q[3].unpack(v11);
#pragma unroll
      for (int i = 0; i < 8; ++i)
        o[i] = ly0 * (lx0 * v00[i] + lx1 * v01[i]) + ly1 * (lx0 * v10[i] + lx1 * v11[i]);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = 0.f;
    }
    store8(out + static_cast<size_t>(row[u]) * C + c0, o);
  }
}

// LayerNorm over the channels of every real pixel (token).  LPT = C/8 lanes share a token,
// 16 bytes per lane; the PM rows are streamed contiguously and pad rows are skipped (they
// stay zero in `out`).  models/unet_cond.py:37, 39, 48.
template <typename T, int LPT>
__global__ void __launch_bounds__(256, 3)
layernorm_kernel(const T* __restrict__ in, const float* __restrict__ w, const float* __restrict__ bias,
                 Geo g, T* __restrict__ out) {
  pdl_trigger();
  pdl_wait();
  constexpr int C = LPT * 8;
  constexpr int TPB = 256 / LPT;  // tokens (PM rows) per block iteration
  const int sub = threadIdx.x % LPT;
  const int c0 = sub * 8;
  float wv[8], bv[8];
  {
    const float4 wa = *reinterpret_cast<const float4*>(w + c0), wb = *reinterpret_cast<const float4*>(w + c0 + 4);
    const float4 ba = *reinterpret_cast<const float4*>(bias + c0), bb = *reinterpret_cast<const float4*>(bias + c0 + 4);
    wv[0] = wa.x; wv[1] = wa.y; wv[2] = wa.z; wv[3] = wa.w; wv[4] = wb.x; wv[5] = wb.y; wv[6] = wb.z; wv[7] = wb.w;
    bv[0] = ba.x; bv[1] = ba.y; bv[2] = ba.z; bv[3] = ba.w; bv[4] = bb.x; bv[5] = bb.y; bv[6] = bb.z; bv[7] = bb.w;
  }
  constexpr int U = 8;
  for (int r0 = blockIdx.x * TPB * U + threadIdx.x / LPT; r0 < g.rows_valid; r0 += gridDim.x * TPB * U) {
    Raw8<T> xr[U];
    bool ok[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int r = r0 + u * TPB;
      ok[u] = false;
      if (r < g.rows_valid) ok[u] = geo_is_pixel(g, r - fdiv(r, g.mS, g.S) * g.S);
      if (ok[u]) xr[u].load(in + static_cast<size_t>(r) * C + c0);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      // all lanes take part in the shuffles; pad rows just do not store
      float v[8];
      if (ok[u]) xr[u].unpack(v);
      else {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = 0.f;
      }
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) s += v[i];
#pragma unroll
      for (int o = LPT / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      const float mean = s * (1.0f / C);
      float q = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) { const float d = v[i] - mean; q += d * d; }
#pragma unroll
      for (int o = LPT / 2; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
      const float rstd = rsqrtf(q * (1.0f / C) + 1e-5f);
      if (ok[u]) {
        float o8[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) o8[i] = (v[i] - mean) * rstd * wv[i] + bv[i];
        store8(out + static_cast<size_t>(r0 + u * TPB) * C + c0, o8);
      }
    }
  }
}

// Attention core: softmax(q k^T / sqrt(D)) v per (sample, head) over the real pixels.
// qkv is the in-proj output [rows, 3E] (q | k | v).  K and V of the head live in shared
// memory as fp32; one thread owns one query and runs an online softmax over key blocks.
// models/unet_cond.py:36, 49 (nn.MultiheadAttention, 4 heads, batch_first).
template <typename T, int D>
__global__ void __launch_bounds__(128)
attention_kernel(const T* __restrict__ qkv, Geo g, int E, T* __restrict__ out) {
  pdl_trigger();
  pdl_wait();
  extern __shared__ float kv[];
  const int L = g.H * g.W;
  float* Ks = kv;
  float* Vs = kv + static_cast<size_t>(L) * D;
  const int b = blockIdx.z, h = blockIdx.y;
  const int ld = 3 * E;
  constexpr int VPT = D / 8;
  for (int idx = threadIdx.x; idx < L * VPT; idx += blockDim.x) {
    const int pix = idx / VPT, dv = (idx % VPT) * 8;
    const size_t ro = static_cast<size_t>(geo_row_of_pixel(g, b, pix)) * ld + h * D + dv;
    load8(qkv + ro + E, Ks + pix * D + dv);
    load8(qkv + ro + 2 * E, Vs + pix * D + dv);
  }
  __syncthreads();
  const int qi = blockIdx.x * blockDim.x + threadIdx.x;
  if (qi >= L) return;
  const int qrow = geo_row(g, b, qi / g.W, qi % g.W);
  float q[D], acc[D];
  {
    const T* qp = qkv + static_cast<size_t>(qrow) * ld + h * D;
    const float sc = rsqrtf(static_cast<float>(D)) * 1.4426950408889634f;  // 1/sqrt(D) * log2(e)
#pragma unroll
    for (int d0 = 0; d0 < D; d0 += 8) {
      load8(qp + d0, q + d0);
#pragma unroll
      for (int i = 0; i < 8; ++i) { q[d0 + i] *= sc; acc[d0 + i] = 0.f; }
    }
  }
  float m = -INFINITY, l = 0.f;
  constexpr int KB = 8;
  for (int j0 = 0; j0 < L; j0 += KB) {
    float s[KB];
    float bm = -INFINITY;
#pragma unroll
    for (int jj = 0; jj < KB; ++jj) {
      const int j = j0 + jj;
      float dot = -INFINITY;
      if (j < L) {
        const float4* kp = reinterpret_cast<const float4*>(Ks + j * D);
        dot = 0.f;
#pragma unroll
        for (int d4 = 0; d4 < D / 4; ++d4) {
          const float4 kk = kp[d4];
          dot = fmaf(q[4 * d4], kk.x, dot); dot = fmaf(q[4 * d4 + 1], kk.y, dot);
          dot = fmaf(q[4 * d4 + 2], kk.z, dot); dot = fmaf(q[4 * d4 + 3], kk.w, dot);
        }
      }
      s[jj] = dot;
      bm = fmaxf(bm, dot);
    }
    const float mn = fmaxf(m, bm);
    const float corr = exp2f(m - mn);
    l *= corr;
#pragma unroll
    for (int d = 0; d < D; ++d) acc[d] *= corr;
#pragma unroll
    for (int jj = 0; jj < KB; ++jj) {
      const int j = j0 + jj;
      if (j < L) {
        const float pj = exp2f(s[jj] - mn);
        l += pj;
        const float4* vp = reinterpret_cast<const float4*>(Vs + j * D);
#pragma unroll
        for (int d4 = 0; d4 < D / 4; ++d4) {
          const float4 vv = vp[d4];
          acc[4 * d4] = fmaf(pj, vv.x, acc[4 * d4]); acc[4 * d4 + 1] = fmaf(pj, vv.y, acc[4 * d4 + 1]);
          acc[4 * d4 + 2] = fmaf(pj, vv.z, acc[4 * d4 + 2]); acc[4 * d4 + 3] = fmaf(pj, vv.w, acc[4 * d4 + 3]);
        }
      }
    }
    m = mn;
  }
  const float inv = 1.0f / l;
  T* op = out + static_cast<size_t>(qrow) * E + h * D;
#pragma unroll
  for (int d0 = 0; d0 < D; d0 += 8) {
    float o[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i] = acc[d0 + i] * inv;
    store8(op + d0, o);
  }
}

// ---------------------------------------------------------------------------
// Tail of one reverse step, fused: the `out` 1x1 conv (64 -> 4, bias) on both CFG halves,
// eps = eps_u + g (eps_c - eps_u), the DDPM posterior mean and the noise injection.
// unet_cond.py:153; diff.py:141-162.  Thread = one latent pixel (4 channels).
// feat rows [0, B) are the null-label pass, rows [B, 2B) the class-label pass.
// ---------------------------------------------------------------------------
struct StepParams {
  const float* x_in;       // (B,4,H,W) fp32
  float* x_out;            // may alias x_in
  const long long* t_vec;  // (B,) current timestep per sample, 1..T
  const float* coef;       // (T,3): (1-a)/sqrt(1-abar), sqrt(a), sigma
  int T;
  int slot_t0;             // noise / dump slot of a step = (slot_t0 - t) / t_stride
  int t_stride;            // timesteps per loop iteration (1 = the reference's schedule)
  float guidance;
  const float* noise;       // injected noise or null (then Philox)
  long long noise_step_stride;  // elements between consecutive steps' noise (0: single step)
  unsigned long long seed;
  unsigned long long sample_offset;
  float* eps_out;          // optional (B,4,H,W) or, with dump stride, (T,B,4,H,W)
  float* x_dump;           // optional (steps,B,4,H,W): x_t before the step
  long long dump_step_stride;
  const float* w_out;      // [4][64]
  const float* b_out;      // [4]
};

template <typename T>
__global__ void __launch_bounds__(128)
cfg_step_kernel(const T* __restrict__ feat, Geo g, int B, const StepParams* __restrict__ pp) {
  pdl_trigger();
  pdl_wait();
  // The 128 feature rows (64 channels) of this CTA's pixels, both guidance halves, are staged through
  // shared memory: eight lanes fetch one row's 16-byte chunks, so a warp instruction covers four
  // whole rows (a thread reading its own row touched 32 different 128-byte lines per instruction);
  // chunk c of row r sits at chunk slot c ^ (r & 7), which makes thread r's reads of its own row
  // conflict-free.  Every load of the CTA is issued before the first one is used.
  constexpr int kRowElems = 64;
  constexpr int kChunks = kRowElems * static_cast<int>(sizeof(T)) / 16;   // 16-byte chunks per row: 8 (bf16) / 16 (fp32)
  constexpr int kE = 16 / static_cast<int>(sizeof(T));                    // elements per chunk
  constexpr int kBuf = sizeof(T) == 2 ? 2 : 1;   // bf16: both halves in flight at once; fp32 (32 KB per half): one after the other
  __shared__ __align__(16) uint8_t rows[kBuf][128 * kChunks * 16];
  __shared__ __align__(16) float ws[kLatC * 64];
  __shared__ float bs[kLatC];
  const int HW = g.H * g.W;
  const int b = blockIdx.y;
  const int tid = threadIdx.x;
  const int pix0 = blockIdx.x * 128;
  const int pix = pix0 + tid;
  // thread -> (row, chunk) of its k-th staged vector: 128 rows x kChunks chunks over 128 threads
  auto load_half = [&](int half, uint4* st) {
#pragma unroll
    for (int k = 0; k < kChunks; ++k) {
      const int v = tid + k * 128, rl = v / kChunks, c = v - rl * kChunks;
      const int pq = pix0 + rl;
      if (pq < HW) {
        const int yq = fdiv(pq, g.mW, g.W), xq = pq - yq * g.W;
        st[k] = *(reinterpret_cast<const uint4*>(feat + static_cast<size_t>(geo_row(g, b + half * B, yq, xq)) * kRowElems) + c);
      }
    }
  };
  auto store_half = [&](int buf, const uint4* st) {
#pragma unroll
    for (int k = 0; k < kChunks; ++k) {
      const int v = tid + k * 128, rl = v / kChunks, c = v - rl * kChunks;
      if (pix0 + rl < HW) *reinterpret_cast<uint4*>(rows[buf] + rl * (kChunks * 16) + ((c ^ (rl & 7)) << 4)) = st[k];
    }
  };
  // the out conv of this thread's pixel from its staged row: k ascending, as one thread per row did
  auto dot_half = [&](int buf, float* e) {
#pragma unroll
    for (int ch = 0; ch < kChunks; ++ch) {
      const uint8_t* src = rows[buf] + tid * (kChunks * 16) + ((ch ^ (tid & 7)) << 4);
      float v[kE];
      if constexpr (sizeof(T) == 2) {
        load8(reinterpret_cast<const T*>(src), v);
      } else {
        const float4 a = *reinterpret_cast<const float4*>(src);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
      }
#pragma unroll
      for (int c = 0; c < kLatC; ++c) {   // (weights as 16-byte broadcast reads; per output the same ascending-k FMA chain)
        float wv[kE];
#pragma unroll
        for (int i = 0; i < kE; i += 4) {
          const float4 w4 = *reinterpret_cast<const float4*>(ws + c * 64 + ch * kE + i);
          wv[i] = w4.x; wv[i + 1] = w4.y; wv[i + 2] = w4.z; wv[i + 3] = w4.w;
        }
#pragma unroll
        for (int i = 0; i < kE; ++i) e[c] = fmaf(v[i], wv[i], e[c]);
      }
    }
  };
  uint4 stage[kBuf][kChunks];
  load_half(0, stage[0]);
  if constexpr (kBuf == 2) load_half(1, stage[1]);
  const StepParams p = *pp;
  for (int i = tid; i < kLatC * 64; i += 128) ws[i] = p.w_out[i];
  if (tid < kLatC) bs[tid] = p.b_out[tid];
  // the step's scalars and this pixel's x_t: a chain of dependent loads (pp -> t -> coefficients) that would
  // otherwise start only after the dot products
  const long long t = p.t_vec[b];
  // the Python layer asserts 1 <= t <= T like the reference (diff.py:140); the row index is clamped
  // as well so that a raw C-ABI caller with a bad t cannot read outside the table
  const long long tr = t < 1 ? 0 : (t > p.T ? p.T - 1 : t - 1);
  const float c1 = p.coef[tr * 3 + 0], sqrt_a = p.coef[tr * 3 + 1], sigma = p.coef[tr * 3 + 2];
  const size_t base = static_cast<size_t>(b) * kLatC * HW + min(pix, HW - 1);
  float xin[kLatC];
#pragma unroll
  for (int c = 0; c < kLatC; ++c) xin[c] = p.x_in[base + c * HW];
  store_half(0, stage[0]);
  if constexpr (kBuf == 2) store_half(1, stage[1]);
  __syncthreads();
  float eu[kLatC], ec[kLatC];
#pragma unroll
  for (int c = 0; c < kLatC; ++c) { eu[c] = bs[c]; ec[c] = bs[c]; }
  if constexpr (kBuf == 2) {
    if (pix < HW) { dot_half(0, eu); dot_half(1, ec); }
  } else {
    if (pix < HW) dot_half(0, eu);
    load_half(1, stage[0]);
    __syncthreads();            // every thread has read its row of the first half
    store_half(0, stage[0]);
    __syncthreads();
    if (pix < HW) dot_half(0, ec);
  }
  if (pix >= HW) return;
  const int slot = (p.slot_t0 - static_cast<int>(t)) / p.t_stride;
  float nz[kLatC] = {0.f, 0.f, 0.f, 0.f};
  if (t != 1) {
    if (p.noise) {
      const float* np = p.noise + p.noise_step_stride * slot + static_cast<size_t>(b) * kLatC * HW + pix;
#pragma unroll
      for (int c = 0; c < kLatC; ++c) nz[c] = np[c * HW];
    } else {
      philox_normal4(p.seed, p.sample_offset + b, static_cast<uint32_t>(t), static_cast<uint32_t>(pix), nz);
    }
  }
#pragma unroll
  for (int c = 0; c < kLatC; ++c) {
    const float xv = xin[c];
    const float eps = eu[c] + p.guidance * (ec[c] - eu[c]);
    if (p.x_dump) p.x_dump[p.dump_step_stride * slot + base + c * HW] = xv;
    if (p.eps_out) p.eps_out[p.dump_step_stride * slot + base + c * HW] = eps;
    const float mu = (xv - c1 * eps) / sqrt_a;
    p.x_out[base + c * HW] = mu + nz[c] * sigma;
  }
}

// `out` conv alone (eager forward): eps (B,4,H,W) fp32.
template <typename T>
__global__ void __launch_bounds__(128)
out_conv_kernel(const T* __restrict__ feat, Geo g, const float* __restrict__ w_out,
                const float* __restrict__ b_out, float* __restrict__ eps) {
  __shared__ float ws[kLatC * 64];
  for (int i = threadIdx.x; i < kLatC * 64; i += blockDim.x) ws[i] = w_out[i];
  __syncthreads();
  const int HW = g.H * g.W;
  const int b = blockIdx.y;
  const int pix = blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= HW) return;
  float e[kLatC];
#pragma unroll
  for (int c = 0; c < kLatC; ++c) e[c] = b_out[c];
  const T* f = feat + static_cast<size_t>(geo_row_of_pixel(g, b, pix)) * 64;
#pragma unroll
  for (int k0 = 0; k0 < 64; k0 += 8) {
    float v[8];
    load8(f + k0, v);
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int c = 0; c < kLatC; ++c) e[c] = fmaf(v[i], ws[c * 64 + k0 + i], e[c]);
  }
#pragma unroll
  for (int c = 0; c < kLatC; ++c) eps[(static_cast<size_t>(b) * kLatC + c) * HW + pix] = e[c];
}

// GeomHead: mean over pixels -> Linear(64,256) -> SiLU -> Linear(256,12).
// unet_cond_geom.py:8-23.  One CTA per sample.
template <typename T>
__global__ void __launch_bounds__(256)
geom_head_kernel(const T* __restrict__ feat, Geo g, const float* __restrict__ w1t, const float* __restrict__ b1,
                 const float* __restrict__ w2, const float* __restrict__ b2, int out_dim, float* __restrict__ geom) {
  __shared__ float part[4][64];
  __shared__ float gmean[64];
  __shared__ float hid[256];
  const int b = blockIdx.x, tid = threadIdx.x;
  const int c = tid & 63, grp = tid >> 6;
  const int HW = g.H * g.W;
  float s = 0.f;
  for (int pix = grp; pix < HW; pix += 4)
    s += to_f32(feat[static_cast<size_t>(geo_row_of_pixel(g, b, pix)) * 64 + c]);
  part[grp][c] = s;
  __syncthreads();
  if (tid < 64) gmean[tid] = (part[0][tid] + part[1][tid] + part[2][tid] + part[3][tid]) / static_cast<float>(HW);
  __syncthreads();
  float h = b1[tid];
  for (int i = 0; i < 64; ++i) h = fmaf(gmean[i], w1t[i * 256 + tid], h);
  hid[tid] = silu(h);
  __syncthreads();
  if (tid < out_dim) {
    float o = b2[tid];
    for (int i = 0; i < 256; ++i) o = fmaf(hid[i], w2[tid * 256 + i], o);
    geom[b * out_dim + tid] = o;
  }
}

// t_vec[b] -= t_stride (end of a captured step).
static __global__ void advance_t_kernel(long long* t_vec, int B, const StepParams* __restrict__ pp) {
  pdl_trigger();
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < B) t_vec[i] -= pp->t_stride;
}
static __global__ void fill_t_kernel(long long* t_vec, int B, long long t) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < B) t_vec[i] = t;
}

// Standalone Philox fill (B,4,HW) for step t: the exact values cfg_step_kernel would draw.
static __global__ void philox_fill_kernel(float* out, int B, int HW, unsigned long long seed,
                                   unsigned long long sample_offset, unsigned int step) {
  const int b = blockIdx.y;
  const int pix = blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= HW) return;
  float nz[4];
  philox_normal4(seed, sample_offset + b, step, static_cast<uint32_t>(pix), nz);
#pragma unroll
  for (int c = 0; c < 4; ++c) out[(static_cast<size_t>(b) * 4 + c) * HW + pix] = nz[c];
}

// ---------------------------------------------------------------------------
// VAE decoder GroupNorm(8, C) + GELU, two passes.  models/vae.py:37-47.
// Pass 1: per (sample, chunk, group) partial sums.  blockDim 256 is a multiple of the
// vectors per row, so every thread stays inside one group.
// ---------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
gn8_stats_kernel(const T* __restrict__ raw, Geo g, int C, double2* __restrict__ partial, int phases = 1) {
  __shared__ double red[2][256];
  const int b = blockIdx.y, tid = threadIdx.x;
  const int vpr = C / 8;
  // A sample's PM rows are one contiguous block and its pad rows hold zeros (no producer writes them
  // anything else), which add nothing to either sum: the kernel streams all S rows with no index
  // arithmetic, eight 16-byte loads in flight per thread (the first version decoded a pixel per load
  // and waited for four: 76 registers, 2.2 TB/s, stalled on the loads -- profiles/r1_v7_ncu_full_gn8_stats.md).
  const int total = g.S * vpr;
  const int per = (((total + gridDim.x - 1) / gridDim.x) + 255) / 256 * 256;   // a multiple of 256, hence of vpr
  const int begin = blockIdx.x * per;
  const int end = min(total, begin + per);
  const T* src = raw + static_cast<size_t>(b) * g.S * C;
  float s = 0.f, ss = 0.f;
  constexpr int U = 8;
  for (int v0 = begin + tid; v0 < end; v0 += 256 * U) {
    Raw8<T> xr[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (v0 + u * 256 < end) xr[u].load(src + static_cast<size_t>(v0 + u * 256) * 8);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (v0 + u * 256 < end) {
        float xv[8];
        xr[u].unpack(xv);
#pragma unroll
        for (int i = 0; i < 8; ++i) { s += xv[i]; ss = fmaf(xv[i], xv[i], ss); }
      }
    }
  }
  red[0][tid] = s; red[1][tid] = ss;
  __syncthreads();
  {
    // thread j's vector column is j % vpr (begin is a multiple of vpr); group = column / (vpr/8) --
    // for a phase-planar row (C = phases x channels, see gn8_apply_gelu_d2s_kernel) the column inside
    // its phase slot.  Warp w gathers group w: fixed lane order, then a fixed shuffle tree.
    const int w = tid >> 5, lane = tid & 31, vpp = vpr / phases, vpg = vpp / 8;
    double a = 0.0, q = 0.0;
    for (int j = lane; j < 256; j += 32) {
      if ((((j % vpr) % vpp) / vpg) == w) { a += red[0][j]; q += red[1][j]; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      a += __shfl_xor_sync(0xffffffffu, a, o);
      q += __shfl_xor_sync(0xffffffffu, q, o);
    }
    if (lane == 0) partial[(static_cast<size_t>(b) * gridDim.x + blockIdx.x) * 8 + w] = make_double2(a, q);
  }
}

// Per-group mean / rstd of a sample from the chunk partials of gn8_stats_kernel: warp w (of 8)
// takes group w, its lanes take the chunks (independent loads), then a fixed shuffle tree.
__device__ __forceinline__ void gn8_finish_stats(const double2* __restrict__ partial, int b, int n_chunks, double cnt,
                                                 float* s_mean, float* s_rstd) {
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double a = 0.0, q = 0.0;
  for (int k = lane; k < n_chunks; k += 32) {
    const double2 v = partial[(static_cast<size_t>(b) * n_chunks + k) * 8 + w];
    a += v.x; q += v.y;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  if (lane == 0) {
    const double mean = a / cnt;
    double var = q / cnt - mean * mean;
    if (var < 0.0) var = 0.0;
    s_mean[w] = static_cast<float>(mean);
    s_rstd[w] = static_cast<float>(1.0 / sqrt(var + 1e-5));
  }
  __syncthreads();
}

// Pass 2: normalise + affine + GELU.  Like pass 1 it walks the sample's contiguous PM block (one
// division per 16-byte vector, to tell pad rows from pixels) and keeps the next batch of loads in
// flight while it works on the current one.
// kD2S = false: out has the geometry of raw; pad rows are rewritten as zeros.
// kD2S = true:  raw is PHASE-PLANAR -- the four ConvTranspose2d(4, 2, 1) phase GEMMs of a layer leave
//   through the TMA-store epilogue into one [pixel of the INPUT map][phase (2a + b)][channel] matrix
//   (row width C = 4 x channels) instead of scattering rows of the 2H x 2W map from registers; this
//   pass, which reads and rewrites every element anyway, puts pixel (y, x)'s phase (a, b) at pixel
//   (2y + a, 2x + b) of out (geometry g2).  Both sides move whole 16-byte channel vectors.
// models/vae.py:37-47.
template <typename T, bool kD2S>
__global__ void __launch_bounds__(256)
gn8_apply_gelu_kernel(const T* __restrict__ raw, const double2* __restrict__ partial, int n_chunks_stats,
                      const float* __restrict__ gamma, const float* __restrict__ beta, Geo g, int C,
                      Geo g2, T* __restrict__ out) {
  __shared__ float s_mean[8], s_rstd[8];
  const int b = blockIdx.y, tid = threadIdx.x;
  gn8_finish_stats(partial, b, n_chunks_stats, static_cast<double>(g.H) * g.W * (C / 8), s_mean, s_rstd);
  const int vpr = C / 8, vpp = kD2S ? vpr / 4 : vpr, vpg = vpp / 8;
  const int total = g.S * vpr;
  const int per = (((total + gridDim.x - 1) / gridDim.x) + 255) / 256 * 256;   // a multiple of 256, hence of vpr
  const int begin = blockIdx.x * per;
  const int end = min(total, begin + per);
  const int vsh = 31 - __clz(vpr);   // vpr is a power of two
  // a thread keeps its channel vector (gamma, beta and the group statistics folded into one multiply-add)
  const int col = tid & (vpr - 1);
  const int phase = col / vpp, vc = col - phase * vpp, c0 = vc * 8;
  const int Cout = kD2S ? C / 4 : C;
  const float mean = s_mean[vc / vpg], rstd = s_rstd[vc / vpg];
  float ga[8], be[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { ga[i] = gamma[c0 + i] * rstd; be[i] = beta[c0 + i] - mean * ga[i]; }
  const T* src = raw + static_cast<size_t>(b) * g.S * C;
  T* dst = out + static_cast<size_t>(b) * g2.S * Cout + c0;
  const int oy = phase >> 1, ox = phase & 1;
  constexpr int U = 4;
  Raw8<T> cur[U], nxt[U];
  int v0 = begin + tid;
#pragma unroll
  for (int u = 0; u < U; ++u)
    if (v0 + u * 256 < end) cur[u].load(src + static_cast<size_t>(v0 + u * 256) * 8);
  for (; v0 < end; v0 += 256 * U) {
    const int vn = v0 + 256 * U;
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (vn + u * 256 < end) nxt[u].load(src + static_cast<size_t>(vn + u * 256) * 8);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int v = v0 + u * 256;
      if (v < end) {
        const int rr = v >> vsh;
        const int yy = fdiv(rr, g.mWp, g.Wp), x = rr - yy * g.Wp;
        const bool pixel = yy >= 1 && x != g.W;
        float xv[8];
        cur[u].unpack(xv);
        if (pixel) {
#pragma unroll
          for (int i = 0; i < 8; ++i) xv[i] = fmaf(xv[i], ga[i], be[i]);
          gelu_inplace<T, 8>(xv);
        }
        if constexpr (kD2S) {
          if (pixel) store8(dst + static_cast<size_t>((2 * yy - 1 + oy) * g2.Wp + 2 * x + ox) * Cout, xv);
        } else {
          if (!pixel) {
#pragma unroll
            for (int i = 0; i < 8; ++i) xv[i] = 0.f;
          }
          store8(dst + static_cast<size_t>(rr) * Cout, xv);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) cur[u] = nxt[u];
  }
}

// ---------------------------------------------------------------------------
// Layout conversion for the op-level test entry points and debug taps.
// ---------------------------------------------------------------------------
template <typename T>
__global__ void nchw_to_pm_kernel(const float* __restrict__ in, Geo g, int C, T* __restrict__ out) {
  const long long total = static_cast<long long>(g.n) * C * g.H * g.W;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= total) return;
  // iterate in PM order (channel fastest) for coalesced writes
  const int c = static_cast<int>(i % C);
  long long pix = i / C;
  const int x = static_cast<int>(pix % g.W); pix /= g.W;
  const int y = static_cast<int>(pix % g.H);
  const int b = static_cast<int>(pix / g.H);
  out[static_cast<size_t>(geo_row(g, b, y, x)) * C + c] =
      from_f32<T>(in[((static_cast<size_t>(b) * C + c) * g.H + y) * g.W + x]);
}
template <typename T>
__global__ void pm_to_nchw_kernel(const T* __restrict__ in, Geo g, int C, int ld, int col0, float* __restrict__ out) {
  const long long total = static_cast<long long>(g.n) * C * g.H * g.W;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int c = static_cast<int>(i % C);
  long long pix = i / C;
  const int x = static_cast<int>(pix % g.W); pix /= g.W;
  const int y = static_cast<int>(pix % g.H);
  const int b = static_cast<int>(pix / g.H);
  out[((static_cast<size_t>(b) * C + c) * g.H + y) * g.W + x] =
      to_f32(in[static_cast<size_t>(geo_row(g, b, y, x)) * ld + col0 + c]);
}

// ---------------------------------------------------------------------------
// VAE encoder (models/vae.py:16-31, 51-62).
//
// GroupNorm(8)+GELU whose consumer is a 4x4 / stride-2 / pad-1 convolution: the result is written
// "space-to-depth with a one-pixel offset".  Position (py, px) of the half-resolution pixel
// matrix -- ALL (H/2+1) x (W/2+1) positions of a sample, the pad column and pad row included,
// PM row py * Wp + px of the sample -- holds the 2x2 block of source pixels (2py-1+sy, 2px-1+sx),
// sy, sx in {0,1}, as 4C channels ((sy*2+sx)*C + c), zeros outside the image.  Output pixel
// (oy, ox) of the strided conv then reads source rows 2oy-1 .. 2oy+2 = blocks oy and oy+1: a
// 2x2-tap, stride-1 convolution with K = 4C per tap.  The output pixel's own PM row is
// (oy+1) * Wp + ox (the pad row leads), so the tap shifts are {-Wp, -Wp+1, 0, 1} -- which is what
// the tcgen05 GEMM already runs: no strided gathers, no zero-padded K.
// ---------------------------------------------------------------------------
// enc.0: 3x3 / pad 1 convolution 3 -> 64 with bias, straight from the fp32 NCHW image into a PM
// activation.  K = 27 is far too short for a tcgen05 pipeline; on the CUDA cores the kernel is
// bound by FFMA issue and by the shared-memory reads of the weights, so one thread computes TWO
// horizontally adjacent pixels x 64 channels with packed FFMA2 (fma.rn.f32x2): every weight pair
// read from shared memory feeds two instructions, each of which does two multiply-adds.
template <typename T>
__global__ void __launch_bounds__(128)
enc_conv_in_kernel(const float* __restrict__ x, const float* __restrict__ w_oihw, const float* __restrict__ bias,
                   Geo g, T* __restrict__ out) {
  __shared__ __align__(16) float ws[27 * 64];
  for (int i = threadIdx.x; i < 27 * 64; i += blockDim.x) {
    const int k = i / 64, co = i % 64;  // k = ci*9 + ky*3 + kx
    ws[i] = w_oihw[co * 27 + k];
  }
  __syncthreads();
  const int be = blockIdx.y;
  const int half_w = g.W / 2;
  const int pp = blockIdx.x * blockDim.x + threadIdx.x;
  if (pp >= g.H * half_w) return;
  const int y = pp / half_w, x0 = 2 * (pp - y * half_w);
  const float* xs = x + static_cast<size_t>(be) * 3 * g.H * g.W;
  float in[3][3][4];   // channel, kernel row, column x0-1 .. x0+2
#pragma unroll
  for (int ci = 0; ci < 3; ++ci)
#pragma unroll
    for (int ky = 0; ky < 3; ++ky)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int iy = y + ky - 1, ix = x0 + j - 1;
        in[ci][ky][j] = (iy >= 0 && iy < g.H && ix >= 0 && ix < g.W) ? xs[(ci * g.H + iy) * g.W + ix] : 0.f;
      }
  float2 acc0[32], acc1[32];
#pragma unroll
  for (int c = 0; c < 32; ++c) acc0[c] = acc1[c] = make_float2(bias[2 * c], bias[2 * c + 1]);
  const float2* ws2 = reinterpret_cast<const float2*>(ws);
#pragma unroll
  for (int ci = 0; ci < 3; ++ci)
#pragma unroll
    for (int ky = 0; ky < 3; ++ky)
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const int k = ci * 9 + ky * 3 + kx;
        const float2 v0 = make_float2(in[ci][ky][kx], in[ci][ky][kx]);
        const float2 v1 = make_float2(in[ci][ky][kx + 1], in[ci][ky][kx + 1]);
#pragma unroll
        for (int c = 0; c < 32; ++c) {
          const float2 w2 = ws2[k * 32 + c];
          acc0[c] = __ffma2_rn(v0, w2, acc0[c]);
          acc1[c] = __ffma2_rn(v1, w2, acc1[c]);
        }
      }
  T* op = out + static_cast<size_t>(geo_row(g, be, y, x0)) * 64;   // the two pixels are adjacent PM rows
#pragma unroll
  for (int c = 0; c < 32; c += 4) {
    const float a[8] = {acc0[c].x, acc0[c].y, acc0[c + 1].x, acc0[c + 1].y,
                        acc0[c + 2].x, acc0[c + 2].y, acc0[c + 3].x, acc0[c + 3].y};
    const float b[8] = {acc1[c].x, acc1[c].y, acc1[c + 1].x, acc1[c + 1].y,
                        acc1[c + 2].x, acc1[c + 2].y, acc1[c + 3].x, acc1[c + 3].y};
    store8(op + 2 * c, a);
    store8(op + 64 + 2 * c, b);
  }
}

// enc.0 on the warp-level tensor-core path (mma.sync m16n8k8, TF32 operands, fp32 accumulate):
// K = 3 channels x 9 taps = 27, padded to 32, is far too short for a tcgen05 pipeline, and the
// CUDA-core version above is bound by FFMA issue.  One warp = tiles of 16 horizontally adjacent
// pixels x 64 output channels; the B fragments (the whole 32 x 64 weight matrix) sit in shared
// memory in fragment order, and the im2col A fragments are gathered straight
// from the fp32 image (L1-resident: every input value is read by nine taps).  TF32 keeps 10
// mantissa bits of the image and the weights: below the bf16 rounding of the output.
__device__ __forceinline__ uint32_t to_tf32(float v) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
  return r;
}
template <typename T>
__global__ void __launch_bounds__(256, 2)
enc_conv_in_mma_kernel(const float* __restrict__ x, const float* __restrict__ w_oihw, const float* __restrict__ bias,
                       Geo g, T* __restrict__ out) {
  const int be = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int gq = lane >> 2, t = lane & 3;   // fragment row group / thread in group
  // B fragments in shared memory, one 8-byte entry per (k step, n tile, lane), read back conflict-free:
  // b0 = W[k = 8 ks + t][n = 8 nt + gq], b1 = W[k + 4][n]; W[k][n] = w[n][k], zero for k >= 27.
  // (Held in registers they cost 64 per thread and leave one CTA per SM.)
  __shared__ uint2 bsm[4 * 8 * 32];
  for (int e = threadIdx.x; e < 4 * 8 * 32; e += 256) {
    const int el = e & 31, nt = (e >> 5) & 7, ks = e >> 8;
    const int k = ks * 8 + (el & 3), n = nt * 8 + (el >> 2);
    bsm[e] = make_uint2(to_tf32(k < 27 ? w_oihw[n * 27 + k] : 0.f), to_tf32(k + 4 < 27 ? w_oihw[n * 27 + k + 4] : 0.f));
  }
  __syncthreads();
  // this thread's eight K indices (j = 2 ks + h  ->  k = 8 ks + t + 4 h) as image offsets
  int koff[8], kdy[8], kdx[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int k = (j >> 1) * 8 + t + (j & 1) * 4;
    const int ci = k / 9, ky = (k - ci * 9) / 3, kx = k - ci * 9 - ky * 3;
    kdy[j] = k < 27 ? ky - 1 : 1 << 20;      // (out of range for every pixel -> contributes zero)
    kdx[j] = kx - 1;
    koff[j] = k < 27 ? ci * g.H * g.W + (ky - 1) * g.W + (kx - 1) : 0;
  }
  float2 bia[8];
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) bia[nt] = make_float2(bias[nt * 8 + 2 * t], bias[nt * 8 + 2 * t + 1]);
  const float* xs = x + static_cast<size_t>(be) * 3 * g.H * g.W;
  const int tiles_per_row = g.W / 16;
  const int n_tiles = g.H * tiles_per_row;
  auto gather = [&](int tile, float (&v)[16]) {
    const int y = tile / tiles_per_row, xa = (tile - y * tiles_per_row) * 16 + gq, xb = xa + 8;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int iy = y + kdy[j];
      const bool row_ok = iy >= 0 && iy < g.H;
      const int ixa = xa + kdx[j], ixb = xb + kdx[j];
      v[2 * j] = (row_ok && ixa >= 0 && ixa < g.W) ? __ldg(xs + koff[j] + y * g.W + xa) : 0.f;
      v[2 * j + 1] = (row_ok && ixb >= 0 && ixb < g.W) ? __ldg(xs + koff[j] + y * g.W + xb) : 0.f;
    }
  };
  for (int tile = blockIdx.x * 8 + warp; tile < n_tiles; tile += gridDim.x * 8) {
    const int y = tile / tiles_per_row, xa = (tile - y * tiles_per_row) * 16 + gq;
    float cur[16];
    gather(tile, cur);
    uint32_t af[4][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      af[j >> 1][(j & 1) * 2] = to_tf32(cur[2 * j]);           // a0 / a2: row gq
      af[j >> 1][(j & 1) * 2 + 1] = to_tf32(cur[2 * j + 1]);   // a1 / a3: row gq + 8
    }
    float acc[8][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) { acc[nt][0] = acc[nt][2] = bia[nt].x; acc[nt][1] = acc[nt][3] = bia[nt].y; }
#pragma unroll
    for (int ks = 0; ks < 4; ++ks)
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        const uint2 bb = bsm[(ks * 8 + nt) * 32 + lane];
        asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(acc[nt][0]), "+f"(acc[nt][1]), "+f"(acc[nt][2]), "+f"(acc[nt][3])
                     : "r"(af[ks][0]), "r"(af[ks][1]), "r"(af[ks][2]), "r"(af[ks][3]), "r"(bb.x), "r"(bb.y));
      }
    T* pa = out + static_cast<size_t>(geo_row(g, be, y, xa)) * 64 + 2 * t;
    T* pb = pa + 8 * 64;   // pixel xb = xa + 8: eight PM rows further
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      if constexpr (sizeof(T) == 2) {
        *reinterpret_cast<__nv_bfloat162*>(pa + nt * 8) = __floats2bfloat162_rn(acc[nt][0], acc[nt][1]);
        *reinterpret_cast<__nv_bfloat162*>(pb + nt * 8) = __floats2bfloat162_rn(acc[nt][2], acc[nt][3]);
      } else {
        *reinterpret_cast<float2*>(pa + nt * 8) = make_float2(acc[nt][0], acc[nt][1]);
        *reinterpret_cast<float2*>(pb + nt * 8) = make_float2(acc[nt][2], acc[nt][3]);
      }
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
gn8_apply_gelu_s2d_kernel(const T* __restrict__ raw, const double2* __restrict__ partial, int n_chunks_stats,
                          const float* __restrict__ gamma, const float* __restrict__ beta, Geo g, int C,
                          Geo g2, T* __restrict__ out) {
  __shared__ float s_mean[8], s_rstd[8];
  const int b = blockIdx.y, tid = threadIdx.x;
  gn8_finish_stats(partial, b, n_chunks_stats, static_cast<double>(g.H) * g.W * (C / 8), s_mean, s_rstd);
  // One thread = one output position x one 8-channel vector: the position is decoded once and
  // its four source pixels are loaded together (ncu on the first version, one source pixel per
  // thread iteration: 235 warp instructions per vector, 51 % issue utilisation).  256 is a
  // multiple of the vectors per pixel, so a thread's channel vector -- and with it its gamma /
  // beta / group statistics -- is the same in every iteration.
  const int vpr = C / 8, vpg = vpr / 8;
  const int vsh = 31 - __clz(vpr);   // vpr is a power of two (C = 64 .. 256)
  const int total = g2.S * vpr;
  const int per = (total + gridDim.x - 1) / gridDim.x;
  const int begin = blockIdx.x * per;
  const int end = min(total, begin + per);
  const int vc = tid & (vpr - 1), c0 = vc * 8;
  const float mean = s_mean[vc / vpg], rstd = s_rstd[vc / vpg];
  float ga[8], be[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { ga[i] = gamma[c0 + i] * rstd; be[i] = beta[c0 + i] - mean * ga[i]; }
  const T* src = raw + static_cast<size_t>(b) * g.S * C + c0;
  T* dst_b = out + static_cast<size_t>(b) * g2.S * (4 * C) + c0;
  for (int w = (begin & ~(vpr - 1)) + tid; w < end; w += 256) {
    if (w < begin) continue;
    const int pos = w >> vsh;
    const int py = fdiv(pos, g2.mWp, g2.Wp), px = pos - py * g2.Wp;
    const int y0 = 2 * py - 1, x0 = 2 * px - 1;
    const int r00 = (y0 + 1) * g.Wp + x0;      // PM row of source pixel (y0, x0) inside the sample
    float xv[4][8];
    bool inside[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int y = y0 + (q >> 1), x = x0 + (q & 1);
      inside[q] = y >= 0 && y < g.H && x >= 0 && x < g.W;
      if (inside[q]) load8(src + (r00 + (q >> 1) * g.Wp + (q & 1)) * C, xv[q]);
    }
    T* dp = dst_b + pos * (4 * C);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
#pragma unroll
      for (int i = 0; i < 8; ++i) xv[q][i] = inside[q] ? fmaf(xv[q][i], ga[i], be[i]) : 0.f;   // (GELU(0) = 0)
      gelu_inplace<T, 8>(xv[q]);
      store8(dp + q * C, xv[q]);
    }
  }
}

// to_mu / to_logvar (1x1 convs, 256 -> 4 each), clamp, reparameterisation with the caller's
// eps, latent scaling and the per-sample KL term (models/vae.py:53-62).  gridDim.y CTAs of 16
// warps per sample, one warp per pixel (lanes split the channels); the KL sum is accumulated in
// float64 in a fixed order: per lane, per warp, per CTA (kl_part), then vae_enc_kl_finish_kernel:
// kl_sample[b] = 0.5 * sum(exp(logvar) + mu^2 - 1 - logvar) / kl_div.
template <typename T>
__global__ void __launch_bounds__(512)
vae_enc_head_kernel(const T* __restrict__ h, Geo g, int C, const float* __restrict__ w_mu,
                    const float* __restrict__ b_mu, const float* __restrict__ w_lv,
                    const float* __restrict__ b_lv, const float* __restrict__ eps, float scale,
                    float* __restrict__ z, float* __restrict__ mu_out, float* __restrict__ logvar_out,
                    double* __restrict__ kl_part) {
  __shared__ double s_kl[16];
  const int b = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int HW = g.H * g.W;
  const int c0 = lane * 8;   // C == 256: eight channels per lane
  float wm[4][8], wl[4][8];
#pragma unroll
  for (int o = 0; o < 4; ++o)
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      wm[o][i] = w_mu[o * C + c0 + i];
      wl[o][i] = w_lv[o * C + c0 + i];
    }
  double kl = 0.0;
  for (int pix = blockIdx.y * 16 + warp; pix < HW; pix += 16 * gridDim.y) {
    float xv[8];
    load8(h + static_cast<size_t>(geo_row_of_pixel(g, b, pix)) * C + c0, xv);
    float acc[8];
#pragma unroll
    for (int o = 0; o < 4; ++o) {
      float a = 0.f, l = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) { a = fmaf(xv[i], wm[o][i], a); l = fmaf(xv[i], wl[o][i], l); }
      acc[o] = a; acc[4 + o] = l;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j)
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], d);
    if (lane < 4) {
      const int o = lane;
      float m = 0.f, lv = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) if (j == o) { m = acc[j]; lv = acc[4 + j]; }
      m += b_mu[o];
      lv = fminf(fmaxf(lv + b_lv[o], -30.0f), 20.0f);
      const size_t idx = (static_cast<size_t>(b) * 4 + o) * HW + pix;
      const float sd = expf(0.5f * lv);
      z[idx] = (m + eps[idx] * sd) * scale;
      if (mu_out) mu_out[idx] = m;
      if (logvar_out) logvar_out[idx] = lv;
      kl += static_cast<double>(expf(lv)) + static_cast<double>(m) * m - 1.0 - lv;
    }
  }
  // lanes 0..3 of each warp hold partial sums: fold them, then the sixteen warps in order
  kl += __shfl_down_sync(0xffffffffu, kl, 2);
  kl += __shfl_down_sync(0xffffffffu, kl, 1);
  if (lane == 0) s_kl[warp] = kl;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < 16; ++w) t += s_kl[w];
    kl_part[b * gridDim.y + blockIdx.y] = t;
  }
}
// the CTAs' partial sums of a sample, added in a fixed order
static __global__ void vae_enc_kl_finish_kernel(const double* __restrict__ kl_part, int parts, int n, float kl_div,
                                                float* __restrict__ kl_sample) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n) return;
  double t = 0.0;
  for (int p = 0; p < parts; ++p) t += kl_part[b * parts + p];
  kl_sample[b] = static_cast<float>(0.5 * t / kl_div);
}

// Diffuser.add_noise (diff.py:18-30): x_t = sqrt(ab) * x_0 + sqrt(1 - ab) * noise with
// ab = alpha_bars[t - 1] per sample.  Every operation is rounded separately, like the three
// torch kernels of the reference, so the result is bit-identical for the same noise.
static __global__ void __launch_bounds__(256)
add_noise_kernel(const float* __restrict__ x0, const long long* __restrict__ t, const float* __restrict__ alpha_bars,
                 int T, const float* __restrict__ noise, long long per_sample, float* __restrict__ x_t) {
  const int b = blockIdx.y;
  long long ti = t[b] - 1;
  ti = ti < 0 ? 0 : (ti >= T ? T - 1 : ti);
  const float ab = alpha_bars[ti];
  const float sa = sqrtf(ab), sb = sqrtf(__fsub_rn(1.0f, ab));
  const long long base = static_cast<long long>(b) * per_sample;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < per_sample;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    x_t[base + i] = __fadd_rn(__fmul_rn(sa, x0[base + i]), __fmul_rn(sb, noise[base + i]));
}

}  // namespace dmb
