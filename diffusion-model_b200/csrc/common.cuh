// Shared device/host helpers for the sm_100a kernels: error plumbing, the padded
// pixel-matrix geometry, bf16/fp32 vector access, and the PTX wrappers for mbarrier,
// TMA (cp.async.bulk.tensor) and tcgen05 (alloc / mma / commit / ld).
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

namespace dmb {

// ---------------------------------------------------------------------------
// Errors.  No exception crosses the C ABI: every entry point catches Error and
// turns it into a status code + dmb_last_error_string().
// ---------------------------------------------------------------------------
struct Error {
  int code;
  std::string msg;
};

[[noreturn]] void fail(int code, const std::string& msg);
void cuda_check(cudaError_t e, const char* what, const char* file, int line);

#define DMB_CUDA(x) ::dmb::cuda_check((x), #x, __FILE__, __LINE__)
#define DMB_CHECK_LAUNCH() ::dmb::cuda_check(cudaPeekAtLastError(), "kernel launch", __FILE__, __LINE__)
#define DMB_REQUIRE(cond, msg)                                        \
  do {                                                                \
    if (!(cond)) ::dmb::fail(2, std::string("invalid argument: ") + (msg)); \
  } while (0)

// ---------------------------------------------------------------------------
// Geometry of a feature map stored as a padded pixel matrix ("PM"):
//   rows  = pixels of all samples, channels contiguous per row (NHWC),
//   every image row carries ONE trailing zero pixel and every image ONE leading
//   zero row, so that row r + dy*Wp + dx is the (dy,dx) neighbour of pixel r and
//   lands on a zero pixel whenever the neighbour is outside the image.
//   One extra zero row closes the last image.  Kernels write real pixels only;
//   pad pixels keep the zeros of the initial memset.
// ---------------------------------------------------------------------------
struct Geo {
  int n;   // samples
  int H, W;
  int Wp;  // W + 1
  int S;   // (H + 1) * Wp rows per sample
  int rows_valid;  // n * S
  int rows_alloc;  // rows_valid + Wp, rounded up to 128
  uint32_t mS, mWp, mW;  // multiply-high magics for division by S, Wp, W (see fdiv)
};

// Exact n / d for 0 <= n < 2^31 with magic = ceil(2^32 / d): the multiply-high estimate is
// q or q + 1 (error n * (magic * d - 2^32) / (d * 2^32) < 1), fixed by one compare.  A few
// integer ops instead of the ~20-instruction emulated division by a run-time divisor.
inline uint32_t fdiv_magic(uint32_t d) { return d <= 1 ? 0u : static_cast<uint32_t>((0x100000000ull + d - 1) / d); }
__device__ __forceinline__ int fdiv(int n, uint32_t magic, int d) {
  if (!magic) return n;
  const int q = static_cast<int>(__umulhi(static_cast<uint32_t>(n), magic));
  return (q * d > n) ? q - 1 : q;
}

inline Geo make_geo(int n, int H, int W) {
  Geo g;
  g.n = n; g.H = H; g.W = W; g.Wp = W + 1; g.S = (H + 1) * (W + 1);
  g.rows_valid = n * g.S;
  g.rows_alloc = ((g.rows_valid + g.Wp + 127) / 128) * 128;
  g.mS = fdiv_magic(g.S); g.mWp = fdiv_magic(g.Wp); g.mW = fdiv_magic(g.W);
  return g;
}
// pixel index (row-major inside the image) -> PM row
__device__ __forceinline__ int geo_row_of_pixel(const Geo& g, int b, int pix) {
  const int y = fdiv(pix, g.mW, g.W);
  return b * g.S + (y + 1) * g.Wp + (pix - y * g.W);
}
// is PM row rr (0 <= rr < S, relative to its sample) a real pixel?
__device__ __forceinline__ bool geo_is_pixel(const Geo& g, int rr) {
  const int yy = fdiv(rr, g.mWp, g.Wp);
  return yy >= 1 && (rr - yy * g.Wp) != g.W;
}

__host__ __device__ inline int geo_row(const Geo& g, int b, int y, int x) {
  return b * g.S + (y + 1) * g.Wp + x;
}

// Decompose a PM row; returns false on pad pixels.
__device__ __forceinline__ bool geo_decode(const Geo& g, int r, int& b, int& y, int& x) {
  if (r < 0 || r >= g.rows_valid) return false;
  b = fdiv(r, g.mS, g.S);
  int rr = r - b * g.S;
  int yy = fdiv(rr, g.mWp, g.Wp);
  x = rr - yy * g.Wp;
  y = yy - 1;
  return yy >= 1 && x < g.W;
}

// ---------------------------------------------------------------------------
// Activation element types.  bf16 mode stores activations as bf16 and feeds
// tcgen05 kind::f16; fp32 mode stores fp32 and feeds kind::tf32.
// ---------------------------------------------------------------------------
template <typename T> struct ElemTraits;
template <> struct ElemTraits<__nv_bfloat16> {
  static constexpr int kChunkK = 64;   // elements per 128-byte K chunk
  static constexpr int kMmaK = 16;     // elements per 32-byte tcgen05 K step
  static constexpr uint32_t kFmt = 1;  // UMMA F16F32Format::BF16
  static constexpr bool kTf32 = false;
};
template <> struct ElemTraits<float> {
  static constexpr int kChunkK = 32;
  static constexpr int kMmaK = 8;
  static constexpr uint32_t kFmt = 2;  // UMMA F16F32Format::TF32
  static constexpr bool kTf32 = true;
};

__device__ __forceinline__ float to_f32(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ float to_f32(float v) { return v; }
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }

// 8 consecutive elements <-> float[8] (16 B for bf16, 32 B for fp32); p must be aligned.
__device__ __forceinline__ void load8(const __nv_bfloat16* p, float* v) {
  uint4 u = *reinterpret_cast<const uint4*>(p);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 f = __bfloat1622float2(h[i]);
    v[2 * i] = f.x; v[2 * i + 1] = f.y;
  }
}
__device__ __forceinline__ void load8(const float* p, float* v) {
  float4 a = *reinterpret_cast<const float4*>(p);
  float4 b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float* v) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
  *reinterpret_cast<uint4*>(p) = u;
}
__device__ __forceinline__ void store8(float* p, const float* v) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}

// 8 consecutive elements kept PACKED in registers (4 regs for bf16, 8 for fp32): streaming
// kernels issue many of these loads back to back to keep >= 100 KB per SM in flight, which is
// what HBM3e needs (latency x bandwidth), and unpack one vector at a time afterwards.
template <typename T> struct Raw8;
template <> struct Raw8<__nv_bfloat16> {
  uint4 u;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { u = *reinterpret_cast<const uint4*>(p); }
  __device__ __forceinline__ void store(__nv_bfloat16* p) const { *reinterpret_cast<uint4*>(p) = u; }
  __device__ __forceinline__ void unpack(float* v) const {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __bfloat1622float2(h[i]);
      v[2 * i] = f.x; v[2 * i + 1] = f.y;
    }
  }
};
template <> struct Raw8<float> {
  float4 a, b;
  __device__ __forceinline__ void load(const float* p) {
    a = *reinterpret_cast<const float4*>(p);
    b = *reinterpret_cast<const float4*>(p + 4);
  }
  __device__ __forceinline__ void store(float* p) const {
    *reinterpret_cast<float4*>(p) = a;
    *reinterpret_cast<float4*>(p + 4) = b;
  }
  __device__ __forceinline__ void unpack(float* v) const {
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
};

// Exact (erf) GELU, as nn.GELU() / F.gelu default.
// ---------------------------------------------------------------------------
// GroupNorm statistics are summed in ONE association order whatever N tile a launch uses (the
// tile is narrowed at small batch), so a sample's result does not depend on the batch it is in
// or on how the batch is sharded: the unit is an aligned group of 16 columns (even and odd
// columns as two sequential chains), units combine as an aligned binary tree (a binary counter:
// level l holds a finished subtree of 2^l units), and left-over subtrees combine from the
// largest to the smallest.
// ---------------------------------------------------------------------------
__device__ __forceinline__ float2 stat_unit16(const float* v) {
  float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll
  for (int j = 0; j < 16; j += 2) {
    s0 += v[j]; q0 = fmaf(v[j], v[j], q0);
    s1 += v[j + 1]; q1 = fmaf(v[j + 1], v[j + 1], q1);
  }
  return make_float2(s0 + s1, q0 + q1);
}
// ---------------------------------------------------------------------------
// Per-sample GroupNorm(1, C) statistics as 64-bit FIXED-POINT sums.  The producing conv adds the
// (sum, sum of squares) of every 16-column unit of every output row into two integers per sample
// (atomics); integer addition is associative, so the totals do not depend on tile shapes, on the
// order in which CTAs retire or on how the batch is sharded -- two runs give the same bits -- and
// the consumer (gn1_apply) reads two numbers instead of re-reducing thousands of row partials in
// every CTA (that prologue was 6 % of a batch-256 step and 10 % of a batch-1 step).
// Scales: 2^20 for sums (|unit sum| < 2^23 is exact to 1e-6), 2^16 for sums of squares; a unit of
// sixteen values of magnitude 1e4 (the first conv of a diverged random-weight trajectory) is
// 1.6e9 * 2^16 = 1e14, and 3136 units per sample stay far below 2^63.
// ---------------------------------------------------------------------------
constexpr float kStatScale1 = 1048576.0f, kStatScale2 = 65536.0f;
constexpr double kStatInv1 = 1.0 / 1048576.0, kStatInv2 = 1.0 / 65536.0;
struct StatAcc {
  long long s1 = 0, s2 = 0;
  __device__ __forceinline__ void add(float2 u) {
    s1 += __float2ll_rn(u.x * kStatScale1);
    s2 += __float2ll_rn(u.y * kStatScale2);
  }
};
// Adds the lanes' accumulators into slots[2 * sample .. +1].  Called by all 32 lanes; `valid`
// lanes carry data for `sample`.  When every valid lane belongs to one sample (the usual case:
// 128 consecutive pixel-matrix rows) the warp reduces first and issues one pair of atomics.
__device__ __forceinline__ void stat_commit(long long* slots, int sample, bool valid, long long s1, long long s2) {
  const unsigned act = __ballot_sync(0xffffffffu, valid);
  if (act == 0u) return;
  const int lead = __ffs(act) - 1;
  const int s0 = __shfl_sync(0xffffffffu, sample, lead);
  if (__all_sync(0xffffffffu, !valid || sample == s0)) {
    if (!valid) { s1 = 0; s2 = 0; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      s1 += __shfl_xor_sync(0xffffffffu, s1, o);
      s2 += __shfl_xor_sync(0xffffffffu, s2, o);
    }
    if ((threadIdx.x & 31) == lead) {
      atomicAdd(reinterpret_cast<unsigned long long*>(slots) + 2 * s0, static_cast<unsigned long long>(s1));
      atomicAdd(reinterpret_cast<unsigned long long*>(slots) + 2 * s0 + 1, static_cast<unsigned long long>(s2));
    }
  } else if (valid) {
    atomicAdd(reinterpret_cast<unsigned long long*>(slots) + 2 * sample, static_cast<unsigned long long>(s1));
    atomicAdd(reinterpret_cast<unsigned long long*>(slots) + 2 * sample + 1, static_cast<unsigned long long>(s2));
  }
}

struct StatTree {
  float2 l0, l1, l2, l3, l4, l5;
  uint32_t n = 0;
  __device__ __forceinline__ static float2 join(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
  __device__ __forceinline__ void add(float2 x) {
    const uint32_t c = n++;
    if (!(c & 1)) { l0 = x; return; }
    x = join(l0, x);
    if (!(c & 2)) { l1 = x; return; }
    x = join(l1, x);
    if (!(c & 4)) { l2 = x; return; }
    x = join(l2, x);
    if (!(c & 8)) { l3 = x; return; }
    x = join(l3, x);
    if (!(c & 16)) { l4 = x; return; }
    l5 = join(l4, x);
  }
  __device__ __forceinline__ float2 total() const {   // n <= 63
    float2 r = make_float2(0.f, 0.f);
    bool has = false;
    if (n & 32) { r = l5; has = true; }
    if (n & 16) { r = has ? join(r, l4) : l4; has = true; }
    if (n & 8) { r = has ? join(r, l3) : l3; has = true; }
    if (n & 4) { r = has ? join(r, l2) : l2; has = true; }
    if (n & 2) { r = has ? join(r, l1) : l1; has = true; }
    if (n & 1) { r = has ? join(r, l0) : l0; }
    return r;
  }
};

__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f)); }
// GELU for bf16 storage: x * sigmoid(u(x)) with u = logit(Phi(x)) fitted by an odd polynomial
// (x * P4(x^2), coefficients pre-multiplied by -log2 e) -- 6 FMA-pipe ops and 2 MUFU ops
// instead of erff's ~35.  Max |error| vs the exact erf form is 4.4e-6 over all x (checked
// against scipy ndtr in fp32), three orders of magnitude below bf16 rounding; the tf32 / fp32
// mode keeps erff.
__device__ __forceinline__ float gelu_fast(float x) {
  const float x2 = x * x;
  float g = -3.4587828849907964e-06f;
  g = fmaf(g, x2, 9.363872959511355e-05f);
  g = fmaf(g, x2, 0.0003210399590898305f);
  g = fmaf(g, x2, -0.10513131320476532f);
  g = fmaf(g, x2, -2.3021068572998047f);
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(g * x));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
  return x * r;
}
// Two elements at once on the packed fp32x2 instructions of sm_100 (FMUL2 / FFMA2 / FADD2): the
// same operations in the same order as gelu_fast (bit-identical results), 8 issue slots + 4 MUFU per
// pair instead of 18 + 4.  The streaming GroupNorm kernels are bound by instruction issue, not by
// HBM (ncu: 55-65 % issue-slot utilisation at 30 % occupancy, profiles/r2_v12_ncu_full_gn1_apply.md).
__device__ __forceinline__ float2 gelu_fast2(float2 x) {
  const float2 x2 = __fmul2_rn(x, x);
  float2 g = make_float2(-3.4587828849907964e-06f, -3.4587828849907964e-06f);
  g = __ffma2_rn(g, x2, make_float2(9.363872959511355e-05f, 9.363872959511355e-05f));
  g = __ffma2_rn(g, x2, make_float2(0.0003210399590898305f, 0.0003210399590898305f));
  g = __ffma2_rn(g, x2, make_float2(-0.10513131320476532f, -0.10513131320476532f));
  g = __ffma2_rn(g, x2, make_float2(-2.3021068572998047f, -2.3021068572998047f));
  const float2 a = __fmul2_rn(g, x);
  float e0, e1, r0, r1;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e0) : "f"(a.x));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e1) : "f"(a.y));
  const float2 d = __fadd2_rn(make_float2(e0, e1), make_float2(1.0f, 1.0f));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r0) : "f"(d.x));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r1) : "f"(d.y));
  return __fmul2_rn(x, make_float2(r0, r1));
}
// Four elements at once with ONE reciprocal: the streaming GroupNorm kernels sit on the MUFU pipe
// (ncu: XU 50-60 % at 30 % occupancy, the highest pipe; ex2 + rcp = 2 MUFU per element at 16 per
// clock per SM).  With d_i = 1 + 2^(t_i):  R = 1 / (d0 d1 d2 d3),  1/d0 = R d1 (d2 d3), ... --
// 5 MUFU per four elements instead of 8, the extra multiplies on the packed fp32x2 pipe.  t is
// clamped at 30 (d <= 2^30 + 1, the product stays below 2^121; beyond it x / d < |x| 1e-9, which
// every caller rounds to zero or adds to a residual of order 1).  The quotients differ from
// gelu_fast's by fp32 rounding only (three roundings instead of one): far below the bf16 rounding
// of the stored result -- but not bit-identical with it, so a kernel uses one form throughout.
#ifndef DMB_GELU4
#define DMB_GELU4 1   // 0: every GELU of the bf16 mode takes its own reciprocal (gelu_fast / gelu_fast2)
#endif
constexpr bool kGelu4 = DMB_GELU4 != 0;
__device__ __forceinline__ void gelu_fast4(float2& xa, float2& xb) {
  const float2 k0 = make_float2(-3.4587828849907964e-06f, -3.4587828849907964e-06f);
  const float2 k1 = make_float2(9.363872959511355e-05f, 9.363872959511355e-05f);
  const float2 k2 = make_float2(0.0003210399590898305f, 0.0003210399590898305f);
  const float2 k3 = make_float2(-0.10513131320476532f, -0.10513131320476532f);
  const float2 k4 = make_float2(-2.3021068572998047f, -2.3021068572998047f);
  const float2 xa2 = __fmul2_rn(xa, xa), xb2 = __fmul2_rn(xb, xb);
  float2 ga = __ffma2_rn(k0, xa2, k1), gb = __ffma2_rn(k0, xb2, k1);
  ga = __ffma2_rn(ga, xa2, k2); gb = __ffma2_rn(gb, xb2, k2);
  ga = __ffma2_rn(ga, xa2, k3); gb = __ffma2_rn(gb, xb2, k3);
  ga = __ffma2_rn(ga, xa2, k4); gb = __ffma2_rn(gb, xb2, k4);
  const float2 ta = __fmul2_rn(ga, xa), tb = __fmul2_rn(gb, xb);
  float e0, e1, e2, e3, R;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e0) : "f"(fminf(ta.x, 30.f)));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e1) : "f"(fminf(ta.y, 30.f)));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e2) : "f"(fminf(tb.x, 30.f)));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e3) : "f"(fminf(tb.y, 30.f)));
  const float2 one = make_float2(1.0f, 1.0f);
  const float2 da = __fadd2_rn(make_float2(e0, e1), one), db = __fadd2_rn(make_float2(e2, e3), one);
  const float2 m = __fmul2_rn(da, db);                       // (d0 d2, d1 d3)
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(R) : "f"(m.x * m.y));
  const float2 q = __fmul2_rn(make_float2(R, R), make_float2(m.y, m.x));   // (1 / (d0 d2), 1 / (d1 d3))
  xa = __fmul2_rn(xa, __fmul2_rn(q, db));                    // 1/d0 = q.x d2, 1/d1 = q.y d3
  xb = __fmul2_rn(xb, __fmul2_rn(q, da));                    // 1/d2 = q.x d0, 1/d3 = q.y d1
}
template <typename T> __device__ __forceinline__ float gelu_for(float x) {
  if constexpr (sizeof(T) == 2) return gelu_fast(x);
  else return gelu_erf(x);
}
// GELU of N (a multiple of 4) values in place: bf16 storage mode four at a time, fp32 mode the erf form
template <typename T, int N> __device__ __forceinline__ void gelu_inplace(float* v) {
  if constexpr (sizeof(T) == 2) {
#pragma unroll
    for (int j = 0; j < N; j += 4) {
      float2 a = make_float2(v[j], v[j + 1]), b = make_float2(v[j + 2], v[j + 3]);
      if constexpr (kGelu4) { gelu_fast4(a, b); } else { a = gelu_fast2(a); b = gelu_fast2(b); }
      v[j] = a.x; v[j + 1] = a.y; v[j + 2] = b.x; v[j + 3] = b.y;
    }
  } else {
#pragma unroll
    for (int j = 0; j < N; ++j) v[j] = gelu_erf(v[j]);
  }
}
__device__ __forceinline__ float silu(float x) { return x / (1.0f + expf(-x)); }

// ---------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------
// Programmatic dependent launch.  Every kernel of the step loop starts with pdl_trigger() (the
// next kernel of the stream may be scheduled as soon as all CTAs of this one are resident or
// done) and calls pdl_wait() before its first global-memory access (returns once the previous
// kernel has completed and its writes are visible).  Without the launch attribute both are no-ops.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "elect.sync _|p, 0xffffffff;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(pred));
  return pred != 0;
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_async_proxy() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a pipeline bug must end in a trap (a CUDA error the host reports), never
// in a hung GPU.  try_wait sleeps in hardware, so the bound is ~seconds, far above any
// legitimate wait.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
#pragma unroll 1
  for (uint32_t spin = 0; spin < (1u << 24); ++spin) {
    if (mbar_try_wait(bar, parity)) return;
  }
  __trap();
}

// 2D tiled TMA load global -> shared, completing on an mbarrier.  Coordinates are signed;
// out-of-range rows arrive as zeros (that is what implements the conv halo).
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
// 3D variants: (channel, PM row inside the sample, sample) -- a box of {64 ch, 1 row, 128 samples}
// is the A tile "pixel q of 128 consecutive samples" of the pixel-major conv mode (gemm_tc.cuh).
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* map, const void* smem_src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}
// 2D tiled TMA store shared -> global (bulk async-group completion).  The smem tile must be
// in the swizzle layout the tensor map names; rows / columns outside the tensor are clipped.
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, const void* smem_src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// wait until at most N committed store groups still READ their shared-memory source
template <int N>
__device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void tma_store_wait_all() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
// named barrier among `count` threads (ids 1..15; 0 is __syncthreads)
__device__ __forceinline__ void named_bar_sync(int id, int count) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}

// tcgen05 / TMEM
// Whole-warp (.sync.aligned) TMEM column allocation; ncols is a power of two >= 32.
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_out, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_out)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem]^T, both operands K-major, issued by ONE thread.
template <bool kTf32>
__device__ __forceinline__ void umma_ss(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  if constexpr (kTf32) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}
// All previously issued MMAs of this thread arrive on the mbarrier when they retire
// (implies tcgen05.fence::before_thread_sync).
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}


// ---------------------------------------------------------------------------
// CTA-pair (cta_group::2) variants: two CTAs of a cluster on the two SMs of one TPC drive
// one 256-row MMA; each CTA holds its 128 rows of A, half of the B tile and its 128 lanes of
// the accumulator.  Only the leader (cluster rank 0) issues MMAs and commits.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of the same shared-memory variable in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_shared(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load issued by either CTA of the pair; the bytes are accounted on the mbarrier at
// `bar_cluster_addr` (the leader's barrier, a shared::cluster address).
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* map, uint32_t bar_cluster_addr,
                                                 int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d_pair(void* smem_dst, const CUtensorMap* map, uint32_t bar_cluster_addr,
                                                 int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar_cluster_addr), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* smem_out, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_out)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
template <bool kTf32>
__device__ __forceinline__ void umma_ss_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  if constexpr (kTf32) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}
// arrives on the barrier at this shared-memory offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      ::"r"(smem_u32(bar)), "h"(static_cast<uint16_t>(3))
      : "memory");
}

// 32 lanes x 32 fp32 columns of the accumulator -> 32 registers per thread (lane = row).
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// 16 registers per thread -> 32 lanes x 16 columns of TMEM (lane = row)
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const float* v) {
  const uint32_t* r = reinterpret_cast<const uint32_t*>(v);
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
        "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Shared-memory matrix descriptor for a K-major tile whose rows are 128 bytes and whose
// 8-row groups are 1024 bytes apart, 128-byte swizzle (what a {128 B x rows} TMA box with
// CU_TENSOR_MAP_SWIZZLE_128B writes).  Fields per the sm_100 UMMA descriptor: start>>4
// [0,14), LBO>>4 [16,30), SBO>>4 [32,46), version=1 [46,48), layout=SWIZZLE_128B(2) [61,64).
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr >> 4) & 0x3FFF);
  d |= static_cast<uint64_t>(1) << 16;
  d |= static_cast<uint64_t>(1024 >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}
// Instruction descriptor: fp32 accumulate, A/B format fmt (1 = bf16, 2 = tf32), both K-major.
__host__ __device__ inline uint32_t umma_idesc(uint32_t fmt, int M, int N) {
  return (1u << 4) | (fmt << 7) | (fmt << 10) | (static_cast<uint32_t>(N >> 3) << 17) |
         (static_cast<uint32_t>(M >> 4) << 24);
}

}  // namespace dmb
