// Engine implementation (see engine.cuh).  One Engine<T> per precision mode:
//   T = __nv_bfloat16 : bf16 activations / weights, tcgen05 kind::f16, fp32 accumulate
//   T = float         : fp32 activations / weights, tcgen05 kind::tf32, fp32 accumulate
#include "engine.cuh"
#include "attention_tc.cuh"
#include "attention_mma.cuh"
#include "attn_tail.cuh"

#include <type_traits>

#include <algorithm>
#include <cmath>
#include <cstring>

namespace dmb {

[[noreturn]] void fail(int code, const std::string& msg) { throw Error{code, msg}; }

void cuda_check(cudaError_t e, const char* what, const char* file, int line) {
  if (e != cudaSuccess) {
    fail(3, std::string("CUDA error: ") + cudaGetErrorString(e) + " in " + what + " at " + file + ":" +
                std::to_string(line));
  }
}

namespace {

// ---------------------------------------------------------------------------
// TMA descriptor encoding through the driver entry point (no link-time libcuda).
// ---------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    DMB_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres));
    if (!p || qres != cudaDriverEntryPointSuccess) fail(3, "cuTensorMapEncodeTiled entry point not available");
    fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// 2D row-major [rows, inner] tensor, box = {128 bytes of inner, box_rows}, 128B swizzle.
template <typename T>
CUtensorMap make_tmap(const void* base, uint64_t inner, uint64_t rows, uint32_t box_rows) {
  CUtensorMap m;
  memset(&m, 0, sizeof(m));
  const cuuint64_t gdim[2] = {inner, rows};
  const cuuint64_t gstride[1] = {inner * sizeof(T)};
  const cuuint32_t box[2] = {static_cast<cuuint32_t>(ElemTraits<T>::kChunkK), box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUtensorMapDataType dt = sizeof(T) == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  CUresult r = encode_tiled_fn()(&m, dt, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) fail(3, "cuTensorMapEncodeTiled failed with code " + std::to_string(static_cast<int>(r)));
  return m;
}

// 3D view of a PM activation: (channel, PM row inside the sample, sample), box = {128 bytes of
// channels, 1 row, 128 samples}: the A / output tile of the pixel-major conv mode (gemm_tc.cuh).
template <typename T>
CUtensorMap make_tmap_samples(const void* base, uint64_t C, uint64_t S, uint64_t n) {
  CUtensorMap m;
  memset(&m, 0, sizeof(m));
  const cuuint64_t gdim[3] = {C, S, n};
  const cuuint64_t gstride[2] = {C * sizeof(T), S * C * sizeof(T)};
  const cuuint32_t box[3] = {static_cast<cuuint32_t>(ElemTraits<T>::kChunkK), 1, static_cast<cuuint32_t>(kTileM)};
  const cuuint32_t estr[3] = {1, 1, 1};
  const CUtensorMapDataType dt = sizeof(T) == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  CUresult r = encode_tiled_fn()(&m, dt, 3, const_cast<void*>(base), gdim, gstride, box, estr,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) fail(3, "cuTensorMapEncodeTiled (3D) failed with code " + std::to_string(static_cast<int>(r)));
  return m;
}

// ---------------------------------------------------------------------------
// Device memory owned by the engine.
// ---------------------------------------------------------------------------
struct Pool {
  std::vector<void*> ptrs;
  size_t bytes = 0;
  void* alloc(size_t n, bool zero = true) {
    void* p = nullptr;
    n = (n + 255) / 256 * 256;
    DMB_CUDA(cudaMalloc(&p, n));
    if (zero) DMB_CUDA(cudaMemset(p, 0, n));
    ptrs.push_back(p);
    bytes += n;
    return p;
  }
  template <typename U> U* alloc_n(size_t n, bool zero = true) { return static_cast<U*>(alloc(n * sizeof(U), zero)); }
  void release() {
    for (void* p : ptrs) cudaFree(p);
    ptrs.clear();
    bytes = 0;
  }
  ~Pool() { release(); }
};

bool g_allow_bn256 = true;   // 128 x 256 tiles where C_out allows (DMB200_BN256=0 turns it off)
int pick_block_n(int N) {
  if (g_allow_bn256 && N % 256 == 0) return 256;
  if (N % 128 == 0) return 128;
  if (N % 64 == 0) return 64;
  if (N % 32 == 0) return 32;
  return 16;
}

uint32_t tmem_cols_for(int block_n) {
  uint32_t need = 2u * block_n, c = 32;
  while (c < need) c <<= 1;
  return c;
}

// Small repack kernels (run once at weight load).
template <typename T>
__global__ void repack_conv3x3_kernel(const float* __restrict__ w /*[Co][Ci][3][3]*/, int Co, int Ci, int rows_per_tap,
                                      T* __restrict__ out /*[9][rows_per_tap][Ci]*/) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long total = 9LL * Co * Ci;
  if (i >= total) return;
  const int ci = static_cast<int>(i % Ci);
  const int co = static_cast<int>((i / Ci) % Co);
  const int tap = static_cast<int>(i / (static_cast<long long>(Ci) * Co));
  out[(static_cast<size_t>(tap) * rows_per_tap + co) * Ci + ci] = from_f32<T>(w[(static_cast<size_t>(co) * Ci + ci) * 9 + tap]);
}
// ConvTranspose2d(4, stride 2, pad 1) weight [Ci][Co][4][4] -> [phase(a,b)][tap(u,v)][Co][Ci].
// Output pixel (2y+a, 2x+b) = sum over input pixels (y+dy, x+dx): a=0: (ky=1,dy=0),(ky=3,dy=-1);
// a=1: (ky=0,dy=+1),(ky=2,dy=0); same for b / kx / dx.  Tap index u selects the first / second pair.
template <typename T>
__global__ void repack_convT_kernel(const float* __restrict__ w, int Ci, int Co, T* __restrict__ out) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long total = 16LL * Co * Ci;
  if (i >= total) return;
  const int ci = static_cast<int>(i % Ci);
  const int co = static_cast<int>((i / Ci) % Co);
  const int tv = static_cast<int>((i / (static_cast<long long>(Ci) * Co)) % 2);
  const int tu = static_cast<int>((i / (2LL * Ci * Co)) % 2);
  const int pb = static_cast<int>((i / (4LL * Ci * Co)) % 2);
  const int pa = static_cast<int>(i / (8LL * Ci * Co));
  const int ky = (pa == 0) ? (tu == 0 ? 1 : 3) : (tu == 0 ? 0 : 2);
  const int kx = (pb == 0) ? (tv == 0 ? 1 : 3) : (tv == 0 ? 0 : 2);
  out[i] = from_f32<T>(w[((static_cast<size_t>(ci) * Co + co) * 4 + ky) * 4 + kx]);
}
// Conv2d(4, stride 2, pad 1) weight [Co][Ci][4][4] -> [tap(u,v)][Co][(sy*2+sx)*Ci + ci] with
// ky = 2u + sy, kx = 2v + sx: the 2x2-tap form over the offset space-to-depth input
// (kernels.cuh, gn8_apply_gelu_s2d_kernel).
template <typename T>
__global__ void repack_conv4s2_kernel(const float* __restrict__ w, int Co, int Ci, T* __restrict__ out) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long total = 16LL * Co * Ci;
  if (i >= total) return;
  const int ci = static_cast<int>(i % Ci);
  const int q = static_cast<int>((i / Ci) % 4);
  const int co = static_cast<int>((i / (4LL * Ci)) % Co);
  const int tap = static_cast<int>(i / (4LL * Ci * Co));
  const int ky = 2 * (tap >> 1) + (q >> 1), kx = 2 * (tap & 1) + (q & 1);
  out[i] = from_f32<T>(w[((static_cast<size_t>(co) * Ci + ci) * 4 + ky) * 4 + kx]);
}
template <typename T>
__global__ void cast_kernel(const float* __restrict__ in, long long n, T* __restrict__ out) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) out[i] = from_f32<T>(in[i]);
}
__global__ void transpose_kernel(const float* __restrict__ in, int rows, int cols, float* __restrict__ out,
                                 int out_ld, int out_col0) {  // out[c][out_col0 + r] = in[r][c]
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * cols) return;
  const int r = i / cols, c = i % cols;
  out[static_cast<size_t>(c) * out_ld + out_col0 + r] = in[i];
}

inline int cdiv(long long a, long long b) { return static_cast<int>((a + b - 1) / b); }

}  // namespace
}  // namespace dmb

namespace dmb {

template <typename T>
struct GemmW {  // a repacked weight: [ntaps * rows_per_tap, K] K-major, plus its TMA descriptor
  T* w = nullptr;
  int K = 0, N = 0, rows_per_tap = 0, ntaps = 0, bn = 0;
  CUtensorMap tmap_bn[6];  // box = 8 << i rows (8 .. 256): the N tile is narrowed when a launch has
                           // too few tiles to fill the chip; the CTA-pair kernel loads half tiles
  const CUtensorMap& tmap_for(int box_rows) const {
    int i = 0;
    while ((8 << i) < box_rows) ++i;
    return tmap_bn[i];
  }
  const float* bias = nullptr;
};

template <typename T>
struct Act {  // an activation in padded pixel-matrix layout
  T* p = nullptr;
  int C = 0;
  Geo g;
  CUtensorMap tmap;
  CUtensorMap tmap_samples; // 3D (channel, row of the sample, sample) view: pixel-major conv mode on small maps
  bool has_samples_map = false;
  CUtensorMap tmap_halo;    // box of kHaloRows rows: the shared A window of a 3x3 kernel row (gemm_tc.cuh, kRow3)
  long long* sstat = nullptr;  // [2 * n] fixed-point (sum, sumsq) per sample for the following GroupNorm(1, C)
                               // (conv outputs only; a slice of the plan's statistics arena, cleared once per step)
};

using Op = std::function<void(cudaStream_t)>;

// Launch list with a label and the algorithmic work of every entry (for the timing mode).
struct OpList {
  std::vector<Op> fns;
  std::vector<std::string> labels;
  std::vector<double> flops, bytes;
  std::string next_label = "op";
  double next_flops = 0, next_bytes = 0;
  void tag(const std::string& label, double fl = 0, double by = 0) { next_label = label; next_flops = fl; next_bytes = by; }
  void push_back(Op f) {
    fns.push_back(std::move(f)); labels.push_back(next_label); flops.push_back(next_flops); bytes.push_back(next_bytes);
    next_label = "op"; next_flops = 0; next_bytes = 0;
  }
  size_t size() const { return fns.size(); }
};

template <typename T>
struct Plan {
  Pool pool;
  OpList ops;
  std::map<std::string, Act<T>> taps;
  int B = 0;
  bool doubled = false;
  // bound inputs / state (device)
  float* x_state = nullptr;
  long long* t_vec = nullptr;
  long long* y = nullptr;
  float* vals = nullptr;
  float* mask = nullptr;
  DynParams* dyn = nullptr;
  float* emb = nullptr;
  float* embproj = nullptr;
  float* eps = nullptr;   // eager forward output staging (B,4,28,28)
  float* geom = nullptr;  // (B,12)
  Act<T> feat;            // sa6 output
  cudaGraphExec_t exec_one = nullptr;
  cudaGraphExec_t exec_chunk = nullptr;
  cudaGraphExec_t exec_eager = nullptr;   // one eager CFG step (dmb_denoise_cond_step) as a graph
  int chunk = 0;
  int launches_per_step = 0;
  int launches_eager = 0;
  ~Plan() {
    if (exec_eager) cudaGraphExecDestroy(exec_eager);
    if (exec_one) cudaGraphExecDestroy(exec_one);
    if (exec_chunk) cudaGraphExecDestroy(exec_chunk);
  }
};

template <typename T>
class Engine : public IEngine {
 public:
  explicit Engine(int device) : device_(device) {
    DMB_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    DMB_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) fail(4, "this library is built for sm_100a (B200); found sm_" + std::to_string(prop.major) + std::to_string(prop.minor));
    num_sms_ = prop.multiProcessorCount;
    // the embedding kernel runs beside inc / down1 on a second stream (fork / join by events,
    // also inside the captured graph)
    DMB_CUDA(cudaStreamCreateWithFlags(&side_, cudaStreamNonBlocking));
    DMB_CUDA(cudaEventCreateWithFlags(&ev_fork_, cudaEventDisableTiming));
    DMB_CUDA(cudaEventCreateWithFlags(&ev_join_, cudaEventDisableTiming));
    DMB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<T, 1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemMax));
    DMB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<T, 1, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemMax));
    DMB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<T, 1, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemMax));
    DMB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<T, 1, 1, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemMax));
    DMB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<T, 1, 2, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemMax));
    DMB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<T, 1, 4, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemMax));
    {
      const char* pe2 = getenv("DMB200_PDL");
      use_pdl_ = pe2 ? (pe2[0] != '0') : false;
      const char* ae = getenv("DMB200_ACC4");
      use_acc4_ = ae ? (ae[0] != '0') : true;
    }
    DMB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<T, 2, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemMax));
    DMB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<T, 2, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemMax));
    DMB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<T, 2, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemMax));
    DMB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<T, 1, 1, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemMax));
    DMB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<T, 2, 1, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemMax));
    DMB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<T, 1, 2, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemMax));
    DMB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<T, 2, 2, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemMax));
    DMB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<T, 1, 3, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemMax));
    DMB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<T, 2, 3, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemMax));
    DMB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<T, 1, 1, 4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemMax));
    DMB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<T, 1, 2, 4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemMax));
    DMB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<T, 1, 3, 4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemMax));
    DMB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<T, 2, 1, 4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemMax));
    DMB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<T, 2, 2, 4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemMax));
    DMB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<T, 2, 3, 4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemMax));
    {
      const char* pe3 = getenv("DMB200_PIX");
      pix_mode_ = !(pe3 && pe3[0] == '0');
      if (const char* e = getenv("DMB200_PIX_MINKN")) pix_min_kn_ = atoll(e);
      if (const char* e = getenv("DMB200_PIX_FILL")) pix_fill_pct_ = atoi(e);
      if (const char* e = getenv("DMB200_PIX_MAXHW")) pix_max_hw_ = atoi(e);
      const char* re = getenv("DMB200_ROW3");   // 0: nine separate A tiles per K chunk
      row3_ = re ? atoi(re) : 1;
      if (const char* e = getenv("DMB200_ROW3_GROUP")) row3_group_ = std::max(0, std::min(3, atoi(e)));
      if (const char* e = getenv("DMB200_ACC4_CONV")) acc4_conv_ = e[0] != '0';
      if (const char* e = getenv("DMB200_CONVT_PLANAR")) convt_planar_ = e[0] != '0';
      if (const char* e = getenv("DMB200_ENC_IN_MMA")) enc_in_mma_ = atoi(e) != 0;
    }
    const char* pe = getenv("DMB200_PAIRS");
    use_pairs_ = pe ? (pe[0] - '0') : 1;
    const char* ne = getenv("DMB200_NARROW_TILES");
    narrow_tiles_ = ne ? (ne[0] != '0') : true;
    if (ne && atoi(ne) >= 16) min_bn_ = atoi(ne);
    DMB_CUDA(cudaFuncSetAttribute(attention_kernel<T, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024));
    DMB_CUDA(cudaFuncSetAttribute(attention_kernel<T, 32>, cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024));
    DMB_CUDA(cudaFuncSetAttribute(attention_kernel<T, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024));
    if constexpr (std::is_same<T, __nv_bfloat16>::value) {
      DMB_CUDA(cudaFuncSetAttribute(attention_tc_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      DMB_CUDA(cudaFuncSetAttribute(attention_tc_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024));
      DMB_CUDA(cudaFuncSetAttribute(attention_tc_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024));
      DMB_CUDA(cudaFuncSetAttribute(attn_tail256_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, attn_tail256_smem_bytes()));
      DMB_CUDA(cudaFuncSetAttribute(attn_tail_kernel<64, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, attn_tail_smem_bytes<64, 4>()));
      DMB_CUDA(cudaFuncSetAttribute(attn_tail_kernel<128, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, attn_tail_smem_bytes<128, 2>()));
      DMB_CUDA(cudaFuncSetAttribute(ln_qkv_kernel<64, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, attn_tail_smem_bytes<64, 4>()));
      DMB_CUDA(cudaFuncSetAttribute(ln_qkv_kernel<128, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, attn_tail_smem_bytes<128, 2>()));
    }
    {
      const char* le = getenv("DMB200_FUSE_LN1");
      fuse_ln1_ = !(le && le[0] == '0');
      const char* ke = getenv("DMB200_KEEP_TAPS");
      keep_taps_ = ke && ke[0] == '1';
    }
    const char* fe = getenv("DMB200_FUSED_TAIL");
    use_fused_tail_ = !(fe && fe[0] == '0');
    const char* bn = getenv("DMB200_BN256");
    g_allow_bn256 = !(bn && bn[0] == '0');
    const char* env = getenv("DMB200_CUDA_CORE_ATTENTION");
    use_tc_attention_ = !(env && env[0] == '1');
    if (const char* e = getenv("DMB200_ATTN_MMA")) use_mma_attention_ = e[0] != '0';    // 0: the tcgen05 kernel
    if (const char* e = getenv("DMB200_EAGER_GRAPH")) eager_graph_ = e[0] != '0';
    if (const char* e = getenv("DMB200_TAIL256")) use_tail256_ = e[0] != '0';
    if (const char* e = getenv("DMB200_LN_QKV")) use_ln_qkv_ = e[0] != '0';
    if (const char* e = getenv("DMB200_GN_SPLITS")) gn_splits_ = atoi(e);
    if (const char* e = getenv("DMB200_ATTN_MMA_MINL")) attn_mma_min_l_ = std::max(1, atoi(e));
    if (const char* e = getenv("DMB200_GN_INPLACE")) gn_inplace_ = e[0] != '0';
    if (const char* e = getenv("DMB200_ALIAS_ACTS")) alias_acts_ = e[0] != '0';
    if (const char* e = getenv("DMB200_CONVIN_Z")) conv_in_z_ = std::max(0, std::min(7, atoi(e)));
    if (const char* e = getenv("DMB200_ATTN_POLY")) attn_poly_ = e[0] != '0';           // 0: every exp2 on MUFU
    if constexpr (std::is_same<T, float>::value) {
      DMB_CUDA(cudaFuncSetAttribute(attention_mma_f32_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
      DMB_CUDA(cudaFuncSetAttribute(attention_mma_f32_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
      DMB_CUDA(cudaFuncSetAttribute(attention_mma_f32_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    }
    if constexpr (std::is_same<T, __nv_bfloat16>::value) {
      const int am_smem = 110 * 1024;
      if (const char* e = getenv("DMB200_ATTN_VAR")) attn_var_ = atoi(e) & 7;
#define DMB_AM_ATTR(D, P, V) DMB_CUDA(cudaFuncSetAttribute(attention_mma_kernel<D, P, V>, cudaFuncAttributeMaxDynamicSharedMemorySize, am_smem))
      DMB_AM_ATTR(16, -1, 0); DMB_AM_ATTR(16, -1, 1); DMB_AM_ATTR(16, -1, 5); DMB_AM_ATTR(16, 3, 5); DMB_AM_ATTR(16, 3, 1);
      DMB_AM_ATTR(32, -1, 0); DMB_AM_ATTR(32, -1, 1); DMB_AM_ATTR(32, -1, 5); DMB_AM_ATTR(32, 3, 5); DMB_AM_ATTR(32, 3, 1);
      DMB_AM_ATTR(64, -1, 0); DMB_AM_ATTR(64, -1, 1); DMB_AM_ATTR(64, -1, 5); DMB_AM_ATTR(64, 3, 5); DMB_AM_ATTR(64, 3, 1);
#undef DMB_AM_ATTR
    }
  }
  ~Engine() override {
    if (ev_fork_) cudaEventDestroy(ev_fork_);
    if (ev_join_) cudaEventDestroy(ev_join_);
    if (side_) cudaStreamDestroy(side_);
  }

  void load_unet(const TensorMapIn& w) override;
  void load_vae(const TensorMapIn& w) override;
  void unet_forward(const float* x, const long long* t, const long long* y, const float* vals, const float* mask,
                    int B, float* eps, float* geom, cudaStream_t s) override;
  void cfg_step(const float* x_in, const long long* t, const long long* y, int null_label, const float* vals,
                const float* mask, float guidance, const float* coef, int T_steps, const float* noise,
                unsigned long long seed, unsigned long long sample_offset, float* x_out, float* eps_out, int B,
                cudaStream_t s) override;
  void sample_loop(float* x, const long long* y, int null_label, const float* vals, const float* mask,
                   float guidance, const float* coef, int T_steps, int t_start, int t_stride, int n_steps,
                   const float* noise, unsigned long long seed, unsigned long long sample_offset, float* x_dump,
                   float* eps_dump, int B, cudaStream_t s) override;
  void vae_decode(const float* z, int B, float* img_f32, uint8_t* img_u8, cudaStream_t s) override;
  void load_vae_enc(const TensorMapIn& w) override;
  void vae_encode(const float* img, const float* eps, int B, float* z, float* mu, float* logvar, float* kl_sample,
                  cudaStream_t s) override;
  void add_noise(const float* x0, const long long* t, const float* alpha_bars, int T_steps, const float* noise, int B,
                 long long per_sample, float* x_t, cudaStream_t s) override;
  void get_tap(const char* name, float* out, int64_t capacity, int64_t* shape4, cudaStream_t s) override;
  void op_conv3x3(const float* x, const float* w, int B, int Cin, int Cout, int H, int W, float* out,
                  cudaStream_t s) override;
  void op_linear(const float* x, const float* w, const float* bias, int rows, int K, int N, int act, float* out,
                 cudaStream_t s) override;
  void op_convT4x4(const float* x, const float* w, const float* bias, int B, int Cin, int Cout, int H, int W,
                   float* out, cudaStream_t s) override;
  void set_timing(bool on) override {
    for (auto& t : timed_) { cudaEventDestroy(t.e0); cudaEventDestroy(t.e1); }
    timed_.clear();
    timing_ = on;
  }
  std::string timing_json() override;
  long long launches() const override { return launches_; }
  void reset_launches() override { launches_ = 0; }

 private:
  static constexpr int kGemmSmemMax = 227 * 1024 - 1024;  // static smem (barriers, < 300 bytes) comes out of the same 227 KB

  // ---- weights
  const float* f(const std::string& name) const {
    auto it = blob_.find(name);
    if (it == blob_.end()) fail(2, "missing weight tensor '" + name + "'");
    return it->second.first;
  }
  int64_t numel(const std::string& name) const {
    auto it = blob_.find(name);
    if (it == blob_.end()) fail(2, "missing weight tensor '" + name + "'");
    return it->second.second;
  }
  void upload(const TensorMapIn& w, const std::string& prefix, Pool& pool);
  GemmW<T> make_conv3x3_w(const std::string& name, int Co, int Ci, Pool& pool);
  GemmW<T> make_linear_w(const std::string& wname, const std::string& bname, int N, int K, Pool& pool);
  GemmW<T> make_convT_w(const std::string& wname, const std::string& bname, int Ci, int Co, Pool& pool);
  GemmW<T> make_conv4s2_w(const std::string& wname, const std::string& bname, int Ci, int Co, Pool& pool);
  GemmW<T> make_conv3x3_from_ptr(const float* w_dev, int Co, int Ci, Pool& pool);

  // ---- launch helpers
  // Launch of a step-loop kernel.  With programmatic dependent launch (DMB200_PDL=1) the kernel may
  // be scheduled while the previous kernel of the stream drains; every kernel launched through
  // here calls pdl_wait() before it touches global memory (common.cuh).  Off by default: the
  // kernels' own latency chains, not the launch gaps, bound the small-batch step (measured).
  template <typename... KArgs, typename... Args>
  void launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = use_pdl_ ? 1 : 0;
    DMB_CUDA(cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...));
  }
  Act<T> new_act(Pool& pool, int n, int H, int W, int C, bool with_stats);
  // N tile of a launch: the weight's own tile, halved (down to 32) while the launch would leave
  // SMs without a tile -- at small batch the GEMMs are bound by streaming the weights, and more,
  // narrower tiles spread that stream over more SMs.
  int effective_bn(const GemmW<T>& w, int rows_valid) const {
    int bn = w.bn;
    const int m_tiles = cdiv(rows_valid, kTileM);
    while (narrow_tiles_ && bn > min_bn_ && m_tiles * (w.rows_per_tap / bn) < num_sms_) bn >>= 1;
    return bn;
  }
  void gemm(const Act<T>& a, const GemmW<T>& w, int phase, const int* shifts, int ntaps, GemmParams ep,
            cudaStream_t s, const Act<T>* out_act = nullptr);
  Op op_conv(const Act<T>& in, const GemmW<T>& w, const Act<T>& out);
  Op op_lin(const Act<T>& in, const GemmW<T>& w, const Act<T>& out, int act, const Act<T>* residual);
  void build_unet_plan(Plan<T>& p);
  Plan<T>& get_plan(int B, bool doubled);
  void bind_inputs(Plan<T>& p, const float* x, const long long* t, const long long* y, const float* vals,
                   const float* mask, int null_label, cudaStream_t s);
  void run_list(OpList& l, cudaStream_t s) {
    for (size_t i = 0; i < l.fns.size(); ++i) {
      if (timing_) {
        Timed t{l.labels[i], l.flops[i], l.bytes[i], nullptr, nullptr};
        DMB_CUDA(cudaEventCreate(&t.e0));
        DMB_CUDA(cudaEventCreate(&t.e1));
        DMB_CUDA(cudaEventRecord(t.e0, s));
        l.fns[i](s);
        DMB_CUDA(cudaEventRecord(t.e1, s));
        timed_.push_back(t);
      } else {
        l.fns[i](s);
      }
    }
  }
  void run_ops(Plan<T>& p, cudaStream_t s) { run_list(p.ops, s); }
  struct Timed {
    std::string label;
    double flops, bytes;
    cudaEvent_t e0, e1;
  };
  bool timing_ = false;
  std::vector<Timed> timed_;

  int device_ = 0;
  int num_sms_ = 148;
  long long launches_ = 0;
  bool counting_ = true;
  long long* stat_arena_cur_ = nullptr;   // bump cursor into the statistics arena of the plan being built
  size_t stat_arena_left_ = 0;
  bool use_tc_attention_ = true;
  int conv_in_z_ = 0;               // DMB200_CONVIN_Z: CTAs per sample of the U-Net's input conv (0: 7 below 64 samples, else 1)
  bool alias_acts_ = true;          // DMB200_ALIAS_ACTS=0: attention-block tails / ln_qkv write separate tensors instead of storing over their input tiles
  bool gn_inplace_ = true;          // DMB200_GN_INPLACE=0: GroupNorm apply writes a separate tensor instead of rewriting the conv output in place
  int attn_mma_min_l_ = 9;          // DMB200_ATTN_MMA_MINL=17: 3x3 maps (L = 9) on the CUDA-core kernel again (17.3 vs 13.4 us at batch 256)
  int gn_splits_ = 0;               // DMB200_GN_SPLITS: CTAs per sample of gn1_apply (0: by batch, see gn())
  bool use_ln_qkv_ = true;          // DMB200_LN_QKV=0: the attention block's LayerNorm and in-projection as separate launches everywhere
  bool use_tail256_ = true;         // DMB200_TAIL256=0: the C = 256 attention blocks keep their separate Linear / LayerNorm launches
  bool eager_graph_ = true;         // DMB200_EAGER_GRAPH=0: dmb_denoise_cond_step launches its kernels one by one
  bool use_mma_attention_ = true;   // warp-level MMA attention (attention_mma.cuh) for maps > 3x3 in bf16 mode
  int attn_var_ = 1;                // experiment bits of attention_mma_kernel (DMB200_ATTN_VAR)
  bool attn_poly_ = false;          // DMB200_ATTN_POLY=1: a quarter of its exponentials on the FMA pipe (measured slower: issue-bound)
  cudaStream_t side_ = nullptr;
  cudaEvent_t ev_fork_ = nullptr, ev_join_ = nullptr;
  bool use_pdl_ = false;   // DMB200_PDL=1: programmatic dependent launch (measured: no gain at batch 1, -1.6 % at 256)
  bool fuse_ln1_ = true;    // DMB200_FUSE_LN1=0: the attention block's first LayerNorm as its own launch
  bool keep_taps_ = false;  // DMB200_KEEP_TAPS=1: also store the un-normalised stage outputs (down1.. up3 taps)
  long long pix_min_kn_ = 256 * 256;   // pixel-major only for C_in * C_out >= this
  int pix_max_hw_ = 14;                // largest map side that may run pixel-major
  int pix_fill_pct_ = 90;              // narrow the N tile while there are fewer tiles than this share of the SMs
  bool pix_mode_ = true;   // DMB200_PIX=0: small maps keep the padded-row GEMM at every batch size
  int row3_ = 1;   // DMB200_ROW3=0: every tap of a 3x3 conv loads its own A tile
  int row3_group_ = 0;   // DMB200_ROW3_GROUP=1|2|3: kernel-row steps per pipeline stage (0: by N tile)
  bool acc4_conv_ = true;    // DMB200_ACC4_CONV=0: two accumulator stages / epilogue warpgroups for 3x3 convs with N tile <= 64 too
  bool convt_planar_ = true;   // DMB200_CONVT_PLANAR=0: ConvTranspose phases scatter their rows from registers
  bool use_fused_tail_ = true;   // DMB200_FUSED_TAIL=0: out_proj / LayerNorm / ff1 / ff2 as four launches
  bool use_acc4_ = true;   // DMB200_ACC4=0: two epilogue warpgroups for every launch
  int min_bn_ = 16;
  bool narrow_tiles_ = true;   // DMB200_NARROW_TILES=0: always the weight's own N tile
  int use_pairs_ = 1;   // cta_group::2 GEMM tiles: DMB200_PAIRS=0 off, 1 when there is enough work, 2 always

  Pool wpool_unet_, wpool_vae_;
  std::map<std::string, std::pair<const float*, int64_t>> blob_;  // name -> fp32 device copy
  std::map<std::string, GemmW<T>> gw_;
  EmbWeights embw_{};
  const float* geom_w1t_ = nullptr;
  bool unet_loaded_ = false, vae_loaded_ = false;

  std::map<std::pair<int, bool>, std::unique_ptr<Plan<T>>> plans_;
  Plan<T>* last_plan_ = nullptr;

  struct VaePlan {
    Pool pool;
    OpList ops;
    int B = 0;
    float* z = nullptr;
    float* img_f32 = nullptr;
    uint8_t* img_u8 = nullptr;
  };
  // plans by chunk size: the full chunk (32) and the tails a caller's batch sizes leave behind, so a
  // batch that is not a multiple of 32 does not rebuild a plan (cudaMalloc / cudaFree / device sync)
  // twice per call.  At most kMaxVaePlans are kept; the full chunk is never evicted.
  static constexpr int kMaxVaePlans = 3;
  std::map<int, std::unique_ptr<VaePlan>> vae_plans_;
  VaePlan* vae_plan_ = nullptr;   // the one used last
  void build_vae_plan(VaePlan& p, int B);
  struct VaeEncPlan {
    Pool pool;
    OpList ops;
    int B = 0;
    const float* img = nullptr;
    const float* eps = nullptr;
    float* z = nullptr;
    float* mu = nullptr;
    float* logvar = nullptr;
    float* kl = nullptr;
    std::map<std::string, Act<T>> taps;   // "venc.<layer>": raw conv outputs and plain GN outputs (debug)
  };
  std::map<int, std::unique_ptr<VaeEncPlan>> vae_enc_plans_;
  VaeEncPlan* vae_enc_plan_ = nullptr;   // the one used last (debug taps read it)
  void build_vae_enc_plan(VaeEncPlan& p, int B);
  Pool wpool_vae_enc_;
  bool vae_enc_loaded_ = false;
  bool enc_in_mma_ = true;   // DMB200_ENC_IN_MMA=0: enc.0 on the CUDA cores (FFMA2) instead of mma.sync TF32

  void count(int n = 1) {
    if (counting_) launches_ += n;
  }
};

// ---------------------------------------------------------------------------
// Weights
// ---------------------------------------------------------------------------
template <typename T>
void Engine<T>::upload(const TensorMapIn& w, const std::string& prefix, Pool& pool) {
  size_t total = 0;
  for (auto& kv : w) total += (static_cast<size_t>(kv.second.numel) + 63) / 64 * 64;
  float* base = pool.alloc_n<float>(total);
  size_t off = 0;
  for (auto& kv : w) {
    DMB_CUDA(cudaMemcpy(base + off, kv.second.ptr, kv.second.numel * sizeof(float), cudaMemcpyDefault));
    blob_[prefix + kv.first] = {base + off, kv.second.numel};
    off += (static_cast<size_t>(kv.second.numel) + 63) / 64 * 64;
  }
}

template <typename T>
GemmW<T> Engine<T>::make_conv3x3_from_ptr(const float* w_dev, int Co, int Ci, Pool& pool) {
  GemmW<T> g;
  g.K = Ci; g.N = Co; g.ntaps = 9;
  g.bn = pick_block_n(Co);
  g.rows_per_tap = (Co + g.bn - 1) / g.bn * g.bn;
  DMB_REQUIRE(Ci % ElemTraits<T>::kChunkK == 0, "conv input channels must be a multiple of the K chunk");
  g.w = pool.alloc_n<T>(static_cast<size_t>(9) * g.rows_per_tap * Ci);
  const long long total = 9LL * Co * Ci;
  repack_conv3x3_kernel<T><<<cdiv(total, 256), 256>>>(w_dev, Co, Ci, g.rows_per_tap, g.w);
  DMB_CHECK_LAUNCH();
  for (int i = 0; i < 6; ++i) g.tmap_bn[i] = make_tmap<T>(g.w, Ci, static_cast<uint64_t>(9) * g.rows_per_tap, 8 << i);
  return g;
}

template <typename T>
GemmW<T> Engine<T>::make_conv3x3_w(const std::string& name, int Co, int Ci, Pool& pool) {
  DMB_REQUIRE(numel(name) == 9LL * Co * Ci, "unexpected size of " + name);
  return make_conv3x3_from_ptr(f(name), Co, Ci, pool);
}

template <typename T>
GemmW<T> Engine<T>::make_linear_w(const std::string& wname, const std::string& bname, int N, int K, Pool& pool) {
  DMB_REQUIRE(numel(wname) == static_cast<int64_t>(N) * K, "unexpected size of " + wname);
  GemmW<T> g;
  g.K = K; g.N = N; g.ntaps = 1;
  g.bn = pick_block_n(N);
  g.rows_per_tap = (N + g.bn - 1) / g.bn * g.bn;
  g.w = pool.alloc_n<T>(static_cast<size_t>(g.rows_per_tap) * K);
  cast_kernel<T><<<cdiv(static_cast<long long>(N) * K, 256), 256>>>(f(wname), static_cast<long long>(N) * K, g.w);
  DMB_CHECK_LAUNCH();
  for (int i = 0; i < 6; ++i) g.tmap_bn[i] = make_tmap<T>(g.w, K, g.rows_per_tap, 8 << i);
  if (!bname.empty()) {
    DMB_REQUIRE(numel(bname) == N, "unexpected size of " + bname);
    g.bias = f(bname);
  }
  return g;
}

template <typename T>
GemmW<T> Engine<T>::make_convT_w(const std::string& wname, const std::string& bname, int Ci, int Co, Pool& pool) {
  DMB_REQUIRE(numel(wname) == 16LL * Co * Ci, "unexpected size of " + wname);
  GemmW<T> g;
  g.K = Ci; g.N = Co; g.ntaps = 4;
  g.bn = pick_block_n(Co);
  g.rows_per_tap = Co;
  DMB_REQUIRE(Co % g.bn == 0, "conv-transpose output channels must be a multiple of the N tile");
  g.w = pool.alloc_n<T>(16ull * Co * Ci);
  repack_convT_kernel<T><<<cdiv(16LL * Co * Ci, 256), 256>>>(f(wname), Ci, Co, g.w);
  DMB_CHECK_LAUNCH();
  for (int i = 0; i < 6; ++i) g.tmap_bn[i] = make_tmap<T>(g.w, Ci, 16ull * Co, 8 << i);
  if (!bname.empty()) g.bias = f(bname);
  return g;
}

template <typename T>
GemmW<T> Engine<T>::make_conv4s2_w(const std::string& wname, const std::string& bname, int Ci, int Co, Pool& pool) {
  DMB_REQUIRE(numel(wname) == 16LL * Co * Ci, "unexpected size of " + wname);
  GemmW<T> g;
  g.K = 4 * Ci; g.N = Co; g.ntaps = 4;
  g.bn = pick_block_n(Co);
  g.rows_per_tap = Co;
  DMB_REQUIRE(Co % g.bn == 0, "strided-conv output channels must be a multiple of the N tile");
  g.w = pool.alloc_n<T>(16ull * Co * Ci);
  repack_conv4s2_kernel<T><<<cdiv(16LL * Co * Ci, 256), 256>>>(f(wname), Co, Ci, g.w);
  DMB_CHECK_LAUNCH();
  for (int i = 0; i < 6; ++i) g.tmap_bn[i] = make_tmap<T>(g.w, 4 * Ci, 4ull * Co, 8 << i);
  if (!bname.empty()) g.bias = f(bname);
  return g;
}

}  // namespace dmb

namespace dmb {

static const char* kResBlocks[] = {"inc", "down1.maxpool_conv.1", "down1.maxpool_conv.2", "down2.maxpool_conv.1",
                                   "down2.maxpool_conv.2", "down3.maxpool_conv.1", "down3.maxpool_conv.2", "bot1",
                                   "bot2", "bot3", "up1.conv.0", "up1.conv.1", "up2.conv.0", "up2.conv.1",
                                   "up3.conv.0", "up3.conv.1"};
// (C_in, C_mid, C_out) of the sixteen ResBlocks, models/unet_cond.py:131-153.
static const int kResDims[][3] = {{4, 64, 64},      {64, 64, 64},     {64, 128, 128},  {128, 128, 128},
                                  {128, 256, 256},  {256, 256, 256},  {256, 256, 256}, {256, 512, 512},
                                  {512, 512, 512},  {512, 256, 256},  {512, 512, 512}, {512, 256, 128},
                                  {256, 256, 256},  {256, 128, 64},   {128, 128, 128}, {128, 64, 64}};
static const char* kAttn[] = {"sa1", "sa2", "sa3", "sa4", "sa5", "sa6"};
static const int kAttnC[] = {128, 256, 256, 128, 64, 64};
static const char* kEmbLayers[] = {"down1", "down2", "down3", "up1", "up2", "up3"};
static const int kEmbC[] = {128, 256, 256, 128, 64, 64};

template <typename T>
void Engine<T>::load_unet(const TensorMapIn& w) {
  DMB_CUDA(cudaSetDevice(device_));
  DMB_CUDA(cudaDeviceSynchronize());
  plans_.clear();
  last_plan_ = nullptr;
  // drop the U-Net's entries only: the VAE decoder ("vae.") and encoder ("venc.") keep theirs
  auto is_vae = [](const std::string& k) { return k.rfind("vae.", 0) == 0 || k.rfind("venc.", 0) == 0; };
  for (auto it = gw_.begin(); it != gw_.end();) {
    if (is_vae(it->first)) ++it; else it = gw_.erase(it);
  }
  for (auto it = blob_.begin(); it != blob_.end();) {
    if (is_vae(it->first)) ++it; else it = blob_.erase(it);
  }
  wpool_unet_.release();
  unet_loaded_ = false;
  upload(w, "", wpool_unet_);
  Pool& pool = wpool_unet_;

  for (int i = 0; i < 16; ++i) {
    const std::string p = kResBlocks[i];
    const int ci = kResDims[i][0], cm = kResDims[i][1], co = kResDims[i][2];
    if (i > 0) gw_[p + ".double_conv.0.weight"] = make_conv3x3_w(p + ".double_conv.0.weight", cm, ci, pool);
    else DMB_REQUIRE(numel(p + ".double_conv.0.weight") == 9LL * cm * ci, "unexpected size of inc conv");
    gw_[p + ".double_conv.3.weight"] = make_conv3x3_w(p + ".double_conv.3.weight", co, cm, pool);
    DMB_REQUIRE(numel(p + ".double_conv.1.weight") == cm && numel(p + ".double_conv.4.bias") == co, "GroupNorm size");
  }
  for (int i = 0; i < 6; ++i) {
    const std::string p = kAttn[i];
    const int C = kAttnC[i];
    gw_[p + ".qkv"] = make_linear_w(p + ".mha.in_proj_weight", p + ".mha.in_proj_bias", 3 * C, C, pool);
    gw_[p + ".proj"] = make_linear_w(p + ".mha.out_proj.weight", p + ".mha.out_proj.bias", C, C, pool);
    gw_[p + ".ff1"] = make_linear_w(p + ".ff_self.1.weight", p + ".ff_self.1.bias", C, C, pool);
    gw_[p + ".ff2"] = make_linear_w(p + ".ff_self.3.weight", p + ".ff_self.3.bias", C, C, pool);
  }
  // Embedding path: small fp32 matrices, transposed so that thread o reads column o coalesced.
  float* w1t = pool.alloc_n<float>(24 * 256);
  float* w2t = pool.alloc_n<float>(256 * 256);
  float* projt = pool.alloc_n<float>(256 * kProjCols);
  float* projb = pool.alloc_n<float>(kProjCols);
  float* invf = pool.alloc_n<float>(128);
  transpose_kernel<<<cdiv(256 * 24, 256), 256>>>(f("cond_mlp.0.weight"), 256, 24, w1t, 256, 0);
  transpose_kernel<<<cdiv(256 * 256, 256), 256>>>(f("cond_mlp.2.weight"), 256, 256, w2t, 256, 0);
  int col = 0;
  for (int i = 0; i < 6; ++i) {
    const std::string p = std::string(kEmbLayers[i]) + ".emb_layer.1";
    DMB_REQUIRE(numel(p + ".weight") == static_cast<int64_t>(kEmbC[i]) * 256, "unexpected size of " + p);
    transpose_kernel<<<cdiv(kEmbC[i] * 256, 256), 256>>>(f(p + ".weight"), kEmbC[i], 256, projt, kProjCols, col);
    DMB_CUDA(cudaMemcpy(projb + col, f(p + ".bias"), kEmbC[i] * sizeof(float), cudaMemcpyDeviceToDevice));
    col += kEmbC[i];
  }
  DMB_CHECK_LAUNCH();
  {
    // inv_freq = 1 / 10000^(2j/256) evaluated in fp32 the way torch does (pow then reciprocal);
    // the host shim overrides it with torch's own values through the "__inv_freq" entry.
    float h[128];
    for (int j = 0; j < 128; ++j) h[j] = 1.0f / powf(10000.0f, static_cast<float>(2 * j) / 256.0f);
    auto it = blob_.find("__inv_freq");
    if (it != blob_.end()) DMB_CUDA(cudaMemcpy(invf, it->second.first, 128 * sizeof(float), cudaMemcpyDeviceToDevice));
    else DMB_CUDA(cudaMemcpy(invf, h, sizeof(h), cudaMemcpyHostToDevice));
  }
  DMB_REQUIRE(numel("class_emb.weight") == 4 * 256, "class_emb must be (4,256)");
  embw_.class_emb = f("class_emb.weight");
  embw_.w1t = w1t; embw_.b1 = f("cond_mlp.0.bias");
  embw_.w2t = w2t; embw_.b2 = f("cond_mlp.2.bias");
  embw_.inv_freq = invf; embw_.projt = projt; embw_.projb = projb;
  float* g1t = pool.alloc_n<float>(64 * 256);
  transpose_kernel<<<cdiv(256 * 64, 256), 256>>>(f("geom_head.mlp.0.weight"), 256, 64, g1t, 256, 0);
  DMB_CHECK_LAUNCH();
  geom_w1t_ = g1t;
  DMB_REQUIRE(numel("out.weight") == 4 * 64 && numel("out.bias") == 4, "out conv must be 64->4");
  DMB_CUDA(cudaDeviceSynchronize());
  unet_loaded_ = true;
}

template <typename T>
void Engine<T>::load_vae(const TensorMapIn& w) {
  DMB_CUDA(cudaSetDevice(device_));
  DMB_CUDA(cudaDeviceSynchronize());
  vae_plans_.clear();
  vae_plan_ = nullptr;
  for (auto it = blob_.begin(); it != blob_.end();) {
    if (it->first.rfind("vae.", 0) == 0) it = blob_.erase(it); else ++it;
  }
  for (auto it = gw_.begin(); it != gw_.end();) {
    if (it->first.rfind("vae.", 0) == 0) it = gw_.erase(it); else ++it;
  }
  wpool_vae_.release();
  vae_loaded_ = false;
  upload(w, "vae.", wpool_vae_);
  Pool& pool = wpool_vae_;
  DMB_REQUIRE(numel("vae.dec.0.weight") == 256 * 4 * 9, "dec.0 must be conv 4->256");
  gw_["vae.dec.3"] = make_convT_w("vae.dec.3.weight", "vae.dec.3.bias", 256, 256, pool);
  gw_["vae.dec.6"] = make_conv3x3_w("vae.dec.6.weight", 128, 256, pool);
  gw_["vae.dec.6"].bias = f("vae.dec.6.bias");
  gw_["vae.dec.9"] = make_convT_w("vae.dec.9.weight", "vae.dec.9.bias", 128, 128, pool);
  gw_["vae.dec.12"] = make_conv3x3_w("vae.dec.12.weight", 64, 128, pool);
  gw_["vae.dec.12"].bias = f("vae.dec.12.bias");
  gw_["vae.dec.15"] = make_convT_w("vae.dec.15.weight", "vae.dec.15.bias", 64, 64, pool);
  gw_["vae.dec.18"] = make_conv3x3_w("vae.dec.18.weight", 3, 64, pool);
  gw_["vae.dec.18"].bias = f("vae.dec.18.bias");
  DMB_CUDA(cudaDeviceSynchronize());
  vae_loaded_ = true;
}

}  // namespace dmb

namespace dmb {

// ---------------------------------------------------------------------------
// Launch helpers
// ---------------------------------------------------------------------------
template <typename T>
Act<T> Engine<T>::new_act(Pool& pool, int n, int H, int W, int C, bool with_stats) {
  Act<T> a;
  a.C = C;
  a.g = make_geo(n, H, W);
  a.p = pool.alloc_n<T>(static_cast<size_t>(a.g.rows_alloc) * C);
  if (C % ElemTraits<T>::kChunkK == 0) {
    a.tmap = make_tmap<T>(a.p, C, a.g.rows_alloc, kTileM);
    a.tmap_halo = make_tmap<T>(a.p, C, a.g.rows_alloc, kHaloRows);
    if (H * W <= 196) {
      a.tmap_samples = make_tmap_samples<T>(a.p, C, a.g.S, n);
      a.has_samples_map = true;
    }
  }
  if (with_stats) {
    const size_t need = 2 * static_cast<size_t>(n);
    if (stat_arena_left_ < need) fail(1, "internal error: GroupNorm statistics arena exhausted");
    a.sstat = stat_arena_cur_;
    stat_arena_cur_ += need;
    stat_arena_left_ -= need;
  }
  return a;
}

// One launch of the tcgen05 GEMM.  `ep` carries the epilogue fields; the problem fields are
// filled here from the activation, the weight and the tap list.  `out_act` (optional) is the
// output activation when the tile can leave through a TMA store (identity row map, 128-byte
// column blocks); otherwise the epilogue stores per thread.
template <typename T>
void Engine<T>::gemm(const Act<T>& a, const GemmW<T>& w, int phase, const int* shifts, int ntaps, GemmParams ep,
                     cudaStream_t s, const Act<T>* out_act) {
  DMB_REQUIRE(a.C == w.K, "GEMM K mismatch");
  DMB_REQUIRE(ntaps <= kMaxTaps, "too many taps");
  GemmParams p = ep;
  p.num_taps = ntaps;
  for (int i = 0; i < ntaps; ++i) {
    p.a_shift[i] = shifts[i];
    p.b_row[i] = (phase * ntaps + i) * w.rows_per_tap;
  }
  p.k_chunks = w.K / ElemTraits<T>::kChunkK;
  // Pixel-major mode for 3x3 convs on small maps at large batch (gemm_tc.cuh, PixTile)
  const bool conv3 = ntaps == 9 && shifts[1] == shifts[0] + 1 && shifts[2] == shifts[1] + 1 &&
                     shifts[4] == shifts[3] + 1 && shifts[5] == shifts[4] + 1 && shifts[7] == shifts[6] + 1 &&
                     shifts[8] == shifts[7] + 1 && shifts[3] == shifts[0] + a.g.Wp && shifts[6] == shifts[3] + a.g.Wp;
  const int HW = a.g.H * a.g.W;
  bool pix = pix_mode_ && conv3 && a.g.H <= pix_max_hw_ && a.g.W <= pix_max_hw_ && HW <= 196 && a.g.n >= 2 * kTileM && a.has_samples_map && out_act &&
             out_act->has_samples_map && ep.epi == EPI_STORE && ep.map_mode == 0 && w.bn % ElemTraits<T>::kChunkK == 0 &&
             out_act->C == w.rows_per_tap && ep.out == out_act->p;
  int bn = effective_bn(w, a.g.rows_valid);
  if (pix && static_cast<long long>(w.K) * w.N < pix_min_kn_) pix = false;   // short K loops: per-tile overheads win
  if (pix) {   // tiles = pixels x batch tiles x N tiles
    bn = w.bn;
    while (narrow_tiles_ && bn > 64 && HW * cdiv(a.g.n, kTileM) * (w.rows_per_tap / bn) * 100 < num_sms_ * pix_fill_pct_) bn >>= 1;
  }
  p.block_n = bn;
  p.n_tiles = w.rows_per_tap / bn;
  // CTA pairs (256-row tiles) when there is enough work to keep every pair busy
  const int m_tiles1 = pix ? HW * cdiv(a.g.n, kTileM) : cdiv(a.g.rows_valid, kTileM);
  // (short K loops -- the Linear layers -- are epilogue-bound and lose on the pair's cross-CTA hand-shakes)
  const bool pair = bn >= 32 && ((use_pairs_ == 1 && ntaps * p.k_chunks >= 8 && (m_tiles1 / 2) * p.n_tiles >= num_sms_ / 2) ||
                                   (use_pairs_ == 2 && m_tiles1 >= 2));
  const int ctas = pair ? 2 : 1;
  p.m_tiles = cdiv(a.g.rows_valid, kTileM * ctas);
  p.pix = pix ? 1 : 0;
  if (pix) {
    p.pix_m_tiles = cdiv(a.g.n, kTileM * ctas);
    p.m_tiles = HW * p.pix_m_tiles;
    // pixels with the most neighbours (the longest K loops) first
    int k = 0;
    for (int want = 9; want >= 1; --want)
      for (int pi = 0; pi < HW; ++pi) {
        const int py = pi / a.g.W, px = pi % a.g.W;
        const int nb = (3 - (py == 0) - (py == a.g.H - 1)) * (3 - (px == 0) - (px == a.g.W - 1));
        if (nb == want) p.pix_order[k++] = static_cast<unsigned char>(pi);
      }
  }
  p.tma_store = (out_act && ep.epi == EPI_STORE && ep.map_mode == 0 && bn % ElemTraits<T>::kChunkK == 0 &&
                 out_act->C % ElemTraits<T>::kChunkK == 0 && out_act->C >= ep.out_col0 + p.n_tiles * bn &&
                 ep.out_col0 % ElemTraits<T>::kChunkK == 0 &&
                 out_act->g.rows_alloc == a.g.rows_alloc && ep.out == out_act->p && ep.ldc == out_act->C)
                    ? 1 : 0;
  if (pix) DMB_REQUIRE(p.tma_store == 1, "pixel-major conv needs the TMA-store epilogue");
  const int pair_bytes = kTileM * 128 + (bn / ctas) * 128;   // one (tap, K chunk) operand pair
  // Linear layers (one tap, short K, N tile <= 128): four accumulator stages / epilogue warpgroups
  // ... and the narrow 3x3 convs (N tile <= 64: 576 cycles of MMA per 128-row tile against an epilogue of a few
  // thousand per warpgroup; 64->64 @28x28 at batch 256: 37.4 -> 31.9 us, same bits)
  const bool acc4_conv = acc4_conv_ && row3_ && !pix && conv3 && bn <= 64 && p.tma_store;
  const bool acc4 = use_acc4_ && ((!pair && ntaps == 1 && bn <= 128 && p.tma_store) || acc4_conv);
  const int n_acc = acc4 ? 4 : 2;
  const int staging = p.tma_store ? gemm_staging_bytes(n_acc) : 0;
  const int ring = kGemmSmemMax - 1024 - staging;
  // One barrier hand-shake costs the issuing threads ~500 cycles (measured): group enough
  // operand pairs per stage that their MMAs (4 x bn/2 cycles per pair) cover it, while
  // keeping at least 3 stages in flight.
  const int n_iter = (pix ? 1 : ntaps) * p.k_chunks;   // (pixel-major: the shortest K loop bounds the group)
  int group = std::max(1, std::min(std::min(n_iter, kMaxGroup), cdiv(256, bn)));
  if (group == 3) group = 2;
  while (group > 1 && ring / (group * pair_bytes) < 3) group >>= 1;
  {
    static const char* ge = getenv("DMB200_GEMM_GROUP");  // experiment override
    if (ge) group = std::max(1, std::min(std::min(n_iter, kMaxGroup), atoi(ge)));
    if (group == 3) group = 2;
    while (group > 1 && ring / (group * pair_bytes) < 2) group >>= 1;
  }
  // 3x3 convs: one 130-row A window per (kernel row, K chunk) shared by the row's three taps,
  // when at least three such stages fit (any N tile for CTA pairs, <= 128 otherwise)
  const int row3_stage = static_cast<int>(kHaloBytes) + 3 * (bn / ctas) * 128;
  const bool row3 = !pix && row3_ && ntaps == 9 && (!acc4 || acc4_conv) && ring / row3_stage >= 3 && shifts[1] == shifts[0] + 1 &&
                    shifts[2] == shifts[1] + 1 && shifts[4] == shifts[3] + 1 && shifts[5] == shifts[4] + 1 &&
                    shifts[7] == shifts[6] + 1 && shifts[8] == shifts[7] + 1;
  // kRow3: steps per stage (gemm_tc.cuh).  A step is 12 MMAs = 3 bn cycles; one mbarrier hand-shake costs ~500,
  // so narrow layers take 3 (bn <= 64) or 2 (bn = 128) steps per stage while at least three stages still fit.
  // (Weights resident in shared memory for the whole launch were measured twice, alone and with this grouping:
  // slower -- the A ring loses the stages the weights take; profiles/r2_experiment_resident_weights.patch.)
  if (row3) {
    const int n3 = 3 * p.k_chunks;
    const int want = row3_group_ > 0 ? row3_group_ : std::min(3, cdiv(512, 3 * bn));
    int cand[4], nc = 0;   // step counts in order of preference (n3 = 3 k_chunks: 2 divides it only for even k_chunks)
    if (want == 2 && n3 % 2 != 0) cand[nc++] = 3;
    for (int g = want; g >= 1; --g)
      if (n3 % g == 0) cand[nc++] = g;
    group = 1;
    for (int i = 0; i < nc; ++i)
      if (ring / (cand[i] * row3_stage) >= 3) { group = cand[i]; break; }
  }
  p.group = group;
  const int stage_bytes = row3 ? group * row3_stage : group * pair_bytes;
  p.stages = std::min(kMaxStages, ring / stage_bytes);
  p.idesc = umma_idesc(ElemTraits<T>::kFmt, kTileM * ctas, bn);
  p.tmem_cols = tmem_cols_for(bn * n_acc / 2);
  p.geo = a.g;
  p.n_valid = w.N;
  const int smem = p.stages * stage_bytes + staging + 1024;
  const CUtensorMap& tout = pix ? out_act->tmap_samples : (p.tma_store ? out_act->tmap : a.tmap);
  const CUtensorMap& ta = pix ? a.tmap_samples : a.tmap;
  if (pair) {
    const int clusters = std::min(p.m_tiles * p.n_tiles, num_sms_ / 2);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(2 * clusters);
    cfg.blockDim = dim3(kGemmThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = use_pdl_ ? 2 : 1;
    if (acc4) cfg.blockDim = dim3(gemm_threads(4));
    if (row3 && acc4 && group == 3) DMB_CUDA(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<T, 2, 3, 4, true>, a.tmap_halo, w.tmap_for(bn / 2), tout, p));
    else if (row3 && acc4 && group == 2) DMB_CUDA(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<T, 2, 2, 4, true>, a.tmap_halo, w.tmap_for(bn / 2), tout, p));
    else if (row3 && acc4) DMB_CUDA(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<T, 2, 1, 4, true>, a.tmap_halo, w.tmap_for(bn / 2), tout, p));
    else if (row3 && group == 3) DMB_CUDA(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<T, 2, 3, 2, true>, a.tmap_halo, w.tmap_for(bn / 2), tout, p));
    else if (row3 && group == 2) DMB_CUDA(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<T, 2, 2, 2, true>, a.tmap_halo, w.tmap_for(bn / 2), tout, p));
    else if (row3) DMB_CUDA(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<T, 2, 1, 2, true>, a.tmap_halo, w.tmap_for(bn / 2), tout, p));
    else if (group == 1) DMB_CUDA(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<T, 2, 1>, ta, w.tmap_for(bn / 2), tout, p));
    else if (group == 2) DMB_CUDA(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<T, 2, 2>, ta, w.tmap_for(bn / 2), tout, p));
    else DMB_CUDA(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<T, 2, 4>, ta, w.tmap_for(bn / 2), tout, p));
  } else {
    const int grid = std::min(p.m_tiles * p.n_tiles, num_sms_);
    if (acc4 && row3) {
      constexpr int kT4 = gemm_threads(4);
      if (group == 3) launch(gemm_tc_kernel<T, 1, 3, 4, true>, grid, kT4, smem, s, a.tmap_halo, w.tmap_for(bn), tout, p);
      else if (group == 2) launch(gemm_tc_kernel<T, 1, 2, 4, true>, grid, kT4, smem, s, a.tmap_halo, w.tmap_for(bn), tout, p);
      else launch(gemm_tc_kernel<T, 1, 1, 4, true>, grid, kT4, smem, s, a.tmap_halo, w.tmap_for(bn), tout, p);
    } else if (acc4) {
      constexpr int kT4 = gemm_threads(4);
      if (group == 1) launch(gemm_tc_kernel<T, 1, 1, 4>, grid, kT4, smem, s, ta, w.tmap_for(bn), tout, p);
      else if (group == 2) launch(gemm_tc_kernel<T, 1, 2, 4>, grid, kT4, smem, s, ta, w.tmap_for(bn), tout, p);
      else launch(gemm_tc_kernel<T, 1, 4, 4>, grid, kT4, smem, s, ta, w.tmap_for(bn), tout, p);
    } else if (row3) {
      if (group == 3) launch(gemm_tc_kernel<T, 1, 3, 2, true>, grid, kGemmThreads, smem, s, a.tmap_halo, w.tmap_for(bn), tout, p);
      else if (group == 2) launch(gemm_tc_kernel<T, 1, 2, 2, true>, grid, kGemmThreads, smem, s, a.tmap_halo, w.tmap_for(bn), tout, p);
      else launch(gemm_tc_kernel<T, 1, 1, 2, true>, grid, kGemmThreads, smem, s, a.tmap_halo, w.tmap_for(bn), tout, p);
    } else
    if (group == 1) launch(gemm_tc_kernel<T, 1, 1>, grid, kGemmThreads, smem, s, ta, w.tmap_for(bn), tout, p);
    else if (group == 2) launch(gemm_tc_kernel<T, 1, 2>, grid, kGemmThreads, smem, s, ta, w.tmap_for(bn), tout, p);
    else launch(gemm_tc_kernel<T, 1, 4>, grid, kGemmThreads, smem, s, ta, w.tmap_for(bn), tout, p);
  }
  DMB_CHECK_LAUNCH();
  count();
}

template <typename T>
Op Engine<T>::op_conv(const Act<T>& in, const GemmW<T>& w, const Act<T>& out) {
  return [this, in, w, out](cudaStream_t s) {
    int shifts[9];
    for (int ky = 0; ky < 3; ++ky)
      for (int kx = 0; kx < 3; ++kx) shifts[ky * 3 + kx] = (ky - 1) * in.g.Wp + (kx - 1);
    GemmParams ep{};
    ep.epi = EPI_STORE;
    ep.out = out.p; ep.ldc = out.C;
    ep.bias = w.bias;
    ep.sample_stats = out.sstat;
    gemm(in, w, 0, shifts, 9, ep, s, &out);
  };
}

template <typename T>
Op Engine<T>::op_lin(const Act<T>& in, const GemmW<T>& w, const Act<T>& out, int act, const Act<T>* residual) {
  const T* res = residual ? residual->p : nullptr;
  return [this, in, w, out, act, res](cudaStream_t s) {
    int shift0 = 0;
    GemmParams ep{};
    ep.epi = EPI_STORE;
    ep.out = out.p; ep.ldc = out.C;
    ep.bias = w.bias; ep.act = act; ep.residual = res;
    gemm(in, w, 0, &shift0, 1, ep, s, &out);
  };
}

// ---------------------------------------------------------------------------
// U-Net plan: the launch list of one forward (models/unet_cond_geom.py:52-76) for a fixed
// effective batch.  doubled = the two CFG passes as one batch of 2B: samples [0,B) carry
// the null label, [B,2B) the class label (diff.py:147-150).
// ---------------------------------------------------------------------------
template <typename T>
void Engine<T>::build_unet_plan(Plan<T>& P) {
  Pool& pool = P.pool;
  const int B = P.B, n = P.doubled ? 2 * B : B;
  P.x_state = pool.alloc_n<float>(static_cast<size_t>(B) * kLatC * kLatH * kLatW);
  P.t_vec = pool.alloc_n<long long>(B);
  P.y = pool.alloc_n<long long>(B);
  P.vals = pool.alloc_n<float>(B * kCondDim);
  P.mask = pool.alloc_n<float>(B * kCondDim);
  P.dyn = pool.alloc_n<DynParams>(1);
  P.emb = pool.alloc_n<float>(static_cast<size_t>(n) * kEmbDim);
  P.embproj = pool.alloc_n<float>(static_cast<size_t>(n) * kProjCols);
  P.eps = pool.alloc_n<float>(static_cast<size_t>(B) * kLatC * kLatH * kLatW);
  P.geom = pool.alloc_n<float>(B * 12);
  auto& ops = P.ops;

  {  // GroupNorm statistics arena: two 64-bit fixed-point sums per sample for each of the 32 convs that
     // feed a GroupNorm(1, C); one memset at the head of every step clears all of them
    constexpr int kStatActs = 40;
    const size_t words = static_cast<size_t>(kStatActs) * 2 * n;
    long long* arena = pool.alloc_n<long long>(words);
    stat_arena_cur_ = arena;
    stat_arena_left_ = words;
    ops.tag("stat clear");
    ops.push_back([arena, words](cudaStream_t s) { DMB_CUDA(cudaMemsetAsync(arena, 0, words * sizeof(long long), s)); });
  }

  {  // embedding + six projections
    const EmbWeights ew = embw_;
    Plan<T>* pp = &P;
    const int doubled = P.doubled ? 1 : 0;
    ops.tag("emb");
    ops.push_back([this, ew, pp, n, B, doubled](cudaStream_t s) {
      // nothing before down1's embedding add reads the result: fork to the side stream
      DMB_CUDA(cudaEventRecord(ev_fork_, s));
      DMB_CUDA(cudaStreamWaitEvent(side_, ev_fork_, 0));
      emb_kernel<<<dim3(cdiv(B, kEmbSamples), kEmbSplit), 256, 0, side_>>>(ew, pp->t_vec, pp->y, pp->dyn->flags, pp->vals,
                                                                          pp->mask, B, doubled, pp->emb, pp->embproj);
      DMB_CHECK_LAUNCH();
      DMB_CUDA(cudaEventRecord(ev_join_, side_));
      count();
    });
  }

  bool emb_joined = false;
  // `ln_name` (GN_ADD_EMB only): the LayerNorm of the attention block that follows is applied in
  // the same kernel and the returned activation is LayerNorm(x); `pre_tap` then holds the
  // un-normalised x when the context keeps taps (DMB200_KEEP_TAPS=1, the parity tests).
  Act<T> pre_tap;
  bool have_pre_tap = false, last_gn_fused_ln = false;
  auto gn = [&](const Act<T>& raw, const std::string& gname, int mode, const Act<T>* residual, int emb_off,
                int n_out = 0, const std::string& ln_name = std::string()) {
    if (n_out == 0) n_out = raw.g.n;
    // In place where nothing else reads the un-normalised conv output (no debug taps, no fan-out of a shared
    // guidance prefix): a thread rewrites exactly the 16-byte vectors it has just read (pad rows are zeros on
    // both sides), so the pass touches half the cache lines and the next conv finds more of its input in L2.
    const bool inplace = gn_inplace_ && !keep_taps_ && n_out == raw.g.n;
    Act<T> out = inplace ? raw : new_act(pool, n_out, raw.g.H, raw.g.W, raw.C, false);
    // fused where the launch is latency-bound (small maps / small batch: one launch and one pass
    // less); on large activations the two streaming kernels are faster than the fused one, whose
    // per-token shuffles make it issue-bound (measured at batch 256: 57 vs 34 + 15 us at 28x28)
    const bool fuse_ln = !ln_name.empty() &&
                         static_cast<size_t>(n_out) * raw.g.H * raw.g.W * raw.C * sizeof(T) <= (4u << 20);
    const float* ln_w = fuse_ln ? f(ln_name + ".weight") : nullptr;
    const float* ln_b = fuse_ln ? f(ln_name + ".bias") : nullptr;
    T* pre = nullptr;
    have_pre_tap = false;
    last_gn_fused_ln = fuse_ln;
    if (fuse_ln && keep_taps_) {
      pre_tap = new_act(pool, n_out, raw.g.H, raw.g.W, raw.C, false);
      pre = pre_tap.p;
      have_pre_tap = true;
    }
    const float* gamma = f(gname + ".weight");
    const float* beta = f(gname + ".bias");
    const T* res = residual ? residual->p : nullptr;
    const float* embproj = P.embproj;
    const int total_vec = raw.g.H * raw.g.W * raw.C / 8;
    // enough CTAs to fill the chip at small batch, few and long CTAs at large batch (measured at batch 256 with
    // 1 / 2 / 3 / 6 / 13 CTAs per sample: step 2.445 / 2.433 / 2.455 / 2.510 / 2.594 ms -- the prologue of a CTA
    // costs more than the last partial wave)
    const int want = (6 * num_sms_ + n_out - 1) / n_out;
    int splits = std::max(1, std::min(std::min(want, 16), total_vec / 512));
    if (gn_splits_ > 0) splits = std::max(1, std::min(gn_splits_, total_vec / 256));   // experiment override
    if (mode == GN_ADD_EMB && !emb_joined) {   // first consumer of the embedding projections: join the side stream
      emb_joined = true;
      ops.tag("emb join");
      ops.push_back([this](cudaStream_t s) { DMB_CUDA(cudaStreamWaitEvent(s, ev_join_, 0)); });
    }
    ops.tag("gn1_apply " + std::to_string(raw.C) + " @" + std::to_string(raw.g.H) + " m" + std::to_string(mode), 0,
            static_cast<double>(n_out) * total_vec * 8 * sizeof(T) * (residual ? 3 : 2));
    ops.push_back([this, raw, out, gamma, beta, mode, res, embproj, emb_off, splits, fuse_ln, ln_w, ln_b, pre](cudaStream_t s) {
      const dim3 grid(splits, out.g.n);
      if (fuse_ln) {
        launch(gn1_apply_kernel<T, GN_ADD_EMB, true>, grid, 256, 0, s, raw.p, raw.sstat,
               gamma, beta, res, embproj, emb_off, out.g, raw.C, out.p, raw.g.n,
               1.0 / (static_cast<double>(raw.g.H) * raw.g.W * raw.C), ln_w, ln_b, pre);
        DMB_CHECK_LAUNCH();
        count();
        return;
      }
#define DMB_GN_ARGS raw.p, raw.sstat, gamma, beta, res, embproj, emb_off, out.g, raw.C, out.p, raw.g.n, 1.0 / (static_cast<double>(raw.g.H) * raw.g.W * raw.C), static_cast<const float*>(nullptr), static_cast<const float*>(nullptr), static_cast<T*>(nullptr)
      if (mode == GN_GELU) launch(gn1_apply_kernel<T, GN_GELU>, grid, 256, 0, s, DMB_GN_ARGS);
      else if (mode == GN_RES_GELU) launch(gn1_apply_kernel<T, GN_RES_GELU>, grid, 256, 0, s, DMB_GN_ARGS);
      else if (mode == GN_ADD_EMB) launch(gn1_apply_kernel<T, GN_ADD_EMB>, grid, 256, 0, s, DMB_GN_ARGS);
      else launch(gn1_apply_kernel<T, GN_NONE>, grid, 256, 0, s, DMB_GN_ARGS);
#undef DMB_GN_ARGS
      DMB_CHECK_LAUNCH();
      count();
    });
    return out;
  };
  auto conv = [&](const Act<T>& in, const std::string& wname) {
    const GemmW<T>& w = gw_.at(wname);
    Act<T> out = new_act(pool, in.g.n, in.g.H, in.g.W, w.N, true);
    ops.tag("conv3x3 " + std::to_string(w.K) + "->" + std::to_string(w.N) + " @" + std::to_string(in.g.H),
            2.0 * in.g.n * in.g.H * in.g.W * 9.0 * w.K * w.N);
    ops.push_back(op_conv(in, w, out));
    return out;
  };
  // ResBlock (models/unet_cond.py:10-30); last_mode is what follows the second GroupNorm.
  auto resblock = [&](const Act<T>& in, const std::string& p, bool residual, int last_mode, int emb_off,
                      int n_out = 0, const std::string& ln_name = std::string()) {
    Act<T> r1 = conv(in, p + ".double_conv.0.weight");
    Act<T> h = gn(r1, p + ".double_conv.1", GN_GELU, nullptr, 0);
    Act<T> r2 = conv(h, p + ".double_conv.3.weight");
    return gn(r2, p + ".double_conv.4", residual ? GN_RES_GELU : last_mode, residual ? &in : nullptr, emb_off, n_out,
              ln_name);
  };
  // x_is_normed: the producing GroupNorm kernel already applied this block's first LayerNorm
  auto attention = [&](const Act<T>& x, const std::string& p, bool x_is_normed) {
    const int C = x.C, H = x.g.H, W = x.g.W, nn = x.g.n;
    auto ln = [&](const Act<T>& in, const std::string& name) {
      Act<T> out = new_act(pool, nn, H, W, C, false);
      const float* wv = f(name + ".weight");
      const float* bv = f(name + ".bias");
      ops.tag("layernorm", 0, 2.0 * in.g.n * in.g.H * in.g.W * C * sizeof(T));
      ops.push_back([this, in, out, wv, bv, C](cudaStream_t s) {
        // C/8 lanes per token; each block walks rows with a grid stride
        const int tpb = 256 / (C / 8);
        const int grid = std::max(1, std::min(cdiv(in.g.rows_valid, tpb * 8), num_sms_ * 8));
        if (C == 64) launch(layernorm_kernel<T, 8>, grid, 256, 0, s, in.p, wv, bv, in.g, out.p);
        else if (C == 128) launch(layernorm_kernel<T, 16>, grid, 256, 0, s, in.p, wv, bv, in.g, out.p);
        else launch(layernorm_kernel<T, 32>, grid, 256, 0, s, in.p, wv, bv, in.g, out.p);
        DMB_CHECK_LAUNCH();
        count();
      });
      return out;
    };
    // LayerNorm + in-projection in one kernel (ln_qkv_kernel) where the LayerNorm is still its own pass
    bool lnqkv = false;
    if constexpr (std::is_same<T, __nv_bfloat16>::value)
      lnqkv = !x_is_normed && use_ln_qkv_ && use_fused_tail_ && (C == 64 || C == 128);
    // (ln_qkv_kernel normalises a tile in shared memory and stores it over the rows it loaded: x_ln takes x's
    // buffer -- nothing reads the un-normalised x after this block's LayerNorm -- unless debug taps keep x)
    Act<T> xl = x_is_normed ? x : (lnqkv ? (alias_acts_ && !keep_taps_ ? x : new_act(pool, nn, H, W, C, false)) : ln(x, p + ".ln"));
    Act<T> qkv = new_act(pool, nn, H, W, 3 * C, false);
    const double tok = static_cast<double>(nn) * H * W;
    auto lin_tag = [&](int K, int N) {
      ops.tag("linear " + std::to_string(K) + "->" + std::to_string(N) + " @" + std::to_string(H), 2.0 * tok * K * N);
    };
    if constexpr (std::is_same<T, __nv_bfloat16>::value) {
      if (lnqkv) {
        const GemmW<T>& wq = gw_.at(p + ".qkv");
        LnQkvParams lp{};
        lp.geo = x.g;
        lp.m_tiles = cdiv(x.g.rows_valid, kTileM);
        lp.ln_w = f(p + ".ln.weight");
        lp.ln_b = f(p + ".ln.bias");
        lp.bqkv = wq.bias;
        DMB_REQUIRE(lp.bqkv && wq.rows_per_tap == 3 * C, "in-projection: bias and an unpadded [3C, C] weight");
        ops.tag("ln_qkv " + std::to_string(C) + " @" + std::to_string(H), 2.0 * tok * C * 3 * C, 5.0 * tok * C * sizeof(T));
        ops.push_back([this, x, xl, qkv, wq, lp, C](cudaStream_t s) {
          const int grid = std::min(lp.m_tiles, num_sms_);
          if (C == 64)
            launch(ln_qkv_kernel<64, 4>, grid, kTailThreads, attn_tail_smem_bytes<64, 4>(), s, x.tmap, wq.tmap_for(64), xl.tmap, qkv.tmap, lp);
          else
            launch(ln_qkv_kernel<128, 2>, grid, kTailThreads, attn_tail_smem_bytes<128, 2>(), s, x.tmap, wq.tmap_for(128), xl.tmap, qkv.tmap, lp);
          DMB_CHECK_LAUNCH();
          count();
        });
      }
    }
    if (!lnqkv) {
      lin_tag(C, 3 * C);
      ops.push_back(op_lin(xl, gw_.at(p + ".qkv"), qkv, 0, nullptr));
    }
    Act<T> av = new_act(pool, nn, H, W, C, false);
    ops.tag("attention L=" + std::to_string(H * W) + " d=" + std::to_string(C / 4), 4.0 * nn * H * W * H * W * C);
    ops.push_back([this, qkv, av, C](cudaStream_t s) {
      const int L = qkv.g.H * qkv.g.W, D = C / 4;
      if constexpr (std::is_same<T, __nv_bfloat16>::value) {
        if (use_mma_attention_ && use_tc_attention_ && L >= attn_mma_min_l_) {
          // one CTA per (sample, head) with up to 7 warps, each owning 16-row query tiles; when the
          // batch is too small to fill the chip twice the tiles are spread over more CTAs
          const int ntiles = cdiv(L, 16);
          const int nw = std::min(7, ntiles);
          const int splits = std::max(1, std::min(cdiv(ntiles, nw), cdiv(2 * num_sms_, 4 * qkv.g.n)));
          const int tpc = cdiv(ntiles, splits);
          const dim3 grid(cdiv(ntiles, tpc), 4, qkv.g.n);
          const size_t smem = attention_mma_smem_bytes(L, D);
          const int threads = 32 * std::min(nw, tpc);
#define DMB_AM_LAUNCH(P, V)                                                                                         \
  do {                                                                                                              \
    if (D == 16) launch(attention_mma_kernel<16, P, V>, grid, threads, smem, s, qkv.p, qkv.g, C, av.p, tpc);       \
    else if (D == 32) launch(attention_mma_kernel<32, P, V>, grid, threads, smem, s, qkv.p, qkv.g, C, av.p, tpc);  \
    else launch(attention_mma_kernel<64, P, V>, grid, threads, smem, s, qkv.p, qkv.g, C, av.p, tpc);               \
  } while (0)
          if (attn_poly_ && attn_var_ == 5) DMB_AM_LAUNCH(3, 5);
          else if (attn_poly_) DMB_AM_LAUNCH(3, 1);
          else if (attn_var_ == 0) DMB_AM_LAUNCH(-1, 0);
          else if (attn_var_ == 5) DMB_AM_LAUNCH(-1, 5);
          else DMB_AM_LAUNCH(-1, 1);
#undef DMB_AM_LAUNCH
          DMB_CHECK_LAUNCH();
          count();
          return;
        }
        if (use_tc_attention_ && L > 16) {   // tiny maps (3x3): a 128-row MMA tile would be > 90 % padding
          // whole (sample, head) per CTA when there are enough of them to fill the chip twice
          // over; otherwise the query blocks are spread over more CTAs
          const int nqb = cdiv(L, kAttnQ);
          // (d_head 16 / 32 run two chains per CTA: at least two query blocks each)
          const int qsplit = std::max(1, std::min(D <= 32 ? cdiv(nqb, 2) : nqb, cdiv(4 * num_sms_, 4 * qkv.g.n)));
          const int qpc = cdiv(nqb, qsplit);
          const dim3 grid(cdiv(nqb, qpc), 4, qkv.g.n);
          static const char* pad_env = getenv("DMB200_ATTN_PAD_SMEM");   // experiment: force 1 CTA / SM
          const size_t smem = attention_tc_smem_bytes(L, D) + ((pad_env && L == 784) ? atoi(pad_env) * 1024 : 0);
          if (D == 16) launch(attention_tc_kernel<16>, grid, kAttnThreads, smem, s, qkv.p, qkv.g, C, av.p, qpc);
          else if (D == 32) launch(attention_tc_kernel<32>, grid, kAttnThreads, smem, s, qkv.p, qkv.g, C, av.p, qpc);
          else launch(attention_tc_kernel<64>, grid, kAttnThreads, smem, s, qkv.p, qkv.g, C, av.p, qpc);
          DMB_CHECK_LAUNCH();
          count();
          return;
        }
      }
      if constexpr (std::is_same<T, float>::value) {
        if (use_mma_attention_ && L > 16) {   // TF32 warp-level MMA attention (attention_mma.cuh)
          const int ntiles = cdiv(L, 16);
          const int nw = std::min(kAfMaxWarps, ntiles);
          const int splits = std::max(1, std::min(cdiv(ntiles, nw), cdiv(2 * num_sms_, 4 * qkv.g.n)));
          const int tpc = cdiv(ntiles, splits);
          const dim3 grid(cdiv(ntiles, tpc), 4, qkv.g.n);
          const size_t smem = attention_mma_f32_smem_bytes(L, D);
          const int threads = 32 * std::min(nw, tpc);
          if (D == 16) launch(attention_mma_f32_kernel<16>, grid, threads, smem, s, qkv.p, qkv.g, C, av.p, tpc);
          else if (D == 32) launch(attention_mma_f32_kernel<32>, grid, threads, smem, s, qkv.p, qkv.g, C, av.p, tpc);
          else launch(attention_mma_f32_kernel<64>, grid, threads, smem, s, qkv.p, qkv.g, C, av.p, tpc);
          DMB_CHECK_LAUNCH();
          count();
          return;
        }
      }
      const int threads = std::min(128, (L + 31) / 32 * 32);
      const dim3 grid(cdiv(L, threads), 4, qkv.g.n);
      const size_t smem = static_cast<size_t>(L) * D * 2 * sizeof(float);
      if (D == 16) launch(attention_kernel<T, 16>, grid, threads, smem, s, qkv.p, qkv.g, C, av.p);
      else if (D == 32) launch(attention_kernel<T, 32>, grid, threads, smem, s, qkv.p, qkv.g, C, av.p);
      else launch(attention_kernel<T, 64>, grid, threads, smem, s, qkv.p, qkv.g, C, av.p);
      DMB_CHECK_LAUNCH();
      count();
    });
    if constexpr (std::is_same<T, __nv_bfloat16>::value) {
      if (use_fused_tail_ && use_tail256_ && C == 256) {
        // the same chain with one weight matrix resident at a time (attn_tail256_kernel); the block's output
        // takes the attention output's buffer (a tile is stored over the rows it was loaded from)
        Act<T> out = alias_acts_ ? av : new_act(pool, nn, H, W, C, false);
        const GemmW<T>&wp = gw_.at(p + ".proj"), &w1 = gw_.at(p + ".ff1"), &w2 = gw_.at(p + ".ff2");
        TailParams tp{};
        tp.geo = av.g;
        tp.m_tiles = cdiv(av.g.rows_valid, kTileM);
        tp.bp = wp.bias; tp.b1 = w1.bias; tp.b2 = w2.bias;
        tp.ln_w = f(p + ".ff_self.0.weight");
        tp.ln_b = f(p + ".ff_self.0.bias");
        DMB_REQUIRE(tp.bp && tp.b1 && tp.b2, "attention-block linears carry biases");
        ops.tag("attn_tail 256 @" + std::to_string(H), 6.0 * tok * C * C, 3.0 * tok * C * sizeof(T));
        ops.push_back([this, av, xl, out, wp, w1, w2, tp](cudaStream_t s) {
          const int grid = std::min(tp.m_tiles, num_sms_);
          launch(attn_tail256_kernel, grid, kTail256Threads, attn_tail256_smem_bytes(), s, av.tmap, wp.tmap_for(256),
                 w1.tmap_for(256), w2.tmap_for(256), out.tmap, reinterpret_cast<const __nv_bfloat16*>(xl.p), tp);
          DMB_CHECK_LAUNCH();
          count();
        });
        return out;
      }
      if (use_fused_tail_ && (C == 64 || C == 128)) {
        // out_proj + residual, LayerNorm, ff1 + GELU, ff2 + residual on chip (attn_tail.cuh), stored over av
        Act<T> out = alias_acts_ ? av : new_act(pool, nn, H, W, C, false);
        const GemmW<T>&wp = gw_.at(p + ".proj"), &w1 = gw_.at(p + ".ff1"), &w2 = gw_.at(p + ".ff2");
        TailParams tp{};
        tp.geo = av.g;
        tp.m_tiles = cdiv(av.g.rows_valid, kTileM);
        tp.bp = wp.bias; tp.b1 = w1.bias; tp.b2 = w2.bias;
        tp.ln_w = f(p + ".ff_self.0.weight");
        tp.ln_b = f(p + ".ff_self.0.bias");
        DMB_REQUIRE(tp.bp && tp.b1 && tp.b2, "attention-block linears carry biases");
        ops.tag("attn_tail " + std::to_string(C) + " @" + std::to_string(H), 6.0 * tok * C * C, 3.0 * tok * C * sizeof(T));
        ops.push_back([this, av, xl, out, wp, w1, w2, tp, C](cudaStream_t s) {
          const int grid = std::min(tp.m_tiles, num_sms_);
          if (C == 64)
            launch(attn_tail_kernel<64, 4>, grid, kTailThreads, attn_tail_smem_bytes<64, 4>(), s, av.tmap, xl.tmap, wp.tmap_for(64),
                   w1.tmap_for(64), w2.tmap_for(64), out.tmap, tp);
          else
            launch(attn_tail_kernel<128, 2>, grid, kTailThreads, attn_tail_smem_bytes<128, 2>(), s, av.tmap, xl.tmap, wp.tmap_for(128),
                   w1.tmap_for(128), w2.tmap_for(128), out.tmap, tp);
          DMB_CHECK_LAUNCH();
          count();
        });
        return out;
      }
    }
    Act<T> o = new_act(pool, nn, H, W, C, false);
    lin_tag(C, C);
    ops.push_back(op_lin(av, gw_.at(p + ".proj"), o, 0, &xl));
    Act<T> h = ln(o, p + ".ff_self.0");
    Act<T> f1 = new_act(pool, nn, H, W, C, false);
    lin_tag(C, C);
    ops.push_back(op_lin(h, gw_.at(p + ".ff1"), f1, 1, nullptr));
    Act<T> out = new_act(pool, nn, H, W, C, false);
    lin_tag(C, C);
    ops.push_back(op_lin(f1, gw_.at(p + ".ff2"), out, 0, &o));
    return out;
  };
  auto down = [&](const Act<T>& x, const std::string& p, int emb_off, const std::string& ln_name, int n_out = 0) {
    Act<T> pooled = new_act(pool, x.g.n, x.g.H / 2, x.g.W / 2, x.C, false);
    ops.tag("maxpool", 0, 1.25 * pooled.g.n * x.g.H * x.g.W * x.C * sizeof(T));
    ops.push_back([this, x, pooled](cudaStream_t s) {
      const int per_sample = pooled.g.H * pooled.g.W * (x.C / 8);
      launch(maxpool2_kernel<T>, dim3(cdiv(per_sample, 256), pooled.g.n), 256, 0, s, x.p, x.g, pooled.g, x.C, pooled.p);
      DMB_CHECK_LAUNCH();
      count();
    });
    Act<T> a = resblock(pooled, p + ".maxpool_conv.1", true, GN_NONE, 0);
    return resblock(a, p + ".maxpool_conv.2", false, GN_ADD_EMB, emb_off, n_out, ln_name);
  };
  auto up = [&](const Act<T>& x, const Act<T>& skip, const std::string& p, int emb_off, const std::string& ln_name) {
    Act<T> cat = new_act(pool, x.g.n, skip.g.H, skip.g.W, skip.C + x.C, false);
    ops.tag("upcat", 0, 2.0 * cat.g.n * cat.g.H * cat.g.W * cat.C * sizeof(T));
    ops.push_back([this, x, skip, cat](cudaStream_t s) {
      // (vectors per sample of each part / kUpcatU, rounded up to whole pixels)
      auto quarter_of = [&](int c) {
        const int vpr = c / 8;
        return static_cast<int>(cdiv(cdiv(static_cast<long long>(cat.g.H) * cat.g.W * vpr, kUpcatU), vpr) * vpr);
      };
      const int qa = quarter_of(skip.C), qb = quarter_of(x.C);
      const int blocks_a = cdiv(qa, 256), blocks_b = cdiv(qb, 256);
      launch(upcat_kernel<T>, dim3(blocks_a + blocks_b, cat.g.n), 256, 0, s, skip.p, skip.C, skip.g.n, x.p, x.g, x.C, cat.g,
             cat.p, blocks_a, qa, qb);
      DMB_CHECK_LAUNCH();
      count();
    });
    Act<T> a = resblock(cat, p + ".conv.0", true, GN_NONE, 0);
    return resblock(a, p + ".conv.1", false, GN_ADD_EMB, emb_off, 0, ln_name);
  };

  // Classifier-free guidance runs the conditional and the unconditional sample of a pair
  // through the same weights from the same x_t; they differ only through the embedding, which
  // first enters at the end of down1 (unet_cond.py:71-74: x = maxpool_conv(x); x + emb).  inc and
  // down1's convolutions are therefore evaluated ONCE per pair (n0 = B samples) and fan out to
  // 2B at down1's embedding add; up3 reads the shared inc output as its skip tensor.
  const int n0 = B;
  // inc: the 4->64 conv runs on CUDA cores straight from the fp32 latent.
  Act<T> r0 = new_act(pool, n0, kLatH, kLatW, 64, true);
  {
    const float* w0 = f("inc.double_conv.0.weight");
    Plan<T>* pp = &P;
    ops.tag("conv_in 4->64 @28", 2.0 * n0 * kLatH * kLatW * 36.0 * 64);
    ops.push_back([this, r0, w0, pp, B](cudaStream_t s) {
      if constexpr (std::is_same<T, __nv_bfloat16>::value)
        launch(conv_in_mma_kernel, dim3(r0.g.n, 1, conv_in_z_ ? conv_in_z_ : (r0.g.n >= 64 ? 1 : 7)), kConvInThreads, 0, s, pp->x_state, B, w0, nullptr, 1.0f, r0.g, 64,
                                                                      r0.p, r0.sstat);
      else
        conv_in_kernel<T><<<dim3(cdiv(kLatH * kLatW, 128), r0.g.n, 1), 128, 0, s>>>(pp->x_state, B, w0, nullptr, 1.0f,
                                                                                    r0.g, 64, r0.p, r0.sstat);
      DMB_CHECK_LAUNCH();
      count();
    });
  }
  Act<T> h0 = gn(r0, "inc.double_conv.1", GN_GELU, nullptr, 0);
  Act<T> r1 = conv(h0, "inc.double_conv.3.weight");
  Act<T> x1 = gn(r1, "inc.double_conv.4", GN_NONE, nullptr, 0);
  P.taps["inc"] = x1;
  // each down / up stage ends in GroupNorm + embedding add and feeds an attention block whose
  // first LayerNorm is applied by the same kernel (fuse_ln1_)
  auto lnn = [&](const char* sa) { return fuse_ln1_ ? std::string(sa) + ".ln" : std::string(); };
  bool normed = false;   // did the stage's last GroupNorm kernel apply the LayerNorm?
  auto stage_tap = [&](const char* name, const Act<T>& stage_out) {
    normed = last_gn_fused_ln;
    if (!normed) P.taps[name] = stage_out;
    else if (have_pre_tap) P.taps[name] = pre_tap;
  };
  Act<T> x2 = down(x1, "down1", 0, lnn("sa1"), n);   stage_tap("down1", x2);
  x2 = attention(x2, "sa1", normed);                 P.taps["sa1"] = x2;
  Act<T> x3 = down(x2, "down2", 128, lnn("sa2"));    stage_tap("down2", x3);
  x3 = attention(x3, "sa2", normed);                 P.taps["sa2"] = x3;
  Act<T> x4 = down(x3, "down3", 384, lnn("sa3"));    stage_tap("down3", x4);
  x4 = attention(x4, "sa3", normed);                 P.taps["sa3"] = x4;
  x4 = resblock(x4, "bot1", false, GN_NONE, 0);  P.taps["bot1"] = x4;
  x4 = resblock(x4, "bot2", false, GN_NONE, 0);  P.taps["bot2"] = x4;
  x4 = resblock(x4, "bot3", false, GN_NONE, 0);  P.taps["bot3"] = x4;
  Act<T> u = up(x4, x3, "up1", 640, lnn("sa4"));     stage_tap("up1", u);
  u = attention(u, "sa4", normed);                   P.taps["sa4"] = u;
  u = up(u, x2, "up2", 768, lnn("sa5"));             stage_tap("up2", u);
  u = attention(u, "sa5", normed);                   P.taps["sa5"] = u;
  u = up(u, x1, "up3", 832, lnn("sa6"));             stage_tap("up3", u);
  u = attention(u, "sa6", normed);                   P.taps["sa6"] = u;
  P.feat = u;
}

}  // namespace dmb

namespace dmb {

template <typename T>
Plan<T>& Engine<T>::get_plan(int B, bool doubled) {
  if (!unet_loaded_) fail(5, "U-Net weights have not been loaded (dmb_load_unet_weights)");
  DMB_REQUIRE(B >= 1, "batch must be >= 1");
  auto key = std::make_pair(B, doubled);
  auto it = plans_.find(key);
  if (it == plans_.end()) {
    // Activations of a plan are private buffers; keep at most two plans alive.
    while (plans_.size() >= 2) {
      DMB_CUDA(cudaDeviceSynchronize());
      auto victim = plans_.begin();
      if (victim->second.get() == last_plan_ && plans_.size() > 1) ++victim;
      if (victim->second.get() == last_plan_) last_plan_ = nullptr;
      plans_.erase(victim);
    }
    std::unique_ptr<Plan<T>> p(new Plan<T>());
    p->B = B;
    p->doubled = doubled;
    build_unet_plan(*p);
    DMB_CUDA(cudaDeviceSynchronize());  // allocation memsets are done before first use
    it = plans_.emplace(key, std::move(p)).first;
  }
  last_plan_ = it->second.get();
  return *it->second;
}

template <typename T>
void Engine<T>::bind_inputs(Plan<T>& p, const float* x, const long long* t, const long long* y, const float* vals,
                            const float* mask, int null_label, cudaStream_t s) {
  const size_t lat = static_cast<size_t>(p.B) * kLatC * kLatH * kLatW * sizeof(float);
  if (x) DMB_CUDA(cudaMemcpyAsync(p.x_state, x, lat, cudaMemcpyDefault, s));
  if (t) DMB_CUDA(cudaMemcpyAsync(p.t_vec, t, p.B * sizeof(long long), cudaMemcpyDefault, s));
  DMB_REQUIRE(y != nullptr, "class labels y are required");
  DMB_CUDA(cudaMemcpyAsync(p.y, y, p.B * sizeof(long long), cudaMemcpyDefault, s));
  const bool has_cond = vals && mask;
  if (has_cond) {
    DMB_CUDA(cudaMemcpyAsync(p.vals, vals, p.B * kCondDim * sizeof(float), cudaMemcpyDefault, s));
    DMB_CUDA(cudaMemcpyAsync(p.mask, mask, p.B * kCondDim * sizeof(float), cudaMemcpyDefault, s));
  }
  const int flags[2] = {null_label, has_cond ? 1 : 0};
  // pageable source: the runtime stages it before returning, so the stack buffer is safe
  DMB_CUDA(cudaMemcpyAsync(p.dyn->flags, flags, sizeof(flags), cudaMemcpyHostToDevice, s));
}

template <typename T>
void Engine<T>::unet_forward(const float* x, const long long* t, const long long* y, const float* vals,
                             const float* mask, int B, float* eps, float* geom, cudaStream_t s) {
  DMB_CUDA(cudaSetDevice(device_));
  Plan<T>& p = get_plan(B, false);
  bind_inputs(p, x, t, y, vals, mask, 0, s);
  run_ops(p, s);
  const Act<T> feat = p.feat;
  if (eps) {
    out_conv_kernel<T><<<dim3(cdiv(kLatH * kLatW, 128), B), 128, 0, s>>>(feat.p, feat.g, f("out.weight"), f("out.bias"), eps);
    DMB_CHECK_LAUNCH();
    count();
  }
  if (geom) {
    geom_head_kernel<T><<<B, 256, 0, s>>>(feat.p, feat.g, geom_w1t_, f("geom_head.mlp.0.bias"),
                                          f("geom_head.mlp.2.weight"), f("geom_head.mlp.2.bias"), 12, geom);
    DMB_CHECK_LAUNCH();
    count();
  }
}

template <typename T>
void Engine<T>::cfg_step(const float* x_in, const long long* t, const long long* y, int null_label,
                         const float* vals, const float* mask, float guidance, const float* coef, int T_steps,
                         const float* noise, unsigned long long seed, unsigned long long sample_offset,
                         float* x_out, float* eps_out, int B, cudaStream_t s) {
  DMB_CUDA(cudaSetDevice(device_));
  Plan<T>& p = get_plan(B, true);
  bind_inputs(p, x_in, t, y, vals, mask, null_label, s);
  StepParams sp{};
  sp.x_in = p.x_state; sp.x_out = x_out; sp.t_vec = p.t_vec; sp.coef = coef; sp.T = T_steps; sp.slot_t0 = 0;
  sp.t_stride = 1;
  sp.guidance = guidance; sp.noise = noise; sp.noise_step_stride = 0; sp.seed = seed;
  sp.sample_offset = sample_offset; sp.eps_out = eps_out; sp.x_dump = nullptr; sp.dump_step_stride = 0;
  sp.w_out = f("out.weight"); sp.b_out = f("out.bias");
  DMB_CUDA(cudaMemcpyAsync(&p.dyn->step, &sp, sizeof(sp), cudaMemcpyHostToDevice, s));
  // One eager step = ~100 launches; generate_steps.py calls this once per timestep from Python
  // (generate_steps.py:180-189), so outside the per-kernel timing mode the step runs as ONE graph
  // launch (same kernels, same order: everything that varies is read from device memory).
  if (!timing_ && eager_graph_) {
    if (!p.exec_eager) {
      cudaStream_t cs;
      DMB_CUDA(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
      cudaGraph_t graph = nullptr;
      const long long before = launches_;
      try {
        DMB_CUDA(cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal));
        run_ops(p, cs);
        launch(cfg_step_kernel<T>, dim3(cdiv(kLatH * kLatW, 128), B), 128, 0, cs, p.feat.p, p.feat.g, B, &p.dyn->step);
        DMB_CHECK_LAUNCH();
        count();
        DMB_CUDA(cudaStreamEndCapture(cs, &graph));
        DMB_CUDA(cudaGraphInstantiate(&p.exec_eager, graph, 0));
      } catch (...) {
        launches_ = before;
        if (graph) cudaGraphDestroy(graph);
        cudaStreamDestroy(cs);
        throw;
      }
      p.launches_eager = static_cast<int>(launches_ - before);
      launches_ = before;
      cudaGraphDestroy(graph);
      cudaStreamDestroy(cs);
    }
    DMB_CUDA(cudaGraphLaunch(p.exec_eager, s));
    launches_ += p.launches_eager;
    return;
  }
  run_ops(p, s);
  OpList tail;
  tail.tag("cfg_step (out conv + CFG + DDPM)", 0, static_cast<double>(B) * kLatH * kLatW * (2 * 64 * sizeof(T) + 4 * 4 * 3));
  Plan<T>* pp = &p;
  tail.push_back([this, pp, B](cudaStream_t cs) {
    launch(cfg_step_kernel<T>, dim3(cdiv(kLatH * kLatW, 128), B), 128, 0, cs, pp->feat.p, pp->feat.g, B, &pp->dyn->step);
    DMB_CHECK_LAUNCH();
    count();
  });
  run_list(tail, s);
}

// The T-step loop (diff.py:332-344) as CUDA-graph replays: one graph holds `chunk` whole
// steps (embedding ... U-Net ... fused CFG/DDPM update, t -= 1); everything that varies
// between calls is read from device memory, so the graph is captured once per batch size.
template <typename T>
void Engine<T>::sample_loop(float* x, const long long* y, int null_label, const float* vals, const float* mask,
                            float guidance, const float* coef, int T_steps, int t_start, int t_stride, int n_steps,
                            const float* noise, unsigned long long seed, unsigned long long sample_offset,
                            float* x_dump, float* eps_dump, int B, cudaStream_t s) {
  DMB_CUDA(cudaSetDevice(device_));
  DMB_REQUIRE(t_start >= 1 && t_start <= T_steps, "t_start must be in [1, T]");
  DMB_REQUIRE(t_stride >= 1, "t_stride must be >= 1");
  DMB_REQUIRE(n_steps >= 1 && static_cast<long long>(n_steps - 1) * t_stride < t_start,
              "the visited timesteps t_start - k * t_stride must stay >= 1");
  Plan<T>& p = get_plan(B, true);
  bind_inputs(p, x, nullptr, y, vals, mask, null_label, s);
  fill_t_kernel<<<cdiv(B, 256), 256, 0, s>>>(p.t_vec, B, t_start);
  DMB_CHECK_LAUNCH();
  const long long lat = static_cast<long long>(B) * kLatC * kLatH * kLatW;
  StepParams sp{};
  sp.x_in = p.x_state; sp.x_out = p.x_state; sp.t_vec = p.t_vec; sp.coef = coef; sp.T = T_steps;
  sp.slot_t0 = t_start; sp.t_stride = t_stride; sp.guidance = guidance; sp.noise = noise; sp.noise_step_stride = lat; sp.seed = seed;
  sp.sample_offset = sample_offset; sp.eps_out = eps_dump; sp.x_dump = x_dump; sp.dump_step_stride = lat;
  sp.w_out = f("out.weight"); sp.b_out = f("out.bias");
  DMB_CUDA(cudaMemcpyAsync(&p.dyn->step, &sp, sizeof(sp), cudaMemcpyHostToDevice, s));

  auto one_step = [&](cudaStream_t cs) {
    run_ops(p, cs);
    launch(cfg_step_kernel<T>, dim3(cdiv(kLatH * kLatW, 128), B), 128, 0, cs, p.feat.p, p.feat.g, B, &p.dyn->step);
    DMB_CHECK_LAUNCH();
    launch(advance_t_kernel, cdiv(B, 256), 256, 0, cs, p.t_vec, B, &p.dyn->step);
    DMB_CHECK_LAUNCH();
    count(2);
  };
  auto capture = [&](int steps) {
    cudaStream_t cs;
    DMB_CUDA(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    const long long before = launches_;  // kernels recorded into the graph are counted per replay
    try {
      DMB_CUDA(cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal));
      for (int i = 0; i < steps; ++i) one_step(cs);
      DMB_CUDA(cudaStreamEndCapture(cs, &graph));
      DMB_CUDA(cudaGraphInstantiate(&exec, graph, 0));
    } catch (...) {
      launches_ = before;
      if (graph) cudaGraphDestroy(graph);
      cudaStreamDestroy(cs);
      throw;
    }
    p.launches_per_step = static_cast<int>((launches_ - before) / steps);
    launches_ = before;
    cudaGraphDestroy(graph);
    cudaStreamDestroy(cs);
    return exec;
  };
  const int chunk = 10;
  if (!p.exec_one) p.exec_one = capture(1);
  if (n_steps >= chunk && !p.exec_chunk) {
    p.exec_chunk = capture(chunk);
    p.chunk = chunk;
  }
  int left = n_steps;
  while (left >= chunk && p.exec_chunk) {
    DMB_CUDA(cudaGraphLaunch(p.exec_chunk, s));
    launches_ += static_cast<long long>(chunk) * p.launches_per_step;
    left -= chunk;
  }
  while (left > 0) {
    DMB_CUDA(cudaGraphLaunch(p.exec_one, s));
    launches_ += p.launches_per_step;
    --left;
  }
  DMB_CUDA(cudaMemcpyAsync(x, p.x_state, lat * sizeof(float), cudaMemcpyDefault, s));
}

}  // namespace dmb

namespace dmb {

// ---------------------------------------------------------------------------
// VAE decoder plan (models/vae.py:35-49, 64-69) for a chunk of n samples.
// ---------------------------------------------------------------------------
template <typename T>
void Engine<T>::build_vae_plan(VaePlan& P, int n) {
  Pool& pool = P.pool;
  P.B = n;
  auto& ops = P.ops;
  VaePlan* pp = &P;

  // planar: `raw` is the phase-planar output of a ConvTranspose (convT below): four phase slots per
  // pixel of the input map; the normalised activation is written at 2H x 2W
  auto gn8 = [&](const Act<T>& raw, const std::string& name, bool planar = false) {
    Act<T> out = planar ? new_act(pool, raw.g.n, raw.g.H * 2, raw.g.W * 2, raw.C / 4, false)
                        : new_act(pool, raw.g.n, raw.g.H, raw.g.W, raw.C, false);
    const int total_vec = raw.g.H * raw.g.W * raw.C / 8;
    const int chunks = std::max(1, std::min(32, total_vec / 4096));
    double2* partial = pool.alloc_n<double2>(static_cast<size_t>(raw.g.n) * chunks * 8);
    const float* gamma = f(name + ".weight");
    const float* beta = f(name + ".bias");
    ops.tag("vae gn8+gelu", 0, 3.0 * raw.g.n * total_vec * 8 * sizeof(T));
    ops.push_back([this, raw, out, chunks, partial, gamma, beta, planar](cudaStream_t s) {
      gn8_stats_kernel<T><<<dim3(chunks, raw.g.n), 256, 0, s>>>(raw.p, raw.g, raw.C, partial, planar ? 4 : 1);
      DMB_CHECK_LAUNCH();
      if (planar)
        gn8_apply_gelu_kernel<T, true><<<dim3(chunks, raw.g.n), 256, 0, s>>>(raw.p, partial, chunks, gamma, beta, raw.g,
                                                                            raw.C, out.g, out.p);
      else
        gn8_apply_gelu_kernel<T, false><<<dim3(chunks, raw.g.n), 256, 0, s>>>(raw.p, partial, chunks, gamma, beta, raw.g,
                                                                             raw.C, out.g, out.p);
      DMB_CHECK_LAUNCH();
      count(2);
    });
    return out;
  };
  auto conv = [&](const Act<T>& in, const std::string& wname) {
    const GemmW<T>& w = gw_.at(wname);
    Act<T> out = new_act(pool, in.g.n, in.g.H, in.g.W, w.N, false);
    ops.tag("vae conv3x3 " + std::to_string(w.K) + "->" + std::to_string(w.N) + " @" + std::to_string(in.g.H),
            2.0 * in.g.n * in.g.H * in.g.W * 9.0 * w.K * w.N);
    ops.push_back(op_conv(in, w, out));
    return out;
  };
  // ConvTranspose2d(4, 2, 1) = four phase GEMMs of 2 x 2 taps.  Planar form (default): all four leave
  // through the TMA-store epilogue into one matrix over the INPUT map's pixels, phase (2a + b) at columns
  // (2a + b) C .. + C; the GroupNorm pass that follows moves them to their 2H x 2W positions
  // (kernels.cuh, gn8_apply_gelu_d2s_kernel).  DMB200_CONVT_PLANAR=0: every phase scatters its rows
  // of the 2H x 2W map from registers.
  const bool planar = convt_planar_;
  auto convT = [&](const Act<T>& in, const std::string& wname) {
    const GemmW<T> w = gw_.at(wname);
    Act<T> out = planar ? new_act(pool, in.g.n, in.g.H, in.g.W, 4 * w.N, false)
                        : new_act(pool, in.g.n, in.g.H * 2, in.g.W * 2, w.N, false);
    ops.tag("vae convT4x4 " + std::to_string(w.K) + "->" + std::to_string(w.N) + " @" + std::to_string(in.g.H),
            2.0 * in.g.n * in.g.H * in.g.W * 16.0 * w.K * w.N);
    ops.push_back([this, in, w, out, planar](cudaStream_t s) {
      for (int a = 0; a < 2; ++a)
        for (int b = 0; b < 2; ++b) {
          int shifts[4];
          for (int u = 0; u < 2; ++u)
            for (int v = 0; v < 2; ++v) {
              const int dy = (a == 0) ? (u == 0 ? 0 : -1) : (u == 0 ? 1 : 0);
              const int dx = (b == 0) ? (v == 0 ? 0 : -1) : (v == 0 ? 1 : 0);
              shifts[u * 2 + v] = dy * in.g.Wp + dx;
            }
          GemmParams ep{};
          ep.epi = EPI_STORE;
          ep.out = out.p; ep.ldc = out.C; ep.bias = w.bias;
          if (planar) {
            ep.out_col0 = (a * 2 + b) * w.N;
            gemm(in, w, a * 2 + b, shifts, 4, ep, s, &out);
          } else {
            ep.map_mode = 1; ep.geo_out = out.g; ep.up_a = a; ep.up_b = b;
            gemm(in, w, a * 2 + b, shifts, 4, ep, s);
          }
        }
    });
    return out;
  };

  Act<T> r0 = new_act(pool, n, kLatH, kLatW, 256, false);
  {
    const float* w0 = f("vae.dec.0.weight");
    const float* b0 = f("vae.dec.0.bias");
    ops.tag("vae conv_in 4->256 @28", 2.0 * n * kLatH * kLatW * 36.0 * 256);
    ops.push_back([this, r0, w0, b0, pp, n](cudaStream_t s) {
      if constexpr (std::is_same<T, __nv_bfloat16>::value)
        launch(conv_in_mma_kernel, dim3(n, 4), kConvInThreads, 0, s, pp->z, n, w0, b0, 0.18215f, r0.g, 256, r0.p, nullptr);
      else
        conv_in_kernel<T><<<dim3(cdiv(kLatH * kLatW, 128), n, 4), 128, 0, s>>>(pp->z, n, w0, b0, 0.18215f, r0.g, 256,
                                                                               r0.p, nullptr);
      DMB_CHECK_LAUNCH();
      count();
    });
  }
  Act<T> a = gn8(r0, "vae.dec.1");
  a = gn8(convT(a, "vae.dec.3"), "vae.dec.4", planar);
  a = gn8(conv(a, "vae.dec.6"), "vae.dec.7");
  a = gn8(convT(a, "vae.dec.9"), "vae.dec.10", planar);
  a = gn8(conv(a, "vae.dec.12"), "vae.dec.13");
  a = gn8(convT(a, "vae.dec.15"), "vae.dec.16", planar);
  {
    const GemmW<T> w = gw_.at("vae.dec.18");
    ops.tag("vae conv3x3 64->3 @224 +sigmoid", 2.0 * n * 224.0 * 224 * 9 * 64 * 3);
    ops.push_back([this, a, w, pp](cudaStream_t s) {
      int shifts[9];
      for (int ky = 0; ky < 3; ++ky)
        for (int kx = 0; kx < 3; ++kx) shifts[ky * 3 + kx] = (ky - 1) * a.g.Wp + (kx - 1);
      GemmParams ep{};
      ep.epi = EPI_IMAGE;
      ep.bias = w.bias;
      ep.img_f32 = pp->img_f32;
      ep.img_u8 = pp->img_u8;
      gemm(a, w, 0, shifts, 9, ep, s);
    });
  }
}

template <typename T>
void Engine<T>::vae_decode(const float* z, int B, float* img_f32, uint8_t* img_u8, cudaStream_t s) {
  DMB_CUDA(cudaSetDevice(device_));
  if (!vae_loaded_) fail(5, "VAE decoder weights have not been loaded (dmb_load_vae_dec_weights)");
  DMB_REQUIRE(B >= 1, "batch must be >= 1");
  const int kChunk = 32;
  const size_t lat = static_cast<size_t>(kLatC) * kLatH * kLatW;
  const size_t img = static_cast<size_t>(3) * 224 * 224;
  for (int b0 = 0; b0 < B; b0 += kChunk) {
    const int n = std::min(kChunk, B - b0);
    auto found = vae_plans_.find(n);
    if (found == vae_plans_.end()) {
      DMB_CUDA(cudaStreamSynchronize(s));
      for (auto it = vae_plans_.begin(); static_cast<int>(vae_plans_.size()) >= kMaxVaePlans && it != vae_plans_.end();)
        it = it->first != kChunk ? vae_plans_.erase(it) : std::next(it);
      found = vae_plans_.emplace(n, std::unique_ptr<VaePlan>(new VaePlan())).first;
      build_vae_plan(*found->second, n);
      DMB_CUDA(cudaDeviceSynchronize());
    }
    vae_plan_ = found->second.get();
    VaePlan& p = *vae_plan_;
    p.z = const_cast<float*>(z) + b0 * lat;
    p.img_f32 = img_f32 ? img_f32 + b0 * img : nullptr;
    p.img_u8 = img_u8 ? img_u8 + b0 * img : nullptr;
    run_list(p.ops, s);
  }
}

// ---------------------------------------------------------------------------
// VAE encoder (models/vae.py:16-31, 51-62).  Weights arrive under their
// state_dict names ("enc.0.weight", ..., "to_mu.weight", "to_logvar.bias") and are kept under
// the prefix "venc." so that reloading the decoder does not drop them.
// ---------------------------------------------------------------------------
template <typename T>
void Engine<T>::load_vae_enc(const TensorMapIn& w) {
  DMB_CUDA(cudaSetDevice(device_));
  DMB_CUDA(cudaDeviceSynchronize());
  vae_enc_plans_.clear();
  vae_enc_plan_ = nullptr;
  for (auto it = blob_.begin(); it != blob_.end();) {
    if (it->first.rfind("venc.", 0) == 0) it = blob_.erase(it); else ++it;
  }
  for (auto it = gw_.begin(); it != gw_.end();) {
    if (it->first.rfind("venc.", 0) == 0) it = gw_.erase(it); else ++it;
  }
  wpool_vae_enc_.release();
  vae_enc_loaded_ = false;
  upload(w, "venc.", wpool_vae_enc_);
  Pool& pool = wpool_vae_enc_;
  DMB_REQUIRE(numel("venc.enc.0.weight") == 64 * 3 * 9, "enc.0 must be conv 3->64");
  DMB_REQUIRE(numel("venc.to_mu.weight") == 4 * 256 && numel("venc.to_logvar.weight") == 4 * 256,
              "to_mu / to_logvar must be 1x1 convs 256->4");
  gw_["venc.enc.3"] = make_conv4s2_w("venc.enc.3.weight", "venc.enc.3.bias", 64, 64, pool);
  gw_["venc.enc.6"] = make_conv3x3_w("venc.enc.6.weight", 128, 64, pool);
  gw_["venc.enc.6"].bias = f("venc.enc.6.bias");
  gw_["venc.enc.9"] = make_conv4s2_w("venc.enc.9.weight", "venc.enc.9.bias", 128, 128, pool);
  gw_["venc.enc.12"] = make_conv3x3_w("venc.enc.12.weight", 256, 128, pool);
  gw_["venc.enc.12"].bias = f("venc.enc.12.bias");
  gw_["venc.enc.15"] = make_conv4s2_w("venc.enc.15.weight", "venc.enc.15.bias", 256, 256, pool);
  DMB_CUDA(cudaDeviceSynchronize());
  vae_enc_loaded_ = true;
}

template <typename T>
void Engine<T>::build_vae_enc_plan(VaeEncPlan& P, int n) {
  Pool& pool = P.pool;
  P.B = n;
  auto& ops = P.ops;
  VaeEncPlan* pp = &P;

  // GroupNorm(8)+GELU; s2d = written as the offset space-to-depth input of the strided conv
  auto gn8 = [&](const Act<T>& raw, const std::string& name, bool s2d) {
    Act<T> out = s2d ? new_act(pool, raw.g.n, raw.g.H / 2, raw.g.W / 2, 4 * raw.C, false)
                     : new_act(pool, raw.g.n, raw.g.H, raw.g.W, raw.C, false);
    const int total_vec = raw.g.H * raw.g.W * raw.C / 8;
    const int chunks = std::max(1, std::min(96, total_vec / 4096));
    double2* partial = pool.alloc_n<double2>(static_cast<size_t>(raw.g.n) * chunks * 8);
    const float* gamma = f(name + ".weight");
    const float* beta = f(name + ".bias");
    ops.tag(s2d ? "vae-enc gn8+gelu (space-to-depth)" : "vae-enc gn8+gelu", 0,
            3.0 * raw.g.n * total_vec * 8 * sizeof(T));
    ops.push_back([this, raw, out, chunks, partial, gamma, beta, s2d](cudaStream_t s) {
      gn8_stats_kernel<T><<<dim3(chunks, raw.g.n), 256, 0, s>>>(raw.p, raw.g, raw.C, partial);
      DMB_CHECK_LAUNCH();
      if (s2d)
        gn8_apply_gelu_s2d_kernel<T><<<dim3(chunks, raw.g.n), 256, 0, s>>>(raw.p, partial, chunks, gamma, beta,
                                                                          raw.g, raw.C, out.g, out.p);
      else
        gn8_apply_gelu_kernel<T, false><<<dim3(chunks, raw.g.n), 256, 0, s>>>(raw.p, partial, chunks, gamma, beta, raw.g,
                                                                             raw.C, out.g, out.p);
      DMB_CHECK_LAUNCH();
      count(2);
    });
    return out;
  };
  auto conv = [&](const Act<T>& in, const std::string& wname) {
    const GemmW<T>& w = gw_.at(wname);
    Act<T> out = new_act(pool, in.g.n, in.g.H, in.g.W, w.N, false);
    ops.tag("vae-enc conv3x3 " + std::to_string(w.K) + "->" + std::to_string(w.N) + " @" + std::to_string(in.g.H),
            2.0 * in.g.n * in.g.H * in.g.W * 9.0 * w.K * w.N);
    ops.push_back(op_conv(in, w, out));
    return out;
  };
  // 4x4 / stride 2 / pad 1 as a 2x2-tap GEMM over the space-to-depth input (K = 4 C_in per tap)
  auto conv4s2 = [&](const Act<T>& in, const std::string& wname) {
    const GemmW<T> w = gw_.at(wname);
    Act<T> out = new_act(pool, in.g.n, in.g.H, in.g.W, w.N, false);
    ops.tag("vae-enc conv4x4s2 " + std::to_string(w.K / 4) + "->" + std::to_string(w.N) + " @" +
                std::to_string(in.g.H * 2),
            2.0 * in.g.n * in.g.H * in.g.W * 4.0 * w.K * w.N);
    ops.push_back([this, in, w, out](cudaStream_t s) {
      // output pixel (oy, ox) sits at PM row (oy + 1) * Wp + ox (leading pad row); block (py, px) at py * Wp + px
      const int shifts[4] = {-in.g.Wp, -in.g.Wp + 1, 0, 1};
      GemmParams ep{};
      ep.epi = EPI_STORE;
      ep.out = out.p; ep.ldc = out.C; ep.bias = w.bias;
      gemm(in, w, 0, shifts, 4, ep, s, &out);
    });
    return out;
  };

  Act<T> r0 = new_act(pool, n, 224, 224, 64, false);
  {
    const float* w0 = f("venc.enc.0.weight");
    const float* b0 = f("venc.enc.0.bias");
    ops.tag("vae-enc conv_in 3->64 @224", 2.0 * n * 224.0 * 224 * 27 * 64);
    ops.push_back([this, r0, w0, b0, pp, n](cudaStream_t s) {
      if (enc_in_mma_) enc_conv_in_mma_kernel<T><<<dim3(49, n), 256, 0, s>>>(pp->img, w0, b0, r0.g, r0.p);
      else enc_conv_in_kernel<T><<<dim3(cdiv(224 * 112, 128), n), 128, 0, s>>>(pp->img, w0, b0, r0.g, r0.p);
      DMB_CHECK_LAUNCH();
      count();
    });
  }
  auto tap = [&](const char* name, const Act<T>& v) { P.taps[name] = v; return v; };
  tap("venc.0", r0);
  Act<T> a = gn8(r0, "venc.enc.1", true);
  a = tap("venc.5", gn8(tap("venc.3", conv4s2(a, "venc.enc.3")), "venc.enc.4", false));
  a = gn8(tap("venc.6", conv(a, "venc.enc.6")), "venc.enc.7", true);
  a = tap("venc.11", gn8(tap("venc.9", conv4s2(a, "venc.enc.9")), "venc.enc.10", false));
  a = gn8(tap("venc.12", conv(a, "venc.enc.12")), "venc.enc.13", true);
  a = tap("venc.17", gn8(tap("venc.15", conv4s2(a, "venc.enc.15")), "venc.enc.16", false));
  {
    const float* wm = f("venc.to_mu.weight");
    const float* bm = f("venc.to_mu.bias");
    const float* wl = f("venc.to_logvar.weight");
    const float* bl = f("venc.to_logvar.bias");
    ops.tag("vae-enc to_mu/to_logvar + reparameterise + KL", 2.0 * n * kLatH * kLatW * 256 * 8);
    constexpr int kParts = 4;   // CTAs per sample
    double* kl_part = pool.alloc_n<double>(static_cast<size_t>(n) * kParts);
    ops.push_back([this, a, wm, bm, wl, bl, pp, n, kl_part](cudaStream_t s) {
      vae_enc_head_kernel<T><<<dim3(n, kParts), 512, 0, s>>>(a.p, a.g, a.C, wm, bm, wl, bl, pp->eps, 0.18215f,
                                                            pp->z, pp->mu, pp->logvar, kl_part);
      DMB_CHECK_LAUNCH();
      vae_enc_kl_finish_kernel<<<cdiv(n, 32), 32, 0, s>>>(kl_part, kParts, n, 224.0f * 224.0f, pp->kl);
      DMB_CHECK_LAUNCH();
      count(2);
    });
  }
}

template <typename T>
void Engine<T>::vae_encode(const float* img, const float* eps, int B, float* z, float* mu, float* logvar,
                           float* kl_sample, cudaStream_t s) {
  DMB_CUDA(cudaSetDevice(device_));
  if (!vae_enc_loaded_) fail(5, "VAE encoder weights have not been loaded (dmb_load_vae_enc_weights)");
  DMB_REQUIRE(B >= 1, "batch must be >= 1");
  const int kChunk = 32;
  const size_t lat = static_cast<size_t>(kLatC) * kLatH * kLatW;
  const size_t im = static_cast<size_t>(3) * 224 * 224;
  for (int b0 = 0; b0 < B; b0 += kChunk) {
    const int n = std::min(kChunk, B - b0);
    auto found = vae_enc_plans_.find(n);
    if (found == vae_enc_plans_.end()) {
      DMB_CUDA(cudaStreamSynchronize(s));
      for (auto it = vae_enc_plans_.begin();
           static_cast<int>(vae_enc_plans_.size()) >= kMaxVaePlans && it != vae_enc_plans_.end();)
        it = it->first != kChunk ? vae_enc_plans_.erase(it) : std::next(it);
      found = vae_enc_plans_.emplace(n, std::unique_ptr<VaeEncPlan>(new VaeEncPlan())).first;
      build_vae_enc_plan(*found->second, n);
      DMB_CUDA(cudaDeviceSynchronize());
    }
    vae_enc_plan_ = found->second.get();
    VaeEncPlan& p = *vae_enc_plan_;
    p.img = img + b0 * im;
    p.eps = eps + b0 * lat;
    p.z = z + b0 * lat;
    p.mu = mu ? mu + b0 * lat : nullptr;
    p.logvar = logvar ? logvar + b0 * lat : nullptr;
    p.kl = kl_sample + b0;
    run_list(p.ops, s);
  }
}

template <typename T>
void Engine<T>::add_noise(const float* x0, const long long* t, const float* alpha_bars, int T_steps, const float* noise,
                          int B, long long per_sample, float* x_t, cudaStream_t s) {
  DMB_CUDA(cudaSetDevice(device_));
  DMB_REQUIRE(B >= 1 && per_sample >= 1 && T_steps >= 1, "B, per_sample and T must be >= 1");
  const int gx = static_cast<int>(std::min<long long>(cdiv(per_sample, 256), 64));
  add_noise_kernel<<<dim3(gx, B), 256, 0, s>>>(x0, t, alpha_bars, T_steps, noise, per_sample, x_t);
  DMB_CHECK_LAUNCH();
  count();
}

// ---------------------------------------------------------------------------
// Debug tap: a named intermediate activation of the last plan, as (n, C, H, W) fp32.
// ---------------------------------------------------------------------------
template <typename T>
void Engine<T>::get_tap(const char* name, float* out, int64_t capacity, int64_t* shape4, cudaStream_t s) {
  const std::string nm(name);
  if (nm.rfind("venc.", 0) == 0) {   // VAE encoder: output of enc.<i> of the last chunk (index of nn.Sequential)
    if (!vae_enc_plan_) fail(5, "no VAE encode has been run yet");
    auto it = vae_enc_plan_->taps.find(nm);
    if (it == vae_enc_plan_->taps.end()) fail(2, "unknown tap '" + nm + "'");
    const Act<T>& a = it->second;
    shape4[0] = a.g.n; shape4[1] = a.C; shape4[2] = a.g.H; shape4[3] = a.g.W;
    const long long total = static_cast<long long>(a.g.n) * a.C * a.g.H * a.g.W;
    DMB_REQUIRE(capacity >= total, "tap buffer too small");
    pm_to_nchw_kernel<T><<<cdiv(total, 256), 256, 0, s>>>(a.p, a.g, a.C, a.C, 0, out);
    DMB_CHECK_LAUNCH();
    return;
  }
  if (!last_plan_) fail(5, "no forward has been run yet");
  Plan<T>& p = *last_plan_;
  const int n = p.doubled ? 2 * p.B : p.B;
  if (nm == "emb" || nm == "embproj") {
    const int cols = nm == "emb" ? kEmbDim : kProjCols;
    shape4[0] = n; shape4[1] = cols; shape4[2] = 1; shape4[3] = 1;
    DMB_REQUIRE(capacity >= static_cast<int64_t>(n) * cols, "tap buffer too small");
    DMB_CUDA(cudaMemcpyAsync(out, nm == "emb" ? p.emb : p.embproj, static_cast<size_t>(n) * cols * sizeof(float),
                             cudaMemcpyDefault, s));
    return;
  }
  auto it = p.taps.find(nm);
  if (it == p.taps.end()) fail(2, "unknown tap '" + nm + "'");
  const Act<T>& a = it->second;
  shape4[0] = a.g.n; shape4[1] = a.C; shape4[2] = a.g.H; shape4[3] = a.g.W;
  const long long total = static_cast<long long>(a.g.n) * a.C * a.g.H * a.g.W;
  DMB_REQUIRE(capacity >= total, "tap buffer too small");
  pm_to_nchw_kernel<T><<<cdiv(total, 256), 256, 0, s>>>(a.p, a.g, a.C, a.C, 0, out);
  DMB_CHECK_LAUNCH();
}

}  // namespace dmb

namespace dmb {

// ---------------------------------------------------------------------------
// Op-level entry points (unit tests of the tensor-core kernel through the C ABI).
// Inputs / outputs are plain fp32 NCHW (or row-major matrices); the conversion to and
// from the padded pixel-matrix layout happens here.
// ---------------------------------------------------------------------------
template <typename T>
void Engine<T>::op_conv3x3(const float* x, const float* w, int B, int Cin, int Cout, int H, int W, float* out,
                           cudaStream_t s) {
  DMB_CUDA(cudaSetDevice(device_));
  DMB_REQUIRE(Cin % ElemTraits<T>::kChunkK == 0, "Cin must be a multiple of the K chunk (64 bf16 / 32 fp32)");
  Pool pool;
  float* wd = pool.alloc_n<float>(9ull * Cout * Cin);
  DMB_CUDA(cudaMemcpy(wd, w, 9ull * Cout * Cin * sizeof(float), cudaMemcpyDefault));
  GemmW<T> gw = make_conv3x3_from_ptr(wd, Cout, Cin, pool);
  Act<T> in = new_act(pool, B, H, W, Cin, false);
  const int cpad = gw.rows_per_tap;
  // (the epilogue's GroupNorm statistics run as in the network: a private two-words-per-sample arena)
  stat_arena_cur_ = pool.alloc_n<long long>(2 * static_cast<size_t>(B));
  stat_arena_left_ = 2 * static_cast<size_t>(B);
  DMB_CUDA(cudaMemset(stat_arena_cur_, 0, 2 * static_cast<size_t>(B) * sizeof(long long)));
  Act<T> o = new_act(pool, B, H, W, cpad, true);
  stat_arena_left_ = 0;
  DMB_CUDA(cudaDeviceSynchronize());
  const long long tin = static_cast<long long>(B) * Cin * H * W;
  nchw_to_pm_kernel<T><<<cdiv(tin, 256), 256, 0, s>>>(x, in.g, Cin, in.p);
  DMB_CHECK_LAUNCH();
  gw.N = cpad;  // store the padded columns too; only [0, Cout) are read back
  op_conv(in, gw, o)(s);
  const long long tout = static_cast<long long>(B) * Cout * H * W;
  pm_to_nchw_kernel<T><<<cdiv(tout, 256), 256, 0, s>>>(o.p, o.g, Cout, cpad, 0, out);
  DMB_CHECK_LAUNCH();
  DMB_CUDA(cudaStreamSynchronize(s));
}

// out[rows, N] = act(x[rows, K] @ w[N, K]^T + bias); rows are laid out as a 1-pixel-high map.
template <typename T>
void Engine<T>::op_linear(const float* x, const float* w, const float* bias, int rows, int K, int N, int act,
                          float* out, cudaStream_t s) {
  DMB_CUDA(cudaSetDevice(device_));
  DMB_REQUIRE(K % ElemTraits<T>::kChunkK == 0, "K must be a multiple of the K chunk (64 bf16 / 32 fp32)");
  Pool pool;
  GemmW<T> gw;
  gw.K = K; gw.N = N; gw.ntaps = 1; gw.bn = pick_block_n(N);
  gw.rows_per_tap = (N + gw.bn - 1) / gw.bn * gw.bn;
  float* wd = pool.alloc_n<float>(static_cast<size_t>(N) * K);
  DMB_CUDA(cudaMemcpy(wd, w, static_cast<size_t>(N) * K * sizeof(float), cudaMemcpyDefault));
  gw.w = pool.alloc_n<T>(static_cast<size_t>(gw.rows_per_tap) * K);
  cast_kernel<T><<<cdiv(static_cast<long long>(N) * K, 256), 256>>>(wd, static_cast<long long>(N) * K, gw.w);
  DMB_CHECK_LAUNCH();
  for (int i = 0; i < 6; ++i) gw.tmap_bn[i] = make_tmap<T>(gw.w, K, gw.rows_per_tap, 8 << i);
  float* bd = nullptr;
  if (bias) {
    bd = pool.alloc_n<float>(gw.rows_per_tap);
    DMB_CUDA(cudaMemcpy(bd, bias, N * sizeof(float), cudaMemcpyDefault));
  }
  gw.bias = bd;
  // rows as an (n=1, H=1, W=rows) map: x is (1, K, 1, rows) in NCHW terms = K-major transposed,
  // so convert from the row-major matrix directly with a (rows, K) -> PM copy.
  Act<T> in = new_act(pool, 1, 1, rows, K, false);
  Act<T> o = new_act(pool, 1, 1, rows, gw.rows_per_tap, false);
  DMB_CUDA(cudaDeviceSynchronize());
  // row i of the matrix -> PM row geo_row(0, 0, i); contiguous, so a strided 2D cast does it.
  float* xd = pool.alloc_n<float>(static_cast<size_t>(rows) * K);
  DMB_CUDA(cudaMemcpyAsync(xd, x, static_cast<size_t>(rows) * K * sizeof(float), cudaMemcpyDefault, s));
  cast_kernel<T><<<cdiv(static_cast<long long>(rows) * K, 256), 256, 0, s>>>(
      xd, static_cast<long long>(rows) * K, in.p + static_cast<size_t>(geo_row(in.g, 0, 0, 0)) * K);
  DMB_CHECK_LAUNCH();
  GemmW<T> gw2 = gw;
  gw2.N = gw.rows_per_tap;
  op_lin(in, gw2, o, act, nullptr)(s);
  // read back [rows, N] from PM rows (ld = rows_per_tap)
  float* od = pool.alloc_n<float>(static_cast<size_t>(rows) * N);
  // view as NCHW with C = N, H = 1, W = rows gives (N, rows); we want (rows, N): use a transposing readback
  pm_to_nchw_kernel<T><<<cdiv(static_cast<long long>(rows) * N, 256), 256, 0, s>>>(o.p, o.g, N, gw.rows_per_tap, 0, od);
  DMB_CHECK_LAUNCH();
  // od is (N, rows); transpose into out (rows, N)
  float* ot = pool.alloc_n<float>(static_cast<size_t>(rows) * N);
  transpose_kernel<<<cdiv(static_cast<long long>(rows) * N, 256), 256, 0, s>>>(od, N, rows, ot, N, 0);
  DMB_CHECK_LAUNCH();
  DMB_CUDA(cudaMemcpyAsync(out, ot, static_cast<size_t>(rows) * N * sizeof(float), cudaMemcpyDefault, s));
  DMB_CUDA(cudaStreamSynchronize(s));
}

template <typename T>
void Engine<T>::op_convT4x4(const float* x, const float* w, const float* bias, int B, int Cin, int Cout, int H,
                            int W, float* out, cudaStream_t s) {
  DMB_CUDA(cudaSetDevice(device_));
  DMB_REQUIRE(Cin % ElemTraits<T>::kChunkK == 0, "Cin must be a multiple of the K chunk (64 bf16 / 32 fp32)");
  Pool pool;
  GemmW<T> gw;
  gw.K = Cin; gw.N = Cout; gw.ntaps = 4; gw.bn = pick_block_n(Cout); gw.rows_per_tap = Cout;
  DMB_REQUIRE(Cout % gw.bn == 0, "Cout must be a multiple of the N tile");
  float* wd = pool.alloc_n<float>(16ull * Cout * Cin);
  DMB_CUDA(cudaMemcpy(wd, w, 16ull * Cout * Cin * sizeof(float), cudaMemcpyDefault));
  gw.w = pool.alloc_n<T>(16ull * Cout * Cin);
  repack_convT_kernel<T><<<cdiv(16LL * Cout * Cin, 256), 256>>>(wd, Cin, Cout, gw.w);
  DMB_CHECK_LAUNCH();
  for (int i = 0; i < 6; ++i) gw.tmap_bn[i] = make_tmap<T>(gw.w, Cin, 16ull * Cout, 8 << i);
  float* bd = nullptr;
  if (bias) {
    bd = pool.alloc_n<float>(Cout);
    DMB_CUDA(cudaMemcpy(bd, bias, Cout * sizeof(float), cudaMemcpyDefault));
  }
  gw.bias = bd;
  Act<T> in = new_act(pool, B, H, W, Cin, false);
  Act<T> o = new_act(pool, B, 2 * H, 2 * W, Cout, false);
  DMB_CUDA(cudaDeviceSynchronize());
  nchw_to_pm_kernel<T><<<cdiv(static_cast<long long>(B) * Cin * H * W, 256), 256, 0, s>>>(x, in.g, Cin, in.p);
  DMB_CHECK_LAUNCH();
  for (int a = 0; a < 2; ++a)
    for (int b = 0; b < 2; ++b) {
      int shifts[4];
      for (int u = 0; u < 2; ++u)
        for (int v = 0; v < 2; ++v) {
          const int dy = (a == 0) ? (u == 0 ? 0 : -1) : (u == 0 ? 1 : 0);
          const int dx = (b == 0) ? (v == 0 ? 0 : -1) : (v == 0 ? 1 : 0);
          shifts[u * 2 + v] = dy * in.g.Wp + dx;
        }
      GemmParams ep{};
      ep.epi = EPI_STORE;
      ep.out = o.p; ep.ldc = o.C; ep.bias = gw.bias;
      ep.map_mode = 1; ep.geo_out = o.g; ep.up_a = a; ep.up_b = b;
      gemm(in, gw, a * 2 + b, shifts, 4, ep, s);
    }
  pm_to_nchw_kernel<T><<<cdiv(static_cast<long long>(B) * Cout * 4 * H * W, 256), 256, 0, s>>>(o.p, o.g, Cout, Cout, 0, out);
  DMB_CHECK_LAUNCH();
  DMB_CUDA(cudaStreamSynchronize(s));
}

template <typename T>
std::string Engine<T>::timing_json() {
  DMB_CUDA(cudaDeviceSynchronize());
  struct Agg { int count = 0; double ms = 0, flops = 0, bytes = 0; };
  std::map<std::string, Agg> agg;
  std::vector<std::string> order;
  for (auto& t : timed_) {
    float ms = 0.f;
    DMB_CUDA(cudaEventElapsedTime(&ms, t.e0, t.e1));
    if (!agg.count(t.label)) order.push_back(t.label);
    Agg& a = agg[t.label];
    a.count += 1; a.ms += ms; a.flops += t.flops; a.bytes += t.bytes;
  }
  std::string out = "[";
  for (size_t i = 0; i < order.size(); ++i) {
    const Agg& a = agg[order[i]];
    char buf[512];
    snprintf(buf, sizeof(buf), "%s{\"label\": \"%s\", \"count\": %d, \"ms\": %.6f, \"flops\": %.6e, \"bytes\": %.6e}",
             i ? ", " : "", order[i].c_str(), a.count, a.ms, a.flops, a.bytes);
    out += buf;
  }
  out += "]";
  return out;
}

template class Engine<__nv_bfloat16>;
template class Engine<float>;

IEngine* make_engine_bf16(int device) { return new Engine<__nv_bfloat16>(device); }
IEngine* make_engine_f32(int device) { return new Engine<float>(device); }

}  // namespace dmb
