// extern "C" boundary: argument validation, exception -> status code translation.
#include "../../include/dmb200.h"

#include <cstring>
#include <atomic>
#include <cstdio>
#include <mutex>
#include <string>
#include <thread>
#include <vector>
#include <zlib.h>

#include "engine.cuh"
#include "score.cuh"

struct dmb_ctx {
  dmb::IEngine* eng;
  int device;
};

namespace {
thread_local std::string g_last_error;

template <typename F>
int guarded(F&& fn) {
  try {
    fn();
    return DMB_OK;
  } catch (const dmb::Error& e) {
    g_last_error = e.msg;
    return e.code ? e.code : 1;
  } catch (const std::exception& e) {
    g_last_error = std::string("internal error: ") + e.what();
    return 1;
  } catch (...) {
    g_last_error = "internal error";
    return 1;
  }
}

dmb::TensorMapIn collect(int n, const char* const* names, const void* const* ptrs, const int64_t* numels) {
  DMB_REQUIRE(n > 0 && names && ptrs && numels, "empty weight list");
  dmb::TensorMapIn m;
  for (int i = 0; i < n; ++i) {
    DMB_REQUIRE(names[i] && ptrs[i] && numels[i] > 0, "null entry in weight list");
    m[names[i]] = dmb::TensorRef{ptrs[i], numels[i]};
  }
  return m;
}
inline cudaStream_t st(void* s) { return static_cast<cudaStream_t>(s); }
inline const long long* ll(const int64_t* p) { return reinterpret_cast<const long long*>(p); }
}  // namespace

extern "C" {

int dmb_abi_version(void) { return 1; }
const char* dmb_last_error_string(void) { return g_last_error.c_str(); }

int dmb_ctx_create(int device, int precision, dmb_ctx** out) {
  return guarded([&] {
    DMB_REQUIRE(out != nullptr, "out is NULL");
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) dmb::fail(4, "no CUDA device available: this library has no CPU fallback");
    DMB_REQUIRE(device >= 0 && device < count, "device ordinal out of range");
    dmb::IEngine* eng = nullptr;
    if (precision == DMB_PRECISION_BF16) eng = dmb::make_engine_bf16(device);
    else if (precision == DMB_PRECISION_TF32) eng = dmb::make_engine_f32(device);
    else dmb::fail(2, "invalid argument: unknown precision mode");
    *out = new dmb_ctx{eng, device};
  });
}

int dmb_ctx_destroy(dmb_ctx* ctx) {
  return guarded([&] {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    delete ctx->eng;
    delete ctx;
  });
}

int dmb_load_unet_weights(dmb_ctx* ctx, int n, const char* const* names, const void* const* ptrs,
                          const int64_t* numels) {
  return guarded([&] {
    DMB_REQUIRE(ctx, "ctx is NULL");
    ctx->eng->load_unet(collect(n, names, ptrs, numels));
  });
}

int dmb_load_vae_dec_weights(dmb_ctx* ctx, int n, const char* const* names, const void* const* ptrs,
                             const int64_t* numels) {
  return guarded([&] {
    DMB_REQUIRE(ctx, "ctx is NULL");
    ctx->eng->load_vae(collect(n, names, ptrs, numels));
  });
}

int dmb_unet_forward(dmb_ctx* ctx, const float* x, const int64_t* t, const int64_t* y, const float* vals,
                     const float* mask, int B, float* eps_out, float* geom_out, void* stream) {
  return guarded([&] {
    DMB_REQUIRE(ctx && x && t && y, "ctx, x, t and y are required");
    DMB_REQUIRE((vals == nullptr) == (mask == nullptr), "cond_vals and cond_mask must be given together");
    ctx->eng->unet_forward(x, ll(t), ll(y), vals, mask, B, eps_out, geom_out, st(stream));
  });
}

int dmb_denoise_cond_step(dmb_ctx* ctx, const float* x_in, const int64_t* t, const int64_t* y, int null_label,
                          const float* vals, const float* mask, float guidance_scale, const float* coef, int T,
                          const float* noise, uint64_t seed, uint64_t sample_offset, float* x_out,
                          float* eps_out, int B, void* stream) {
  return guarded([&] {
    DMB_REQUIRE(ctx && x_in && t && y && coef && x_out, "ctx, x_in, t, y, coef and x_out are required");
    DMB_REQUIRE((vals == nullptr) == (mask == nullptr), "cond_vals and cond_mask must be given together");
    DMB_REQUIRE(null_label >= 0 && null_label <= 3, "null_label must index class_emb (0..3)");
    DMB_REQUIRE(T >= 1, "T must be >= 1");
    ctx->eng->cfg_step(x_in, ll(t), ll(y), null_label, vals, mask, guidance_scale, coef, T, noise, seed,
                       sample_offset, x_out, eps_out, B, st(stream));
  });
}

int dmb_sample_loop(dmb_ctx* ctx, float* x, const int64_t* y, int null_label, const float* vals,
                    const float* mask, float guidance_scale, const float* coef, int T, int t_start, int n_steps,
                    const float* noise, uint64_t seed, uint64_t sample_offset, float* x_dump, float* eps_dump,
                    int B, void* stream) {
  return guarded([&] {
    DMB_REQUIRE(ctx && x && y && coef, "ctx, x, y and coef are required");
    DMB_REQUIRE((vals == nullptr) == (mask == nullptr), "cond_vals and cond_mask must be given together");
    DMB_REQUIRE(null_label >= 0 && null_label <= 3, "null_label must index class_emb (0..3)");
    ctx->eng->sample_loop(x, ll(y), null_label, vals, mask, guidance_scale, coef, T, t_start, 1, n_steps, noise, seed,
                          sample_offset, x_dump, eps_dump, B, st(stream));
  });
}

int dmb_sample_loop_strided(dmb_ctx* ctx, float* x, const int64_t* y, int null_label, const float* vals,
                            const float* mask, float guidance_scale, const float* coef, int T, int t_start,
                            int t_stride, int n_steps, const float* noise, uint64_t seed, uint64_t sample_offset,
                            float* x_dump, float* eps_dump, int B, void* stream) {
  return guarded([&] {
    DMB_REQUIRE(ctx && x && y && coef, "ctx, x, y and coef are required");
    DMB_REQUIRE((vals == nullptr) == (mask == nullptr), "cond_vals and cond_mask must be given together");
    DMB_REQUIRE(null_label >= 0 && null_label <= 3, "null_label must index class_emb (0..3)");
    ctx->eng->sample_loop(x, ll(y), null_label, vals, mask, guidance_scale, coef, T, t_start, t_stride, n_steps, noise,
                          seed, sample_offset, x_dump, eps_dump, B, st(stream));
  });
}

int dmb_vae_decode(dmb_ctx* ctx, const float* z, int B, float* img_f32, uint8_t* img_u8, void* stream) {
  return guarded([&] {
    DMB_REQUIRE(ctx && z, "ctx and z are required");
    DMB_REQUIRE(img_f32 || img_u8, "at least one output buffer is required");
    ctx->eng->vae_decode(z, B, img_f32, img_u8, st(stream));
  });
}

int dmb_load_vae_enc_weights(dmb_ctx* ctx, int n, const char* const* names, const void* const* ptrs,
                             const int64_t* numels) {
  return guarded([&] {
    DMB_REQUIRE(ctx, "ctx is NULL");
    ctx->eng->load_vae_enc(collect(n, names, ptrs, numels));
  });
}

int dmb_vae_encode(dmb_ctx* ctx, const float* img, const float* eps, int B, float* z, float* mu, float* logvar,
                   float* kl_sample, void* stream) {
  return guarded([&] {
    DMB_REQUIRE(ctx && img && eps && z && kl_sample, "ctx, img, eps, z and kl_sample are required");
    ctx->eng->vae_encode(img, eps, B, z, mu, logvar, kl_sample, st(stream));
  });
}

int dmb_add_noise(dmb_ctx* ctx, const float* x0, const int64_t* t, const float* alpha_bars, int T,
                  const float* noise, int B, int64_t per_sample, float* x_t, void* stream) {
  return guarded([&] {
    DMB_REQUIRE(ctx && x0 && t && alpha_bars && noise && x_t, "ctx, x0, t, alpha_bars, noise and x_t are required");
    ctx->eng->add_noise(x0, ll(t), alpha_bars, T, noise, B, static_cast<long long>(per_sample), x_t, st(stream));
  });
}

namespace {
constexpr int kReconThreads = 1024;
// one CTA per sample; thread t takes float4 vectors t, t + 1024, ... (then the scalar tail), sums in
// double, and the CTA combines the 1024 sums with a fixed shuffle + shared-memory tree
__global__ void __launch_bounds__(kReconThreads)
recon_error_kernel(const float* __restrict__ a, const float* __restrict__ b, long long per_sample,
                   double* __restrict__ sq_sum, double* __restrict__ abs_sum) {
  __shared__ double red[2][kReconThreads / 32];
  const float* pa = a + static_cast<size_t>(blockIdx.x) * per_sample;
  const float* pb = b + static_cast<size_t>(blockIdx.x) * per_sample;
  double sq = 0.0, ab = 0.0;
  const bool vec = (per_sample % 4 == 0) && ((reinterpret_cast<uintptr_t>(pa) | reinterpret_cast<uintptr_t>(pb)) % 16 == 0);
  const long long nvec = vec ? per_sample / 4 : 0;
  for (long long i = threadIdx.x; i < nvec; i += kReconThreads) {
    const float4 x = reinterpret_cast<const float4*>(pa)[i], y = reinterpret_cast<const float4*>(pb)[i];
    const float d0 = x.x - y.x, d1 = x.y - y.y, d2 = x.z - y.z, d3 = x.w - y.w;
    sq += static_cast<double>(d0) * d0 + static_cast<double>(d1) * d1 + static_cast<double>(d2) * d2 + static_cast<double>(d3) * d3;
    ab += static_cast<double>(fabsf(d0)) + fabsf(d1) + fabsf(d2) + fabsf(d3);
  }
  for (long long i = nvec * 4 + threadIdx.x; i < per_sample; i += kReconThreads) {
    const float d = pa[i] - pb[i];
    sq += static_cast<double>(d) * d;
    ab += fabsf(d);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    sq += __shfl_xor_sync(0xffffffffu, sq, o);
    ab += __shfl_xor_sync(0xffffffffu, ab, o);
  }
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = sq; red[1][threadIdx.x >> 5] = ab; }
  __syncthreads();
  if (threadIdx.x < 32) {
    sq = red[0][threadIdx.x]; ab = red[1][threadIdx.x];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      sq += __shfl_xor_sync(0xffffffffu, sq, o);
      ab += __shfl_xor_sync(0xffffffffu, ab, o);
    }
    if (threadIdx.x == 0) {
      if (sq_sum) sq_sum[blockIdx.x] = sq;
      if (abs_sum) abs_sum[blockIdx.x] = ab;
    }
  }
}
}  // namespace

int dmb_recon_error(dmb_ctx* ctx, const float* a, const float* b, int B, int64_t per_sample,
                    double* sq_sum, double* abs_sum, void* stream) {
  return guarded([&] {
    DMB_REQUIRE(ctx && a && b && (sq_sum || abs_sum), "ctx, a, b and one of sq_sum / abs_sum are required");
    DMB_REQUIRE(B >= 1 && per_sample >= 1, "B >= 1 and per_sample >= 1 are required");
    DMB_CUDA(cudaSetDevice(ctx->device));
    recon_error_kernel<<<B, kReconThreads, 0, st(stream)>>>(a, b, static_cast<long long>(per_sample), sq_sum, abs_sum);
    DMB_CHECK_LAUNCH();
  });
}

int dmb_score_masks(dmb_ctx* ctx, const uint8_t* pred, int pred_channels, const uint8_t* gt, int gt_channels,
                    int B, int H, int W, int threshold, int invert, double sigma, int64_t* counts,
                    double* weighted_hit, void* stream) {
  return guarded([&] {
    DMB_REQUIRE(ctx && pred && gt && counts && weighted_hit, "ctx, pred, gt, counts and weighted_hit are required");
    DMB_REQUIRE(B >= 1 && H >= 1 && W >= 1 && H <= dmb::kScoreMaxSide && W <= dmb::kScoreMaxSide,
                "B >= 1 and 1 <= H, W <= 254 are required");
    DMB_REQUIRE((pred_channels == 0 || pred_channels == 1 || pred_channels == 3) &&
                    (gt_channels == 0 || gt_channels == 1 || gt_channels == 3),
                "channels must be 0 (mask), 1 (grey) or 3 (RGB)");
    DMB_REQUIRE(sigma > 0.0, "sigma must be > 0");   // eval_iou_noise.py:197, 225
    DMB_CUDA(cudaSetDevice(ctx->device));
    // function attributes are per device: one flag per ordinal
    static std::atomic<bool> attr_set_dev[64];
    DMB_REQUIRE(ctx->device >= 0 && ctx->device < 64, "device ordinal out of range");
    std::atomic<bool>& attr_set = attr_set_dev[ctx->device];
    if (!attr_set.load()) {
      DMB_CUDA(cudaFuncSetAttribute(dmb::score_masks_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    static_cast<int>(dmb::score_smem_bytes(dmb::kScoreMaxSide, dmb::kScoreMaxSide))));
      // two 100 KB CTAs per SM at 224 x 224: ask for the largest shared-memory carve-out
      DMB_CUDA(cudaFuncSetAttribute(dmb::score_masks_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                    cudaSharedmemCarveoutMaxShared));
      attr_set.store(true);
    }
    dmb::ScoreParams p{};
    p.pred = pred; p.gt = gt; p.pred_channels = pred_channels; p.gt_channels = gt_channels;
    p.B = B; p.H = H; p.W = W; p.threshold = threshold; p.invert = invert; p.sigma = sigma;
    p.counts = reinterpret_cast<long long*>(counts);
    p.weighted_hit = weighted_hit;
    dmb::score_masks_kernel<<<B, dmb::kScoreThreads, dmb::score_smem_bytes(H, W), st(stream)>>>(p);
    DMB_CHECK_LAUNCH();
  });
}

// ---------------------------------------------------------------------------
// Host side of the output path: PNG files from a batch of planar u8 images.
// ---------------------------------------------------------------------------
namespace {
void png_chunk(std::vector<unsigned char>& out, const char* type, const unsigned char* data, size_t n) {
  auto be32 = [&](uint32_t v) { for (int s = 24; s >= 0; s -= 8) out.push_back(static_cast<unsigned char>(v >> s)); };
  be32(static_cast<uint32_t>(n));
  const size_t start = out.size();
  out.insert(out.end(), type, type + 4);
  if (n) out.insert(out.end(), data, data + n);
  be32(static_cast<uint32_t>(crc32(0L, out.data() + start, static_cast<uInt>(n + 4))));
}

// one image (C, H, W) planar u8 -> PNG bytes (8-bit grey or RGB, filter type 0 on every row)
bool png_encode(const unsigned char* img, int C, int H, int W, int level, std::vector<unsigned char>& out) {
  std::vector<unsigned char> raw(static_cast<size_t>(H) * (1 + static_cast<size_t>(W) * C));
  const size_t plane = static_cast<size_t>(H) * W;
  unsigned char* q = raw.data();
  for (int y = 0; y < H; ++y) {
    *q++ = 0;
    const unsigned char* r0 = img + static_cast<size_t>(y) * W;
    if (C == 1) {
      memcpy(q, r0, W);
      q += W;
    } else {
      for (int x = 0; x < W; ++x) { *q++ = r0[x]; *q++ = r0[plane + x]; *q++ = r0[2 * plane + x]; }
    }
  }
  uLongf zn = compressBound(static_cast<uLong>(raw.size()));
  std::vector<unsigned char> z(zn);
  if (compress2(z.data(), &zn, raw.data(), static_cast<uLong>(raw.size()), level) != Z_OK) return false;
  static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
  out.assign(sig, sig + 8);
  unsigned char ihdr[13];
  for (int i = 0; i < 4; ++i) { ihdr[i] = static_cast<unsigned char>(W >> (24 - 8 * i)); ihdr[4 + i] = static_cast<unsigned char>(H >> (24 - 8 * i)); }
  ihdr[8] = 8; ihdr[9] = C == 1 ? 0 : 2; ihdr[10] = 0; ihdr[11] = 0; ihdr[12] = 0;
  png_chunk(out, "IHDR", ihdr, 13);
  png_chunk(out, "IDAT", z.data(), zn);
  png_chunk(out, "IEND", nullptr, 0);
  return true;
}
}  // namespace

int dmb_write_png_batch(const uint8_t* images, int n, int C, int H, int W, const char* const* paths, int threads,
                        int compression_level) {
  return guarded([&] {
    DMB_REQUIRE(images && paths && n >= 0, "images and paths are required");
    DMB_REQUIRE((C == 1 || C == 3) && H >= 1 && W >= 1, "C must be 1 (grey) or 3 (RGB); H, W >= 1");
    DMB_REQUIRE(compression_level >= 0 && compression_level <= 9, "compression_level must be 0..9");
    const int nt = std::max(1, std::min(threads > 0 ? threads : static_cast<int>(std::thread::hardware_concurrency()), n));
    std::atomic<int> next{0};
    std::atomic<int> failed{-1};
    auto work = [&]() {
      std::vector<unsigned char> png;
      for (int i = next.fetch_add(1); i < n; i = next.fetch_add(1)) {
        bool ok = png_encode(images + static_cast<size_t>(i) * C * H * W, C, H, W, compression_level, png);
        if (ok) {
          FILE* f = fopen(paths[i], "wb");
          ok = f && fwrite(png.data(), 1, png.size(), f) == png.size();
          if (f) ok = (fclose(f) == 0) && ok;
        }
        if (!ok) { int exp = -1; failed.compare_exchange_strong(exp, i); }
      }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < nt; ++t) pool.emplace_back(work);
    work();
    for (auto& t : pool) t.join();
    if (failed.load() >= 0) dmb::fail(6, std::string("could not write '") + paths[failed.load()] + "'");
  });
}

int dmb_philox_normal(dmb_ctx* ctx, float* out, int B, uint64_t seed, uint64_t sample_offset, int t,
                      void* stream) {
  return guarded([&] {
    DMB_REQUIRE(ctx && out && B >= 1, "ctx, out and B >= 1 are required");
    DMB_CUDA(cudaSetDevice(ctx->device));
    const int HW = dmb::kLatH * dmb::kLatW;
    dmb::philox_fill_kernel<<<dim3((HW + 127) / 128, B), 128, 0, st(stream)>>>(out, B, HW, seed, sample_offset,
                                                                              static_cast<unsigned int>(t));
    DMB_CHECK_LAUNCH();
  });
}

int64_t dmb_launch_count(dmb_ctx* ctx) { return ctx ? ctx->eng->launches() : -1; }
int dmb_reset_launch_count(dmb_ctx* ctx) {
  return guarded([&] {
    DMB_REQUIRE(ctx, "ctx is NULL");
    ctx->eng->reset_launches();
  });
}

int dmb_set_kernel_timing(dmb_ctx* ctx, int on) {
  return guarded([&] {
    DMB_REQUIRE(ctx, "ctx is NULL");
    ctx->eng->set_timing(on != 0);
  });
}

int dmb_get_kernel_timing(dmb_ctx* ctx, char* buf, int64_t capacity, int64_t* needed) {
  return guarded([&] {
    DMB_REQUIRE(ctx && needed, "ctx and needed are required");
    const std::string js = ctx->eng->timing_json();
    *needed = static_cast<int64_t>(js.size()) + 1;
    if (buf && capacity >= *needed) memcpy(buf, js.c_str(), js.size() + 1);
  });
}

int dmb_debug_get_tap(dmb_ctx* ctx, const char* name, float* out, int64_t capacity, int64_t* shape4,
                      void* stream) {
  return guarded([&] {
    DMB_REQUIRE(ctx && name && out && shape4, "ctx, name, out and shape4 are required");
    ctx->eng->get_tap(name, out, capacity, shape4, st(stream));
  });
}

int dmb_op_conv3x3(dmb_ctx* ctx, const float* x, const float* w, int B, int Cin, int Cout, int H, int W,
                   float* out, void* stream) {
  return guarded([&] {
    DMB_REQUIRE(ctx && x && w && out, "ctx, x, w and out are required");
    DMB_REQUIRE(B >= 1 && Cin >= 1 && Cout >= 1 && H >= 1 && W >= 1, "sizes must be positive");
    ctx->eng->op_conv3x3(x, w, B, Cin, Cout, H, W, out, st(stream));
  });
}

int dmb_op_linear(dmb_ctx* ctx, const float* x, const float* w, const float* bias, int rows, int K, int N,
                  int act, float* out, void* stream) {
  return guarded([&] {
    DMB_REQUIRE(ctx && x && w && out, "ctx, x, w and out are required");
    DMB_REQUIRE(rows >= 1 && K >= 1 && N >= 1, "sizes must be positive");
    ctx->eng->op_linear(x, w, bias, rows, K, N, act, out, st(stream));
  });
}

int dmb_op_conv_transpose4x4(dmb_ctx* ctx, const float* x, const float* w, const float* bias, int B, int Cin,
                             int Cout, int H, int W, float* out, void* stream) {
  return guarded([&] {
    DMB_REQUIRE(ctx && x && w && out, "ctx, x, w and out are required");
    DMB_REQUIRE(B >= 1 && Cin >= 1 && Cout >= 1 && H >= 1 && W >= 1, "sizes must be positive");
    ctx->eng->op_convT4x4(x, w, bias, B, Cin, Cout, H, W, out, st(stream));
  });
}

}  // extern "C"

// ---------------------------------------------------------------------------
// Allocation-free entry point (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA-graph capture, tight serving
// loops, harnesses that forbid it).  A variant of the VAE encoder's KL reduction
// (vae_enc_head_kernel -> kl_part scratch -> vae_enc_kl_finish_kernel, kernels.cuh) whose
// per-CTA partial sums live in a file-scope __device__ array instead of the context's pool.
//
// g_kl_part_static is a DEVICE symbol: in host code its name is only a host-side stand-in, so
// the host never passes it to a kernel.  Once per device, with that device current, the entry
// point asks the runtime for the array's device address (cudaGetSymbolAddress) and hands THAT
// pointer to the kernels.  Every GPU that loads the library holds the array (128 KB) until exit.
// ---------------------------------------------------------------------------
namespace {
constexpr int kKlStaticMaxB = DMB_KL_STATIC_MAX_B;
constexpr int kKlStaticParts = 4;      // CTAs per sample, as in vae_enc_head_kernel's launch
constexpr int kKlStaticThreads = 256;
constexpr int kKlStaticMaxDevices = 64;
static __device__ __align__(256) double g_kl_part_static[kKlStaticMaxB * kKlStaticParts];

std::mutex g_kl_static_mu;
double* g_kl_static_ptr[kKlStaticMaxDevices] = {};   // device address of the array, per device

// One CTA = one quarter of a sample's elements.  Same summand and the same float64 accumulation
// as vae_enc_head_kernel (models/vae.py:61): exp(logvar) + mu^2 - 1 - logvar; fixed order
// (thread stride, lanes, warps), so two calls on the same input give the same bits.  Every
// partial this launch's finish kernel reads is written here first: nothing has to start at zero.
__global__ void __launch_bounds__(kKlStaticThreads)
kl_part_static_kernel(const float* __restrict__ mu, const float* __restrict__ logvar, long long per_sample,
                      double* __restrict__ part) {
  __shared__ double s_w[kKlStaticThreads / 32];
  const int b = blockIdx.x, p = blockIdx.y;
  const long long chunk = (per_sample + kKlStaticParts - 1) / kKlStaticParts;
  const long long lo = p * chunk, hi = lo + chunk < per_sample ? lo + chunk : per_sample;
  const float* m_ = mu + static_cast<size_t>(b) * per_sample;
  const float* l_ = logvar + static_cast<size_t>(b) * per_sample;
  double acc = 0.0;
  for (long long i = lo + threadIdx.x; i < hi; i += kKlStaticThreads) {
    const float m = m_[i], lv = l_[i];
    acc += static_cast<double>(expf(lv)) + static_cast<double>(m) * m - 1.0 - lv;
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, d);
  if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < kKlStaticThreads / 32; ++w) t += s_w[w];
    part[b * kKlStaticParts + p] = t;
  }
}

__global__ void kl_finish_static_kernel(const double* __restrict__ part, int n, float kl_div,
                                        float* __restrict__ kl_sample) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n) return;
  double t = 0.0;
  for (int p = 0; p < kKlStaticParts; ++p) t += part[b * kKlStaticParts + p];
  kl_sample[b] = static_cast<float>(0.5 * t / kl_div);
}

// Set-up, once per device: the device address of g_kl_part_static on `device` (made current).
double* kl_static_scratch(int device) {
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
    dmb::fail(4, "no CUDA device available: this library has no CPU fallback");
  DMB_REQUIRE(device >= 0 && device < count && device < kKlStaticMaxDevices, "device ordinal out of range");
  DMB_CUDA(cudaSetDevice(device));
  std::lock_guard<std::mutex> lock(g_kl_static_mu);
  if (!g_kl_static_ptr[device]) {
    double* ptr = nullptr;
    DMB_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_kl_part_static));
    if (!ptr) dmb::fail(5, "cudaGetSymbolAddress returned a null address for the static scratch array");
    g_kl_static_ptr[device] = ptr;
  }
  return g_kl_static_ptr[device];
}
}  // namespace

extern "C" int dmb_kl_per_sample_static(int device, const float* mu, const float* logvar, int B, int64_t per_sample,
                                        float kl_div, float* kl_sample, void* stream) {
  return guarded([&] {
    DMB_REQUIRE(mu && logvar && kl_sample, "mu, logvar and kl_sample are required");
    DMB_REQUIRE(B >= 1 && per_sample >= 1, "B >= 1 and per_sample >= 1 are required");
    if (B > kKlStaticMaxB)
      dmb::fail(2, "invalid argument: B = " + std::to_string(B) + " exceeds the static scratch array of "
                       "dmb_kl_per_sample_static (DMB_KL_STATIC_MAX_B = " + std::to_string(kKlStaticMaxB) +
                       " samples per call); split the batch or use dmb_vae_encode");
    DMB_REQUIRE(kl_div > 0.0f, "kl_div must be > 0");
    double* part = kl_static_scratch(device);   // device pointer, never the host-side symbol
    kl_part_static_kernel<<<dim3(B, kKlStaticParts), kKlStaticThreads, 0, st(stream)>>>(
        mu, logvar, static_cast<long long>(per_sample), part);
    DMB_CHECK_LAUNCH();
    kl_finish_static_kernel<<<(B + 127) / 128, 128, 0, st(stream)>>>(part, B, kl_div, kl_sample);
    DMB_CHECK_LAUNCH();
  });
}
