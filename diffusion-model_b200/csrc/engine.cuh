// Host-side engine behind the C ABI: device memory, TMA descriptors, repacked weights,
// the per-batch launch plans of the U-Net / VAE decoder and the CUDA-graph sampling loop.
#pragma once

#include <functional>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "common.cuh"
#include "gemm_tc.cuh"
#include "kernels.cuh"

namespace dmb {


struct TensorRef {  // a named fp32 tensor handed over the ABI (host or device memory)
  const void* ptr;
  int64_t numel;
};
using TensorMapIn = std::map<std::string, TensorRef>;

// Parameters that change between calls without changing the launch plan.  They live in
// device memory so that a captured graph stays valid.
struct DynParams {
  StepParams step;
  int flags[2];  // {null_label, has_cond}
};

class IEngine {
 public:
  virtual ~IEngine() {}
  virtual void load_unet(const TensorMapIn& w) = 0;
  virtual void load_vae(const TensorMapIn& w) = 0;
  virtual void unet_forward(const float* x, const long long* t, const long long* y, const float* vals,
                            const float* mask, int B, float* eps, float* geom, cudaStream_t s) = 0;
  virtual void cfg_step(const float* x_in, const long long* t, const long long* y, int null_label,
                        const float* vals, const float* mask, float guidance, const float* coef, int T,
                        const float* noise, unsigned long long seed, unsigned long long sample_offset,
                        float* x_out, float* eps_out, int B, cudaStream_t s) = 0;
  virtual void sample_loop(float* x, const long long* y, int null_label, const float* vals, const float* mask,
                           float guidance, const float* coef, int T, int t_start, int t_stride, int n_steps,
                           const float* noise, unsigned long long seed, unsigned long long sample_offset,
                           float* x_dump, float* eps_dump, int B, cudaStream_t s) = 0;
  virtual void vae_decode(const float* z, int B, float* img_f32, uint8_t* img_u8, cudaStream_t s) = 0;
  // VAE encoder + forward diffusion
  virtual void load_vae_enc(const TensorMapIn& w) = 0;
  virtual void vae_encode(const float* img, const float* eps, int B, float* z, float* mu, float* logvar,
                          float* kl_sample, cudaStream_t s) = 0;
  virtual void add_noise(const float* x0, const long long* t, const float* alpha_bars, int T, const float* noise,
                         int B, long long per_sample, float* x_t, cudaStream_t s) = 0;
  virtual void get_tap(const char* name, float* out, int64_t capacity, int64_t* shape4, cudaStream_t s) = 0;
  virtual void op_conv3x3(const float* x, const float* w, int B, int Cin, int Cout, int H, int W, float* out,
                          cudaStream_t s) = 0;
  virtual void op_linear(const float* x, const float* w, const float* bias, int rows, int K, int N, int act,
                         float* out, cudaStream_t s) = 0;
  virtual void op_convT4x4(const float* x, const float* w, const float* bias, int B, int Cin, int Cout, int H,
                           int W, float* out, cudaStream_t s) = 0;
  virtual void set_timing(bool on) = 0;
  virtual std::string timing_json() = 0;
  virtual long long launches() const = 0;
  virtual void reset_launches() = 0;
};

IEngine* make_engine_bf16(int device);
IEngine* make_engine_f32(int device);

}  // namespace dmb
