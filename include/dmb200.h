/* dmb200 -- C ABI of the B200-native conditional reverse-diffusion sampling path.
 *
 * The reference (S-Taichiii/diffusion-model) is pure Python and has no FFI; the boundary
 * it exposes for this path is its Python module surface.  Each entry point
 * below names the reference function whose arithmetic it replaces (file:line relative to
 * the reference root); `diffusion-model_b200/` holds the same-named Python modules that
 * bind these symbols with ctypes.
 *
 * Conventions
 *   - plain C types only; tensors are raw pointers + sizes, fp32 contiguous NCHW unless
 *     stated; int64 for t / y exactly as torch.long;
 *   - pointers marked [dev] must be device memory of the context's GPU (tensor.data_ptr());
 *     pointers marked [any] may be host or device memory (copied with cudaMemcpyDefault);
 *   - all work is enqueued on `stream` (a cudaStream_t passed as void*, 0 = legacy default
 *     stream); no hidden synchronisation except where stated;
 *   - every function returns 0 on success, non-zero on failure; the message is available
 *     from dmb_last_error_string() (thread-local).  No C++ exception crosses the boundary;
 *   - a context is bound to one GPU and is not thread-safe; use one per device per process;
 *   - the caller owns every input and output buffer; the library owns its context only
 *     (repacked weights, activations, CUDA graphs).  No pointer is retained after a call
 *     returns except weights, which are copied.
 *   - there is no CPU fallback: dmb_ctx_create fails on anything but an sm_100 GPU.
 */
#ifndef DMB200_H_
#define DMB200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dmb_ctx dmb_ctx;

#define DMB_OK 0
#define DMB_PRECISION_BF16 0 /* bf16 activations/weights, tcgen05 kind::f16, fp32 accumulate */
#define DMB_PRECISION_TF32 1 /* fp32 activations/weights, tcgen05 kind::tf32, fp32 accumulate */

int dmb_abi_version(void);
const char* dmb_last_error_string(void);

/* One context per GPU.  precision: DMB_PRECISION_*. */
int dmb_ctx_create(int device, int precision, dmb_ctx** out);
int dmb_ctx_destroy(dmb_ctx* ctx);

/* Weights in the reference's state_dict layout, fp32, by name.
 * Replaces Utils.loadModel's model.load_state_dict(...).to(device) hand-over (utils.py:68-73):
 * U-Net keys of UnetCondWithGeomHead (models/unet_cond_geom.py:31-49, models/unet_cond.py:111-153;
 * 191 tensors, conv OIHW, Linear (out,in), mha.in_proj_weight (3E,E)); optional extra key
 * "__inv_freq" (128,) = the pos_encoding frequencies as torch computes them (unet_cond.py:156).
 * VAE keys "dec.{0..18}.{weight,bias}" (models/vae.py:35-49; 26 tensors, conv-transpose IOHW). */
int dmb_load_unet_weights(dmb_ctx* ctx, int n, const char* const* names, const void* const* ptrs /*[any]*/,
                          const int64_t* numels);
int dmb_load_vae_dec_weights(dmb_ctx* ctx, int n, const char* const* names, const void* const* ptrs /*[any]*/,
                             const int64_t* numels);

/* UnetCondWithGeomHead.forward (models/unet_cond_geom.py:79-100): one eager forward.
 * x (B,4,28,28) [dev]; t,y (B,) int64 [any]; vals,mask (B,12) [any] or both NULL (then no
 * numeric-condition term, unet_cond_geom.py:91); eps_out (B,4,28,28) [dev] or NULL;
 * geom_out (B,12) [dev] or NULL. */
int dmb_unet_forward(dmb_ctx* ctx, const float* x, const int64_t* t, const int64_t* y, const float* vals,
                     const float* mask, int B, float* eps_out, float* geom_out, void* stream);

/* Diffuser.denoise_cond, CFG branch (diff.py:127-162): the two U-Net passes as one doubled
 * batch, eps = eps_u + g (eps_c - eps_u), posterior mean, noise (zeroed where t == 1).
 * coef (T,3) fp32 [dev] = per-t ((1-a)/sqrt(1-abar), sqrt(a), sigma), computed by the host
 * with the reference's own torch expressions (diff.py:14-16, 160-161).
 * noise (B,4,28,28) [dev] injects the step's noise; NULL draws Philox noise keyed by
 * (seed, sample_offset + b, t).  x_out may alias x_in.  eps_out optional. */
int dmb_denoise_cond_step(dmb_ctx* ctx, const float* x_in, const int64_t* t, const int64_t* y, int null_label,
                          const float* vals, const float* mask, float guidance_scale, const float* coef, int T,
                          const float* noise, uint64_t seed, uint64_t sample_offset, float* x_out,
                          float* eps_out, int B, void* stream);

/* The reverse loop of Diffuser.sample_latent_cond (diff.py:327-344) for n_steps steps
 * t = t_start, t_start-1, ..., as CUDA-graph replays.  x (B,4,28,28) [dev] holds x_{t_start}
 * on entry and x_{t_start-n_steps} on return.  noise: NULL (Philox) or (n_steps,B,4,28,28)
 * [dev], entry i used at t = t_start - i.  x_dump / eps_dump: optional (n_steps,B,4,28,28)
 * [dev] receiving x_t BEFORE each step (generate_steps.py:160-178 saves that state) and the
 * guided eps of each step. */
int dmb_sample_loop(dmb_ctx* ctx, float* x, const int64_t* y, int null_label, const float* vals,
                    const float* mask, float guidance_scale, const float* coef, int T, int t_start, int n_steps,
                    const float* noise, uint64_t seed, uint64_t sample_offset, float* x_dump, float* eps_dump,
                    int B, void* stream);

/* Extension (NOT in the reference): the same loop over a strided
 * schedule.  Iteration k runs at t = t_start - k * t_stride (all >= 1) and reads its constants
 * from row t-1 of coef like dmb_sample_loop; the caller fills those rows with the update it
 * wants, e.g. deterministic DDIM (eta = 0) written in the posterior-step form
 * x' = (x - c0 * eps) / c1 + c2 * z  (Diffuser.strided_coefficients).  noise / x_dump / eps_dump
 * are indexed by iteration.  t_stride = 1 is exactly dmb_sample_loop. */
int dmb_sample_loop_strided(dmb_ctx* ctx, float* x, const int64_t* y, int null_label, const float* vals,
                            const float* mask, float guidance_scale, const float* coef, int T, int t_start,
                            int t_stride, int n_steps, const float* noise, uint64_t seed, uint64_t sample_offset,
                            float* x_dump, float* eps_dump, int B, void* stream);

/* VAE.decode (models/vae.py:64-69) + Diffuser.reverse_to_img's x*255 / clamp / uint8 truncation
 * (diff.py:58-61).  z (B,4,28,28) [dev]; img_f32 (B,3,224,224) [dev] in [0,1] or NULL;
 * img_u8 (B,3,224,224) [dev] or NULL. */
int dmb_vae_decode(dmb_ctx* ctx, const float* z, int B, float* img_f32, uint8_t* img_u8, void* stream);

/* VAE encoder weights: the state_dict entries of VAE.enc, VAE.to_mu and
 * VAE.to_logvar under their own names ("enc.0.weight" ... "enc.16.bias", "to_mu.weight",
 * "to_mu.bias", "to_logvar.weight", "to_logvar.bias"), fp32, host or device  (models/vae.py:16-31). */
int dmb_load_vae_enc_weights(dmb_ctx* ctx, int n, const char* const* names, const void* const* ptrs /*[any]*/,
                             const int64_t* numels);

/* VAE.encode (models/vae.py:51-62).  img (B,3,224,224) fp32 [dev]; eps (B,4,28,28) [dev] is the
 * standard-normal draw the reference takes from torch.randn_like(std) (:56) -- the caller owns
 * the generator; z (B,4,28,28) [dev] = (mu + eps * exp(0.5 * logvar)) * 0.18215; mu / logvar
 * (B,4,28,28) [dev] or NULL (logvar after the clamp to [-30, 20], :54); kl_sample (B) [dev] =
 * 0.5 * sum(exp(logvar) + mu^2 - 1 - logvar) / (224 * 224) per sample (:61) -- the reference
 * returns its mean over the batch. */
int dmb_vae_encode(dmb_ctx* ctx, const float* img, const float* eps, int B, float* z, float* mu, float* logvar,
                   float* kl_sample, void* stream);

/* Diffuser.add_noise (diff.py:18-30): x_t = sqrt(ab) * x_0 + sqrt(1 - ab) * noise, ab =
 * alpha_bars[t[b] - 1].  x0 / noise / x_t (B, per_sample) fp32 [dev]; t (B) int64 [dev], 1..T
 * (the caller checks the range, as the reference asserts it on the host); alpha_bars (T) fp32
 * [dev].  Each operation is rounded separately, as in the reference's three torch kernels, so
 * the result is bit-identical for the same noise. */
int dmb_add_noise(dmb_ctx* ctx, const float* x0, const int64_t* t, const float* alpha_bars, int T,
                  const float* noise, int B, int64_t per_sample, float* x_t, void* stream);

/* Reconstruction error of VAE.forward (vae.py:71-76: F.mse_loss(x_recon, x)) and of
 * vae_recon_check.py:29-31 (per-image mean squared / absolute error): per sample, the float64 sums
 * of (a - b)^2 and |a - b| over per_sample elements, accumulated in a fixed order (one CTA per
 * sample: per-thread sums over a fixed stride, then a fixed tree), so two calls give the same bits.
 * a, b (B, per_sample) fp32 [dev]; sq_sum, abs_sum (B) float64 [dev], either may be NULL.
 * The caller divides (mean over the batch: sum(sq_sum) / (B * per_sample)). */
int dmb_recon_error(dmb_ctx* ctx, const float* a, const float* b, int B, int64_t per_sample,
                    double* sq_sum, double* abs_sum, void* stream);

/* Allocation-free KL reduction.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight
 * serving loops, harnesses that forbid it).  A variant of the KL term of dmb_vae_encode
 * (0.5 * sum(exp(logvar) + mu^2 - 1 - logvar) / kl_div per sample, float64 accumulation in a
 * fixed order) that needs no context: its scratch (per-CTA partial sums) is a file-scope
 * __device__ array of DMB_KL_STATIC_MAX_B * 4 doubles (128 KB, held by every GPU that loads the
 * library until exit), and the call makes no cudaMalloc / cudaFree / allocating library call.
 * device: CUDA ordinal (made current).  mu / logvar (B, per_sample) fp32 [dev] (logvar already
 * clamped, as dmb_vae_encode returns it); kl_sample (B) fp32 [dev], caller-allocated.
 * B > DMB_KL_STATIC_MAX_B is rejected with an error (split the batch).  All calls on a device
 * share the one array, which keeps the last call's contents (every entry a call reads is written
 * by that call first, so nothing needs clearing): calls that overlap on the GPU -- different
 * streams, or concurrent graph branches -- race; serialise them on one stream. */
#define DMB_KL_STATIC_MAX_B 4096
int dmb_kl_per_sample_static(int device, const float* mu, const float* logvar, int B, int64_t per_sample,
                             float kl_div, float* kl_sample, void* stream);

/* Scoring of generated images against ground-truth masks on the device:
 * eval_iou_noise.py:77-94 load_binary_mask (PIL "L" conversion + threshold), :162-171 exact
 * Euclidean distance transform, :185-232 gaussian_weighted_recall / far_noise_ratio, :239-272
 * compute_metrics.  pred (B, pc, H, W) u8 [dev], gt (B, gc, H, W) u8 [dev]; channels 3 = RGB image,
 * 1 = grey image (both binarised: L < threshold if invert else L >= threshold), 0 = mask (non-zero
 * = foreground).  H, W <= 254.  counts (B,6) int64 [dev] = inter, union, gt_area, pred_area, fp,
 * far_noise (pred pixels farther than sigma from the ground truth); weighted_hit (B,) f64 [dev] =
 * sum over pred pixels of exp(-d^2 / (2 sigma^2)).  The caller forms the reference's float64
 * ratios from these (iou = inter/union ...), which makes every ratio but gauss_recall bit-exact. */
int dmb_score_masks(dmb_ctx* ctx, const uint8_t* pred, int pred_channels, const uint8_t* gt, int gt_channels,
                    int B, int H, int W, int threshold, int invert, double sigma, int64_t* counts,
                    double* weighted_hit, void* stream);

/* Output path: writes n PNG files from a HOST batch of planar u8 images
 * (n, C, H, W), C = 3 (RGB) or 1 (grey), in parallel on `threads` host threads (<= 0: all).
 * Replaces the per-image `.cpu()` -> ToPILImage -> image.save loop of Diffuser.reverse_to_img +
 * Utils.saveImages (diff.py:58-64, utils.py:216-224) and the per-step saves of
 * generate_steps.py:47-66, 164-167; the caller keeps the reference's file names.  The files
 * decode to exactly the input pixels (8-bit, no interlace); their bytes are not PIL's.
 * Needs no context and no GPU. */
int dmb_write_png_batch(const uint8_t* images, int n, int C, int H, int W, const char* const* paths, int threads,
                        int compression_level);

/* The Philox noise dmb_denoise_cond_step / dmb_sample_loop would draw for step t:
 * out (B,4,28,28) [dev]. */
int dmb_philox_normal(dmb_ctx* ctx, float* out, int B, uint64_t seed, uint64_t sample_offset, int t,
                      void* stream);

/* Number of kernels launched by this context since the last reset (graph replays count the
 * kernels they contain). */
int64_t dmb_launch_count(dmb_ctx* ctx);
int dmb_reset_launch_count(dmb_ctx* ctx);

/* Per-kernel device timing (CUDA events around every launch of the EAGER entry points:
 * dmb_unet_forward, dmb_denoise_cond_step, dmb_vae_decode).  dmb_get_kernel_timing synchronises
 * and writes a JSON array [{label,count,ms,flops,bytes}] (algorithmic FLOPs / bytes per label)
 * into buf; *needed receives the size including the terminator. */
int dmb_set_kernel_timing(dmb_ctx* ctx, int on);
int dmb_get_kernel_timing(dmb_ctx* ctx, char* buf, int64_t capacity, int64_t* needed);

/* ---- test / debug entry points (unit-level parity of the kernels through the same ABI) ---- */

/* Named intermediate activation of the most recent forward / step as (n,C,H,W) fp32 [dev].
 * Names: emb, embproj, inc, down1..3, sa1..6, bot1..3, up1..3.  shape4 receives n,C,H,W. */
int dmb_debug_get_tap(dmb_ctx* ctx, const char* name, float* out, int64_t capacity, int64_t* shape4,
                      void* stream);
/* 3x3 / pad 1 / no-bias convolution (nn.Conv2d, models/unet_cond.py:19): x (B,Cin,H,W) [dev],
 * w (Cout,Cin,3,3) [any], out (B,Cout,H,W) [dev].  Synchronises the stream. */
int dmb_op_conv3x3(dmb_ctx* ctx, const float* x, const float* w, int B, int Cin, int Cout, int H, int W,
                   float* out, void* stream);
/* out (rows,N) = act(x (rows,K) @ w (N,K)^T + bias); act 0 none, 1 GELU(erf).  All [any]
 * except out [dev].  Synchronises the stream. */
int dmb_op_linear(dmb_ctx* ctx, const float* x, const float* w, const float* bias, int rows, int K, int N,
                  int act, float* out, void* stream);
/* ConvTranspose2d(4, stride 2, pad 1) (models/vae.py:38): x (B,Cin,H,W) [dev], w (Cin,Cout,4,4)
 * [any], bias (Cout) [any] or NULL, out (B,Cout,2H,2W) [dev].  Synchronises the stream. */
int dmb_op_conv_transpose4x4(dmb_ctx* ctx, const float* x, const float* w, const float* bias, int B, int Cin,
                             int Cout, int H, int W, float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DMB200_H_ */
