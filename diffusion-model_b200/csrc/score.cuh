// On-GPU scoring of generated images against ground-truth masks:
// eval_iou_noise.py:77-94 (load_binary_mask), :162-171 (_distance_map_to_gt, scipy backend),
// :185-232 (gaussian_weighted_recall, far_noise_ratio), :239-272 (compute_metrics).
//
// One CTA per image pair.  Everything integer is exact: the masks, the five areas, and the
// SQUARED Euclidean distance of every predicted pixel to the nearest ground-truth pixel (exact
// two-pass transform: per-column 1-D distance, then min over the row of dx^2 + g^2).  Only the
// predicted pixels need a distance (both distance-based metrics sum over pred == 1), so the row
// pass is done warp-cooperatively per predicted pixel: 32 lanes split the row, shuffle-min.
// The host turns the counters into the reference's float64 ratios, so iou / gt_iou /
// far_noise_ratio are bit-identical to compute_metrics; gauss_recall differs only by the order
// of a float64 sum.
//
// scipy quirk mirrored on purpose: distance_transform_edt of an input WITHOUT any background
// (ground truth empty) returns sqrt((y+1)^2 + x^2), as if one background pixel sat at
// (-1, 0) (scipy 1.18; pinned by tests/test_oracle_golden.py::test_scoring_oracle_matches_scipy).
// far_noise_ratio reaches that case (gaussian_weighted_recall returns 1.0 before it).
#pragma once

#include "common.cuh"

namespace dmb {

constexpr int kScoreThreads = 256;
constexpr int kScoreMaxSide = 254;   // column distances are kept in one byte (255 = no ground truth in the column)

struct ScoreParams {
  const uint8_t* pred;   // (B, pc, H, W) u8: pc = 3 RGB image, 1 grey image, 0 mask (non-zero = foreground)
  const uint8_t* gt;     // (B, gc, H, W) likewise
  int pred_channels, gt_channels;
  int B, H, W;
  int threshold;         // grey level: foreground = L < threshold (invert) or L >= threshold
  int invert;
  double sigma;
  long long* counts;     // (B, 6): inter, union, gt_area, pred_area, fp, far_noise
  double* weighted_hit;  // (B,): sum over pred of exp(-d^2 / (2 sigma^2))
};

// PIL's RGB -> L (ImagingConvert rgb2l): (R*19595 + G*38470 + B*7471 + 0x8000) >> 16
__device__ __forceinline__ uint32_t pil_luma(uint32_t r, uint32_t g, uint32_t b) {
  return (r * 19595u + g * 38470u + b * 7471u + 0x8000u) >> 16;
}

__device__ __forceinline__ void binarise_plane(const uint8_t* __restrict__ src, int channels, int HW, int threshold,
                                               int invert, uint8_t* __restrict__ dst) {
  // 4 pixels per thread per iteration (HW is a multiple of 4 for every map this path sees; the
  // tail is handled bytewise)
  const bool vec = (HW & 3) == 0 && (reinterpret_cast<uintptr_t>(src) & 3) == 0;
  const int n4 = vec ? HW >> 2 : 0;
  for (int i = threadIdx.x; i < n4; i += kScoreThreads) {
    uint32_t out = 0;
    if (channels == 3) {
      const uint32_t r = *reinterpret_cast<const uint32_t*>(src + 4 * i);
      const uint32_t g = *reinterpret_cast<const uint32_t*>(src + HW + 4 * i);
      const uint32_t b = *reinterpret_cast<const uint32_t*>(src + 2 * HW + 4 * i);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t L = pil_luma((r >> (8 * k)) & 255u, (g >> (8 * k)) & 255u, (b >> (8 * k)) & 255u);
        const bool fg = invert ? (static_cast<int>(L) < threshold) : (static_cast<int>(L) >= threshold);
        out |= (fg ? 1u : 0u) << (8 * k);
      }
    } else {
      const uint32_t v = *reinterpret_cast<const uint32_t*>(src + 4 * i);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int L = static_cast<int>((v >> (8 * k)) & 255u);
        const bool fg = channels == 0 ? (L != 0) : (invert ? (L < threshold) : (L >= threshold));
        out |= (fg ? 1u : 0u) << (8 * k);
      }
    }
    *reinterpret_cast<uint32_t*>(dst + 4 * i) = out;
  }
  for (int i = 4 * n4 + threadIdx.x; i < HW; i += kScoreThreads) {
    int L;
    if (channels == 3) L = static_cast<int>(pil_luma(src[i], src[HW + i], src[2 * HW + i]));
    else L = src[i];
    dst[i] = channels == 0 ? (L != 0) : (invert ? (L < threshold) : (L >= threshold));
  }
}

static __global__ void __launch_bounds__(kScoreThreads)
score_masks_kernel(const ScoreParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ int cnt[5];            // inter, union, gt_area, pred_area, fp
  __shared__ int far_cnt;
  __shared__ double hit_part[kScoreThreads / 32];
  const int b = blockIdx.x;
  const int H = p.H, W = p.W, HW = H * W;
  uint8_t* pm = smem_raw;                       // predicted mask, one byte per pixel
  uint8_t* g = smem_raw + ((HW + 15) & ~15);    // ground-truth mask, then per-column distance to it
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < 5) cnt[tid] = 0;
  if (tid == 0) far_cnt = 0;

  const int pc = p.pred_channels == 0 ? 1 : p.pred_channels, gc = p.gt_channels == 0 ? 1 : p.gt_channels;
  binarise_plane(p.pred + static_cast<size_t>(b) * pc * HW, p.pred_channels, HW, p.threshold, p.invert, pm);
  binarise_plane(p.gt + static_cast<size_t>(b) * gc * HW, p.gt_channels, HW, p.threshold, p.invert, g);
  __syncthreads();

  {  // areas
    int c_inter = 0, c_union = 0, c_gt = 0, c_pred = 0, c_fp = 0;
    for (int i = tid; i < HW; i += kScoreThreads) {
      const int a = g[i], q = pm[i];
      c_inter += a & q; c_union += a | q; c_gt += a; c_pred += q; c_fp += q & (a ^ 1);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      c_inter += __shfl_xor_sync(0xffffffffu, c_inter, o);
      c_union += __shfl_xor_sync(0xffffffffu, c_union, o);
      c_gt += __shfl_xor_sync(0xffffffffu, c_gt, o);
      c_pred += __shfl_xor_sync(0xffffffffu, c_pred, o);
      c_fp += __shfl_xor_sync(0xffffffffu, c_fp, o);
    }
    if (lane == 0) {
      atomicAdd(&cnt[0], c_inter); atomicAdd(&cnt[1], c_union); atomicAdd(&cnt[2], c_gt);
      atomicAdd(&cnt[3], c_pred); atomicAdd(&cnt[4], c_fp);
    }
  }
  __syncthreads();
  const bool gt_empty = cnt[2] == 0;

  // column pass (in place): g[y][x] = distance along the column to the nearest ground-truth pixel
  for (int x = tid; x < W; x += kScoreThreads) {
    int d = 255;
    for (int y = 0; y < H; ++y) {
      d = g[y * W + x] ? 0 : min(d + 1, 255);
      g[y * W + x] = static_cast<uint8_t>(d);
    }
    d = 255;
    for (int y = H - 1; y >= 0; --y) {
      const int down = g[y * W + x];
      d = down == 0 ? 0 : min(d + 1, 255);
      g[y * W + x] = static_cast<uint8_t>(min(d, down));
    }
  }
  __syncthreads();

  // row pass at the predicted pixels; a warp owns 32 consecutive pixels at a time
  const double two_s2 = 2.0 * (p.sigma * p.sigma);
  int far = 0;
  double hit = 0.0;
  for (int base = warp * 32; base < HW; base += kScoreThreads) {
    const int i = base + lane;
    uint32_t todo = __ballot_sync(0xffffffffu, i < HW && pm[i] != 0);
    while (todo) {
      const int src = __ffs(todo) - 1;
      todo &= todo - 1;
      const int pi = base + src;
      const int y = pi / W, x = pi - y * W;
      long long d2;
      if (gt_empty) {
        d2 = static_cast<long long>(y + 1) * (y + 1) + static_cast<long long>(x) * x;
      } else {
        int best = 0x7fffffff;
        const uint8_t* grow = g + y * W;
        for (int xx = lane; xx < W; xx += 32) {
          const int gy = grow[xx];
          if (gy != 255) best = min(best, (x - xx) * (x - xx) + gy * gy);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
        d2 = best;
      }
      if (lane == 0) {
        const double dist = sqrt(static_cast<double>(d2));
        far += dist > p.sigma ? 1 : 0;
        hit += exp(-(dist * dist) / two_s2);   // the reference's expression, in float64
      }
    }
  }
  if (lane == 0) {
    atomicAdd(&far_cnt, far);
    hit_part[warp] = hit;
  }
  __syncthreads();
  if (tid == 0) {
    long long* c = p.counts + static_cast<size_t>(b) * 6;
    c[0] = cnt[0]; c[1] = cnt[1]; c[2] = cnt[2]; c[3] = cnt[3]; c[4] = cnt[4]; c[5] = far_cnt;
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < kScoreThreads / 32; ++w) s += hit_part[w];
    p.weighted_hit[b] = s;
  }
}

inline size_t score_smem_bytes(int H, int W) { return 2 * static_cast<size_t>((H * W + 15) & ~15); }

}  // namespace dmb
