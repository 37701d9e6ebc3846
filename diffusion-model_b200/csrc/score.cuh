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

constexpr int kScoreThreads = 512;   // the per-pixel searches are latency chains: 32 warps per SM (two CTAs)
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

// foreground bytes (0 / 1) of four pixels packed in a word
__device__ __forceinline__ uint32_t fg4(uint32_t r, uint32_t g, uint32_t b, int channels, int threshold, int invert) {
  if (channels == 0) {   // mask bytes: non-zero -> 1 (OR-fold every byte into its lowest bit)
    uint32_t x = r | (r >> 4);
    x |= x >> 2;
    x |= x >> 1;
    return x & 0x01010101u;
  }
  uint32_t out = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int L = channels == 3 ? static_cast<int>(pil_luma((r >> (8 * k)) & 255u, (g >> (8 * k)) & 255u, (b >> (8 * k)) & 255u))
                                : static_cast<int>((r >> (8 * k)) & 255u);
    const bool fg = channels == 0 ? (L != 0) : (invert ? (L < threshold) : (L >= threshold));
    out |= (fg ? 1u : 0u) << (8 * k);
  }
  return out;
}

// One image plane set -> one byte per pixel in shared memory.  16 pixels per thread and load,
// two such vectors in flight per thread and plane (1024 threads per SM): the image is read once from HBM and this is
// the only HBM traffic of the kernel, so the loads have to cover the memory latency.
__device__ __forceinline__ void binarise_plane(const uint8_t* __restrict__ src, int channels, int HW, int threshold,
                                               int invert, uint8_t* __restrict__ dst) {
  const bool vec = (HW & 15) == 0 && (reinterpret_cast<uintptr_t>(src) & 15) == 0;
  const int n16 = vec ? HW >> 4 : 0;
  constexpr int U = 2;
  for (int i0 = threadIdx.x; i0 < n16; i0 += U * kScoreThreads) {
    uint4 r[U], g[U], b[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int i = i0 + u * kScoreThreads;
      if (i < n16) {
        r[u] = *reinterpret_cast<const uint4*>(src + 16 * i);
        if (channels == 3) {
          g[u] = *reinterpret_cast<const uint4*>(src + HW + 16 * i);
          b[u] = *reinterpret_cast<const uint4*>(src + 2 * HW + 16 * i);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int i = i0 + u * kScoreThreads;
      if (i < n16) {
        uint4 o;
        o.x = fg4(r[u].x, g[u].x, b[u].x, channels, threshold, invert);
        o.y = fg4(r[u].y, g[u].y, b[u].y, channels, threshold, invert);
        o.z = fg4(r[u].z, g[u].z, b[u].z, channels, threshold, invert);
        o.w = fg4(r[u].w, g[u].w, b[u].w, channels, threshold, invert);
        *reinterpret_cast<uint4*>(dst + 16 * i) = o;
      }
    }
  }
  for (int i = 16 * n16 + threadIdx.x; i < HW; i += kScoreThreads) {
    int L;
    if (channels == 3) L = static_cast<int>(pil_luma(src[i], src[HW + i], src[2 * HW + i]));
    else L = src[i];
    dst[i] = channels == 0 ? (L != 0) : (invert ? (L < threshold) : (L >= threshold));
  }
}

static __global__ void __launch_bounds__(kScoreThreads, 2)
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

  {  // areas: 16 pixels per load, the 0 / 1 bytes counted with popc
    int c_inter = 0, c_union = 0, c_gt = 0, c_pred = 0;
    const int n16 = HW >> 4;
    for (int i = tid; i < n16; i += kScoreThreads) {
      const uint4 a = *reinterpret_cast<const uint4*>(g + 16 * i);
      const uint4 q = *reinterpret_cast<const uint4*>(pm + 16 * i);
      c_inter += __popc(a.x & q.x) + __popc(a.y & q.y) + __popc(a.z & q.z) + __popc(a.w & q.w);
      c_union += __popc(a.x | q.x) + __popc(a.y | q.y) + __popc(a.z | q.z) + __popc(a.w | q.w);
      c_gt += __popc(a.x) + __popc(a.y) + __popc(a.z) + __popc(a.w);
      c_pred += __popc(q.x) + __popc(q.y) + __popc(q.z) + __popc(q.w);
    }
    for (int i = 16 * n16 + tid; i < HW; i += kScoreThreads) {
      const int a = g[i], q = pm[i];
      c_inter += a & q; c_union += a | q; c_gt += a; c_pred += q;
    }
    int c_fp = c_pred - c_inter;   // pred and not gt
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

  // column pass (in place): g[y][x] = distance along the column to the nearest ground-truth pixel.
  // The loads do not depend on the running distance, so eight rows are fetched at a time and
  // only the register chain is serial.
  for (int x = tid; x < W; x += kScoreThreads) {
    uint8_t* col = g + x;
    int d = 255, y = 0;
    for (; y + 8 <= H; y += 8, col += 8 * W) {
      int v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = col[k * W];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        d = v[k] ? 0 : min(d + 1, 255);
        col[k * W] = static_cast<uint8_t>(d);
      }
    }
    for (; y < H; ++y, col += W) {
      d = col[0] ? 0 : min(d + 1, 255);
      col[0] = static_cast<uint8_t>(d);
    }
    d = 255;
    y = H;
    for (; y >= 8; y -= 8) {
      col -= 8 * W;
      int v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = col[(7 - k) * W];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        d = v[k] == 0 ? 0 : min(d + 1, 255);
        col[(7 - k) * W] = static_cast<uint8_t>(min(d, v[k]));
      }
    }
    for (; y > 0; --y) {
      col -= W;
      const int v = col[0];
      d = v == 0 ? 0 : min(d + 1, 255);
      col[0] = static_cast<uint8_t>(min(d, v));
    }
  }
  __syncthreads();

  // row pass at the predicted pixels; a warp owns 32 consecutive pixels at a time: the squared
  // distance of each predicted pixel is found by all 32 lanes (they split the row, shuffle-min)
  // and kept by the pixel's own lane, which then does the float64 part for it
  const double two_s2 = 2.0 * (p.sigma * p.sigma);
  int far = 0;
  double hit = 0.0;
  // (a lane first looks at four pixels at once: most 128-pixel spans of a line drawing are empty)
  const bool words = (HW & 3) == 0;
  const int span = words ? 128 : 32;
  for (int base0 = warp * span; base0 < HW; base0 += (kScoreThreads / 32) * span) {
    uint32_t w4 = 0;
    if (words) {
      const int wi = base0 + 4 * lane;
      if (wi < HW) w4 = *reinterpret_cast<const uint32_t*>(pm + wi);
      if (!__any_sync(0xffffffffu, w4 != 0)) continue;
    }
    for (int sub = 0; sub < (words ? 4 : 1); ++sub) {
      // sub-span `sub`: pixel base0 + 4 * lane + sub belongs to this lane
      const int i = words ? base0 + 4 * lane + sub : base0 + lane;
      const bool mine = words ? ((w4 >> (8 * sub)) & 255u) != 0 : (i < HW && pm[i] != 0);
      uint32_t todo = __ballot_sync(0xffffffffu, mine);
      long long d2_mine = 0;
      while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const int pi = words ? base0 + 4 * src + sub : base0 + src;
        const int y = pi / W, x = pi - y * W;
        long long d2;
        if (gt_empty) {
          d2 = static_cast<long long>(y + 1) * (y + 1) + static_cast<long long>(x) * x;
        } else {
          // the 32 columns around x first: a candidate <= 16^2 there cannot be beaten by a
          // column at least 16 away, which is the common case (predictions lie near the truth)
          int best = 0x7fffffff;
          const uint8_t* grow = g + y * W;
          {
            const int xx = x - 16 + lane;
            if (xx >= 0 && xx < W) {
              const int gy = grow[xx];
              if (gy != 255) best = (x - xx) * (x - xx) + gy * gy;
            }
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
          if (best > 256) {   // (uniform: every lane holds the minimum) -- search the whole row
            for (int xx = lane; xx < W; xx += 32) {
              const int gy = grow[xx];
              if (gy != 255) best = min(best, (x - xx) * (x - xx) + gy * gy);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
          }
          d2 = best;
        }
        if (lane == src) d2_mine = d2;
      }
      if (mine) {
        const double dist = sqrt(static_cast<double>(d2_mine));
        far += dist > p.sigma ? 1 : 0;
        hit += exp(-(dist * dist) / two_s2);   // the reference's expression, in float64
      }
    }
  }
  // fixed-order reduction over the lanes, then over the warps: the sum is deterministic
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    far += __shfl_xor_sync(0xffffffffu, far, o);
    hit += __shfl_xor_sync(0xffffffffu, hit, o);
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
