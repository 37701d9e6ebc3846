// Multi-tap implicit GEMM on tcgen05 tensor cores with TMEM accumulators, fed by TMA.
//
//   D[r, co] = sum_{tap} sum_{ci} A[r + shift(tap), ci] * Wt[tap][co, ci]
//
// A is an activation in padded pixel-matrix layout (common.cuh), so a 3x3 / pad-1
// convolution is 9 taps whose operand tiles are plain 128-row boxes of the same 2D
// tensor at shifted row coordinates: one TMA box per (tap, K chunk); rows outside the
// tensor arrive as zeros.  A 1x1 conv / Linear is the one-tap case; a 4x4 stride-2
// transposed conv is four 4-tap launches (one per output phase) with a scattering row map.
//
// CTA = 384 threads, one CTA (kCtas = 1) or one 2-CTA cluster (kCtas = 2, cta_group::2, a
// 256 x BLOCK_N tile whose B operand is split between the pair) per SM, persistent over tiles:
//   warp 0    TMA producer : warp-uniform loop, elect_one() around the asm only; fills the smem
//                            ring, kGroup (tap, K chunk) operand pairs per stage, signals full[s]
//                            (of the leader CTA for pairs) via complete_tx
//   warp 1    MMA issuer   : tcgen05.mma SS (leader CTA only for pairs), commits to empty[s]
//                            (multicast to both CTAs) and to tmem_full[a]
//   warp 2    TMEM allocator
//   warps 4-7 / 8-11  two epilogue warpgroups, one per TMEM accumulator stage (tile parity):
//                            tcgen05.ld lane == row, bias / GELU / residual, GroupNorm
//                            statistics (fixed-point per-sample sums), then the tile goes through the
//                            warpgroup's own 128B-swizzled staging buffer and out with one TMA
//                            store per 128-byte column block (pad-pixel rows are stored as zeros).
// Two TMEM accumulator stages let the epilogue of tile i overlap the MMAs of tile i+1 and the
// two epilogues overlap each other.  Scattering outputs (conv-transpose phases), the image
// epilogue and tiles narrower than 128 bytes keep per-thread stores.
// Why kGroup: one producer <-> issuer mbarrier round trip costs ~500 cycles of issue-thread
// time whatever the payload (measured), so a stage carries >= ~512 cycles of MMA work.
#pragma once

#include "common.cuh"

namespace dmb {

constexpr int kMaxTaps = 9;
constexpr int kTileM = 128;
constexpr int kMaxStages = 8;
constexpr int kMaxGroup = 4;   // (tap, K chunk) operand pairs per pipeline stage

enum EpiMode : int { EPI_STORE = 0, EPI_IMAGE = 1 };

struct GemmParams {
  int num_taps;
  int a_shift[kMaxTaps];  // row shift of the A box per tap
  int b_row[kMaxTaps];    // first row of the tap's [Cout, Cin] block in the weight matrix
  int k_chunks;           // Cin / chunkK
  int m_tiles, n_tiles, block_n, stages;
  int group;              // (tap, K chunk) operand pairs per pipeline stage: one barrier hand-shake
                          // (~500 cycles of issue-thread time) must cover >= that much MMA time
  uint32_t idesc;
  uint32_t tmem_cols;

  Geo geo;     // geometry of the A rows (and of the output rows under the identity map)
  int epi;     // EpiMode
  void* out;   // T [rows, ldc]
  int ldc;
  int n_valid;            // number of real output columns (<= n_tiles * block_n)
  const float* bias;      // [n_valid] or null
  int act;                // 0 none, 1 GELU(erf)
  const void* residual;   // T [rows, ldc] or null (added after the activation)
  long long* sample_stats;  // [2 * samples] fixed-point (sum, sum of squares) of the output per sample for the
                            // following GroupNorm(1, C) (StatAcc / stat_commit, common.cuh), or null
  int map_mode;           // 0 identity; 1 = out pixel (2y+up_a, 2x+up_b) of geo_out
  Geo geo_out;
  int up_a, up_b;
  int tma_store;          // 1: epilogue stores through smem + TMA (tmap_out); 0: per-thread stores
  int out_col0;           // EPI_STORE: first output column of the launch (the phase slot of a phase-planar
                          // ConvTranspose output, engine.cu build_vae_plan); bias / residual are not shifted
  // pixel-major 3x3 conv (small maps at large batch, see PixTile): tmap_a / tmap_out are the 3D
  // (channel, PM row of the sample, sample) maps, M runs over SAMPLES, a tile belongs to one
  // output pixel and its taps are the pixel's in-image neighbours
  int pix;
  int pix_m_tiles;             // tiles (of kTileM * kCtas samples) along the batch
  unsigned char pix_order[196]; // output pixels (maps up to 14 x 14), those with the most neighbours first (longest tiles first)
  float* img_f32;         // EPI_IMAGE: (n, n_valid, H, W) planar fp32, sigmoid applied
  uint8_t* img_u8;        // EPI_IMAGE: same, x255 clamp truncate
};

constexpr int kGemmThreads = 384;                // kAcc = 2: 4 control warps + 2 epilogue warpgroups
constexpr int kStagingBytes = 2 * kTileM * 128;  // one 128-row x 128-byte sub-tile per epilogue warpgroup
constexpr int gemm_threads(int acc) { return 128 + 128 * acc; }
constexpr int gemm_staging_bytes(int acc) { return acc * kTileM * 128; }

// Pixel-major tile of a 3x3 / pad-1 convolution on a small map.  The PM layout spends
// (H+1)(W+1)/(HW) MMA rows on padding (1.31x at 7x7, 1.78x at 3x3) and every output row runs all
// nine taps, although border pixels have only 4 or 6 neighbours inside the image.  When the
// batch is large the GEMM is turned around: the M dimension runs over SAMPLES, a tile belongs to
// ONE output pixel p, and its K loop visits only the neighbours q of p that exist: A tile = pixel
// q of 128 consecutive samples (a {64 ch, 1 row, 128 samples} box of the 3D tensor map), B tile =
// the weight block of tap q - p.  MMA work = (3H-2)(3W-2) pixel pairs per sample instead of
// 9(H+1)(W+1): 1.6x less at 7x7, 2.9x less at 3x3.
struct PixTile {
  int n_tile, m_tile, py, px, prow;   // prow = PM row of the output pixel inside its sample
  int dy_lo, dx_lo, ndx, ntap;        // neighbours: dy in [dy_lo, dy_lo + ntap / ndx), dx in [dx_lo, dx_lo + ndx)
};
__device__ __forceinline__ PixTile pix_tile(const GemmParams& p, int tile) {
  PixTile t;
  t.n_tile = tile % p.n_tiles;
  const int rest = tile / p.n_tiles;
  t.m_tile = rest % p.pix_m_tiles;
  const int pix = p.pix_order[rest / p.pix_m_tiles];
  t.py = pix / p.geo.W;
  t.px = pix - t.py * p.geo.W;
  t.prow = (t.py + 1) * p.geo.Wp + t.px;
  t.dy_lo = t.py == 0 ? 1 : 0;
  t.dx_lo = t.px == 0 ? 1 : 0;
  const int ndy = 3 - t.dy_lo - (t.py == p.geo.H - 1 ? 1 : 0);
  t.ndx = 3 - t.dx_lo - (t.px == p.geo.W - 1 ? 1 : 0);
  t.ntap = ndy * t.ndx;
  return t;
}

// kCtas == 2: launched as clusters of two CTAs (cta_group::2).  The pair owns a 256-row tile:
// CTA rank r loads rows [m0 + 128 r, +128) of A and rows [n0 + r BLOCK_N/2, +BLOCK_N/2) of the
// weight tile, keeps its 128 accumulator lanes in its own TMEM and runs its own epilogue;
// rank 0 issues the M=256 MMAs and multicasts the commits.  Per SM this halves the B
// operand's shared-memory traffic, which is what bounds the single-CTA kernel.
//
// kAcc = number of TMEM accumulator stages = number of epilogue warpgroups (2 or 4).  Short-K
// launches (the Linear layers) are bound by the epilogue's dependent instruction stream: with
// two warpgroups an SM scheduler holds two epilogue warps and issues in 23-38 % of its cycles
// (ncu, profiles/r1_v3_ncu_full_up_path.md); four warpgroups (N tile <= 128) double that.
//
// kRow3 (3x3 convolutions): the three taps of one kernel row are row shifts by -1 / 0 / +1 of the
// same PM rows, so their A operands overlap in 127 of 128 rows.  A stage then holds ONE 130-row
// A window (rows m0 + dy Wp - 1 ...) and the weight tiles of the row's three taps; the MMAs of
// tap dx read the window at a start address 128 dx bytes further.  The 128B swizzle is an XOR
// on shared-memory ADDRESS bits (TMA wrote the window with the same rule), so the shifted start
// needs nothing else: descriptor base_offset stays 0 (measured on B200: base_offset = dx gives
// wrong results, 0 is exact); for the same reason the windows need 128-byte, not 1024-byte,
// alignment and are packed back to back behind the (1024-aligned) weight tiles.  The A operand crosses the L2 -> SM fabric 3 times per K chunk
// instead of 9, and that fabric, not the tensor pipe, is what bounds these launches (ncu:
// 10-14 TB/s of xbar reads at 55-72 % tensor-pipe activity).
constexpr int kHaloRows = kTileM + 2;
constexpr uint32_t kHaloBytes = kHaloRows * 128;   // windows are packed back to back: the swizzle follows the address
template <typename T, int kCtas, int kGroup, int kAcc = 2, bool kRow3 = false>
__global__ void __launch_bounds__(gemm_threads(kAcc), 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const __grid_constant__ CUtensorMap tmap_out, const GemmParams p) {
  using Tr = ElemTraits<T>;
  constexpr int kSubN = Tr::kChunkK;   // columns per 128-byte sub-tile (64 bf16 / 32 fp32)
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[kMaxStages];
  __shared__ __align__(8) uint64_t empty_bar[kMaxStages];
  __shared__ __align__(8) uint64_t tmem_full_bar[kAcc];
  __shared__ __align__(8) uint64_t tmem_empty_bar[kAcc];
  __shared__ uint32_t tmem_base_smem;

  pdl_trigger();
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // 128B-swizzled operand tiles need 1024-byte alignment.
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  const uint32_t rank = (kCtas == 2) ? cluster_ctarank() : 0u;
  const int num_cl = gridDim.x / kCtas, cl = blockIdx.x / kCtas;   // tile loop runs per cluster
  const uint32_t a_bytes = kTileM * 128;
  const uint32_t b_bytes = static_cast<uint32_t>(p.block_n / kCtas) * 128;
  const uint32_t pair_bytes = a_bytes + b_bytes;                       // one (tap, K chunk) operand pair
  // kRow3: a stage = kGroup consecutive (kernel row, K chunk) steps -- a 130-row A window and the three weight
  // tiles of the row's taps each.  One mbarrier hand-shake costs the issuing warps ~500 cycles and a step is only
  // 12 MMAs (3 N cycles), so narrow layers take 3 or 2 steps per hand-shake: where a launch has few tiles per
  // CTA (batch 1; the 64 -> 3 output conv of the VAE) the chain of hand-shakes is what it waits for.
  const uint32_t step_bytes = kHaloBytes + 3u * b_bytes;
  const uint32_t stage_bytes = kRow3 ? static_cast<uint32_t>(kGroup) * step_bytes : static_cast<uint32_t>(kGroup) * pair_bytes;
  // shared memory: [stage ring][epilogue staging]; kRow3: [staging][weight tiles of all stages][A windows]
  const uint32_t staging_bytes = p.tma_store ? static_cast<uint32_t>(gemm_staging_bytes(kAcc)) : 0u;
  uint8_t* staging = kRow3 ? smem : smem + static_cast<size_t>(p.stages) * stage_bytes;  // 1024-aligned (stage_bytes % 1024 == 0)
  uint8_t* ring_b = smem + staging_bytes;                                          // kRow3: 3 weight tiles per step
  uint8_t* ring_a = ring_b + static_cast<size_t>(p.stages) * kGroup * 3u * b_bytes;  // kRow3: one window per step

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    if (p.tma_store) tma_prefetch_desc(&tmap_out);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < kAcc; ++a) {
      mbar_init(&tmem_full_bar[a], 1);
      mbar_init(&tmem_empty_bar[a], 4 * kCtas);  // one arrive per warp of the epilogue group (of each CTA)
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    if constexpr (kCtas == 2) tmem_alloc_pair(&tmem_base_smem, p.tmem_cols);
    else tmem_alloc(&tmem_base_smem, p.tmem_cols);
  }
  tc_fence_before();
  if constexpr (kCtas == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  pdl_wait();   // barriers, TMEM and descriptors are set up; from here on global memory is touched

  const int total_tiles = p.m_tiles * p.n_tiles;   // (pixel-major: m_tiles = pixels x batch tiles)
  const int n_iter_all = p.num_taps * p.k_chunks;

  // The producer and MMA warps run their loops with ALL 32 lanes (warp-uniform control flow
  // and operands) and elect one lane only around the asynchronous instructions: ptxas then
  // keeps descriptors, barrier and TMEM addresses in uniform registers instead of wrapping
  // every UTMALDG / UTCHMMA in an ELECT + R2UR.BROADCAST waterfall (~60 cycles per MMA).
  if (warp == 0) {
    // ------------------------------------------------------------- TMA producer
    uint32_t stage = 0, phase = 0;
    const uint32_t lead_full0 = (kCtas == 2) ? mapa_shared(smem_u32(&full_bar[0]), 0) : 0u;
    for (int tile = cl; tile < total_tiles; tile += num_cl) {
      int n_tile = tile % p.n_tiles;
      int m_tile = tile / p.n_tiles;
      int n_iter = n_iter_all;
      PixTile pt{};
      if (!kRow3 && p.pix) {
        pt = pix_tile(p, tile);
        n_tile = pt.n_tile; m_tile = pt.m_tile;
        n_iter = pt.ntap * p.k_chunks;
      }
      const int m0 = (m_tile * kCtas + static_cast<int>(rank)) * kTileM;
      const int n0 = n_tile * p.block_n + static_cast<int>(rank) * (p.block_n / kCtas);
      // operand coordinates of a tap: generic = (shifted PM row, weight row); pixel-major = (PM row
      // of the tap's neighbour pixel inside the sample, weight row); the sample coordinate is m0
      auto tap_coords = [&](int tap, int& a_c, int& b_c) {
        if (!kRow3 && p.pix) {
          const int dyi = tap / pt.ndx;
          const int dy = pt.dy_lo + dyi, dx = pt.dx_lo + (tap - dyi * pt.ndx);
          a_c = pt.prow + (dy - 1) * p.geo.Wp + (dx - 1);
          b_c = p.b_row[dy * 3 + dx] + n0;
        } else {
          a_c = m0 + p.a_shift[tap];
          b_c = p.b_row[tap] + n0;
        }
      };
      if constexpr (kRow3) {
        // step = (kernel row dy, K chunk): the 130-row window + the weight tiles of its three taps; kGroup steps per stage
        int dy = 0, kc = 0;
        for (int it0 = 0; it0 < 3 * p.k_chunks; it0 += kGroup) {
          uint8_t* sa0 = ring_a + stage * (kGroup * kHaloBytes);
          uint8_t* sb0 = ring_b + stage * (kGroup * 3u) * b_bytes;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          const bool issue = elect_one();
          if (issue && (kCtas == 1 || rank == 0)) mbar_expect_tx(&full_bar[stage], stage_bytes * kCtas);
#pragma unroll
          for (int gi = 0; gi < kGroup; ++gi) {
            const int a_row = m0 + p.a_shift[3 * dy + 1] - 1;
            const int k_col = kc * Tr::kChunkK;
            const int b0 = p.b_row[3 * dy] + n0, b1 = p.b_row[3 * dy + 1] + n0, b2 = p.b_row[3 * dy + 2] + n0;
            if (++kc == p.k_chunks) { kc = 0; ++dy; }
            uint8_t* sa = sa0 + gi * kHaloBytes;
            uint8_t* sb = sb0 + gi * 3u * b_bytes;
            if (issue) {
              if constexpr (kCtas == 2) {
                const uint32_t lead_bar = lead_full0 + stage * 8;
                tma_load_2d_pair(sa, &tmap_a, lead_bar, k_col, a_row);
                tma_load_2d_pair(sb, &tmap_b, lead_bar, k_col, b0);
                tma_load_2d_pair(sb + b_bytes, &tmap_b, lead_bar, k_col, b1);
                tma_load_2d_pair(sb + 2 * b_bytes, &tmap_b, lead_bar, k_col, b2);
              } else {
                tma_load_2d(sa, &tmap_a, &full_bar[stage], k_col, a_row);
                tma_load_2d(sb, &tmap_b, &full_bar[stage], k_col, b0);
                tma_load_2d(sb + b_bytes, &tmap_b, &full_bar[stage], k_col, b1);
                tma_load_2d(sb + 2 * b_bytes, &tmap_b, &full_bar[stage], k_col, b2);
              }
            }
          }
          __syncwarp();
          if (++stage == static_cast<uint32_t>(p.stages)) { stage = 0; phase ^= 1; }
        }
      }
      // (tap, K chunk) pairs are walked in order; the per-tap row offsets are looked up when
      // the tap changes, i.e. long before the next barrier wake-up needs them
      int tap = 0, kc = 0;
      int a_cur, b_cur;
      tap_coords(0, a_cur, b_cur);
      const bool pix = !kRow3 && p.pix;
      if constexpr (!kRow3)
      for (int it0 = 0; it0 < n_iter; it0 += kGroup) {
        int a_row[kGroup], b_row[kGroup], k_col[kGroup];
#pragma unroll
        for (int gi = 0; gi < kGroup; ++gi) {
          a_row[gi] = a_cur; b_row[gi] = b_cur; k_col[gi] = kc * Tr::kChunkK;
          if (it0 + gi < n_iter && ++kc == p.k_chunks) {
            kc = 0;
            ++tap;
            if (tap * p.k_chunks < n_iter) tap_coords(tap, a_cur, b_cur);
          }
        }
        const int cnt = min(kGroup, n_iter - it0);
        uint8_t* sa = smem + stage * stage_bytes;
        mbar_wait(&empty_bar[stage], phase ^ 1);
        if (elect_one()) {
          if (kCtas == 1 || rank == 0) mbar_expect_tx(&full_bar[stage], static_cast<uint32_t>(cnt) * pair_bytes * kCtas);
#pragma unroll
          for (int gi = 0; gi < kGroup; ++gi) {
            if (gi < cnt) {
              uint8_t* sp = sa + gi * pair_bytes;
              if constexpr (kCtas == 2) {
                // both CTAs' bytes are accounted on the leader's barrier
                const uint32_t lead_bar = lead_full0 + stage * 8;
                if (pix) tma_load_3d_pair(sp, &tmap_a, lead_bar, k_col[gi], a_row[gi], m0);
                else tma_load_2d_pair(sp, &tmap_a, lead_bar, k_col[gi], a_row[gi]);
                tma_load_2d_pair(sp + a_bytes, &tmap_b, lead_bar, k_col[gi], b_row[gi]);
              } else {
                if (pix) tma_load_3d(sp, &tmap_a, &full_bar[stage], k_col[gi], a_row[gi], m0);
                else tma_load_2d(sp, &tmap_a, &full_bar[stage], k_col[gi], a_row[gi]);
                tma_load_2d(sp + a_bytes, &tmap_b, &full_bar[stage], k_col[gi], b_row[gi]);
              }
            }
          }
        }
        __syncwarp();
        if (++stage == static_cast<uint32_t>(p.stages)) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------- MMA issuer
    if (rank == 0) {
      uint32_t stage = 0, phase = 0;
      uint32_t tile_count = 0;
      for (int tile = cl; tile < total_tiles; tile += num_cl, ++tile_count) {
        const uint32_t as = tile_count % kAcc;
        const uint32_t aph = (tile_count / kAcc) & 1;
        mbar_wait(&tmem_empty_bar[as], aph ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * p.block_n;
        const int n_iter = (!kRow3 && p.pix) ? pix_tile(p, tile).ntap * p.k_chunks : n_iter_all;
        if constexpr (kRow3) {
          const int n3 = 3 * p.k_chunks;   // a multiple of kGroup (host)
          for (int it0 = 0; it0 < n3; it0 += kGroup) {
            const uint32_t sa0 = smem_u32(ring_a + stage * (kGroup * kHaloBytes));
            const uint32_t sb0 = smem_u32(ring_b + stage * (kGroup * 3u) * b_bytes);
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            const bool issue = elect_one();
#pragma unroll
            for (int gi = 0; gi < kGroup; ++gi) {
              const uint64_t adesc0 = umma_desc_sw128(sa0 + gi * kHaloBytes);
              const uint64_t bdesc0 = umma_desc_sw128(sb0 + gi * 3u * b_bytes);
              const uint64_t bstep = static_cast<uint64_t>(b_bytes >> 4);
              if (issue) {
#pragma unroll
                for (int dx = 0; dx < 3; ++dx) {
                  // window rows dx .. dx + 127: start address + 128 dx bytes (16-byte units)
                  const uint64_t ad = adesc0 + static_cast<uint64_t>(dx * 8);
#pragma unroll
                  for (int k = 0; k < 4; ++k) {
                    const uint32_t acc = (it0 + gi > 0 || dx > 0 || k > 0) ? 1u : 0u;
                    if constexpr (kCtas == 2) umma_ss_pair<Tr::kTf32>(d_tmem, ad + (k * 2), bdesc0 + dx * bstep + (k * 2), p.idesc, acc);
                    else umma_ss<Tr::kTf32>(d_tmem, ad + (k * 2), bdesc0 + dx * bstep + (k * 2), p.idesc, acc);
                  }
                }
              }
            }
            if (issue) {
              if constexpr (kCtas == 2) umma_commit_pair(&empty_bar[stage]); else umma_commit(&empty_bar[stage]);
              if (it0 + kGroup >= n3) {
                if constexpr (kCtas == 2) umma_commit_pair(&tmem_full_bar[as]); else umma_commit(&tmem_full_bar[as]);
              }
            }
            __syncwarp();
            if (++stage == static_cast<uint32_t>(p.stages)) { stage = 0; phase ^= 1; }
          }
        }
        if constexpr (!kRow3)
        for (int it0 = 0; it0 < n_iter; it0 += kGroup) {
          const int cnt = min(kGroup, n_iter - it0);
          const uint32_t sa = smem_u32(smem + stage * stage_bytes);
          const uint64_t adesc0 = umma_desc_sw128(sa);
          const uint64_t bdesc0 = umma_desc_sw128(sa + a_bytes);
          const uint64_t pstep = static_cast<uint64_t>(pair_bytes >> 4);   // start-address field is in 16-byte units
          const bool last = it0 + cnt >= n_iter;
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
#pragma unroll
            for (int gi = 0; gi < kGroup; ++gi) {
              if (gi < cnt) {
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  // advance 32 bytes (one MMA K step) inside the 128-byte swizzle row
                  const uint32_t acc = (it0 + gi > 0 || k > 0) ? 1u : 0u;
                  if constexpr (kCtas == 2)
                    umma_ss_pair<Tr::kTf32>(d_tmem, adesc0 + gi * pstep + (k * 2), bdesc0 + gi * pstep + (k * 2), p.idesc, acc);
                  else
                    umma_ss<Tr::kTf32>(d_tmem, adesc0 + gi * pstep + (k * 2), bdesc0 + gi * pstep + (k * 2), p.idesc, acc);
                }
              }
            }
            if constexpr (kCtas == 2) umma_commit_pair(&empty_bar[stage]); else umma_commit(&empty_bar[stage]);
            if (last) {
              if constexpr (kCtas == 2) umma_commit_pair(&tmem_full_bar[as]); else umma_commit(&tmem_full_bar[as]);
            }
          }
          __syncwarp();
          if (++stage == static_cast<uint32_t>(p.stages)) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------- epilogue
    // Two independent warpgroups: group g drains accumulator stage g (every other tile), with
    // its own staging buffer, named barrier and TMA-store thread, so the latency chains of
    // two tiles (TMEM load -> math -> smem -> TMA store) overlap.
    const int q = warp & 3;            // TMEM lane quadrant this warp may read
    const int wg = (warp - 4) >> 2;    // epilogue warpgroup = accumulator stage it serves
    const int row_in_tile = q * 32 + lane;
    const bool store_thread = (threadIdx.x & 127) == 0;
    const int bar_id = 1 + wg;
    T* out = reinterpret_cast<T*>(p.out);
    const T* residual = reinterpret_cast<const T*>(p.residual);
    uint8_t* buf = staging + wg * (kTileM * 128);
    const uint32_t as = static_cast<uint32_t>(wg);
    auto release_tmem = [&]() {  // this warp has read all it needs of accumulator `as`
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (kCtas == 2) mbar_arrive_cluster(mapa_shared(smem_u32(&tmem_empty_bar[as]), 0));
        else mbar_arrive(&tmem_empty_bar[as]);
      }
    };
    uint32_t tile_count = 0;
    for (int tile = cl; tile < total_tiles; tile += num_cl, ++tile_count) {
      if ((tile_count % kAcc) != as) continue;
      int n_tile = tile % p.n_tiles;
      int m_tile = tile / p.n_tiles;
      int prow = 0;
      const bool pix = !kRow3 && p.pix;
      if (pix) {
        const PixTile pt = pix_tile(p, tile);
        n_tile = pt.n_tile; m_tile = pt.m_tile; prow = pt.prow;
      }
      const int n0 = n_tile * p.block_n;
      const int m0 = (m_tile * kCtas + static_cast<int>(rank)) * kTileM;
      int r = m0 + row_in_tile;
      int b = 0, y = 0, x = 0;
      bool valid;
      if (pix) {   // rows of the tile are samples; r = the output pixel's PM row
        valid = r < p.geo.n;
        b = r;
        r = r * p.geo.S + prow;
      } else {
        valid = geo_decode(p.geo, r, b, y, x);
      }
      StatAcc sacc;   // this row's share of its sample's GroupNorm statistics (all columns of the tile)

      const uint32_t aph = (tile_count / kAcc) & 1;
      mbar_wait(&tmem_full_bar[as], aph);
      tc_fence_after();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * p.block_n;
      // GroupNorm statistics: every aligned 16-column unit of the row is summed in fp32 (stat_unit16),
      // converted to fixed point and added as integers, so the per-sample totals do not depend on the
      // N tile of the launch, on the tile order or on the batch (common.cuh, StatAcc).

      if (p.tma_store) {
        // kP columns are in registers at a time: a whole 128-byte sub-tile with two epilogue
        // warpgroups, half of one (bf16) with four, whose 640-thread CTA leaves 96 registers
        // per thread.
        constexpr int kP = (kAcc == 4) ? 32 : kSubN;
        constexpr int kHalves = kSubN / kP;
        constexpr int kVecs = kP * static_cast<int>(sizeof(T)) / 16;   // 16-byte staging chunks per step
        const int n_sub = p.block_n / kSubN;
        for (int k = 0; k < n_sub; ++k) {
#pragma unroll
          for (int h = 0; h < kHalves; ++h) {
            float v[kP];
            tmem_ld32(t_row + k * kSubN + h * kP, v);
            if constexpr (kP == 64) tmem_ld32(t_row + k * kSubN + 32, v + 32);
            tmem_ld_wait();
            if (k == n_sub - 1 && h == kHalves - 1) release_tmem();  // accumulator fully read: hand it back to the MMA issuer
            const int col0 = n0 + k * kSubN + h * kP;
            if (valid) {
              if (p.bias) {
#pragma unroll
                for (int j = 0; j < kP; j += 4) {
                  const float4 bv = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + j));
                  v[j] += bv.x; v[j + 1] += bv.y; v[j + 2] += bv.z; v[j + 3] += bv.w;
                }
              }
              if (p.act == 1) {
                gelu_inplace<T, kP>(v);
              }
              if (residual) {
                const T* rp = residual + static_cast<size_t>(r) * p.ldc + col0;
                Raw8<T> rr[kP / 8];
#pragma unroll
                for (int j = 0; j < kP / 8; ++j) rr[j].load(rp + 8 * j);
#pragma unroll
                for (int j = 0; j < kP / 8; ++j) {
                  float rv[8];
                  rr[j].unpack(rv);
#pragma unroll
                  for (int i = 0; i < 8; ++i) v[8 * j + i] += rv[i];
                }
              }
              if (p.sample_stats) {
#pragma unroll
                for (int j = 0; j < kP / 16; ++j) sacc.add(stat_unit16(v + 16 * j));
              }
            } else {
#pragma unroll
              for (int j = 0; j < kP; ++j) v[j] = 0.f;  // pad pixels stay zero
            }
            if (h == 0) {
              if (store_thread) tma_store_wait_read<0>();  // the previous store of this group has read the buffer
              named_bar_sync(bar_id, 128);
            }
            {
              uint8_t* rowp = buf + row_in_tile * 128;
              const int sw = row_in_tile & 7;
              if constexpr (sizeof(T) == 2) {
#pragma unroll
                for (int c = 0; c < kVecs; ++c) {
                  uint4 u;
                  __nv_bfloat162* hp = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
                  for (int i = 0; i < 4; ++i) hp[i] = __floats2bfloat162_rn(v[c * 8 + 2 * i], v[c * 8 + 2 * i + 1]);
                  *reinterpret_cast<uint4*>(rowp + (((h * kVecs + c) ^ sw) << 4)) = u;
                }
              } else {
#pragma unroll
                for (int c = 0; c < kVecs; ++c) {
                  *reinterpret_cast<float4*>(rowp + (((h * kVecs + c) ^ sw) << 4)) =
                      make_float4(v[c * 4], v[c * 4 + 1], v[c * 4 + 2], v[c * 4 + 3]);
                }
              }
            }
          }
          fence_async_proxy();
          named_bar_sync(bar_id, 128);
          if (store_thread) {
            if (pix) tma_store_3d(&tmap_out, buf, n0 + k * kSubN, prow, m0);
            else tma_store_2d(&tmap_out, buf, p.out_col0 + n0 + k * kSubN, m0);
            tma_store_commit();
          }
        }
        if (p.sample_stats) stat_commit(p.sample_stats, b, valid, sacc.s1, sacc.s2);
      } else if (p.epi == EPI_IMAGE) {
        float v[16];
        tmem_ld16(t_row, v);
        tmem_ld_wait();
        release_tmem();
        if (valid) {
          const size_t plane = static_cast<size_t>(p.geo.H) * p.geo.W;
          for (int c = 0; c < p.n_valid; ++c) {
            float s = v[c] + (p.bias ? __ldg(p.bias + c) : 0.f);
            s = 1.0f / (1.0f + expf(-s));
            const size_t o = (static_cast<size_t>(b) * p.n_valid + c) * plane + static_cast<size_t>(y) * p.geo.W + x;
            if (p.img_f32) p.img_f32[o] = s;
            if (p.img_u8) p.img_u8[o] = static_cast<uint8_t>(fminf(fmaxf(s * 255.0f, 0.f), 255.f));
          }
        }
      } else {
        int out_r = r;
        if (p.map_mode == 1 && valid) out_r = geo_row(p.geo_out, b, 2 * y + p.up_a, 2 * x + p.up_b);
        for (int c0 = 0; c0 < p.block_n; c0 += 32) {
          float v[32];
          if (p.block_n - c0 >= 32) {
            tmem_ld32(t_row + c0, v);
          } else {
            tmem_ld16(t_row + c0, v);
#pragma unroll
            for (int j = 16; j < 32; ++j) v[j] = 0.f;
          }
          tmem_ld_wait();
          const int ncols = min(32, p.block_n - c0);
          if (valid) {
            const int col0 = n0 + c0;
            if (p.bias) {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (j < ncols) v[j] += __ldg(p.bias + col0 + j);
            }
            if (p.act == 1) {
              gelu_inplace<T, 32>(v);
            }
            if (residual) {
              const T* rp = residual + static_cast<size_t>(r) * p.ldc + col0;
#pragma unroll
              for (int j = 0; j < 32; j += 8) {
                if (j < ncols) {
                  float rv[8];
                  load8(rp + j, rv);
#pragma unroll
                  for (int i = 0; i < 8; ++i) v[j + i] += rv[i];
                }
              }
            }
            if (p.sample_stats) {   // block_n is a multiple of 16: whole units only
              sacc.add(stat_unit16(v));
              if (ncols > 16) sacc.add(stat_unit16(v + 16));
            }
            T* op = out + static_cast<size_t>(out_r) * p.ldc + p.out_col0 + col0;
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              if (j < ncols) store8(op + j, v + j);
            }
          }
        }
        release_tmem();
        if (p.sample_stats) stat_commit(p.sample_stats, b, valid, sacc.s1, sacc.s2);
      }
    }
    if (store_thread) tma_store_wait_all<0>();
  }

  tc_fence_before();
  if constexpr (kCtas == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    if constexpr (kCtas == 2) tmem_dealloc_pair(tmem_base, p.tmem_cols);
    else tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

}  // namespace dmb
