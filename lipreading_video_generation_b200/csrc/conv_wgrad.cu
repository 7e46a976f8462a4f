// conv_wgrad.cu -- convolution weight gradient on tcgen05 tensor cores.
//
//   dw[co][tap][ci] += sum_pixels dy[pix][co] * X[pix + tap][ci]
//
// GEMM view per (tap, co-tile, ci-tile):  M = 128 output channels, N = 64 or 128 input channels,
// K = pixels, walked 128 pixels per pipeline stage (8 tcgen05.mma of K=16) and split across CTAs
// (grid.y); partial sums are reduced with fp32 atomics into dw.
//
// Both operands are "MN-major" for the tensor core: the TMA boxes land as [128 pixels][64 channels]
// (pixel = K index is the slow axis, channel = M/N index is contiguous), described to tcgen05 by an
// MN-major SWIZZLE_128B descriptor (SBO = 1024 B to the next 8 pixels, LBO = 16 KB to the next
// 64-channel block).  The X box is shifted by the tap and zero-filled out of bounds exactly as in
// the forward kernel, so padding and stride-2 phases are handled identically.
//
// Cout == 64 ("swap" mode): dy fills only half of an M = 128 tile, and these layers are bound by the L2 -> smem
// traffic per stage (48 KB per 256 tensor-pipe cycles), not by the MMA.  The operands swap roles instead: two
// shifted X blocks -- consecutive (tap, 64-channel block) pairs -- form the M = 128 operand and the single dy
// block is the N = 64 operand, D[(x block, ci)][co]: the same bytes per stage carry twice the useful FLOPs.
#include "common.cuh"
#include "tmap.cuh"
#include "umma.cuh"

namespace b2d {
using namespace umma;

struct WgradMaps {
  CUtensorMap x[4];
  CUtensorMap x2;
  CUtensorMap dy;
};

struct WgradParams {
  int taps, ksize, stride;
  int C1, Ctot;          // channels of source 1, total input channels
  int HoWo, rows, imgs;
  int pixel_tiles;       // ceil(N*Ho*Wo / 128)
  int Cout;
  int co_tiles, ci_tiles, BNci;
  int co_blocks;         // 64-channel dy blocks loaded per stage
  int swap;              // Cout == 64: M = two X blocks, N = dy
  int cib, xb_total;     // swap mode: 64-channel X blocks per tap, taps * cib
  float* dw;
};

constexpr int WG_THREADS = 192;
constexpr int WG_STAGES = 3;
constexpr int BLK = 128 * 128;  // one [128 pixels][64 ch] block

__global__ void __launch_bounds__(WG_THREADS)
conv_wgrad_kernel(const __grid_constant__ WgradMaps maps, const WgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int nb = p.BNci / 64;
  const uint32_t stage_bytes = (uint32_t)(2 + nb) * BLK;
  uint64_t* full_bar = (uint64_t*)(smem + (size_t)WG_STAGES * stage_bytes);
  uint64_t* empty_bar = full_bar + WG_STAGES;
  uint64_t* acc_bar = empty_bar + WG_STAGES;
  uint32_t* tmem_slot = (uint32_t*)(acc_bar + 1);

  // work item decode: blockIdx.x = (tap * co_tiles + co_t) * ci_tiles + ci_t
  // (swap mode: blockIdx.x = pair of consecutive X blocks, X block = tap * cib + ci block)
  int w = blockIdx.x;
  const int ci_t = w % p.ci_tiles; w /= p.ci_tiles;
  const int co_t = w % p.co_tiles; w /= p.co_tiles;
  const int tap = w;
  const int co0 = p.swap ? 0 : co_t * 128;
  const int ci0 = ci_t * p.BNci;
  const int split = blockIdx.y, nsplit = gridDim.y;
  const int my_iters = (p.pixel_tiles - split + nsplit - 1) / nsplit;

  if (threadIdx.x == 0) {
    for (int s = 0; s < WG_STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(acc_bar, 1);
    fence_barrier_init();
  }
  const uint32_t tmem_cols = p.BNci <= 64 ? 64 : 128;
  if (warp == 1) tmem_alloc(tmem_slot, tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      const int pad = p.ksize >> 1;
      // tap -> (row offset, column offset, stride-2 phase map)
      auto tap_geom = [&](int t, int& hoff, int& wc, int& ph) {
        const int dy_ = t / p.ksize, dx_ = t - dy_ * p.ksize;
        ph = 0;
        if (p.stride == 1) { hoff = dy_ - pad; wc = dx_ - pad; }
        else { ph = ((dy_ + 1) & 1) * 2 + ((dx_ + 1) & 1); hoff = (dy_ == 0 ? -1 : 0); wc = (dx_ == 0 ? -1 : 0); }
      };
      int hoff[2], wc[2], ph[2], cx[2];
      if (p.swap) {
        for (int j = 0; j < 2; ++j) {
          int xb = 2 * (int)blockIdx.x + j;
          if (xb >= p.xb_total) xb = p.xb_total - 1;    // odd block count: the last pair repeats its block
          tap_geom(xb / p.cib, hoff[j], wc[j], ph[j]);
          cx[j] = (xb % p.cib) * 64;
        }
      } else {
        tap_geom(tap, hoff[0], wc[0], ph[0]);
      }
      auto load_x = [&](uint8_t* dst, uint64_t* bar, int c, int j, int h0, int img0) {
        if (c < p.C1) tma_load_4d(dst, &maps.x[ph[j]], bar, c, wc[j], h0 + hoff[j], img0);
        else          tma_load_4d(dst, &maps.x2, bar, c - p.C1, wc[j], h0 + hoff[j], img0);
      };
      for (int it = 0; it < my_iters; ++it) {
        const int tile = split + it * nsplit;
        int img0, h0;
        if (p.HoWo >= 128) { const int tpi = p.HoWo >> 7; img0 = tile / tpi; h0 = (tile - img0 * tpi) * p.rows; }
        else { img0 = tile * p.imgs; h0 = 0; }
        const int s = it % WG_STAGES;
        const uint32_t par = (uint32_t)((it / WG_STAGES) & 1);
        mbar_wait(&empty_bar[s], par ^ 1u);
        uint8_t* base = smem + (size_t)s * stage_bytes;
        mbar_expect_tx(&full_bar[s], stage_bytes);
        if (p.swap) {
          // A: two shifted X blocks; B: the dy block
          load_x(base, &full_bar[s], cx[0], 0, h0, img0);
          load_x(base + BLK, &full_bar[s], cx[1], 1, h0, img0);
          tma_load_4d(base + 2 * BLK, &maps.dy, &full_bar[s], 0, 0, h0, img0);
        } else {
          // A: two 64-channel dy blocks; B: shifted X blocks
          tma_load_4d(base, &maps.dy, &full_bar[s], co0, 0, h0, img0);
          tma_load_4d(base + BLK, &maps.dy, &full_bar[s], co0 + 64, 0, h0, img0);
          for (int j = 0; j < nb; ++j) load_x(base + (2 + j) * BLK, &full_bar[s], ci0 + j * 64, 0, h0, img0);
        }
      }
    }
  } else if (warp == 1) {
    const uint32_t idesc = make_idesc(128, p.BNci, 1, 1, FMT_F16, FMT_F16);  // both operands MN-major
    for (int it = 0; it < my_iters; ++it) {
      const int s = it % WG_STAGES;
      const uint32_t par = (uint32_t)((it / WG_STAGES) & 1);
      mbar_wait(&full_bar[s], par);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t a_addr = smem_u32(smem + (size_t)s * stage_bytes);
        const uint32_t b_addr = a_addr + 2 * BLK;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          mma_f16_ss(tmem_base, make_sdesc_sw128(a_addr + k * 2048, BLK, 1024),
                      make_sdesc_sw128(b_addr + k * 2048, BLK, 1024), idesc, (uint32_t)((it | k) != 0));
        }
        mma_commit(&empty_bar[s]);
        if (it == my_iters - 1) mma_commit(acc_bar);
      }
      __syncwarp();
    }
  } else {
    const int q = warp & 3;
    const int row = q * 32 + lane;  // accumulator row: output channel, or (X block, input channel) in swap mode
    if (my_iters > 0) {
      mbar_wait(acc_bar, 0);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
      const long long Ktot = (long long)p.taps * p.Ctot;
      if (p.swap) {
        const int xb = 2 * (int)blockIdx.x + (row >> 6);
        const bool valid = xb < p.xb_total;
        // dw[co][tap][ci]: lanes hold consecutive ci -> each atomic instruction covers 128 contiguous bytes
        float* dst = p.dw + (long long)(xb / p.cib) * p.Ctot + (xb % p.cib) * 64 + (row & 63);
        for (int c0 = 0; c0 < 64; c0 += 16) {
          uint32_t r[16];
          __syncwarp();
          tmem_ld16(taddr + (uint32_t)c0, r);
          tmem_ld_wait();
          if (valid) {
#pragma unroll
            for (int j = 0; j < 16; ++j) atomicAdd(dst + (long long)(c0 + j) * Ktot, __uint_as_float(r[j]));
          }
        }
      } else {
        const int co = co0 + row;
        const bool valid = co < p.Cout;
        for (int c0 = 0; c0 < p.BNci; c0 += 16) {
          uint32_t r[16];
          __syncwarp();
          tmem_ld16(taddr + (uint32_t)c0, r);
          tmem_ld_wait();
          if (valid) {
            float* dst = p.dw + (long long)co * Ktot + (long long)tap * p.Ctot + ci0 + c0;
#pragma unroll
            for (int j = 0; j < 16; j += 4)
              red_add_f32x4(dst + j, __uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]),
                            __uint_as_float(r[j + 3]));
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
}

// ---------------------------------------------------------------------------------------------------------------
// 3x3 stride-1 weight gradient from HALOED X tiles.
//
// The kernel above loads a shifted X box per tap: nine taps = nine copies of nearly the same pixels, and at 42 B/clk/SM
// of L2 -> shared-memory bandwidth (all SMs streaming) that traffic, not the tensor pipe, bounds it (32 MAC/B with a
// 128 x 128 tile, 21 MAC/B when Cout = 64).  Here a pixel tile's X is loaded ONCE with its halo, (rows + 2) x (Wo + 2)
// pixels of 64 channels, and every tap is a shifted VIEW of that tile: an MN-major SWIZZLE_128B operand may start at any
// 128-byte row, because the swizzle is a function of the absolute shared-memory address (scripts/micro/umma_shift.cu:
// start addresses shifted by 0..16 rows and M = 128 operands made of two overlapping views are exact).  A K = 16
// slice is 16 consecutive pixels of one image row, so its rows are contiguous in the haloed tile too.
//
//   MODE 1 (Cout == 64): CTA = one 64-channel block of X, all nine taps.  A (M = 128) = TWO taps' views of X -- the
//     descriptor's LBO is the byte distance between the two views -- B (N = 64) = the dy block; five MMAs per slice,
//     accumulator rows = (tap parity, ci), columns = co, 5 x 64 TMEM columns.  49 KB per 128 pixels for 4.7 M MACs
//     (94 MAC/B, was 21).
//   MODE 0 (Cout % 128 == 0): CTA = (kernel row, 128 output channels, 64 or 128 input channels): A (M = 128) = the two dy
//     blocks, B = the view of tap (ky, kx) of an X tile with a COLUMN halo only (its rows are already those of ky),
//     three MMAs per slice into three accumulators.  66 KB per 128 pixels for 6.3 M MACs (96 MAC/B, was 32).
// (Nine taps of a 128 x 64 tile would need 576 TMEM columns, hence the kernel-row split in MODE 0.)
struct HaloMaps {
  CUtensorMap x, x2, dy;
};

struct HaloParams {
  int C1, Ctot, Cout;
  int R, tiles_per_img, pixel_tiles;
  int pitch;             // Wo + 2: pixels per row of the haloed tile
  int spr_shift;         // log2(Wo / 16): K = 16 slices per image row
  int co_tiles, ci_tiles, nb;
  uint32_t xblk_bytes;   // one haloed 64-channel X block in shared memory, rounded up to 1024 B
  uint32_t stage_bytes, tx_bytes;
  int stages;
  float* dw;
};

constexpr int WGH_MAX_STAGES = 4;

template <int MODE>
__global__ void __launch_bounds__(WG_THREADS)
conv_wgrad_halo_kernel(const __grid_constant__ HaloMaps maps, const HaloParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  uint64_t* full_bar = (uint64_t*)(smem + (size_t)p.stages * p.stage_bytes);
  uint64_t* empty_bar = full_bar + WGH_MAX_STAGES;
  uint64_t* acc_bar = empty_bar + WGH_MAX_STAGES;
  uint32_t* tmem_slot = (uint32_t*)(acc_bar + 1);

  int ky = 0, co0 = 0, ci0 = 0;
  if (MODE == 0) {
    int w = blockIdx.x;
    ci0 = (w % p.ci_tiles) * p.nb * 64; w /= p.ci_tiles;
    co0 = (w % p.co_tiles) * 128;
    ky = w / p.co_tiles;
  } else {
    ci0 = (int)blockIdx.x * 64;
  }
  const int BN = MODE == 0 ? p.nb * 64 : 64;
  const int split = blockIdx.y, nsplit = gridDim.y;
  const int my_iters = (p.pixel_tiles - split + nsplit - 1) / nsplit;
  // shared-memory offsets inside a stage
  const uint32_t dy_off = MODE == 0 ? 0u : p.xblk_bytes;
  const uint32_t x_off = MODE == 0 ? 2u * BLK : 0u;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(acc_bar, 1);
    fence_barrier_init();
  }
  const uint32_t tmem_cols = (MODE == 0 && BN == 64) ? 256 : 512;
  if (warp == 1) tmem_alloc(tmem_slot, tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      for (int it = 0; it < my_iters; ++it) {
        const int tile = split + it * nsplit;
        const int img = tile / p.tiles_per_img;
        const int h0 = (tile - img * p.tiles_per_img) * p.R;
        const int s = it % p.stages;
        const uint32_t par = (uint32_t)((it / p.stages) & 1);
        mbar_wait(&empty_bar[s], par ^ 1u);
        uint8_t* base = smem + (size_t)s * p.stage_bytes;
        mbar_expect_tx(&full_bar[s], p.tx_bytes);
        if (MODE == 0) {
          tma_load_4d(base, &maps.dy, &full_bar[s], co0, 0, h0, img);
          tma_load_4d(base + BLK, &maps.dy, &full_bar[s], co0 + 64, 0, h0, img);
          for (int j = 0; j < p.nb; ++j) {
            const int c = ci0 + j * 64;
            uint8_t* dst = base + x_off + (size_t)j * p.xblk_bytes;
            if (c < p.C1) tma_load_4d(dst, &maps.x, &full_bar[s], c, -1, h0 + ky - 1, img);
            else          tma_load_4d(dst, &maps.x2, &full_bar[s], c - p.C1, -1, h0 + ky - 1, img);
          }
        } else {
          if (ci0 < p.C1) tma_load_4d(base, &maps.x, &full_bar[s], ci0, -1, h0 - 1, img);
          else            tma_load_4d(base, &maps.x2, &full_bar[s], ci0 - p.C1, -1, h0 - 1, img);
          tma_load_4d(base + dy_off, &maps.dy, &full_bar[s], 0, 0, h0, img);
        }
      }
    }
  } else if (warp == 1) {
    // One issuing thread, ring slot / phase by counting, descriptors by addition on the 32-bit words (hi = stride byte
    // offset 1024 + version + swizzle mode, constant; lo = address >> 4 | LBO >> 4 << 16): the same finding as in
    // conv_gemm.cu -- ~100 instructions of descriptor assembly per K = 16 slice cost more issue time than the slice's
    // three to five MMAs take on the tensor pipe (162-270 clk).
    if (lane == 0) {
      const uint32_t idesc = make_idesc(128, BN, 1, 1, FMT_F16, FMT_F16);  // both operands MN-major
      const int spr_mask = (1 << p.spr_shift) - 1;
      const uint32_t hi = (uint32_t)(make_sdesc_sw128(0, 0, 1024) >> 32);
      auto desc = [hi](uint32_t lo) { return ((uint64_t)hi << 32) | lo; };
      const uint32_t ring = smem_u32(smem);
      const uint32_t pitch16 = (uint32_t)p.pitch * 8u;                     // one pixel row of the haloed tile, in 16-byte units
      const uint32_t lbo_blk = (uint32_t)(BLK >> 4) << 16, lbo_x = (p.xblk_bytes >> 4) << 16;
      int s = 0;
      uint32_t par = 0;
      for (int it = 0; it < my_iters; ++it) {
        mbar_wait(&full_bar[s], par);
        tc_fence_after();
        const uint32_t base = ring + (uint32_t)s * p.stage_bytes;
        const uint32_t dy_lo = ((base + dy_off) >> 4) | lbo_blk;
        const uint32_t x16 = (base + x_off) >> 4;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int r = j >> p.spr_shift, c0 = (j & spr_mask) << 4;
          const uint32_t acc = (uint32_t)((it | j) != 0);
          const uint32_t xv = x16 + (uint32_t)r * pitch16 + (uint32_t)c0 * 8u;   // the slice's 16 pixels in the tile
          if (MODE == 0) {
            const uint64_t a = desc(dy_lo + (uint32_t)j * 128u);
#pragma unroll
            for (int kx = 0; kx < 3; ++kx)
              mma_f16_ss(tmem_base + (uint32_t)(kx * BN), a, desc((xv + (uint32_t)kx * 8u) | lbo_x), idesc, acc);
          } else {
            const uint64_t b = desc(dy_lo + (uint32_t)j * 128u);
            // views of taps t = 3 * ky + kx at xv + ky * row + kx * 128 B; pairs (0,1) (2,3) (4,5) (6,7) (8,8):
            // LBO = byte distance between the pair's two views
            mma_f16_ss(tmem_base + 0,   desc(xv | (8u << 16)), b, idesc, acc);
            mma_f16_ss(tmem_base + 64,  desc((xv + 16u) | ((pitch16 - 16u) << 16)), b, idesc, acc);
            mma_f16_ss(tmem_base + 128, desc((xv + pitch16 + 8u) | (8u << 16)), b, idesc, acc);
            mma_f16_ss(tmem_base + 192, desc((xv + 2u * pitch16) | (8u << 16)), b, idesc, acc);
            mma_f16_ss(tmem_base + 256, desc(xv + 2u * pitch16 + 16u), b, idesc, acc);
          }
        }
        mma_commit(&empty_bar[s]);
        if (it == my_iters - 1) mma_commit(acc_bar);
        if (++s == p.stages) { s = 0; par ^= 1u; }
      }
    }
    __syncwarp();
  } else {
    const int q = warp & 3;
    const int row = q * 32 + lane;
    if (my_iters > 0) {
      mbar_wait(acc_bar, 0);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
      const long long Ktot = 9LL * p.Ctot;
      if (MODE == 0) {
        float* dst_row = p.dw + (long long)(co0 + row) * Ktot + (long long)(ky * 3) * p.Ctot + ci0;
        for (int kx = 0; kx < 3; ++kx)
          for (int c0 = 0; c0 < BN; c0 += 16) {
            uint32_t r[16];
            __syncwarp();
            tmem_ld16(taddr + (uint32_t)(kx * BN + c0), r);
            tmem_ld_wait();
            float* dst = dst_row + (long long)kx * p.Ctot + c0;
#pragma unroll
            for (int j = 0; j < 16; j += 4)
              red_add_f32x4(dst + j, __uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]),
                            __uint_as_float(r[j + 3]));
          }
      } else {
        // rows = (tap parity, ci): lanes hold consecutive ci -> each atomic instruction covers 128 contiguous bytes
        for (int pr = 0; pr < 5; ++pr) {
          const int tap = 2 * pr + (row >> 6);
          const bool valid = tap < 9;
          float* dst = p.dw + (long long)tap * p.Ctot + ci0 + (row & 63);
          for (int c0 = 0; c0 < 64; c0 += 16) {
            uint32_t r[16];
            __syncwarp();
            tmem_ld16(taddr + (uint32_t)(pr * 64 + c0), r);
            tmem_ld_wait();
            if (valid) {
#pragma unroll
              for (int j = 0; j < 16; ++j) atomicAdd(dst + (long long)(c0 + j) * Ktot, __uint_as_float(r[j]));
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
}

static bool is_pow2w(int v) { return v > 0 && (v & (v - 1)) == 0; }

static int make_map4(CUtensorMap* m, const void* base, int C, long long ld, int N, int H, int W, int phase_stride,
                     int Wo, int rows, int imgs) {
  const uint64_t dims[4] = {(uint64_t)C, (uint64_t)(W / phase_stride), (uint64_t)(H / phase_stride), (uint64_t)N};
  const uint64_t str[3] = {(uint64_t)ld * 2 * phase_stride, (uint64_t)ld * 2 * W * phase_stride,
                           (uint64_t)ld * 2 * W * H};
  const uint32_t box[4] = {64u, (uint32_t)Wo, (uint32_t)rows, (uint32_t)imgs};
  return make_tmap_16b(m, base, 4, dims, str, box);
}

static int make_map4_box(CUtensorMap* m, const void* base, int C, long long ld, int N, int H, int W, int box_w, int box_h) {
  const uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)N};
  const uint64_t str[3] = {(uint64_t)ld * 2, (uint64_t)ld * 2 * W, (uint64_t)ld * 2 * W * H};
  const uint32_t box[4] = {64u, (uint32_t)box_w, (uint32_t)box_h, 1u};
  return make_tmap_16b(m, base, 4, dims, str, box);
}

// 3x3 stride-1, planes of >= 128 pixels, rows of >= 16: the haloed-tile kernels
static int wgrad_halo(const b2d_wgrad_args* a, int C2, int Ho, int Wo, int R, void* stream) {
  const int mode = a->Cout == 64 ? 1 : 0;
  const int Rh = mode ? R + 2 : R;
  HaloMaps maps;
  memset(&maps, 0, sizeof(maps));
  int rc = make_map4_box(&maps.x, a->x, a->C1, a->ldx, a->N, a->H, a->W, Wo + 2, Rh);
  if (rc) return rc;
  if (C2) { rc = make_map4_box(&maps.x2, a->x2, C2, a->ldx2, a->N, a->H, a->W, Wo + 2, Rh); if (rc) return rc; }
  else maps.x2 = maps.x;
  rc = make_map4_box(&maps.dy, a->dy, a->Cout, a->lddy, a->N, Ho, Wo, Wo, R);
  if (rc) return rc;
  HaloParams p;
  memset(&p, 0, sizeof(p));
  p.C1 = a->C1; p.Ctot = a->C1 + C2; p.Cout = a->Cout;
  p.R = R; p.tiles_per_img = Ho / R; p.pixel_tiles = a->N * p.tiles_per_img;
  p.pitch = Wo + 2;
  p.spr_shift = 0;
  while ((16 << p.spr_shift) < Wo) ++p.spr_shift;
  const uint32_t xbox = (uint32_t)Rh * (uint32_t)(Wo + 2) * 128u;
  p.xblk_bytes = (xbox + 1023u) & ~1023u;
  int items;
  if (mode == 0) {
    p.nb = (a->C1 % 128 == 0 && C2 % 128 == 0) ? 2 : 1;     // a 128-wide ci tile must not straddle the two sources
    p.co_tiles = a->Cout / 128;
    p.ci_tiles = p.Ctot / (64 * p.nb);
    p.stage_bytes = 2u * BLK + (uint32_t)p.nb * p.xblk_bytes;
    p.tx_bytes = 2u * BLK + (uint32_t)p.nb * xbox;
    items = 3 * p.co_tiles * p.ci_tiles;
  } else {
    p.nb = 1; p.co_tiles = 1; p.ci_tiles = p.Ctot / 64;
    p.stage_bytes = p.xblk_bytes + (uint32_t)BLK;
    p.tx_bytes = xbox + (uint32_t)BLK;
    items = p.ci_tiles;
  }
  const size_t smem_max = 227 * 1024;
  const size_t fixed = (2 * WGH_MAX_STAGES + 1) * 8 + 16 + 1024;
  p.stages = (int)((smem_max - fixed) / p.stage_bytes);
  if (p.stages > WGH_MAX_STAGES) p.stages = WGH_MAX_STAGES;
  B2D_REQUIRE(p.stages >= 2, B2D_ERR_SHAPE, "wgrad: a %d-byte stage does not fit twice in shared memory", (int)p.stage_bytes);
  p.dw = a->dw;
  int nsplit = num_sms() / items;           // one CTA per SM (shared memory, TMEM): one wave
  if (nsplit > p.pixel_tiles) nsplit = p.pixel_tiles;
  if (nsplit < 1) nsplit = 1;
  const size_t smem = (size_t)p.stages * p.stage_bytes + fixed;
  dim3 grid((unsigned)items, (unsigned)nsplit);
  if (mode == 0) {
    B2D_SET_SMEM_ONCE((conv_wgrad_halo_kernel<0>), (int)smem_max);
    conv_wgrad_halo_kernel<0><<<grid, WG_THREADS, smem, (cudaStream_t)stream>>>(maps, p);
  } else {
    B2D_SET_SMEM_ONCE((conv_wgrad_halo_kernel<1>), (int)smem_max);
    conv_wgrad_halo_kernel<1><<<grid, WG_THREADS, smem, (cudaStream_t)stream>>>(maps, p);
  }
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

}  // namespace b2d

using namespace b2d;

extern "C" int b2d_conv2d_wgrad(const b2d_wgrad_args* a, void* stream) {
  B2D_REQUIRE(a && a->x && a->dy && a->dw, B2D_ERR_SHAPE, "wgrad: null argument");
  B2D_REQUIRE(aligned16(a->dw), B2D_ERR_ALIGN, "wgrad: dw must be 16-byte aligned (vector accumulation)");
  B2D_REQUIRE(a->ksize == 1 || a->ksize == 3, B2D_ERR_SHAPE, "wgrad: ksize %d", a->ksize);
  B2D_REQUIRE(a->stride == 1 || (a->stride == 2 && a->ksize == 3), B2D_ERR_SHAPE, "wgrad: stride %d", a->stride);
  const int C2 = a->x2 ? a->C2 : 0;
  B2D_REQUIRE(a->C1 > 0 && a->C1 % 64 == 0 && C2 % 64 == 0, B2D_ERR_SHAPE, "wgrad: channels (%d,%d) not multiples of 64",
              a->C1, C2);
  B2D_REQUIRE(a->stride == 1 || C2 == 0, B2D_ERR_SHAPE, "wgrad: two-source input needs stride 1");
  B2D_REQUIRE(a->Cout > 0 && a->Cout % 64 == 0, B2D_ERR_SHAPE, "wgrad: Cout=%d must be a multiple of 64", a->Cout);
  B2D_REQUIRE(a->H % a->stride == 0 && a->W % a->stride == 0, B2D_ERR_SHAPE, "wgrad: odd plane");
  const int Ho = a->H / a->stride, Wo = a->W / a->stride;
  B2D_REQUIRE(is_pow2w(Wo) && Wo <= 128, B2D_ERR_SHAPE, "wgrad: output width %d must be a power of two <= 128", Wo);
  int rows = 128 / Wo, imgs = 1;
  if (rows > Ho) {
    B2D_REQUIRE((128 % (Ho * Wo)) == 0, B2D_ERR_SHAPE, "wgrad: plane %dx%d does not tile 128 pixels", Ho, Wo);
    rows = Ho; imgs = 128 / (Ho * Wo);
  } else {
    B2D_REQUIRE(Ho % rows == 0, B2D_ERR_SHAPE, "wgrad: height %d not a multiple of %d", Ho, rows);
  }
  static const bool halo_on = [] { const char* e = getenv("B2D_WGRAD_HALO"); return !(e && e[0] == '0'); }();
  if (halo_on && a->ksize == 3 && a->stride == 1 && imgs == 1 && Wo >= 16 && (a->Cout == 64 || a->Cout % 128 == 0))
    return wgrad_halo(a, C2, Ho, Wo, rows, stream);
  WgradMaps maps;
  memset(&maps, 0, sizeof(maps));
  int rc;
  if (a->stride == 1) {
    rc = make_map4(&maps.x[0], a->x, a->C1, a->ldx, a->N, a->H, a->W, 1, Wo, rows, imgs);
    if (rc) return rc;
    for (int i = 1; i < 4; ++i) maps.x[i] = maps.x[0];
  } else {
    for (int pa = 0; pa < 2; ++pa)
      for (int pb = 0; pb < 2; ++pb) {
        const __half* base = (const __half*)a->x + ((long long)pa * a->W + pb) * a->ldx;
        rc = make_map4(&maps.x[pa * 2 + pb], base, a->C1, a->ldx, a->N, a->H, a->W, 2, Wo, rows, imgs);
        if (rc) return rc;
      }
  }
  if (C2) { rc = make_map4(&maps.x2, a->x2, C2, a->ldx2, a->N, a->H, a->W, 1, Wo, rows, imgs); if (rc) return rc; }
  else maps.x2 = maps.x[0];
  rc = make_map4(&maps.dy, a->dy, a->Cout, a->lddy, a->N, Ho, Wo, 1, Wo, rows, imgs);
  if (rc) return rc;

  WgradParams p;
  p.taps = a->ksize * a->ksize; p.ksize = a->ksize; p.stride = a->stride;
  p.C1 = a->C1; p.Ctot = a->C1 + C2;
  p.HoWo = Ho * Wo; p.rows = rows; p.imgs = imgs;
  const long long total_pixels = (long long)a->N * Ho * Wo;
  p.pixel_tiles = (int)((total_pixels + 127) / 128);
  p.Cout = a->Cout;
  p.co_tiles = (a->Cout + 127) / 128;
  p.co_blocks = 2;
  p.swap = a->Cout == 64;
  B2D_REQUIRE(p.swap || a->Cout % 128 == 0, B2D_ERR_SHAPE, "wgrad: Cout=%d must be 64 or a multiple of 128", a->Cout);
  // a 128-wide ci tile must not straddle the two sources unless both are multiples of 128
  p.BNci = (!p.swap && a->C1 % 128 == 0 && C2 % 128 == 0) ? 128 : 64;
  p.ci_tiles = p.Ctot / p.BNci;
  p.cib = p.Ctot / 64;
  p.xb_total = p.taps * p.cib;
  p.dw = a->dw;
  const int items = p.swap ? (p.xb_total + 1) / 2 : p.taps * p.co_tiles * p.ci_tiles;
  int nsplit = (2 * num_sms() + items - 1) / items;
  if (nsplit > p.pixel_tiles) nsplit = p.pixel_tiles;
  if (nsplit < 1) nsplit = 1;
  const size_t smem = (size_t)WG_STAGES * 4 * BLK + (2 * WG_STAGES + 1) * 8 + 16 + 1024;
  B2D_SET_SMEM_ONCE((conv_wgrad_kernel), (int)smem);
  dim3 grid((unsigned)items, (unsigned)nsplit);
  conv_wgrad_kernel<<<grid, WG_THREADS, smem, (cudaStream_t)stream>>>(maps, p);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}
