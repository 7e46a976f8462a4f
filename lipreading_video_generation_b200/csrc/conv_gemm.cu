// conv_gemm.cu -- implicit-GEMM convolution (3x3 pad 1 stride 1/2, 1x1) on tcgen05 tensor cores.
//
//   M = 128 output pixels per CTA (a rectangle of full rows, possibly spanning images),
//   N = BN output channels (16..256), K = taps * (C1 + C2) walked in 64-channel chunks.
//
// A tiles come straight from the pixel-major activation tensor with a 4-D TMA box
// (64 ch, Wo, rows, imgs) whose start coordinate is shifted by the tap; out-of-bounds rows /
// columns are zero-filled by TMA, which IS the convolution's zero padding.  Stride-2 convs
// use four "phase" tensor maps (even/odd rows x even/odd columns of the input).
// B tiles (weights [Cout][K], K-major) use a 2-D TMA box.  Both land in 128B-swizzled smem,
// consumed by tcgen05.mma (M=128, N=BN, K=16) with the fp32 accumulator in TMEM.
//
// Warp roles (320 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer,
// warps 2..9 = epilogue (TMEM -> registers -> bias/emb/residual -> bf16 / fp32-NCHW global): two warps per TMEM lane
// quarter, one taking the even and one the odd 16-column chunks of the tile (the fused GroupNorm epilogues are ~2300
// warp instructions per 128 x 64 tile: four warps per CTA left them latency bound, profiles/r2_groupnorm_fused.md).
//
// HALO variant (3x3, stride 1, planes of 16k x 8m pixels, TMA-store epilogue): the 128 pixels of a tile are a PATCH of
// 16 image rows x 8 pixels, and its input is loaded ONCE per 64-channel block with its halo -- a (64 ch, 10, 18, 1) box,
// 23 KB -- instead of nine shifted 16 KB boxes.  Every tap is then a view of that tile: in the K-major SWIZZLE_128B
// layout the 8 pixels of image row g sit at tile rows (g + ky) * 10 + kx .., i.e. the operand's sixteen 8-row groups are
// 1280 B apart (stride byte offset 1280) and it starts at row ky * 10 + kx -- legal because the swizzle is a function of
// the absolute shared-memory address (scripts/micro/umma_shift_k.cu: 0 mismatches for every start row and pitch).
// The round-2 launch list showed why: at 64 input channels the per-tap kernel moves 216 KB from L2 per 128 x 64 tile
// = 48 B/clk/SM, the L2 -> SM ceiling, for 1152 clk of tensor work (585 TFLOP/s).  With 64 input channels the CTA's
// 64 x 576 weight block is also loaded once and stays in shared memory (one CTA per SM, tile_n fixed per CTA); wider
// inputs stream the weight blocks through their own ring.  One CTA per SM, 16 epilogue warps, and the epilogue's
// residual / GroupNorm-input tile of the NEXT tile is fetched while the current one is computed (two staging slots).
#include "common.cuh"
#include "tmap.cuh"
#include "umma.cuh"
#include <mutex>

namespace b2d {
using namespace umma;

__device__ __forceinline__ void named_bar_sync(int id, int count) {
  asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ void red_shared_add_f32(uint32_t saddr, float v) {
  asm volatile("red.shared.add.f32 [%0], %1;" :: "r"(saddr), "f"(v) : "memory");
}

struct ConvMaps {
  CUtensorMap a[4];  // source 1 (4 phases when stride == 2, else a[0])
  CUtensorMap a2;    // source 2 (stride 1 only)
  CUtensorMap b;     // weights
  CUtensorMap o;     // output   (N,Ho,Wo,ldo): box (64, Wo, rows, imgs)   [TMA-store epilogue]
  CUtensorMap r;     // residual (N,Ho,Wo,ldres): same box; EPI_GNB: the GroupNorm input x1
  CUtensorMap r2;    // EPI_GNB: the GroupNorm input x2 (second tensor of a skip concatenation)
};

// Epilogue variants (template parameter of the kernel; the plain one carries none of the extra code):
//   EPI_PLAIN  bias + emb + residual
//   EPI_STATS  the same, plus per-(image, channel pair) sum / sum of squares of the output for the GroupNorm that
//              consumes it (b2d_conv_args.gn_stats_out)
//   EPI_GNB    data-gradient convolution whose output is dL/dy of y = act(GroupNorm(x)): stores dyh = dy * act'(z) * mask
//              and accumulates (sum dyh, sum dyh * x) per (image, channel) (b2d_conv_args.gnb_*)
constexpr int EPI_PLAIN = 0, EPI_STATS = 1, EPI_GNB = 2;

struct ConvParams {
  int taps, ksize, stride;
  int kc1, kc2;        // 64-channel chunks of source 1 / source 2
  int HoWo, Wo;        // output plane
  int rows, imgs;      // box: rows per tile, images per tile
  int total_pixels;    // N*Ho*Wo
  int n_images;
  int tma_out;         // epilogue stages the tile in smem and uses TMA for the residual load and the output store
  int m_tiles, n_tiles; // persistent tile loop: tile = tile_m * n_tiles + tile_n
  int Cout, BN;
  int stages;
  int fmt;             // storage format of x / residual / out: FMT_F16 (forward) or FMT_F16 (gradient = dgrad)
  uint32_t tmem_cols;
  const float* bias;
  const float* emb; long long ldemb;
  const __half* residual; long long ldres;
  __half* out; long long ldo;
  float* out_nchw;
  // EPI_STATS
  double* stats;           // [N][Cout/2][2]
  // EPI_GNB
  int gnb_C1, gnb_cpg, gnb_act;
  const float* gnb_mean; const float* gnb_rstd; const float* gnb_gamma; const float* gnb_beta;
  const uint8_t* gnb_mask; float gnb_keep_scale;
  float* gnb_ab;           // [N][Cout][2]
  // HALO variant
  int Ho;
  int b_resident;      // the CTA's weight block (9 taps x BN x 64) is loaded once and stays in shared memory
  int na, nb;          // slots of the haloed-input ring (<= 4) / of the streamed weight ring (<= 12)
  int ns;              // staging slots of the epilogue (1..3); 3: the residual / GroupNorm input runs a tile ahead
  uint32_t b_bytes;    // bytes of the weight area (resident block or nb slots)
};

constexpr int CONV_EPI_WARPS = 8;          // two per TMEM lane quarter: even / odd 16-column chunks of the tile
constexpr int CONV_EPI_THREADS = CONV_EPI_WARPS * 32;
constexpr int CONV_THREADS = 64 + CONV_EPI_THREADS;
constexpr int EB_FLOATS = 512;              // images per tile x BN <= 512 (bias + emb staging)
constexpr int SACC_FLOATS = 512;            // fused GroupNorm sums of one tile: images per tile x BN x 2 <= 512
constexpr int A_TILE_BYTES = 128 * 128;  // 128 pixels x 64 bf16
constexpr int HALO_PW = 8, HALO_PH = 16;                     // patch: 16 image rows x 8 pixels
constexpr int HALO_PITCH = HALO_PW + 2, HALO_ROWS = HALO_PH + 2;
constexpr uint32_t HALO_DATA_BYTES = HALO_ROWS * HALO_PITCH * 128;   // 23040: what one haloed load delivers
constexpr uint32_t HALO_A_BYTES = 23 * 1024;                 // slot size (1024-byte aligned slots)
constexpr int HALO_EPI_WARPS = 16;
constexpr int HALO_THREADS = 64 + HALO_EPI_WARPS * 32;

// Sum 16 per-lane values over the 32 lanes of a warp with 16 shuffles (reduce-scatter): on return v[0] of lane l holds
// the warp total of value index ((l>>4)&1)*8 + ((l>>3)&1)*4 + ((l>>2)&1)*2 + ((l>>1)&1) (lanes l and l^1 hold the same).
__device__ __forceinline__ void warp_reduce_scatter16(float* v, int lane) {
#define B2D_RS_STEP(HALF, BIT)                                                      \
  {                                                                                 \
    const bool up = (lane & BIT) != 0;                                              \
    _Pragma("unroll") for (int i = 0; i < HALF; ++i) {                              \
      const float send = up ? v[i] : v[i + HALF];                                   \
      const float keep = up ? v[i + HALF] : v[i];                                   \
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, BIT);                        \
    }                                                                               \
  }
  B2D_RS_STEP(8, 16) B2D_RS_STEP(4, 8) B2D_RS_STEP(2, 4) B2D_RS_STEP(1, 2)
  v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
}
// The same for 32 values with 31 shuffles: lane l ends with the total of value index l.
__device__ __forceinline__ void warp_reduce_scatter32(float* v, int lane) {
  B2D_RS_STEP(16, 16) B2D_RS_STEP(8, 8) B2D_RS_STEP(4, 4) B2D_RS_STEP(2, 2) B2D_RS_STEP(1, 1)
#undef B2D_RS_STEP
}

template <int EPI, bool HALO>
__global__ void __launch_bounds__(HALO ? HALO_THREADS : CONV_THREADS, HALO ? 1 : 2)
conv_gemm_kernel(const __grid_constant__ ConvMaps maps, const ConvParams p) {
  constexpr int EPI_WARPS = HALO ? HALO_EPI_WARPS : CONV_EPI_WARPS;
  constexpr int EPI_THREADS = EPI_WARPS * 32;
  constexpr int NTHREADS = 64 + EPI_THREADS;
  constexpr int CSPLIT = EPI_WARPS / 4;      // epilogue warps per TMEM lane quarter: warp takes chunks cc % CSPLIT == csel
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t b_tile_bytes = (uint32_t)p.BN * 128u;
  const uint32_t stage_bytes = A_TILE_BYTES + b_tile_bytes;
  // operand area: [stages x (A tile | B tile)], HALO: [na haloed input slots][weight area]
  const uint32_t ring_bytes = HALO ? (uint32_t)p.na * HALO_A_BYTES + p.b_bytes : (uint32_t)p.stages * stage_bytes;
  uint64_t* bars = (uint64_t*)(smem + ring_bytes);
  uint64_t* full_bar = bars;                              // HALO: a_full[4]
  uint64_t* empty_bar = HALO ? bars + 4 : bars + p.stages;   // HALO: a_empty[4]
  uint64_t* b_full = bars + 8;                            // HALO only: [12]
  uint64_t* b_empty = bars + 20;                          // HALO only: [12]
  uint64_t* acc_full = HALO ? bars + 32 : empty_bar + p.stages;   // [2] accumulator buffer b holds a finished tile
  uint64_t* acc_empty = acc_full + 2;          // [2] epilogue has drained buffer b
  uint32_t* tmem_slot = HALO ? (uint32_t*)(bars + 40) : (uint32_t*)(acc_empty + 2);
  const uint32_t eb_base = HALO ? smem_u32(smem + ring_bytes + 512)
                                : smem_u32(tmem_slot + 4);   // float [2][EB_FLOATS]: bias + emb of the current tile
  // residual / GroupNorm-input tile landed in the staging buffer; HALO: one barrier per staging slot
  uint64_t* res_full = HALO ? bars + 36 : (uint64_t*)(tmem_slot + 2);
  // staging: [128 pix][64 ch] fp16 blocks, 1024-aligned; 2 blocks, HALO: ns slots x (BN / 64) blocks
  uint8_t* stage_out = smem + (size_t)ring_bytes + 1024 + 2 * EB_FLOATS * 4;
  const int nblk_stage = p.BN >> 6;
  const uint32_t sacc_u32 = smem_u32(stage_out + (HALO ? p.ns * nblk_stage : 2) * A_TILE_BYTES);   // float [SACC_FLOATS]
  if (EPI != EPI_PLAIN) {
    for (int i = threadIdx.x; i < SACC_FLOATS; i += NTHREADS) sts32f(sacc_u32 + (uint32_t)i * 4u, 0.f);
  }

  // Persistent CTA: tiles tile0, tile0 + gridDim.x, ...; the TMA ring and the two TMEM accumulator buffers run
  // across tile boundaries, so tile i's epilogue overlaps tile i+1's loads and MMAs (the non-persistent
  // version spent ~22k clk per CTA on a 1.2k-clk mainloop: profiles/r1_conv.md).
  const int total_tiles = p.m_tiles * p.n_tiles;
  const int chunks = p.kc1 + p.kc2;
  const int total_iters = p.taps * chunks;

  if (threadIdx.x == 0) {
    if (HALO) {
      for (int s = 0; s < 4; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
      for (int s = 0; s < 12; ++s) { mbar_init(&b_full[s], 1); mbar_init(&b_empty[s], 1); }
      for (int s = 0; s < 3; ++s) mbar_init(&res_full[s], 1);
    } else {
      for (int s = 0; s < p.stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
      mbar_init(res_full, 1);
    }
    mbar_init(&acc_full[0], 1); mbar_init(&acc_full[1], 1);
    mbar_init(&acc_empty[0], EPI_THREADS); mbar_init(&acc_empty[1], EPI_THREADS);
    fence_barrier_init();
    prefetch_tmap(&maps.a[0]);
    prefetch_tmap(&maps.b);
  }
  if (warp == 1) tmem_alloc(tmem_slot, 2 * p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int halo_tpi = p.HoWo >> 7, halo_pw = p.Wo >> 3;     // HALO: patches per image / per patch row

  if (warp == 0 && HALO) {
    // ------------------------------------------------------------ TMA producer, haloed patches
    if (elect_one()) {
      uint8_t* b_area = smem + (size_t)p.na * HALO_A_BYTES;
      const int my_tiles = (int)blockIdx.x < total_tiles ? (total_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
      const int total_seq = my_tiles * chunks;               // (tile, 64-channel block) pairs of this CTA, in MMA order
      // haloed input of pair `seq` -> slot seq % na
      int a_slot = 0, a_ti = 0, a_ch = 0;                    // next haloed load: ring slot, tile index, channel block
      uint32_t a_par = 1;
      auto issue_a = [&]() {
        const int tile = (int)blockIdx.x + a_ti * (int)gridDim.x;
        const int tile_m = tile / p.n_tiles;
        const int img0 = tile_m / halo_tpi, t = tile_m - img0 * halo_tpi;
        const int py = t / halo_pw, px = t - py * halo_pw;
        const int h0 = py * HALO_PH, w0 = px * HALO_PW;
        mbar_wait(&empty_bar[a_slot], a_par);
        mbar_expect_tx(&full_bar[a_slot], HALO_DATA_BYTES);
        uint8_t* a_dst = smem + (size_t)a_slot * HALO_A_BYTES;
        if (a_ch < p.kc1) tma_load_4d(a_dst, &maps.a[0], &full_bar[a_slot], a_ch * 64, w0 - 1, h0 - 1, img0);
        else              tma_load_4d(a_dst, &maps.a2, &full_bar[a_slot], (a_ch - p.kc1) * 64, w0 - 1, h0 - 1, img0);
        if (++a_slot == p.na) { a_slot = 0; a_par ^= 1u; }
        if (++a_ch == chunks) { a_ch = 0; ++a_ti; }
      };
      if (total_seq > 0) {
        const int n0_first = ((int)blockIdx.x % p.n_tiles) * p.BN;
        if (p.b_resident) {                                   // tile_n is the same for every tile of this CTA
          mbar_expect_tx(&b_full[0], 9u * b_tile_bytes);
          for (int tap = 0; tap < 9; ++tap)
            tma_load_2d(b_area + (size_t)tap * b_tile_bytes, &maps.b, &b_full[0], tap * 64, n0_first);
        }
        issue_a();
      }
      int bs = 0, ch = 0, ti = 0;
      uint32_t bpar = 1;
      int n0 = total_seq > 0 ? (((int)blockIdx.x) % p.n_tiles) * p.BN : 0;
      for (int seq = 0; seq < total_seq; ++seq) {
        // the NEXT pair's input goes out before this pair's nine weight blocks: the producer is in order, and behind the
        // weight ring (which the MMAs drain slot by slot) it would leave with only a few taps to spare
        if (seq + 1 < total_seq) issue_a();
        if (!p.b_resident) {
          int kcol = ch * 64;
          for (int tap = 0; tap < 9; ++tap, kcol += chunks * 64) {
            mbar_wait(&b_empty[bs], bpar);
            mbar_expect_tx(&b_full[bs], b_tile_bytes);
            tma_load_2d(b_area + (size_t)bs * b_tile_bytes, &maps.b, &b_full[bs], kcol, n0);
            if (++bs == p.nb) { bs = 0; bpar ^= 1u; }
          }
        }
        if (++ch == chunks) {
          ch = 0; ++ti;
          n0 = (((int)blockIdx.x + ti * (int)gridDim.x) % p.n_tiles) * p.BN;
        }
      }
    }
  } else if (warp == 1 && HALO) {
    // ------------------------------------------------------------ MMA issuer, haloed patches
    // One thread, no division, descriptors by addition: the first version of this loop (ring slots by `%`, an election
    // and ~50 uniform-datapath instructions of descriptor assembly per tap) took ~810 clk per tap for 256 clk of
    // tensor work -- the issuing warp, not the operands, bounded the kernel (ncu: tensor pipe 32 % active, the MMA
    // warp's samples on integer division; profiles/r2_conv_halo.md).
    if (lane == 0) {
      const uint32_t idesc = make_idesc(128, p.BN, 0, 0, FMT_F16, FMT_F16);
      const uint32_t b_area = smem_u32(smem + (size_t)p.na * HALO_A_BYTES);
      const uint32_t a_area = smem_u32(smem);
      // descriptor = hi word (stride byte offset, version, swizzle mode: constant) | lo word (address >> 4, LBO = 16 B)
      const uint32_t a_hi = (uint32_t)(make_sdesc_sw128(0, 16, HALO_PITCH * 128) >> 32);
      const uint32_t b_hi = (uint32_t)(make_sdesc_sw128(0, 16, 1024) >> 32);
      const uint32_t lo_c = (uint32_t)make_sdesc_sw128(0, 16, 1024);            // the LBO field
      auto desc = [](uint32_t hi, uint32_t lo) { return ((uint64_t)hi << 32) | lo; };
      int as = 0, bs = 0, lt = 0;
      uint32_t apar = 0, bpar = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++lt) {
        const int buf = lt & 1;
        mbar_wait(&acc_empty[buf], (uint32_t)(((lt >> 1) & 1) ^ 1));
        if (p.b_resident && lt == 0) mbar_wait(&b_full[0], 0u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)buf * p.tmem_cols;
        for (int ch = 0; ch < chunks; ++ch) {
          mbar_wait(&full_bar[as], apar);
          tc_fence_after();
          const uint32_t a_lo = lo_c + ((a_area + (uint32_t)as * HALO_A_BYTES) >> 4);
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) {
            uint32_t b_lo;
            if (p.b_resident) {
              b_lo = lo_c + ((b_area + (uint32_t)tap * b_tile_bytes) >> 4);
            } else {
              mbar_wait(&b_full[bs], bpar);
              tc_fence_after();
              b_lo = lo_c + ((b_area + (uint32_t)bs * b_tile_bytes) >> 4);
            }
            const uint32_t a_tap = a_lo + (uint32_t)(((tap / 3) * HALO_PITCH + tap % 3) * 128 >> 4);   // the shifted view
#pragma unroll
            for (int k = 0; k < 4; ++k)
              mma_f16_ss(d_tmem, desc(a_hi, a_tap + 2 * k), desc(b_hi, b_lo + 2 * k), idesc, (uint32_t)((ch | tap | k) != 0));
            if (!p.b_resident) {
              mma_commit(&b_empty[bs]);
              if (++bs == p.nb) { bs = 0; bpar ^= 1u; }
            }
          }
          mma_commit(&empty_bar[as]);
          if (++as == p.na) { as = 0; apar ^= 1u; }
        }
        mma_commit(&acc_full[buf]);
      }
    }
    __syncwarp();
  } else if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (elect_one()) {
      int s = 0;
      uint32_t par = 1;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int tile_m = tile / p.n_tiles;
      const int n0 = (tile - tile_m * p.n_tiles) * p.BN;
      int kcol = 0;                                       // K column of the weight block: (tap * chunks + ch) * 64
      int img0, h0;
      if (p.HoWo >= 128) {
        const int tpi = p.HoWo >> 7;
        img0 = tile_m / tpi;
        h0 = (tile_m - img0 * tpi) * p.rows;
      } else {
        img0 = tile_m * p.imgs;
        h0 = 0;
      }
      const int pad = p.ksize >> 1;
      int dy = 0, dx = 0;
      for (int tap = 0; tap < p.taps; ++tap) {
        if (tap) { if (++dx == p.ksize) { dx = 0; ++dy; } }
        int hc, wc, ph = 0;
        if (p.stride == 1) {
          hc = h0 + dy - pad;
          wc = dx - pad;
        } else {  // stride 2, k = 3, pad = 1: input row 2*ho + dy - 1
          const int pa = (dy + 1) & 1, pb = (dx + 1) & 1;
          ph = pa * 2 + pb;
          hc = h0 + (dy == 0 ? -1 : 0);
          wc = (dx == 0 ? -1 : 0);
        }
        for (int ch = 0; ch < chunks; ++ch, kcol += 64) {
          mbar_wait(&empty_bar[s], par);
          uint8_t* a_dst = smem + (size_t)s * stage_bytes;
          uint8_t* b_dst = a_dst + A_TILE_BYTES;
          mbar_expect_tx(&full_bar[s], stage_bytes);
          if (ch < p.kc1) tma_load_4d(a_dst, &maps.a[ph], &full_bar[s], ch * 64, wc, hc, img0);
          else            tma_load_4d(a_dst, &maps.a2, &full_bar[s], (ch - p.kc1) * 64, wc, hc, img0);
          tma_load_2d(b_dst, &maps.b, &full_bar[s], kcol, n0);
          if (++s == p.stages) { s = 0; par ^= 1u; }      // ring slot and phase by counting: no division per block
        }
      }
      }  // tile
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    // one thread; ring slot / phase by counting and descriptors by addition (see the haloed issuer above: division and
    // descriptor assembly per block cost more issue time than the block's four MMAs take on the tensor pipe)
    if (lane == 0) {
      const uint32_t idesc = make_idesc(128, p.BN, 0, 0, FMT_F16, FMT_F16);
      const uint32_t hi = (uint32_t)(make_sdesc_sw128(0, 16, 1024) >> 32);
      const uint32_t lo_c = (uint32_t)make_sdesc_sw128(0, 16, 1024);
      const uint32_t ring = smem_u32(smem);
      int s = 0, lt = 0;
      uint32_t par = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++lt) {
        const int buf = lt & 1;
        mbar_wait(&acc_empty[buf], (uint32_t)(((lt >> 1) & 1) ^ 1));   // epilogue drained this buffer (2 tiles ago)
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)buf * p.tmem_cols;
        for (int it = 0; it < total_iters; ++it) {
          mbar_wait(&full_bar[s], par);
          tc_fence_after();
          const uint32_t a_lo = lo_c + ((ring + (uint32_t)s * stage_bytes) >> 4);
          const uint32_t b_lo = a_lo + (A_TILE_BYTES >> 4);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            mma_f16_ss(d_tmem, ((uint64_t)hi << 32) | (a_lo + 2 * k), ((uint64_t)hi << 32) | (b_lo + 2 * k), idesc,
                       (uint32_t)((it | k) != 0));
          mma_commit(&empty_bar[s]);  // frees the smem stage once these MMAs have read it
          if (++s == p.stages) { s = 0; par ^= 1u; }
        }
        mma_commit(&acc_full[buf]);
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------ epilogue (warps 2..9)
    // Per tile: (1) stage bias + emb for the tile's images in smem and issue ALL residual loads of this thread's
    // row into registers BEFORE waiting for the accumulator, so their HBM latency hides behind the tile's MMAs
    // (the first version loaded them 16 columns at a time after the wait: 62 % of its samples were long_sb);
    // (2) TMEM -> registers -> + (bias + emb) -> + residual -> fp16 / fp32-NCHW global.
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    const int csel = (warp - 2) >> 2;       // this warp takes the 16-column chunks cc with cc % CSPLIT == csel
    const int row = q * 32 + lane;
    const int et = threadIdx.x - 64;        // 0..EPI_THREADS-1 among the epilogue threads
    const bool has_res = (EPI == EPI_GNB) || (p.residual != nullptr);
    // HALO: tile origin (w0, h0, img0) of a patch tile and the staged TMA loads of a tile's residual / GroupNorm input
    auto halo_origin = [&](int tile_, int& n0_, int& w0_, int& h0_, int& img0_) {
      const int tm = tile_ / p.n_tiles;
      n0_ = (tile_ - tm * p.n_tiles) * p.BN;
      img0_ = tm / halo_tpi;
      const int t = tm - img0_ * halo_tpi, py = t / halo_pw;
      h0_ = py * HALO_PH;
      w0_ = (t - py * halo_pw) * HALO_PW;
    };
    auto issue_res_loads = [&](uint8_t* dst, uint64_t* bar, int n0_, int w0_, int h0_, int img0_) {
      const int nb_ = p.BN >> 6;
      mbar_expect_tx(bar, (uint32_t)nb_ * A_TILE_BYTES);
      for (int b = 0; b < nb_; ++b) {
        const int c = n0_ + b * 64;
        if (EPI != EPI_GNB || c < p.gnb_C1) tma_load_4d(dst + b * A_TILE_BYTES, &maps.r, bar, c, w0_, h0_, img0_);
        else                                tma_load_4d(dst + b * A_TILE_BYTES, &maps.r2, bar, c - p.gnb_C1, w0_, h0_, img0_);
      }
    };
    if (HALO && has_res && p.ns == 3 && et == 0 && (int)blockIdx.x < total_tiles) {   // first tile; later ones run a tile ahead
      int n0_, w0_, h0_, img0_;
      halo_origin(blockIdx.x, n0_, w0_, h0_, img0_);
      issue_res_loads(stage_out, &res_full[0], n0_, w0_, h0_, img0_);
    }
    const bool has_eb = (p.bias != nullptr) || (p.emb != nullptr);
    const int imgs_in_tile = p.HoWo >= 128 ? 1 : p.imgs;
    int lt = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++lt) {
    const int buf = lt & 1;
    const int tile_m = tile / p.n_tiles;
    const int n0 = (tile - tile_m * p.n_tiles) * p.BN;
    int img0 = 0, h0 = 0, w0 = 0;                      // tile origin for the TMA boxes (as in the producer)
    long long pix;
    if (HALO) {
      int n0_;
      halo_origin(tile, n0_, w0, h0, img0);
      pix = ((long long)img0 * p.Ho + h0 + (row >> 3)) * p.Wo + w0 + (row & 7);
    } else {
      pix = (long long)tile_m * 128 + row;
      if (p.HoWo >= 128) { const int tpi = p.HoWo >> 7; img0 = tile_m / tpi; h0 = (tile_m - img0 * tpi) * p.rows; }
      else { img0 = tile_m * p.imgs; h0 = 0; }
    }
    const bool valid = HALO ? true : pix < p.total_pixels;
    const int n_img = HALO ? img0 : valid ? (int)(pix / p.HoWo) : 0;
    const int hw = valid ? (int)(pix - (long long)n_img * p.HoWo) : 0;
    const int img_first = HALO ? img0 : (int)(((long long)tile_m * 128) / p.HoWo);
    const int sslot = HALO ? lt % p.ns : 0;           // staging slot of this tile
    uint8_t* const so_ptr = stage_out + (size_t)sslot * nblk_stage * A_TILE_BYTES;
    const uint32_t eb_u32 = eb_base + (uint32_t)buf * (EB_FLOATS * 4u);
    if (EPI == EPI_GNB) {
      // per (image of the tile, channel): z = x * rg + bb with rg = rstd * gamma, bb = beta - mean * rg
      for (int idx = et; idx < imgs_in_tile * p.BN; idx += EPI_THREADS) {
        const int im = idx / p.BN, c = idx - im * p.BN;
        const int col = n0 + c, img = min(img_first + im, p.n_images - 1);
        const int g = col / p.gnb_cpg;
        const float rs = __ldg(p.gnb_rstd + img * 32 + g), mu = __ldg(p.gnb_mean + img * 32 + g);
        const float rg = rs * __ldg(p.gnb_gamma + col);
        sts32f(eb_u32 + (uint32_t)idx * 8u, rg);
        sts32f(eb_u32 + (uint32_t)idx * 8u + 4u, fmaf(-mu, rg, __ldg(p.gnb_beta + col)));
      }
    } else if (has_eb) {
      for (int idx = et; idx < imgs_in_tile * p.BN; idx += EPI_THREADS) {
        const int im = idx / p.BN, c = idx - im * p.BN;
        const int col = n0 + c, img = img_first + im;
        float v = 0.f;
        if (col < p.Cout) {
          if (p.bias) v = __ldg(p.bias + col);
          if (p.emb && img < p.n_images) v += __ldg(p.emb + (long long)img * p.ldemb + col);
        }
        sts32f(eb_u32 + (uint32_t)idx * 4u, v);
      }
    }
    const int nblk = p.BN >> 6;
    if (!HALO && p.tma_out && et == 0) {
      tma_store_wait_read();                           // previous tile's store has finished reading the staging buffer
      // residual, or the GroupNorm input x = [x1 | x2], 64-channel block by block
      if (has_res) issue_res_loads(stage_out, res_full, n0, 0, h0, img0);
    }
    if (HALO && et == 0) {
      // Staging slots rotate; one bulk-store group is committed per tile.  With two or more slots the previous tile's
      // store may still be draining while this tile is computed (only older stores must have finished reading); with
      // three, the NEXT tile's residual / GroupNorm input is fetched now, a whole tile ahead of its use, into the slot
      // the store of two tiles ago has left.
      if (p.ns >= 2) tma_store_wait_read_1(); else tma_store_wait_read();
      if (has_res) {
        if (p.ns == 3) {
          const int next = tile + (int)gridDim.x;
          if (next < total_tiles) {
            int n0_, w0_, h0_, img0_;
            const int ns_ = (lt + 1) % 3;
            halo_origin(next, n0_, w0_, h0_, img0_);
            issue_res_loads(stage_out + (size_t)ns_ * nblk_stage * A_TILE_BYTES, &res_full[ns_], n0_, w0_, h0_, img0_);
          }
        } else {
          issue_res_loads(so_ptr, &res_full[sslot], n0, w0, h0, img0);
        }
      }
    }
    uint4 mbits = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);   // EPI_GNB: dropout keep bits of this row
    if (EPI == EPI_GNB && p.gnb_mask != nullptr && valid) {
      const uint8_t* mp = p.gnb_mask + (pix * p.Cout + n0) / 8;                      // 8-byte aligned (n0 % 64 == 0)
      const uint2 lo = __ldg(reinterpret_cast<const uint2*>(mp));
      mbits.x = lo.x; mbits.y = lo.y;
      if (p.BN > 64) { const uint2 hi = __ldg(reinterpret_cast<const uint2*>(mp + 8)); mbits.z = hi.x; mbits.w = hi.y; }
    }
    const bool fast_cols = (n0 + p.BN <= p.Cout);      // whole tile inside the logical channels: vector path
    // direct path (outputs that are not whole 64-channel blocks): residual segments by 16-byte loads inside the chunk
    // loop (the staged path, which serves every other layer, fetches the residual tile by TMA)
    const bool res_vec = !p.tma_out && p.residual && valid && fast_cols && ((p.ldres & 7) == 0) && ((((uintptr_t)p.residual) & 15) == 0);
    named_bar_sync(1, EPI_THREADS);                                 // eb staging visible to the epilogue warps
    mbar_wait(&acc_full[buf], (uint32_t)((lt >> 1) & 1));
    tc_fence_after();
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)buf * p.tmem_cols;
    const uint32_t eb_row = eb_u32 + (uint32_t)((n_img - img_first) * p.BN) * 4u;
    if (p.tma_out) {
      // ---- staged path: residual arrives by TMA in the 128B-swizzled tile layout, results go back to the same
      // smem slots (conflict-free 16-byte accesses) and leave with one TMA store per 64-channel block
      const uint32_t so_u32 = smem_u32(so_ptr);
      if (has_res) {
        if (HALO) mbar_wait(&res_full[sslot], (uint32_t)((lt / p.ns) & 1));
        else      mbar_wait(res_full, (uint32_t)(lt & 1));
      }
      // fused GroupNorm sums: this warp's 32 rows belong to one image (Ho*Wo % 32 == 0, checked on the host)
      const long long wpix = (long long)tile_m * 128 + q * 32;
      const bool wvalid = HALO ? true : wpix < p.total_pixels;
      const int wslot = (!HALO && wvalid) ? (int)(wpix / p.HoWo) - img_first : 0;
      // a real loop over this warp's chunks: unrolled eight-fold the three epilogue variants came to 30 k SASS
      // instructions (480 KB of code), most of it never executed by a given warp
#pragma unroll 1
      for (int cc = csel; cc < 8; cc += CSPLIT) {
        const int c0 = cc * 16;
        if (c0 < p.BN) {
          uint32_t r[16];
          __syncwarp();
          tmem_ld16(taddr + (uint32_t)c0, r);
          tmem_ld_wait();
          float v[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
          if (EPI != EPI_GNB && has_eb) {
#pragma unroll
            for (int j4 = 0; j4 < 4; ++j4) {
              const float4 e = lds128f(eb_row + (uint32_t)(c0 + 4 * j4) * 4u);
              v[4 * j4] += e.x; v[4 * j4 + 1] += e.y; v[4 * j4 + 2] += e.z; v[4 * j4 + 3] += e.w;
            }
          }
          const uint32_t a0 = so_u32 + (uint32_t)(c0 >> 6) * A_TILE_BYTES + sw128_offset((uint32_t)row, (uint32_t)(c0 & 63));
          const uint32_t a1 = so_u32 + (uint32_t)(c0 >> 6) * A_TILE_BYTES + sw128_offset((uint32_t)row, (uint32_t)((c0 & 63) + 8));
          if (EPI == EPI_GNB) {
            // v = dy.  dyh = dy * silu'(z) * mask / (1 - p);  sums of dyh and dyh * x over the warp's 32 pixels
            const uint4 u0 = lds128(a0), u1 = lds128(a1);
            const uint32_t w[8] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w};
            float x[16];
#pragma unroll
            for (int j = 0; j < 8; ++j) { float2 f = unpack_f16x2(w[j]); x[2 * j] = f.x; x[2 * j + 1] = f.y; }
            const uint32_t cst = eb_u32 + (uint32_t)((valid ? n_img - img_first : 0) * p.BN + c0) * 8u;
            const uint32_t mword = cc < 2 ? mbits.x : cc < 4 ? mbits.y : cc < 6 ? mbits.z : mbits.w;
            const uint32_t m16 = (mword >> ((cc & 1) * 16)) & 0xffffu;
            float red[32];
#pragma unroll
            for (int j2 = 0; j2 < 8; ++j2) {
              const float4 c4 = lds128f(cst + (uint32_t)j2 * 16u);      // (rg, bb) of channels 2*j2, 2*j2 + 1
#pragma unroll
              for (int h = 0; h < 2; ++h) {
                const int j = 2 * j2 + h;
                float dh = v[j];
                if (p.gnb_act) {
                  const float z = fmaf(x[j], h ? c4.z : c4.x, h ? c4.w : c4.y);
                  const float sg = fast_sigmoid(z);
                  dh *= sg * fmaf(z, 1.0f - sg, 1.0f);
                }
                dh = ((m16 >> j) & 1u) ? dh * p.gnb_keep_scale : 0.f;
                v[j] = dh;
                red[2 * j] = wvalid ? dh : 0.f;
                red[2 * j + 1] = wvalid ? dh * x[j] : 0.f;
              }
            }
            warp_reduce_scatter32(red, lane);                           // lane l: channel c0 + l/2, quantity l & 1
            red_shared_add_f32(sacc_u32 + (uint32_t)((wslot * p.BN + c0) * 2 + lane) * 4u, red[0]);
          } else if (p.residual) {
            const uint4 u0 = lds128(a0), u1 = lds128(a1);
            const uint32_t w[8] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w};
#pragma unroll
            for (int j = 0; j < 8; ++j) { float2 f = unpack_f16x2(w[j]); v[2 * j] += f.x; v[2 * j + 1] += f.y; }
          }
          if (EPI == EPI_STATS) {
            float red[16];                                              // 8 channel pairs x (sum, sum of squares)
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              red[2 * j] = wvalid ? v[2 * j] + v[2 * j + 1] : 0.f;
              red[2 * j + 1] = wvalid ? fmaf(v[2 * j], v[2 * j], v[2 * j + 1] * v[2 * j + 1]) : 0.f;
            }
            warp_reduce_scatter16(red, lane);
            if ((lane & 1) == 0) {
              const int vi = ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
              red_shared_add_f32(sacc_u32 + (uint32_t)(wslot * p.BN + c0 + vi) * 4u, red[0]);
            }
          }
          uint4 o0, o1;
          o0.x = pack_f16x2(v[0], v[1]);   o0.y = pack_f16x2(v[2], v[3]);
          o0.z = pack_f16x2(v[4], v[5]);   o0.w = pack_f16x2(v[6], v[7]);
          o1.x = pack_f16x2(v[8], v[9]);   o1.y = pack_f16x2(v[10], v[11]);
          o1.z = pack_f16x2(v[12], v[13]); o1.w = pack_f16x2(v[14], v[15]);
          sts128(a0, o0);
          sts128(a1, o1);
        }
      }
      tc_fence_before();
      mbar_arrive(&acc_empty[buf]);
      fence_proxy_async_smem();                        // staged tile visible to the TMA engine
      named_bar_sync(2, EPI_THREADS);
      if (et == 0) {
        for (int b = 0; b < nblk; ++b) tma_store_4d(&maps.o, so_ptr + b * A_TILE_BYTES, n0 + b * 64, w0, h0, img0);
        tma_store_commit();
      }
      if (EPI == EPI_STATS) {
        // sacc[slot][BN] -> stats[img][(n0 + k) / 2][k & 1] (fp64), re-zeroed for the next tile (whose sums start
        // after its named barrier 1, which this thread reaches only after the stores below)
        for (int idx = et; idx < imgs_in_tile * p.BN; idx += EPI_THREADS) {
          const int im = idx / p.BN, k = idx - im * p.BN;
          const int img = img_first + im;
          const uint32_t sa = sacc_u32 + (uint32_t)idx * 4u;
          float val;
          asm volatile("ld.shared.f32 %0, [%1];" : "=f"(val) : "r"(sa));
          sts32f(sa, 0.f);
          if (img < p.n_images) atomicAdd(p.stats + ((long long)img * (p.Cout >> 1) + ((n0 + k) >> 1)) * 2 + (k & 1), (double)val);
        }
      }
      if (EPI == EPI_GNB) {
        for (int idx = et; idx < imgs_in_tile * p.BN * 2; idx += EPI_THREADS) {
          const int im = idx / (p.BN * 2), k = idx - im * p.BN * 2;
          const int img = img_first + im;
          const uint32_t sa = sacc_u32 + (uint32_t)idx * 4u;
          float val;
          asm volatile("ld.shared.f32 %0, [%1];" : "=f"(val) : "r"(sa));
          sts32f(sa, 0.f);
          if (img < p.n_images) atomicAdd(p.gnb_ab + ((long long)img * p.Cout + n0) * 2 + k, val);
        }
      }
      continue;
    }
#pragma unroll 1
    for (int cc = csel; cc < 8; cc += CSPLIT) {
      const int c0 = cc * 16;
      if (c0 < p.BN) {
      uint32_t r[16];
      __syncwarp();
      tmem_ld16(taddr + (uint32_t)c0, r);
      tmem_ld_wait();
      const int col0 = n0 + c0;
      if (valid && col0 < p.Cout) {
      float v[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
      const int ncol = min(16, p.Cout - col0);
      if (has_eb) {
#pragma unroll
        for (int j4 = 0; j4 < 4; ++j4) {
          const float4 e = lds128f(eb_row + (uint32_t)(c0 + 4 * j4) * 4u);   // zero beyond Cout
          v[4 * j4] += e.x; v[4 * j4 + 1] += e.y; v[4 * j4 + 2] += e.z; v[4 * j4 + 3] += e.w;
        }
      }
      if (p.residual) {
        if (res_vec) {
          const __half* rp = p.residual + pix * p.ldres + col0;
          const uint4 u0 = ldg_nc_u4(rp), u1 = ldg_nc_u4(rp + 8);
          const uint32_t w[8] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w};
#pragma unroll
          for (int j = 0; j < 8; ++j) { float2 f = unpack_f16x2(w[j]); v[2 * j] += f.x; v[2 * j + 1] += f.y; }
        } else {
          const __half* rp = p.residual + pix * p.ldres + col0;
          for (int j = 0; j < ncol; ++j) v[j] += __half2float(rp[j]);
        }
      }
      if (p.out) {
        __half* op = p.out + pix * p.ldo + col0;
        if (ncol == 16 && ((((uintptr_t)op) & 15) == 0)) {
          uint4 o0, o1;
          o0.x = pack_f16x2(v[0], v[1]);   o0.y = pack_f16x2(v[2], v[3]);
          o0.z = pack_f16x2(v[4], v[5]);   o0.w = pack_f16x2(v[6], v[7]);
          o1.x = pack_f16x2(v[8], v[9]);   o1.y = pack_f16x2(v[10], v[11]);
          o1.z = pack_f16x2(v[12], v[13]); o1.w = pack_f16x2(v[14], v[15]);
          *reinterpret_cast<uint4*>(op) = o0;
          *reinterpret_cast<uint4*>(op + 8) = o1;
        } else {
          for (int j = 0; j < ncol; ++j) op[j] = __float2half_rn(fminf(fmaxf(v[j], -65504.f), 65504.f));
        }
      }
      if (p.out_nchw) {
        for (int j = 0; j < ncol; ++j)
          p.out_nchw[((long long)n_img * p.Cout + col0 + j) * p.HoWo + hw] = v[j];
      }
      }  // valid
      }  // c0 < BN
    }
    tc_fence_before();
    mbar_arrive(&acc_empty[buf]);   // all of this thread's TMEM reads of the buffer have completed (ld wait above)
    }  // tile
    if (p.tma_out && et == 0) tma_store_wait_all();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 2 * p.tmem_cols);
}

static bool is_pow2(int v) { return v > 0 && (v & (v - 1)) == 0; }

// Build the 4-D activation map (C, W, H, N) with box (64, Wo, rows, imgs).
static int make_act_map(CUtensorMap* m, const void* base, int C, long long ld, int N, int H, int W, int phase_stride,
                        int Wo, int rows, int imgs) {
  // phase_stride == 2: this map views every second row/column starting at `base`
  const uint64_t dims[4] = {(uint64_t)C, (uint64_t)(W / phase_stride), (uint64_t)(H / phase_stride), (uint64_t)N};
  const uint64_t str[3] = {(uint64_t)ld * 2 * phase_stride, (uint64_t)ld * 2 * W * phase_stride,
                           (uint64_t)ld * 2 * W * H};
  const uint32_t box[4] = {64u, (uint32_t)Wo, (uint32_t)rows, (uint32_t)imgs};
  return make_tmap_16b(m, base, 4, dims, str, box);
}

}  // namespace b2d

using namespace b2d;

extern "C" int b2d_conv2d(const b2d_conv_args* a, void* stream) {
  B2D_REQUIRE(a && a->x && a->w, B2D_ERR_SHAPE, "conv2d: null argument");
  B2D_REQUIRE(a->ksize == 1 || a->ksize == 3, B2D_ERR_SHAPE, "conv2d: ksize %d not in {1,3}", a->ksize);
  B2D_REQUIRE(a->stride == 1 || (a->stride == 2 && a->ksize == 3), B2D_ERR_SHAPE, "conv2d: stride %d unsupported",
              a->stride);
  const int C2 = a->x2 ? a->C2 : 0;
  B2D_REQUIRE(a->C1 > 0 && a->C1 % 64 == 0 && C2 % 64 == 0, B2D_ERR_SHAPE,
              "conv2d: channel counts (%d,%d) must be multiples of 64", a->C1, C2);
  B2D_REQUIRE(a->stride == 1 || C2 == 0, B2D_ERR_SHAPE, "conv2d: two-source input needs stride 1");
  B2D_REQUIRE(a->N > 0 && a->H > 0 && a->W > 0 && a->H % a->stride == 0 && a->W % a->stride == 0, B2D_ERR_SHAPE,
              "conv2d: bad spatial size N=%d H=%d W=%d", a->N, a->H, a->W);
  const int Ho = a->H / a->stride, Wo = a->W / a->stride;
  B2D_REQUIRE(is_pow2(Wo) && Wo <= 128, B2D_ERR_SHAPE, "conv2d: output width %d must be a power of two <= 128", Wo);
  int rows = 128 / Wo, imgs = 1;
  if (rows > Ho) {
    B2D_REQUIRE((128 % (Ho * Wo)) == 0, B2D_ERR_SHAPE, "conv2d: plane %dx%d does not tile 128 pixels", Ho, Wo);
    rows = Ho;
    imgs = 128 / (Ho * Wo);
    B2D_REQUIRE(imgs <= 8, B2D_ERR_SHAPE, "conv2d: plane %dx%d too small (more than 8 images per 128-pixel tile)", Ho, Wo);
  } else {
    B2D_REQUIRE(Ho % rows == 0, B2D_ERR_SHAPE, "conv2d: height %d not a multiple of %d tile rows", Ho, rows);
  }
  B2D_REQUIRE(a->Cout > 0 && a->CoutPad >= a->Cout && a->CoutPad % 16 == 0, B2D_ERR_SHAPE,
              "conv2d: Cout=%d CoutPad=%d", a->Cout, a->CoutPad);
  B2D_REQUIRE(a->out || a->out_nchw_f32, B2D_ERR_SHAPE, "conv2d: no output buffer");
  int BN;
  if (a->CoutPad <= 256 && (a->CoutPad <= 128 || a->CoutPad == 256)) BN = a->CoutPad;
  else if (a->CoutPad % 128 == 0) BN = 128;
  else if (a->CoutPad % 64 == 0) BN = 64;
  else return set_err(B2D_ERR_SHAPE, "conv2d: CoutPad=%d cannot be tiled", a->CoutPad);
  if (BN == 256 && a->CoutPad == 256) BN = 128;  // two N tiles: more CTAs, 2 CTAs/SM
  B2D_REQUIRE(BN % 16 == 0 && BN >= 16 && BN <= 256, B2D_ERR_SHAPE, "conv2d: BN=%d", BN);
  const bool tma_out = a->out != nullptr && a->out_nchw_f32 == nullptr && a->Cout % 64 == 0 && BN % 64 == 0 &&
                       a->ldo % 8 == 0 && aligned16(a->out) &&
                       (a->residual == nullptr || (a->ldres % 8 == 0 && aligned16(a->residual)));
  // Haloed patch tiles (see the header): 3x3 stride 1 on planes of 16k x 8m pixels with the TMA-store epilogue.
  // OPT-IN (B2D_CONV_HALO=1, read per call so that tests and benchmarks can switch): measured on the train64 layers
  // (scripts/bench_conv_halo.py, profiles/r2_conv_halo.md) the variant wins 3-10 % on the deep forward layers
  // (>= 192 input channels, BN = 128) and loses wherever the epilogue or HBM bounds the layer -- every 64 x 64 plane and
  // every GroupNorm-backward epilogue -- because one CTA per SM runs one epilogue chain where two CTAs run two.
  bool halo = false;
  {
    const char* e = getenv("B2D_CONV_HALO");
    halo = (e && e[0] == '1') && a->ksize == 3 && a->stride == 1 && tma_out && a->H % HALO_PH == 0 && a->W % HALO_PW == 0;
  }
  const bool b_resident = halo && (a->C1 + C2 == 64);
  if (halo) {
    BN = (a->CoutPad % 128 == 0 && !b_resident) ? 128 : 64;
    rows = HALO_PH; imgs = 1;
  }

  ConvMaps maps;
  memset(&maps, 0, sizeof(maps));
  int rc;
  // box of the output / residual / GroupNorm-input maps: full rows, or the 8 x 16 patch
  const int obox_w = halo ? HALO_PW : Wo;
  if (halo) {
    rc = make_act_map(&maps.a[0], a->x, a->C1, a->ldx, a->N, a->H, a->W, 1, HALO_PITCH, HALO_ROWS, 1);
    if (rc) return rc;
    for (int i = 1; i < 4; ++i) maps.a[i] = maps.a[0];
  } else if (a->stride == 1) {
    rc = make_act_map(&maps.a[0], a->x, a->C1, a->ldx, a->N, a->H, a->W, 1, Wo, rows, imgs);
    if (rc) return rc;
    for (int i = 1; i < 4; ++i) maps.a[i] = maps.a[0];
  } else {
    for (int pa = 0; pa < 2; ++pa)
      for (int pb = 0; pb < 2; ++pb) {
        const __half* base = (const __half*)a->x + ((long long)pa * a->W + pb) * a->ldx;
        rc = make_act_map(&maps.a[pa * 2 + pb], base, a->C1, a->ldx, a->N, a->H, a->W, 2, Wo, rows, imgs);
        if (rc) return rc;
      }
  }
  if (C2) {
    rc = halo ? make_act_map(&maps.a2, a->x2, C2, a->ldx2, a->N, a->H, a->W, 1, HALO_PITCH, HALO_ROWS, 1)
              : make_act_map(&maps.a2, a->x2, C2, a->ldx2, a->N, a->H, a->W, 1, Wo, rows, imgs);
    if (rc) return rc;
  } else {
    maps.a2 = maps.a[0];
  }
  const int taps = a->ksize * a->ksize;
  const long long Ktot = (long long)taps * (a->C1 + C2);
  {
    const uint64_t dims[2] = {(uint64_t)Ktot, (uint64_t)a->CoutPad};
    const uint64_t str[1] = {(uint64_t)Ktot * 2};
    const uint32_t box[2] = {64u, (uint32_t)BN};
    rc = make_tmap_16b(&maps.b, a->w, 2, dims, str, box);
    if (rc) return rc;
  }

  const bool want_stats = a->gn_stats_out != nullptr;
  const bool want_gnb = a->gnb_ab != nullptr;
  if (want_stats || want_gnb) {
    B2D_REQUIRE(!(want_stats && want_gnb), B2D_ERR_SHAPE, "conv2d: gn_stats_out and gnb_* are exclusive");
    B2D_REQUIRE(tma_out && b2d_conv2d_gn_fusable(Ho, Wo, a->Cout), B2D_ERR_SHAPE,
                "conv2d: fused GroupNorm epilogue needs Cout %% 64 == 0, fp16 out, Ho*Wo %% 32 == 0 (Cout=%d, plane %dx%d)",
                a->Cout, Ho, Wo);
    B2D_REQUIRE((128 / (Ho * Wo) <= 1 ? 1 : 128 / (Ho * Wo)) * BN * 2 <= SACC_FLOATS, B2D_ERR_SHAPE,
                "conv2d: fused GroupNorm epilogue: %d images x %d channels per tile exceed the staging", imgs, BN);
  }
  if (want_gnb) {
    B2D_REQUIRE(a->residual == nullptr && a->bias == nullptr && a->emb == nullptr, B2D_ERR_SHAPE,
                "conv2d: the GroupNorm-backward epilogue takes no bias / emb / residual");
    B2D_REQUIRE(a->gnb_x1 && a->gnb_mean && a->gnb_rstd && a->gnb_gamma && a->gnb_beta, B2D_ERR_SHAPE,
                "conv2d: gnb_* pointers missing");
    const int gC2 = a->Cout - a->gnb_C1;
    B2D_REQUIRE(a->gnb_C1 > 0 && a->gnb_C1 % 64 == 0 && gC2 >= 0 && gC2 % 64 == 0 && (gC2 == 0 || a->gnb_x2), B2D_ERR_SHAPE,
                "conv2d: gnb channel split (%d, %d) must be multiples of 64", a->gnb_C1, gC2);
    B2D_REQUIRE(a->Cout % 32 == 0 && a->gnb_ld1 % 8 == 0 && aligned16(a->gnb_x1) &&
                (gC2 == 0 || (a->gnb_ld2 % 8 == 0 && aligned16(a->gnb_x2))), B2D_ERR_ALIGN, "conv2d: gnb inputs misaligned");
  }
  if (tma_out) {
    rc = make_act_map(&maps.o, a->out, a->Cout, a->ldo, a->N, Ho, Wo, 1, obox_w, rows, imgs);
    if (rc) return rc;
    maps.r = maps.o; maps.r2 = maps.o;
    if (want_gnb) {
      rc = make_act_map(&maps.r, a->gnb_x1, a->gnb_C1, a->gnb_ld1, a->N, Ho, Wo, 1, obox_w, rows, imgs);
      if (rc) return rc;
      if (a->Cout > a->gnb_C1) {
        rc = make_act_map(&maps.r2, a->gnb_x2, a->Cout - a->gnb_C1, a->gnb_ld2, a->N, Ho, Wo, 1, obox_w, rows, imgs);
        if (rc) return rc;
      }
    } else if (a->residual) {
      rc = make_act_map(&maps.r, a->residual, a->Cout, a->ldres, a->N, Ho, Wo, 1, obox_w, rows, imgs);
      if (rc) return rc;
    }
  } else {
    maps.o = maps.a[0]; maps.r = maps.a[0]; maps.r2 = maps.a[0];
  }

  ConvParams p;
  p.tma_out = tma_out ? 1 : 0;
  p.taps = taps; p.ksize = a->ksize; p.stride = a->stride;
  p.kc1 = a->C1 / 64; p.kc2 = C2 / 64;
  p.HoWo = Ho * Wo; p.Wo = Wo; p.rows = rows; p.imgs = imgs;
  p.total_pixels = a->N * Ho * Wo;
  p.n_images = a->N;
  p.Cout = a->Cout; p.BN = BN;
  const uint32_t stage_bytes = A_TILE_BYTES + (uint32_t)BN * 128u;
  // smem per CTA <= 112 KB so that two persistent CTAs share an SM: 72 KB operand ring + 32 KB output staging
  // + 4 KB bias/emb staging + barriers
  int stages = (int)(73728u / stage_bytes);
  if (stages < 2) stages = 2;
  if (stages > 6) stages = 6;
  B2D_REQUIRE(imgs * BN <= EB_FLOATS, B2D_ERR_SHAPE, "conv2d: %d images x %d channels per tile exceed the epilogue staging", imgs, BN);
  p.stages = stages;
  p.tmem_cols = BN <= 32 ? 32 : BN <= 64 ? 64 : BN <= 128 ? 128 : 256;
  p.bias = a->bias; p.emb = a->emb; p.ldemb = a->ldemb;
  p.residual = (const __half*)a->residual; p.ldres = a->ldres;
  p.out = (__half*)a->out; p.ldo = a->ldo;
  p.out_nchw = a->out_nchw_f32;
  p.stats = a->gn_stats_out;
  p.gnb_C1 = a->gnb_C1; p.gnb_cpg = a->Cout / 32 > 0 ? a->Cout / 32 : 1; p.gnb_act = a->gnb_act;
  p.gnb_mean = a->gnb_mean; p.gnb_rstd = a->gnb_rstd; p.gnb_gamma = a->gnb_gamma; p.gnb_beta = a->gnb_beta;
  p.gnb_mask = a->gnb_mask; p.gnb_keep_scale = a->gnb_mask ? a->gnb_keep_scale : 1.0f;
  p.gnb_ab = a->gnb_ab;

  p.Ho = Ho;
  p.b_resident = b_resident ? 1 : 0;
  // shared memory of the haloed variant (<= 227 KB, one CTA per SM): input slots | weights | barriers + bias/emb |
  // staging | GroupNorm sums.  64 input channels: 4 x 23 + 72 (resident weights) + 3 x 16 staging = 220 KB;
  // streamed weights: what the mainloop waits for is the weight ring (L2 latency ~2000 clk at ~60 B/clk/SM wants
  // >= 100 KB in flight), so it gets the room: BN = 64: 3 x 23 + 12 x 8 + 3 x 16; BN = 128: 3 x 23 + 5 x 16 + 2 x 32.
  p.na = b_resident ? 4 : 3;
  p.nb = b_resident ? 0 : (BN == 128 ? 5 : 12);
  p.ns = BN == 128 ? 2 : 3;
  if (halo) {
    const char* e = getenv("B2D_CONV_HALO_CFG");      // "na,nb,ns": developer override for ring-depth experiments
    int v[3];
    if (e && sscanf(e, "%d,%d,%d", &v[0], &v[1], &v[2]) == 3 && !b_resident && BN == 128) { p.na = v[0]; p.nb = v[1]; p.ns = v[2]; }
    B2D_REQUIRE(p.na >= 1 && p.na <= 4 && p.nb >= 0 && p.nb <= 12 && p.ns >= 1 && p.ns <= 3 && (b_resident || p.nb >= 2),
                B2D_ERR_SHAPE, "conv2d: haloed tiles: ring depths (%d, %d, %d)", p.na, p.nb, p.ns);
  }
  p.b_bytes = (uint32_t)(b_resident ? 9 : p.nb) * (uint32_t)BN * 128u;
  const size_t smem = halo ? (size_t)p.na * HALO_A_BYTES + p.b_bytes + 1024 + 2 * EB_FLOATS * 4 +
                                 (size_t)p.ns * (BN / 64) * A_TILE_BYTES + SACC_FLOATS * 4 + 1024
                           : (size_t)stages * stage_bytes + 1024 /* barriers */ + 2 * EB_FLOATS * 4 + 2 * A_TILE_BYTES +
                                 SACC_FLOATS * 4 + 1024;
  static std::once_flag attr_once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(attr_once, [] {
    for (auto fn : {(const void*)conv_gemm_kernel<EPI_PLAIN, false>, (const void*)conv_gemm_kernel<EPI_STATS, false>,
                    (const void*)conv_gemm_kernel<EPI_GNB, false>}) {
      cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
      if (e != cudaSuccess) attr_err = e;
    }
    for (auto fn : {(const void*)conv_gemm_kernel<EPI_PLAIN, true>, (const void*)conv_gemm_kernel<EPI_STATS, true>,
                    (const void*)conv_gemm_kernel<EPI_GNB, true>}) {
      cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
      if (e != cudaSuccess) attr_err = e;
    }
  });
  B2D_CHECK_CUDA(attr_err);
  p.m_tiles = (p.total_pixels + 127) / 128;
  p.n_tiles = (a->CoutPad + BN - 1) / BN;
  const long long tiles = (long long)p.m_tiles * p.n_tiles;
  if (halo) {
    B2D_REQUIRE(smem <= 227 * 1024 && p.total_pixels % 128 == 0, B2D_ERR_SHAPE, "conv2d: haloed tiles: %zu bytes", smem);
    long long ctas = num_sms();                // one persistent CTA per SM (up to 221 KB of shared memory)
    if (ctas > tiles) ctas = tiles;
    if (b_resident) ctas -= ctas % p.n_tiles;  // tile_n = tile % n_tiles is then the same for every tile of a CTA
    dim3 grid((unsigned)ctas);
    if (want_gnb)        conv_gemm_kernel<EPI_GNB, true><<<grid, HALO_THREADS, smem, (cudaStream_t)stream>>>(maps, p);
    else if (want_stats) conv_gemm_kernel<EPI_STATS, true><<<grid, HALO_THREADS, smem, (cudaStream_t)stream>>>(maps, p);
    else                 conv_gemm_kernel<EPI_PLAIN, true><<<grid, HALO_THREADS, smem, (cudaStream_t)stream>>>(maps, p);
    B2D_CHECK_LAUNCH();
    return B2D_OK;
  }
  long long ctas = 2LL * num_sms();            // two persistent CTAs per SM (<= 96 KB ring, <= 256 TMEM columns each)
  if (ctas > tiles) ctas = tiles;
  dim3 grid((unsigned)ctas);
  if (want_gnb)        conv_gemm_kernel<EPI_GNB, false><<<grid, CONV_THREADS, smem, (cudaStream_t)stream>>>(maps, p);
  else if (want_stats) conv_gemm_kernel<EPI_STATS, false><<<grid, CONV_THREADS, smem, (cudaStream_t)stream>>>(maps, p);
  else                 conv_gemm_kernel<EPI_PLAIN, false><<<grid, CONV_THREADS, smem, (cudaStream_t)stream>>>(maps, p);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

extern "C" int b2d_conv2d_gn_fusable(int Ho, int Wo, int Cout) {
  if (Ho <= 0 || Wo <= 0 || Cout <= 0 || Cout % 64 != 0) return 0;
  const int hw = Ho * Wo;
  if (hw % 32 != 0) return 0;
  const int imgs = hw >= 128 ? 1 : 128 / hw;
  const int BN = Cout <= 128 ? Cout : (Cout % 128 == 0 ? 128 : 64);
  return imgs * BN * 2 <= 512 ? 1 : 0;
}
