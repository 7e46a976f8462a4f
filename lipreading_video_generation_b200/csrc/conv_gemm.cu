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
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer,
// warps 2..5 = epilogue (TMEM -> registers -> bias/emb/residual -> bf16 / fp32-NCHW global).
#include "common.cuh"
#include "tmap.cuh"
#include "umma.cuh"

namespace b2d {
using namespace umma;

struct ConvMaps {
  CUtensorMap a[4];  // source 1 (4 phases when stride == 2, else a[0])
  CUtensorMap a2;    // source 2 (stride 1 only)
  CUtensorMap b;     // weights
};

struct ConvParams {
  int taps, ksize, stride;
  int kc1, kc2;        // 64-channel chunks of source 1 / source 2
  int HoWo, Wo;        // output plane
  int rows, imgs;      // box: rows per tile, images per tile
  int total_pixels;    // N*Ho*Wo
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
};

constexpr int CONV_THREADS = 192;
constexpr int A_TILE_BYTES = 128 * 128;  // 128 pixels x 64 bf16

__global__ void __launch_bounds__(CONV_THREADS)
conv_gemm_kernel(const __grid_constant__ ConvMaps maps, const ConvParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t b_tile_bytes = (uint32_t)p.BN * 128u;
  const uint32_t stage_bytes = A_TILE_BYTES + b_tile_bytes;
  uint64_t* full_bar = (uint64_t*)(smem + (size_t)p.stages * stage_bytes);
  uint64_t* empty_bar = full_bar + p.stages;
  uint64_t* acc_full = empty_bar + p.stages;   // [2] accumulator buffer b holds a finished tile
  uint64_t* acc_empty = acc_full + 2;          // [2] epilogue has drained buffer b
  uint32_t* tmem_slot = (uint32_t*)(acc_empty + 2);

  // Persistent CTA: tiles tile0, tile0 + gridDim.x, ...; the TMA ring and the two TMEM accumulator buffers run
  // across tile boundaries, so tile i's epilogue overlaps tile i+1's loads and MMAs (the non-persistent
  // version spent ~22k clk per CTA on a 1.2k-clk mainloop: profiles/r1_conv.md).
  const int total_tiles = p.m_tiles * p.n_tiles;
  const int chunks = p.kc1 + p.kc2;
  const int total_iters = p.taps * chunks;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(&acc_full[0], 1); mbar_init(&acc_full[1], 1);
    mbar_init(&acc_empty[0], 128); mbar_init(&acc_empty[1], 128);
    fence_barrier_init();
    prefetch_tmap(&maps.a[0]);
    prefetch_tmap(&maps.b);
  }
  if (warp == 1) tmem_alloc(tmem_slot, 2 * p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (elect_one()) {
      int it = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int tile_m = tile / p.n_tiles;
      const int n0 = (tile - tile_m * p.n_tiles) * p.BN;
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
      for (int tap = 0; tap < p.taps; ++tap) {
        const int dy = tap / p.ksize, dx = tap - dy * p.ksize;
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
        for (int ch = 0; ch < chunks; ++ch, ++it) {
          const int s = it % p.stages;
          const uint32_t par = (uint32_t)((it / p.stages) & 1);
          mbar_wait(&empty_bar[s], par ^ 1u);
          uint8_t* a_dst = smem + (size_t)s * stage_bytes;
          uint8_t* b_dst = a_dst + A_TILE_BYTES;
          mbar_expect_tx(&full_bar[s], stage_bytes);
          if (ch < p.kc1) tma_load_4d(a_dst, &maps.a[ph], &full_bar[s], ch * 64, wc, hc, img0);
          else            tma_load_4d(a_dst, &maps.a2, &full_bar[s], (ch - p.kc1) * 64, wc, hc, img0);
          tma_load_2d(b_dst, &maps.b, &full_bar[s], (tap * chunks + ch) * 64, n0);
        }
      }
      }  // tile
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    const uint32_t idesc = make_idesc(128, p.BN, 0, 0, FMT_F16, FMT_F16);
    int git = 0, lt = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++lt) {
      const int buf = lt & 1;
      mbar_wait(&acc_empty[buf], (uint32_t)(((lt >> 1) & 1) ^ 1));   // epilogue drained this buffer (2 tiles ago)
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)buf * p.tmem_cols;
      for (int it = 0; it < total_iters; ++it, ++git) {
        const int s = git % p.stages;
        const uint32_t par = (uint32_t)((git / p.stages) & 1);
        mbar_wait(&full_bar[s], par);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t a_addr = smem_u32(smem + (size_t)s * stage_bytes);
          const uint32_t b_addr = a_addr + A_TILE_BYTES;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            mma_f16_ss(d_tmem, make_sdesc_sw128(a_addr + k * 32, 16, 1024),
                       make_sdesc_sw128(b_addr + k * 32, 16, 1024), idesc, (uint32_t)((it | k) != 0));
          }
          mma_commit(&empty_bar[s]);  // frees the smem stage once these MMAs have read it
          if (it == total_iters - 1) mma_commit(&acc_full[buf]);
        }
        __syncwarp();
      }
    }
  } else {
    // ------------------------------------------------------------ epilogue (warps 2..5)
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    const int row = q * 32 + lane;
    int lt = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++lt) {
    const int buf = lt & 1;
    const int tile_m = tile / p.n_tiles;
    const int n0 = (tile - tile_m * p.n_tiles) * p.BN;
    const long long pix = (long long)tile_m * 128 + row;
    const bool valid = pix < p.total_pixels;
    const int n_img = valid ? (int)(pix / p.HoWo) : 0;
    const int hw = valid ? (int)(pix - (long long)n_img * p.HoWo) : 0;
    mbar_wait(&acc_full[buf], (uint32_t)((lt >> 1) & 1));
    tc_fence_after();
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)buf * p.tmem_cols;
    for (int c0 = 0; c0 < p.BN; c0 += 16) {
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
      if (p.bias) {
#pragma unroll
        for (int j = 0; j < 16; ++j) if (j < ncol) v[j] += __ldg(p.bias + col0 + j);
      }
      if (p.emb) {
        const float* e = p.emb + (long long)n_img * p.ldemb + col0;
#pragma unroll
        for (int j = 0; j < 16; ++j) if (j < ncol) v[j] += __ldg(e + j);
      }
      if (p.residual) {
        const __half* rp = p.residual + pix * p.ldres + col0;
        if (ncol == 16 && ((((uintptr_t)rp) & 15) == 0)) {
          uint4 u0 = ldg_nc_u4(rp), u1 = ldg_nc_u4(rp + 8);
          const uint32_t w[8] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w};
#pragma unroll
          for (int j = 0; j < 8; ++j) { float2 f = unpack_f16x2(w[j]); v[2 * j] += f.x; v[2 * j + 1] += f.y; }
        } else {
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
    }
    tc_fence_before();
    mbar_arrive(&acc_empty[buf]);   // all of this thread's TMEM reads of the buffer have completed (ld wait above)
    }  // tile
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

  ConvMaps maps;
  memset(&maps, 0, sizeof(maps));
  int rc;
  if (a->stride == 1) {
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
    rc = make_act_map(&maps.a2, a->x2, C2, a->ldx2, a->N, a->H, a->W, 1, Wo, rows, imgs);
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

  ConvParams p;
  p.taps = taps; p.ksize = a->ksize; p.stride = a->stride;
  p.kc1 = a->C1 / 64; p.kc2 = C2 / 64;
  p.HoWo = Ho * Wo; p.Wo = Wo; p.rows = rows; p.imgs = imgs;
  p.total_pixels = a->N * Ho * Wo;
  p.Cout = a->Cout; p.BN = BN;
  const uint32_t stage_bytes = A_TILE_BYTES + (uint32_t)BN * 128u;
  int stages = (int)(98304u / stage_bytes);  // <= 96 KB of operand ring: two CTAs per SM
  if (stages < 3) stages = 3;
  if (stages > 6) stages = 6;
  p.stages = stages;
  p.tmem_cols = BN <= 32 ? 32 : BN <= 64 ? 64 : BN <= 128 ? 128 : 256;
  p.bias = a->bias; p.emb = a->emb; p.ldemb = a->ldemb;
  p.residual = (const __half*)a->residual; p.ldres = a->ldres;
  p.out = (__half*)a->out; p.ldo = a->ldo;
  p.out_nchw = a->out_nchw_f32;

  const size_t smem = (size_t)stages * stage_bytes + (2 * stages + 4) * 8 + 16 + 1024;
  static size_t smem_set = 0;
  if (smem > smem_set) {
    B2D_CHECK_CUDA(cudaFuncSetAttribute(conv_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    smem_set = 200 * 1024;
  }
  p.m_tiles = (p.total_pixels + 127) / 128;
  p.n_tiles = (a->CoutPad + BN - 1) / BN;
  const long long tiles = (long long)p.m_tiles * p.n_tiles;
  long long ctas = 2LL * num_sms();            // two persistent CTAs per SM (<= 96 KB ring, <= 256 TMEM columns each)
  if (ctas > tiles) ctas = tiles;
  dim3 grid((unsigned)ctas);
  conv_gemm_kernel<<<grid, CONV_THREADS, smem, (cudaStream_t)stream>>>(maps, p);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}
