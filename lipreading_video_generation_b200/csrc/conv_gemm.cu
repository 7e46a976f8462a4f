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
};

constexpr int CONV_EPI_WARPS = 8;          // two per TMEM lane quarter: even / odd 16-column chunks of the tile
constexpr int CONV_EPI_THREADS = CONV_EPI_WARPS * 32;
constexpr int CONV_THREADS = 64 + CONV_EPI_THREADS;
constexpr int EB_FLOATS = 512;              // images per tile x BN <= 512 (bias + emb staging)
constexpr int SACC_FLOATS = 512;            // fused GroupNorm sums of one tile: images per tile x BN x 2 <= 512
constexpr int A_TILE_BYTES = 128 * 128;  // 128 pixels x 64 bf16

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

template <int EPI>
__global__ void __launch_bounds__(CONV_THREADS, 2)
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
  const uint32_t eb_base = smem_u32(tmem_slot + 4);   // float [2][EB_FLOATS]: bias + emb of the current tile
  uint64_t* res_full = (uint64_t*)(tmem_slot + 2);    // residual tile landed in the staging buffer
  uint8_t* stage_out = smem + (size_t)p.stages * stage_bytes + 1024 + 2 * EB_FLOATS * 4;   // 2 x [128 pix][64 ch] fp16, 1024-aligned
  const uint32_t sacc_u32 = smem_u32(stage_out + 2 * A_TILE_BYTES);   // float [SACC_FLOATS]: fused GroupNorm sums of the tile
  if (EPI != EPI_PLAIN) {
    for (int i = threadIdx.x; i < SACC_FLOATS; i += CONV_THREADS) sts32f(sacc_u32 + (uint32_t)i * 4u, 0.f);
  }

  // Persistent CTA: tiles tile0, tile0 + gridDim.x, ...; the TMA ring and the two TMEM accumulator buffers run
  // across tile boundaries, so tile i's epilogue overlaps tile i+1's loads and MMAs (the non-persistent
  // version spent ~22k clk per CTA on a 1.2k-clk mainloop: profiles/r1_conv.md).
  const int total_tiles = p.m_tiles * p.n_tiles;
  const int chunks = p.kc1 + p.kc2;
  const int total_iters = p.taps * chunks;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(&acc_full[0], 1); mbar_init(&acc_full[1], 1);
    mbar_init(&acc_empty[0], CONV_EPI_THREADS); mbar_init(&acc_empty[1], CONV_EPI_THREADS);
    mbar_init(res_full, 1);
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
    // ------------------------------------------------------------ epilogue (warps 2..9)
    // Per tile: (1) stage bias + emb for the tile's images in smem and issue ALL residual loads of this thread's
    // row into registers BEFORE waiting for the accumulator, so their HBM latency hides behind the tile's MMAs
    // (the first version loaded them 16 columns at a time after the wait: 62 % of its samples were long_sb);
    // (2) TMEM -> registers -> + (bias + emb) -> + residual -> fp16 / fp32-NCHW global.
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    const int chalf = (warp - 2) >> 2;      // this warp takes the 16-column chunks cc with (cc & 1) == chalf
    const int row = q * 32 + lane;
    const int et = threadIdx.x - 64;        // 0..255 among the epilogue threads
    const bool has_eb = (p.bias != nullptr) || (p.emb != nullptr);
    const int imgs_in_tile = p.HoWo >= 128 ? 1 : p.imgs;
    int lt = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++lt) {
    const int buf = lt & 1;
    const int tile_m = tile / p.n_tiles;
    const int n0 = (tile - tile_m * p.n_tiles) * p.BN;
    const long long pix = (long long)tile_m * 128 + row;
    const bool valid = pix < p.total_pixels;
    const int n_img = valid ? (int)(pix / p.HoWo) : 0;
    const int hw = valid ? (int)(pix - (long long)n_img * p.HoWo) : 0;
    const int img_first = (int)(((long long)tile_m * 128) / p.HoWo);
    const uint32_t eb_u32 = eb_base + (uint32_t)buf * (EB_FLOATS * 4u);
    if (EPI == EPI_GNB) {
      // per (image of the tile, channel): z = x * rg + bb with rg = rstd * gamma, bb = beta - mean * rg
      for (int idx = et; idx < imgs_in_tile * p.BN; idx += CONV_EPI_THREADS) {
        const int im = idx / p.BN, c = idx - im * p.BN;
        const int col = n0 + c, img = min(img_first + im, p.n_images - 1);
        const int g = col / p.gnb_cpg;
        const float rs = __ldg(p.gnb_rstd + img * 32 + g), mu = __ldg(p.gnb_mean + img * 32 + g);
        const float rg = rs * __ldg(p.gnb_gamma + col);
        sts32f(eb_u32 + (uint32_t)idx * 8u, rg);
        sts32f(eb_u32 + (uint32_t)idx * 8u + 4u, fmaf(-mu, rg, __ldg(p.gnb_beta + col)));
      }
    } else if (has_eb) {
      for (int idx = et; idx < imgs_in_tile * p.BN; idx += CONV_EPI_THREADS) {
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
    int img0, h0;                                      // tile origin for the TMA boxes (as in the producer)
    if (p.HoWo >= 128) { const int tpi = p.HoWo >> 7; img0 = tile_m / tpi; h0 = (tile_m - img0 * tpi) * p.rows; }
    else { img0 = tile_m * p.imgs; h0 = 0; }
    const int nblk = p.BN >> 6;
    if (p.tma_out && et == 0) {
      tma_store_wait_read();                           // previous tile's store has finished reading the staging buffer
      if (EPI == EPI_GNB) {                            // the GroupNorm input x = [x1 | x2], 64-channel block by block
        mbar_expect_tx(res_full, (uint32_t)nblk * A_TILE_BYTES);
        for (int b = 0; b < nblk; ++b) {
          const int c = n0 + b * 64;
          if (c < p.gnb_C1) tma_load_4d(stage_out + b * A_TILE_BYTES, &maps.r, res_full, c, 0, h0, img0);
          else              tma_load_4d(stage_out + b * A_TILE_BYTES, &maps.r2, res_full, c - p.gnb_C1, 0, h0, img0);
        }
      } else if (p.residual) {
        mbar_expect_tx(res_full, (uint32_t)nblk * A_TILE_BYTES);
        for (int b = 0; b < nblk; ++b) tma_load_4d(stage_out + b * A_TILE_BYTES, &maps.r, res_full, n0 + b * 64, 0, h0, img0);
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
    uint4 rres[16];                                    // residual row segment: up to 128 columns of fp16
    const bool res_vec = !p.tma_out && p.residual && valid && fast_cols && ((p.ldres & 7) == 0) && ((((uintptr_t)p.residual) & 15) == 0);
    if (res_vec) {
      const __half* rp = p.residual + pix * p.ldres + n0;
#pragma unroll
      for (int g = 0; g < 16; ++g) if (g * 8 < p.BN && ((g >> 1) & 1) == chalf) rres[g] = ldg_nc_u4(rp + g * 8);
    }
    named_bar_sync(1, CONV_EPI_THREADS);                            // eb staging visible to the four epilogue warps
    mbar_wait(&acc_full[buf], (uint32_t)((lt >> 1) & 1));
    tc_fence_after();
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)buf * p.tmem_cols;
    const uint32_t eb_row = eb_u32 + (uint32_t)((n_img - img_first) * p.BN) * 4u;
    if (p.tma_out) {
      // ---- staged path: residual arrives by TMA in the 128B-swizzled tile layout, results go back to the same
      // smem slots (conflict-free 16-byte accesses) and leave with one TMA store per 64-channel block
      const uint32_t so_u32 = smem_u32(stage_out);
      if (EPI == EPI_GNB || p.residual) mbar_wait(res_full, (uint32_t)(lt & 1));
      // fused GroupNorm sums: this warp's 32 rows belong to one image (Ho*Wo % 32 == 0, checked on the host)
      const long long wpix = (long long)tile_m * 128 + q * 32;
      const bool wvalid = wpix < p.total_pixels;
      const int wslot = wvalid ? (int)(wpix / p.HoWo) - img_first : 0;
#pragma unroll
      for (int cc = 0; cc < 8; ++cc) {
        const int c0 = cc * 16;
        if (c0 < p.BN && (cc & 1) == chalf) {
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
      named_bar_sync(2, CONV_EPI_THREADS);
      if (et == 0) {
        for (int b = 0; b < nblk; ++b) tma_store_4d(&maps.o, stage_out + b * A_TILE_BYTES, n0 + b * 64, 0, h0, img0);
        tma_store_commit();
      }
      if (EPI == EPI_STATS) {
        // sacc[slot][BN] -> stats[img][(n0 + k) / 2][k & 1] (fp64), re-zeroed for the next tile (whose sums start
        // after its named barrier 1, which this thread reaches only after the stores below)
        for (int idx = et; idx < imgs_in_tile * p.BN; idx += CONV_EPI_THREADS) {
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
        for (int idx = et; idx < imgs_in_tile * p.BN * 2; idx += CONV_EPI_THREADS) {
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
#pragma unroll
    for (int cc = 0; cc < 8; ++cc) {
      const int c0 = cc * 16;
      if (c0 < p.BN && (cc & 1) == chalf) {
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
          const uint4 u0 = rres[2 * cc], u1 = rres[2 * cc + 1];
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

  const bool tma_out = a->out != nullptr && a->out_nchw_f32 == nullptr && a->Cout % 64 == 0 && BN % 64 == 0 &&
                       a->ldo % 8 == 0 && aligned16(a->out) &&
                       (a->residual == nullptr || (a->ldres % 8 == 0 && aligned16(a->residual)));
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
    rc = make_act_map(&maps.o, a->out, a->Cout, a->ldo, a->N, Ho, Wo, 1, Wo, rows, imgs);
    if (rc) return rc;
    maps.r = maps.o; maps.r2 = maps.o;
    if (want_gnb) {
      rc = make_act_map(&maps.r, a->gnb_x1, a->gnb_C1, a->gnb_ld1, a->N, Ho, Wo, 1, Wo, rows, imgs);
      if (rc) return rc;
      if (a->Cout > a->gnb_C1) {
        rc = make_act_map(&maps.r2, a->gnb_x2, a->Cout - a->gnb_C1, a->gnb_ld2, a->N, Ho, Wo, 1, Wo, rows, imgs);
        if (rc) return rc;
      }
    } else if (a->residual) {
      rc = make_act_map(&maps.r, a->residual, a->Cout, a->ldres, a->N, Ho, Wo, 1, Wo, rows, imgs);
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

  const size_t smem = (size_t)stages * stage_bytes + 1024 /* barriers */ + 2 * EB_FLOATS * 4 + 2 * A_TILE_BYTES +
                      SACC_FLOATS * 4 + 1024;
  static std::once_flag attr_once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(attr_once, [] {
    for (auto fn : {(const void*)conv_gemm_kernel<EPI_PLAIN>, (const void*)conv_gemm_kernel<EPI_STATS>,
                    (const void*)conv_gemm_kernel<EPI_GNB>}) {
      cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
      if (e != cudaSuccess) attr_err = e;
    }
  });
  B2D_CHECK_CUDA(attr_err);
  p.m_tiles = (p.total_pixels + 127) / 128;
  p.n_tiles = (a->CoutPad + BN - 1) / BN;
  const long long tiles = (long long)p.m_tiles * p.n_tiles;
  long long ctas = 2LL * num_sms();            // two persistent CTAs per SM (<= 96 KB ring, <= 256 TMEM columns each)
  if (ctas > tiles) ctas = tiles;
  dim3 grid((unsigned)ctas);
  if (want_gnb)        conv_gemm_kernel<EPI_GNB><<<grid, CONV_THREADS, smem, (cudaStream_t)stream>>>(maps, p);
  else if (want_stats) conv_gemm_kernel<EPI_STATS><<<grid, CONV_THREADS, smem, (cudaStream_t)stream>>>(maps, p);
  else                 conv_gemm_kernel<EPI_PLAIN><<<grid, CONV_THREADS, smem, (cudaStream_t)stream>>>(maps, p);
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
