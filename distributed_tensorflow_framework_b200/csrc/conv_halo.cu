// K-CONV (halo variant): stride-1 KxK convolution (fprop and data gradient) on NHWC bf16 with C % 64 == 0.
//
// The gather kernel of igemm.cu re-reads every input pixel KH*KW times through L2. Here a CTA owns R whole image
// rows; ONE 4-D TMA load brings the (R+KH-1) x (W+KW-1) x 64-channel halo tile into shared memory (TMA zero-fills
// the SAME padding), laid out as a linear array of 128-byte pixel rows with row pitch Wp = W+KW-1. In that "padded
// linear" space the GEMM row of output pixel (y, x) is m = (y-y0)*Wp + x and the input it needs for filter tap
// (kh, kw) sits at m + kh*Wp + kw: every tap is the SAME shared-memory buffer viewed at a different start address.
// So the A operand of all KH*KW taps is one smem tile, consumed by tcgen05.mma through shifted descriptors; only the
// small weight tiles stream per tap. Rows with x >= W are padding rows of the GEMM and are dropped in the epilogue.
// The data gradient is the same kernel on dy with per-tap-transposed filters and mirrored tap shifts.
//
// Replaces tf.nn.conv2d / its input gradient for 3x3 layers (/root/reference/distribute_tools.py:69-121).
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include "ptx.cuh"
#include "dtf_internal.h"

namespace dtf {
namespace halo {

constexpr int BLOCK_M = 128;
constexpr int NUM_THREADS = 192;  // warps 0-3 epilogue, warp 4 MMA, warp 5 TMA
constexpr int A_STAGES = 2;

struct Params {
  int n_img, H, W, C, n_out;
  int KH, KW;
  int Wp;            // Wt + KW - 1: row pitch of the padded-linear tile space
  int Wt;            // output columns per tile
  int R;             // image rows per tile
  int tiles_x;       // tiles along the image width
  int tiles_per_img;
  int org_x, org_y;  // halo origin relative to the first output row of the tile (fprop: -pad_w, -pad_h)
  int mirror;        // 1: data gradient (tap shifts mirrored)
  int a_stage_bytes;
  int a_tx_bytes;
  int base_off_mode;
  void* out;
  const float* bias;
  int relu;
  float* stats;  // pixels-as-N kernel: [pixel tiles][2][n_out] per-tile column sums / sums of squares (batch norm)
};

template <int BN, int B_STAGES>
__global__ void __launch_bounds__(NUM_THREADS)
conv_halo_kernel(const __grid_constant__ CUtensorMap tma_x, const __grid_constant__ CUtensorMap tma_w, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr int B_BYTES = BN * 64 * 2;
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t a_base = smem_base;
  const uint32_t b_base = smem_base + A_STAGES * p.a_stage_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + A_STAGES * p.a_stage_bytes + B_STAGES * B_BYTES);
  uint64_t* a_full = bars;
  uint64_t* a_empty = a_full + A_STAGES;
  uint64_t* b_full = a_empty + A_STAGES;
  uint64_t* b_empty = b_full + B_STAGES;
  uint64_t* tmem_full = b_empty + B_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile_n = blockIdx.x;
  const int img = blockIdx.y / p.tiles_per_img;
  const int t_in_img = blockIdx.y - img * p.tiles_per_img;
  const int y0 = (t_in_img / p.tiles_x) * p.R;
  const int x0 = (t_in_img % p.tiles_x) * p.Wt;
  const int ntaps = p.KH * p.KW;
  const int ncblk = p.C >> 6;

  if (threadIdx.x == 0) {
    for (int s = 0; s < A_STAGES; ++s) {
      mbar_init(&a_full[s], 1);
      mbar_init(&a_empty[s], 1);
    }
    for (int s = 0; s < B_STAGES; ++s) {
      mbar_init(&b_full[s], 1);
      mbar_init(&b_empty[s], 1);
    }
    mbar_init(tmem_full, 1);
    mbar_fence_init();
  }
  if (warp == 4) tmem_alloc<BN>(tmem_slot);
  if (warp == 5 && lane == 0) {
    tma_prefetch_desc(&tma_x);
    tma_prefetch_desc(&tma_w);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 5) {
    if (lane == 0) {
      int bit = 0;
      for (int cb = 0; cb < ncblk; ++cb) {
        const int as = cb % A_STAGES;
        mbar_wait(&a_empty[as], ((cb / A_STAGES) & 1) ^ 1);
        mbar_arrive_expect_tx(&a_full[as], p.a_tx_bytes);
        tma_load_4d(a_base + as * p.a_stage_bytes, &tma_x, &a_full[as], cb * 64, x0 + p.org_x, y0 + p.org_y, img);
        for (int t = 0; t < ntaps; ++t, ++bit) {
          const int bs = bit % B_STAGES;
          mbar_wait(&b_empty[bs], ((bit / B_STAGES) & 1) ^ 1);
          mbar_arrive_expect_tx(&b_full[bs], B_BYTES);
#pragma unroll
          for (int j = 0; j < BN / 64; ++j)
            tma_load_2d(b_base + bs * B_BYTES + j * 8192, &tma_w, &b_full[bs], tile_n * BN + j * 64, t * p.C + cb * 64);
        }
      }
    }
    __syncwarp();
  } else if (warp == 4) {
    constexpr uint32_t idesc = make_idesc_bf16(BLOCK_M, BN, /*a_mn=*/false, /*b_mn=*/true);
    int bit = 0;
    for (int cb = 0; cb < ncblk; ++cb) {
      const int as = cb % A_STAGES;
      mbar_wait(&a_full[as], (cb / A_STAGES) & 1);
      for (int t = 0; t < ntaps; ++t, ++bit) {
        const int bs = bit % B_STAGES;
        mbar_wait(&b_full[bs], (bit / B_STAGES) & 1);
        tc_fence_after();
        if (lane == 0) {
          int kh = t / p.KW, kw = t - kh * p.KW;
          if (p.mirror) {
            kh = p.KH - 1 - kh;
            kw = p.KW - 1 - kw;
          }
          const uint32_t a_tap = a_base + as * p.a_stage_bytes + (uint32_t)(kh * p.Wp + kw) * 128u;
          const uint32_t boff = (p.base_off_mode & 1) ? ((a_tap >> 7) & 7u) : 0u;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t adesc = make_smem_desc_off(a_tap + k * 32, 16, 1024, boff);
            const uint64_t bdesc = make_smem_desc(b_base + bs * B_BYTES + k * 2048, 8192, 1024);
            umma_bf16(tmem_base, adesc, bdesc, idesc, (bit > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&b_empty[bs]);
          if (t == ntaps - 1) umma_commit(&a_empty[as]);
        }
        __syncwarp();
      }
    }
    if (lane == 0) umma_commit(tmem_full);
    __syncwarp();
  } else {
    // epilogue: GEMM row m -> (y, x) in the padded-linear space of this tile
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    const int m = threadIdx.x;
    const int ry = m / p.Wp;
    const int xx = m - ry * p.Wp;
    const int x = x0 + xx;
    const int y = y0 + ry;
    const bool valid = (ry < p.R) && (xx < p.Wt) && (x < p.W) && (y < p.H);
    const size_t pix = ((size_t)img * p.H + y) * p.W + x;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
#pragma unroll 1
    for (int cc = 0; cc < BN / 32; ++cc) {
      uint32_t v[32];
      tmem_ld32(lane_addr + cc * 32, v);
      tmem_ld_wait();
      const int n0 = tile_n * BN + cc * 32;
      if (valid && n0 < p.n_out) {
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
        if (p.bias != nullptr) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (n0 + j < p.n_out) f[j] += __ldg(p.bias + n0 + j);
        }
        if (p.relu) {
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
        }
        __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + pix * p.n_out + n0;
        if (n0 + 32 <= p.n_out) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            uint4 q;
            q.x = pack_bf16x2(f[j], f[j + 1]);
            q.y = pack_bf16x2(f[j + 2], f[j + 3]);
            q.z = pack_bf16x2(f[j + 4], f[j + 5]);
            q.w = pack_bf16x2(f[j + 6], f[j + 7]);
            *reinterpret_cast<uint4*>(o + j) = q;
          }
        } else {
          for (int j = 0; j < 32; ++j)
            if (n0 + j < p.n_out) o[j] = __float2bfloat16(f[j]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tmem_dealloc<BN>(tmem_base);
  }
}


// Persistent variant: one CTA per SM walks a strided list of (pixel tile, channel tile) items. The accumulator is
// double-buffered in TMEM (2 x BN columns) so the epilogue of item i overlaps the MMAs of item i+1, the halo tile of
// the next item is prefetched into the second A stage while the current one is consumed, and -- when all KH*KW*C/64
// weight tiles of the layer fit in shared memory next to the halo stages (64- and 128-channel layers) -- the weights
// are loaded ONCE per CTA and stay resident, so the steady state streams only activations. Otherwise the weight ring
// keeps running across items. The epilogue goes through a swizzled per-warp staging tile so that every global store
// instruction writes whole 64-byte row segments.
constexpr int MAX_B_STAGES = 40;
constexpr int PERSIST_STAGE_BYTES = 4 * 32 * 64;  // 4 epilogue warps x 32 rows x 32 bf16 columns

struct PersistParams {
  int tiles_n;      // channel tiles
  int total_items;  // pixel tiles x channel tiles
  int nb;           // weight stages in shared memory
  int resident;     // 1: nb == KH*KW*C/64 and every weight tile is loaded once
};

template <int BN>
__global__ void __launch_bounds__(NUM_THREADS, 1)
conv_halo_persistent_kernel(const __grid_constant__ CUtensorMap tma_x, const __grid_constant__ CUtensorMap tma_w,
                            const Params p, const PersistParams pp) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr int B_BYTES = BN * 64 * 2;
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t a_base = smem_base;
  const uint32_t b_base = smem_base + A_STAGES * p.a_stage_bytes;
  const int bar_off = A_STAGES * p.a_stage_bytes + pp.nb * B_BYTES;
  uint64_t* a_full = reinterpret_cast<uint64_t*>(smem + bar_off);
  uint64_t* a_empty = a_full + A_STAGES;
  uint64_t* acc_full = a_empty + A_STAGES;  // [2]
  uint64_t* acc_empty = acc_full + 2;       // [2]
  uint64_t* b_full = acc_empty + 2;         // [MAX_B_STAGES]
  uint64_t* b_empty = b_full + MAX_B_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(b_empty + MAX_B_STAGES);
  uint32_t* tap_off = reinterpret_cast<uint32_t*>(smem + bar_off + 768);  // [KH*KW] <= 49 entries
  uint8_t* stage_base = smem + bar_off + 1024;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ntaps = p.KH * p.KW;
  const int ncblk = p.C >> 6;
  const int nb = pp.nb;
  const bool resident = pp.resident != 0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < A_STAGES; ++s) {
      mbar_init(&a_full[s], 1);
      mbar_init(&a_empty[s], 1);
    }
    for (int s = 0; s < nb; ++s) {
      mbar_init(&b_full[s], 1);
      mbar_init(&b_empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&acc_full[b], 1);
      mbar_init(&acc_empty[b], 4);  // one arrival per epilogue warp
    }
    mbar_fence_init();
  }
  if (warp == 4) tmem_alloc<2 * BN>(tmem_slot);
  if (warp == 5 && lane == 0) {
    tma_prefetch_desc(&tma_x);
    tma_prefetch_desc(&tma_w);
  }
  if (threadIdx.x < ntaps) {  // byte shift of every tap's view of the halo tile (mirrored for the data gradient)
    int kh = threadIdx.x / p.KW, kw = threadIdx.x - kh * p.KW;
    if (p.mirror) {
      kh = p.KH - 1 - kh;
      kw = p.KW - 1 - kw;
    }
    tap_off[threadIdx.x] = (uint32_t)(kh * p.Wp + kw) * 128u;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  auto decode = [&](int item, int& tile_n, int& img, int& y0, int& x0) {
    tile_n = item % pp.tiles_n;
    const int mt = item / pp.tiles_n;
    img = mt / p.tiles_per_img;
    const int tin = mt - img * p.tiles_per_img;
    y0 = (tin / p.tiles_x) * p.R;
    x0 = (tin % p.tiles_x) * p.Wt;
  };

  // The issuing threads run lean loops: ring indices and phases are carried as counters (no div / mod), tap shifts
  // come from a table, and descriptors are a precomputed constant plus the 16-byte-granular address.
  if (warp == 5) {
    if (lane == 0) {
      int as = 0, bs = 0;
      uint32_t a_par = 1, b_par = 1;  // "empty" barriers: the first pass over a ring must not block
      bool first = true;
      for (int item = blockIdx.x; item < pp.total_items; item += gridDim.x) {
        int tile_n, img, y0, x0;
        decode(item, tile_n, img, y0, x0);
        for (int cb = 0; cb < ncblk; ++cb) {
          mbar_wait(&a_empty[as], a_par);
          mbar_arrive_expect_tx(&a_full[as], p.a_tx_bytes);
          tma_load_4d(a_base + as * p.a_stage_bytes, &tma_x, &a_full[as], cb * 64, x0 + p.org_x, y0 + p.org_y, img);
          if (++as == A_STAGES) {
            as = 0;
            a_par ^= 1;
          }
          if (resident && !first) continue;
          int krow = cb * 64;
          for (int t = 0; t < ntaps; ++t, krow += p.C) {
            if (!resident) mbar_wait(&b_empty[bs], b_par);
            mbar_arrive_expect_tx(&b_full[bs], B_BYTES);
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_2d(b_base + bs * B_BYTES + j * 8192, &tma_w, &b_full[bs], tile_n * BN + j * 64, krow);
            if (++bs == nb) {
              bs = 0;
              b_par ^= 1;
            }
          }
        }
        first = false;
      }
    }
    __syncwarp();
  } else if (warp == 4) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BLOCK_M, BN, /*a_mn=*/false, /*b_mn=*/true);
      const uint64_t adesc0 = make_smem_desc(0, 16, 1024);
      const uint64_t bdesc0 = make_smem_desc(0, 8192, 1024);
      int as = 0, bs = 0, buf = 0;
      uint32_t a_par = 0, b_par = 0, acc_par = 1;
      for (int item = blockIdx.x; item < pp.total_items; item += gridDim.x) {
        mbar_wait(&acc_empty[buf], acc_par);  // the epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * BN;
        uint32_t accum = 0;
        for (int cb = 0; cb < ncblk; ++cb) {
          mbar_wait(&a_full[as], a_par);
          const uint32_t a_stage = a_base + as * p.a_stage_bytes;
          for (int t = 0; t < ntaps; ++t) {
            mbar_wait(&b_full[bs], b_par);  // resident: phase 0 completes once, b_par stays 0
            tc_fence_after();
            const uint64_t adesc = adesc0 + ((a_stage + tap_off[t]) >> 4);
            const uint64_t bdesc = bdesc0 + ((b_base + bs * B_BYTES) >> 4);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              umma_bf16(d_tmem, adesc + 2 * k, bdesc + 128 * k, idesc, accum);
              accum = 1;
            }
            if (!resident) umma_commit(&b_empty[bs]);
            if (++bs == nb) {
              bs = 0;
              if (!resident) b_par ^= 1;
            }
          }
          umma_commit(&a_empty[as]);
          if (++as == A_STAGES) {
            as = 0;
            a_par ^= 1;
          }
        }
        umma_commit(&acc_full[buf]);
        if ((buf ^= 1) == 0) acc_par ^= 1;
      }
    }
    __syncwarp();
  } else {
    // epilogue: GEMM row m -> (y, x) in the padded-linear space of the tile; TMEM -> registers -> bias / ReLU ->
    // bf16 -> per-warp staging tile -> 64-byte row segments (8 rows per store instruction)
    uint8_t* stage = stage_base + warp * (32 * 64);
    const int m = threadIdx.x;
    const int ry = m / p.Wp;
    const int xx = m - ry * p.Wp;
    const uint32_t lane_addr = ((uint32_t)(warp * 32) << 16);
    int lt = 0;
    for (int item = blockIdx.x; item < pp.total_items; item += gridDim.x, ++lt) {
      int tile_n, img, y0, x0;
      decode(item, tile_n, img, y0, x0);
      const int buf = lt & 1;
      const int x = x0 + xx, y = y0 + ry;
      const bool valid = (ry < p.R) && (xx < p.Wt) && (x < p.W) && (y < p.H);
      const int pix = valid ? (int)(((long long)img * p.H + y) * p.W + x) : -1;
      mbar_wait(&acc_full[buf], (lt >> 1) & 1);
      tc_fence_after();
#pragma unroll 1
      for (int cc = 0; cc < BN / 32; ++cc) {
        const int n0 = tile_n * BN + cc * 32;
        if (n0 >= p.n_out) break;  // uniform across the CTA
        uint32_t v[32];
        tmem_ld32(tmem_base + buf * BN + lane_addr + cc * 32, v);
        tmem_ld_wait();
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
        if (p.bias != nullptr) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (n0 + j < p.n_out) f[j] += __ldg(p.bias + n0 + j);
        }
        if (p.relu) {
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint4 q;
          q.x = pack_bf16x2(f[c * 8 + 0], f[c * 8 + 1]);
          q.y = pack_bf16x2(f[c * 8 + 2], f[c * 8 + 3]);
          q.z = pack_bf16x2(f[c * 8 + 4], f[c * 8 + 5]);
          q.w = pack_bf16x2(f[c * 8 + 6], f[c * 8 + 7]);
          *reinterpret_cast<uint4*>(stage + lane * 64 + ((c ^ (lane & 3)) << 4)) = q;
        }
        __syncwarp();
        const int chunk = lane & 3;
        const int n = n0 + chunk * 8;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int r = i * 8 + (lane >> 2);
          const int pr = __shfl_sync(0xffffffffu, pix, r);
          const uint4 q = *reinterpret_cast<const uint4*>(stage + r * 64 + ((chunk ^ (r & 3)) << 4));
          if (pr >= 0 && n + 8 <= p.n_out)
            *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.out) + (size_t)pr * p.n_out + n) = q;
        }
        __syncwarp();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[buf]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tmem_dealloc<2 * BN>(tmem_base);
  }
}

// --------------------------------------------------------------------------------------------------------------
// Pixels-as-N variant (the default). A single thread issues every tcgen05.mma of a CTA, and that thread needs a
// fixed number of scalar instructions per MMA; with 64 or 128 output channels as the N dimension an MMA is only
// 32-64 tensor-core cycles long and the issue loop, not the tensor core, sets the pace (ncu: tensor pipe 15-19 %).
// Here the roles are swapped: D^T[co, pixel] = W^T[co, k] * X^T[k, pixel], so A is the weight tile (MN-major,
// M = 128 output channels) and B is the halo tile (K-major, N = 256 padded-linear pixels); every MMA is the
// largest shape (128x256x16, 128 cycles), the per-FLOP instruction count drops 2-4x, and weights stream at half the
// rate. A tap is again just a shifted start address, now of the B descriptor. Whole small images are packed G per
// tile (4-D TMA box over the image index) so 14x14 and 7x7 layers fill the 256-pixel N. The accumulator
// [co lanes x pixel columns] is double-buffered in TMEM (2 x 256 columns); the epilogue transposes it through a
// per-warp staging tile into 64-byte NHWC row segments.
constexpr int PN_PIX = 256;
constexpr int PN_MAX_X_STAGES = 4;
constexpr int PN_STAGE_BYTES = 4 * 32 * 64;  // per warp: 32 pixels x 32 channels bf16
constexpr int PN_TAB_BYTES = 4 * PN_PIX * 4; // per warp: output pixel index (or -1) of every GEMM column

struct PnParams {
  int tiles_co;     // output-channel tiles of MW
  int total_items;  // pixel tiles x channel tiles
  int nb;           // weight stages
  int resident;     // all weight tiles stay in shared memory
  int G;            // images per tile (whole-image tiles only), else 1
  int img_rows;     // padded-linear rows per packed image: (R+KH-1)*Wp
  int groups;       // pixel tiles: n_img*tiles_per_img (G == 1) or ceil(n_img / G)
  int xst;          // halo-tile stages (2 or 3)
  // structurally absent (tap, channel-block) weight tiles (space-to-depth packed filters) are skipped:
  // mask_by 0: none; 1: the mask is chosen by the input-channel block (cb / blk_per_phase);
  // 2: by the output-channel tile (tile_co / blk_per_phase). Bit t of a mask = tap t present.
  int mask_by, blk_per_phase;
  unsigned long long phase_mask[4];
};

// MW: output channels per item, 64 (single 64-channel weight box) or 128. STATS: also emit per-tile batch-norm
// statistics (a separate instantiation: it needs the chunk loop fully unrolled, which costs the plain kernel speed).
template <int MW, bool STATS>
__global__ void __launch_bounds__(NUM_THREADS, 1)
conv_halo_pn_kernel(const __grid_constant__ CUtensorMap tma_x, const __grid_constant__ CUtensorMap tma_w,
                    const Params p, const PnParams pp) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr int W_BYTES = MW * 64 * 2;
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t x_base = smem_base;
  const uint32_t w_base = smem_base + pp.xst * p.a_stage_bytes;
  const int bar_off = pp.xst * p.a_stage_bytes + pp.nb * W_BYTES;
  uint64_t* x_full = reinterpret_cast<uint64_t*>(smem + bar_off);
  uint64_t* x_empty = x_full + PN_MAX_X_STAGES;
  uint64_t* acc_full = x_empty + PN_MAX_X_STAGES;  // [2]
  uint64_t* acc_empty = acc_full + 2;       // [2]
  uint64_t* w_full = acc_empty + 2;         // [MAX_B_STAGES]
  uint64_t* w_empty = w_full + MAX_B_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(w_empty + MAX_B_STAGES);
  uint32_t* tap_off = reinterpret_cast<uint32_t*>(smem + bar_off + 768);
  uint8_t* stage_base = smem + bar_off + 1024;
  int* tab_base = reinterpret_cast<int*>(smem + bar_off + 1024 + PN_STAGE_BYTES);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ntaps = p.KH * p.KW;
  const int ncblk = p.C >> 6;
  const int nb = pp.nb;
  const bool resident = pp.resident != 0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < pp.xst; ++s) {
      mbar_init(&x_full[s], 1);
      mbar_init(&x_empty[s], 1);
    }
    for (int s = 0; s < nb; ++s) {
      mbar_init(&w_full[s], 1);
      mbar_init(&w_empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&acc_full[b], 1);
      mbar_init(&acc_empty[b], 4);
    }
    mbar_fence_init();
  }
  if (warp == 4) tmem_alloc<512>(tmem_slot);
  if (warp == 5 && lane == 0) {
    tma_prefetch_desc(&tma_x);
    tma_prefetch_desc(&tma_w);
  }
  if (threadIdx.x < ntaps) {
    int kh = threadIdx.x / p.KW, kw = threadIdx.x - kh * p.KW;
    if (p.mirror) {
      kh = p.KH - 1 - kh;
      kw = p.KW - 1 - kw;
    }
    tap_off[threadIdx.x] = (uint32_t)(kh * p.Wp + kw) * 128u;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // item -> channel tile, first image of the tile, tile origin
  auto decode = [&](int item, int& tile_co, int& img, int& y0, int& x0) {
    tile_co = item % pp.tiles_co;
    const int mt = item / pp.tiles_co;
    if (pp.G > 1) {
      img = mt * pp.G;
      y0 = 0;
      x0 = 0;
    } else {
      img = mt / p.tiles_per_img;
      const int tin = mt - img * p.tiles_per_img;
      y0 = (tin / p.tiles_x) * p.R;
      x0 = (tin % p.tiles_x) * p.Wt;
    }
  };

  auto tap_mask = [&](int cb, int tile_co) -> unsigned long long {
    if (pp.mask_by == 1) return pp.phase_mask[cb / pp.blk_per_phase];
    if (pp.mask_by == 2) return pp.phase_mask[tile_co / pp.blk_per_phase];
    return ~0ull;
  };

  if (warp == 5) {
    if (lane == 0) {
      int xs = 0, ws = 0;
      uint32_t x_par = 1, w_par = 1;
      bool first = true;
      for (int item = blockIdx.x; item < pp.total_items; item += gridDim.x) {
        int tile_co, img, y0, x0;
        decode(item, tile_co, img, y0, x0);
        for (int cb = 0; cb < ncblk; ++cb) {
          mbar_wait(&x_empty[xs], x_par);
          mbar_arrive_expect_tx(&x_full[xs], p.a_tx_bytes);
          tma_load_4d(x_base + xs * p.a_stage_bytes, &tma_x, &x_full[xs], cb * 64, x0 + p.org_x, y0 + p.org_y, img);
          if (++xs == pp.xst) {
            xs = 0;
            x_par ^= 1;
          }
          if (resident && !first) continue;
          const unsigned long long tmask = tap_mask(cb, tile_co);
          int krow = cb * 64;
          for (int t = 0; t < ntaps; ++t, krow += p.C) {
            if (!((tmask >> t) & 1ull)) continue;
            if (!resident) mbar_wait(&w_empty[ws], w_par);
            mbar_arrive_expect_tx(&w_full[ws], W_BYTES);
#pragma unroll
            for (int j = 0; j < MW / 64; ++j)
              tma_load_2d(w_base + ws * W_BYTES + j * 8192, &tma_w, &w_full[ws], tile_co * MW + j * 64, krow);
            if (++ws == nb) {
              ws = 0;
              w_par ^= 1;
            }
          }
        }
        first = false;
      }
    }
    __syncwarp();
  } else if (warp == 4) {
    // The whole warp walks the loop with warp-uniform state (everything derives from kernel parameters and
    // counters, so descriptors live in uniform registers); one elected lane issues the MMAs and commits.
    // M = 128 even for MW = 64: the upper 64 accumulator lanes then hold products of whatever follows the weight
    // tile in shared memory and are never read.
    constexpr uint32_t idesc = make_idesc_bf16(128, PN_PIX, /*a_mn=*/true, /*b_mn=*/false);
    const uint64_t wdesc0 = make_smem_desc(w_base, 8192, 1024);
    const uint64_t xdesc0 = make_smem_desc(x_base, 16, 1024);
    // tap t -> start-address shift (16-byte units) of the halo tile: +-1 pixel per tap, row wrap every KW taps
    const int sgn = p.mirror ? -1 : 1;
    const int off_first = p.mirror ? ((p.KH - 1) * p.Wp + (p.KW - 1)) * 8 : 0;
    const int off_step = sgn * 8;
    const int off_wrap = sgn * (p.Wp - p.KW + 1) * 8;
    const uint32_t x_stage16 = (uint32_t)p.a_stage_bytes >> 4;
    int xs = 0, ws = 0, buf = 0;
    uint32_t x_par = 0, w_par = 0, acc_par = 1;
    for (int item = blockIdx.x; item < pp.total_items; item += gridDim.x) {
      mbar_wait(&acc_empty[buf], acc_par);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + buf * PN_PIX;
      uint32_t accum = 0;
      const int tile_co = item % pp.tiles_co;
      for (int cb = 0; cb < ncblk; ++cb) {
        mbar_wait(&x_full[xs], x_par);
        const uint64_t xdesc_s = xdesc0 + xs * x_stage16;
        const unsigned long long tmask = tap_mask(cb, tile_co) & ((ntaps < 64) ? ((1ull << ntaps) - 1) : ~0ull);
        const int last_t = 63 - __clzll((long long)tmask);  // the commits ride on the last tap that is executed
        int off = off_first, kw = 0;
        for (int t = 0; t < ntaps; ++t) {
          if (!((tmask >> t) & 1ull)) {  // absent weight tile: only advance the tap shift
            if (++kw == p.KW) {
              kw = 0;
              off += off_wrap;
            } else {
              off += off_step;
            }
            continue;
          }
          mbar_wait(&w_full[ws], w_par);  // resident: phase 0 completes once, w_par stays 0
          tc_fence_after();
          if (elect_one()) {
            const uint64_t wdesc = wdesc0 + (uint32_t)(ws * (W_BYTES >> 4));
            const uint64_t xdesc = xdesc_s + (uint32_t)off;
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_bf16(d_tmem, wdesc + 128 * k, xdesc + 2 * k, idesc, (k > 0) ? 1u : accum);
            if (!resident) umma_commit(&w_empty[ws]);
            if (t == last_t) {
              umma_commit(&x_empty[xs]);
              if (cb == ncblk - 1) umma_commit(&acc_full[buf]);
            }
          }
          __syncwarp();
          accum = 1;
          if (++kw == p.KW) {
            kw = 0;
            off += off_wrap;
          } else {
            off += off_step;
          }
          if (++ws == nb) {
            ws = 0;
            if (!resident) w_par ^= 1;
          }
        }
        if (++xs == pp.xst) {
          xs = 0;
          x_par ^= 1;
        }
      }
      if ((buf ^= 1) == 0) acc_par ^= 1;
    }
  } else {
    // epilogue warp w owns channels [32w, 32w+32) of the item: TMEM lane = channel, column = padded-linear pixel
    // staging tile and pixel table are addressed by 32-bit shared addresses (explicit st.shared / ld.shared: the
    // pointers lose their address space in the carve-up arithmetic and would compile to generic LD.E / ST.E)
    const uint32_t stage = smem_u32(stage_base) + warp * (32 * 64);
    const uint32_t tab = smem_u32(tab_base) + warp * PN_PIX * 4;
    const uint32_t lane_addr = ((uint32_t)(warp * 32) << 16);
    const int m_limit = (pp.G - 1) * pp.img_rows + p.R * p.Wp;  // columns beyond carry no output pixel
    int buf = 0;
    uint32_t acc_par = 0;
    for (int item = blockIdx.x; item < pp.total_items; item += gridDim.x) {
      int tile_co, img, y0, x0;
      decode(item, tile_co, img, y0, x0);
      const int co = tile_co * MW + warp * 32 + lane;
      const bool warp_live = (warp * 32 < MW) && (tile_co * MW + warp * 32 < p.n_out);
      uint32_t vmask[STATS ? PN_PIX / 32 : 1];  // validity of the 32 pixels of every chunk (warp-uniform)
      if (warp_live) {
#pragma unroll(STATS ? PN_PIX / 32 : 1)
        for (int q = 0; q < PN_PIX / 32; ++q) {
          const int m = q * 32 + lane;
          const int g = m / pp.img_rows;
          const int rem = m - g * pp.img_rows;
          const int ry = rem / p.Wp;
          const int xx = rem - ry * p.Wp;
          const int x = x0 + xx, y = y0 + ry, im = img + g;
          const bool valid = (g < pp.G) && (im < p.n_img) && (ry < p.R) && (xx < p.Wt) && (x < p.W) && (y < p.H);
          st_shared_b32(tab + m * 4, valid ? (uint32_t)(int)(((long long)im * p.H + y) * p.W + x) : 0xffffffffu);
          if (STATS) vmask[STATS ? q : 0] = __ballot_sync(0xffffffffu, valid);
        }
        __syncwarp();
      }
      float st_sum = 0.f, st_sq = 0.f;
      const float bias = (warp_live && p.bias != nullptr && co < p.n_out) ? __ldg(p.bias + co) : 0.f;
      mbar_wait(&acc_full[buf], acc_par);
      tc_fence_after();
      if (warp_live) {
#pragma unroll(STATS ? PN_PIX / 32 : 1)
        for (int cc = 0; cc < PN_PIX / 32; ++cc) {
          if (cc * 32 >= m_limit) break;
          uint32_t v[32];
          tmem_ld32(tmem_base + buf * PN_PIX + lane_addr + cc * 32, v);
          tmem_ld_wait();
          const uint32_t st = stage + lane * 2;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            float f = __uint_as_float(v[j]) + bias;
            if (STATS && ((vmask[STATS ? cc : 0] >> j) & 1u)) {  // this thread's channel, valid pixels only
              st_sum += f;
              st_sq = fmaf(f, f, st_sq);
            }
            if (p.relu) f = fmaxf(f, 0.f);
            st_shared_b16(st + j * 64, __bfloat16_as_ushort(__float2bfloat16(f)));  // stage[pixel j][channel lane]
          }
          __syncwarp();
          const int c8 = (lane & 3) * 8;
          const int n = tile_co * MW + warp * 32 + c8;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int r = i * 8 + (lane >> 2);
            const int pr = (int)ld_shared_b32(tab + (cc * 32 + r) * 4);
            const uint4 q = ld_shared_v4(stage + r * 64 + c8 * 2);
            if (pr >= 0 && n + 8 <= p.n_out)
              *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.out) + (size_t)pr * p.n_out + n) = q;
          }
          __syncwarp();
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[buf]);
      if ((buf ^= 1) == 0) acc_par ^= 1;
      if (STATS && warp_live && co < p.n_out) {
        float* o = p.stats + (size_t)(item / pp.tiles_co) * 2 * p.n_out + co;
        o[0] = st_sum;
        o[p.n_out] = st_sq;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

// --------------------------------------------------------------------------------------------------------------
// Filter gradient on halo tiles: dW[tap, ci, co] = sum over pixels p of x[p + shift(tap), ci] * dy[p, co].
// A CTA owns one (64 ci) x (64 co) block of ALL taps and walks a range of image tiles. Per tile one TMA load brings
// the x halo (same box as fprop) and R row loads bring the dy tile into the padded-linear layout (gap positions stay
// zero from the one-time clear). Both operands are MN-major: shared-memory rows are pixels (the reduction dim),
// the 128 bytes of a row are 64 channels. A tap is again only a start-address shift of the x halo; two taps are
// paired into one M=128 MMA by using the shift difference as the descriptor's leading-dimension byte offset, so the
// KH*KW accumulators [128 x 64] live side by side in TMEM for the whole kernel and are reduced into dW with red.add
// once at the end (split over tiles across CTAs).
struct WgradParams {
  int n_img, H, W, C, n_out;
  int KH, KW, Wp, Wt, R, tiles_x, tiles_per_img;
  int org_x, org_y;
  int a_stage_bytes, a_tx_bytes, b_tx_bytes;
  int total_tiles, tiles_per_split;
  float* dw;
};
constexpr int WG_STAGES = 3;
constexpr int WG_B_BYTES = 128 * 128;  // 128 pixel rows x 64 co

__global__ void __launch_bounds__(NUM_THREADS, 1)
wgrad_halo_kernel(const __grid_constant__ CUtensorMap tma_x, const __grid_constant__ CUtensorMap tma_dy,
                  const WgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t a_base = smem_base;
  const uint32_t b_base = smem_base + WG_STAGES * p.a_stage_bytes;
  const int data_bytes = WG_STAGES * (p.a_stage_bytes + WG_B_BYTES);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + data_bytes);
  uint64_t* empty = full + WG_STAGES;
  uint64_t* tmem_full = empty + WG_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);
  uint64_t* pair_desc = reinterpret_cast<uint64_t*>(smem + data_bytes + 128);  // [8]
  uint32_t* pair_off = reinterpret_cast<uint32_t*>(smem + data_bytes + 192);   // [8]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int co0 = blockIdx.x * 64, ci0 = blockIdx.y * 64;
  const int t_begin = blockIdx.z * p.tiles_per_split;
  const int t_end = min(p.total_tiles, t_begin + p.tiles_per_split);
  const int ntaps = p.KH * p.KW;
  const int npairs = (ntaps + 1) >> 1;

  // one-time clear: positions that TMA never writes (row gaps, tails) must read as finite zeros
  for (int i = threadIdx.x * 16; i < data_bytes; i += NUM_THREADS * 16)
    *reinterpret_cast<uint4*>(smem + i) = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  if (threadIdx.x == 0) {
    for (int s = 0; s < WG_STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(tmem_full, 1);
    mbar_fence_init();
  }
  if (warp == 4) tmem_alloc<512>(tmem_slot);
  if (warp == 5 && lane == 0) {
    tma_prefetch_desc(&tma_x);
    tma_prefetch_desc(&tma_dy);
  }
  if (threadIdx.x < npairs) {  // tap pair j: shift of its first tap and the descriptor with LBO = shift difference
    const int t0 = 2 * threadIdx.x, t1 = min(t0 + 1, ntaps - 1);
    const int sh0 = (t0 / p.KW) * p.Wp + (t0 % p.KW);
    const int sh1 = (t1 / p.KW) * p.Wp + (t1 % p.KW);
    pair_off[threadIdx.x] = (uint32_t)sh0 * 128u;
    // odd tail: the second half is a dummy view of the next row
    pair_desc[threadIdx.x] = make_smem_desc(0, (t1 == t0) ? 128u : (uint32_t)(sh1 - sh0) * 128u, 1024);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 5) {
    if (lane == 0) {
      int it = 0;
      for (int t = t_begin; t < t_end; ++t, ++it) {
        const int s = it % WG_STAGES;
        mbar_wait(&empty[s], ((it / WG_STAGES) & 1) ^ 1);
        const int img = t / p.tiles_per_img;
        const int tin = t - img * p.tiles_per_img;
        const int y0 = (tin / p.tiles_x) * p.R, x0 = (tin % p.tiles_x) * p.Wt;
        mbar_arrive_expect_tx(&full[s], p.a_tx_bytes + p.b_tx_bytes);
        tma_load_4d(a_base + s * p.a_stage_bytes, &tma_x, &full[s], ci0, x0 + p.org_x, y0 + p.org_y, img);
        for (int ry = 0; ry < p.R; ++ry)
          tma_load_4d(b_base + s * WG_B_BYTES + ry * p.Wp * 128, &tma_dy, &full[s], co0, x0, y0 + ry, img);
      }
    }
    __syncwarp();
  } else if (warp == 4) {
    // warp-uniform loop; one elected lane issues. Per-pair shift / LBO are read from a small table.
    constexpr uint32_t idesc = make_idesc_bf16(128, 64, /*a_mn=*/true, /*b_mn=*/true);
    const uint64_t bdesc0 = make_smem_desc(b_base, 8192, 1024);
    int s = 0;
    uint32_t ph = 0, accum0 = 0;
    for (int t = t_begin; t < t_end; ++t) {
      mbar_wait(&full[s], ph);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t a_stage = a_base + s * p.a_stage_bytes;
        const uint64_t bdesc = bdesc0 + (uint32_t)(s * (WG_B_BYTES >> 4));
        for (int j = 0; j < npairs; ++j) {
          const uint64_t adesc = pair_desc[j] + ((a_stage + pair_off[j]) >> 4);
          const uint32_t d_tmem = tmem_base + j * 64;
          umma_bf16(d_tmem, adesc, bdesc, idesc, accum0);
#pragma unroll
          for (int k = 1; k < 8; ++k) umma_bf16(d_tmem, adesc + k * 128, bdesc + k * 128, idesc, 1u);
        }
        umma_commit(&empty[s]);
        if (t == t_end - 1) umma_commit(tmem_full);
      }
      __syncwarp();
      accum0 = 1;
      if (++s == WG_STAGES) {
        s = 0;
        ph ^= 1;
      }
    }
    if (t_end <= t_begin && elect_one()) umma_commit(tmem_full);
    __syncwarp();
  } else {
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    const int r = threadIdx.x;  // TMEM lane: rows 0-63 = first tap of the pair, 64-127 = second tap
    const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
    const bool have = t_end > t_begin;
    for (int j = 0; j < npairs; ++j) {
      const int tap = 2 * j + (r >> 6);
      const int ci = ci0 + (r & 63);
#pragma unroll 1
      for (int cc = 0; cc < 2; ++cc) {
        uint32_t v[32];
        tmem_ld32(lane_addr + j * 64 + cc * 32, v);
        tmem_ld_wait();
        if (have && tap < ntaps && ci < p.C) {
          float* o = p.dw + ((size_t)tap * p.C + ci) * p.n_out + co0 + cc * 32;
          if (co0 + cc * 32 + 32 <= p.n_out && (p.n_out & 3) == 0) {
#pragma unroll
            for (int q = 0; q < 32; q += 4)
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};\n" ::"l"(o + q), "f"(__uint_as_float(v[q])),
                           "f"(__uint_as_float(v[q + 1])), "f"(__uint_as_float(v[q + 2])),
                           "f"(__uint_as_float(v[q + 3]))
                           : "memory");
          } else {
            for (int q = 0; q < 32; ++q)
              if (co0 + cc * 32 + q < p.n_out) red_add_f32(o + q, __uint_as_float(v[q]));
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess)
    return nullptr;
  fn = reinterpret_cast<EncodeTiledFn>(sym);
  return fn;
}

template <int BN, int B_STAGES>
static int launch(const CUtensorMap& tx, const CUtensorMap& tw, const Params& p, cudaStream_t stream) {
  const int smem = A_STAGES * p.a_stage_bytes + B_STAGES * BN * 128 + 256 + 1024;
  if (smem > 227 * 1024) return DTF_ERR_TOO_LARGE;
  auto kern = conv_halo_kernel<BN, B_STAGES>;
  static int attr[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (attr[dev & 63] < smem) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return DTF_ERR_CUDA;
    attr[dev & 63] = smem;
  }
  dim3 grid((p.n_out + BN - 1) / BN, p.n_img * p.tiles_per_img);
  kern<<<grid, NUM_THREADS, smem, stream>>>(tx, tw, p);
  return DTF_CHECK_LAUNCH();
}


static int g_halo_persistent = 2;  // 2: pixels-as-N kernel, 1: persistent channel-as-N, 0: one tile per CTA

// Persistent launch: returns DTF_ERR_UNSUPPORTED when the shared-memory plan does not fit (caller falls back).
template <int BN>
static int launch_persistent(const CUtensorMap& tx, const CUtensorMap& tw, const Params& p, cudaStream_t stream) {
  constexpr int B_BYTES = BN * 128;
  const int fixed = 1024 /*alignment*/ + A_STAGES * p.a_stage_bytes + 1024 /*barriers*/ + PERSIST_STAGE_BYTES;
  const int avail = 227 * 1024 - fixed;
  const int total_b = p.KH * p.KW * (p.C >> 6);
  PersistParams pp{};
  pp.tiles_n = (p.n_out + BN - 1) / BN;
  const long long items = (long long)p.n_img * p.tiles_per_img * pp.tiles_n;
  if (items > 0x7fffffffLL || (long long)p.n_img * p.H * p.W > 0x7fffffffLL) return DTF_ERR_UNSUPPORTED;
  pp.total_items = (int)items;
  if (pp.tiles_n == 1 && total_b <= MAX_B_STAGES && total_b * B_BYTES <= avail) {
    pp.resident = 1;
    pp.nb = total_b;
  } else {
    pp.resident = 0;
    pp.nb = avail / B_BYTES;
    if (pp.nb > 8) pp.nb = 8;
    if (pp.nb < 2) return DTF_ERR_UNSUPPORTED;
  }
  const int smem = fixed + pp.nb * B_BYTES;
  auto kern = conv_halo_persistent_kernel<BN>;
  static int attr[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (attr[dev & 63] < smem) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return DTF_ERR_CUDA;
    attr[dev & 63] = smem;
  }
  static int sms[64] = {};
  if (sms[dev & 63] == 0) cudaDeviceGetAttribute(&sms[dev & 63], cudaDevAttrMultiProcessorCount, dev);
  const int grid = pp.total_items < sms[dev & 63] ? pp.total_items : sms[dev & 63];
  kern<<<grid, NUM_THREADS, smem, stream>>>(tx, tw, p, pp);
  return DTF_CHECK_LAUNCH();
}


// Tile plan of the pixels-as-N kernel: Wt x R output pixels per tile with R*(Wt+KW-1) <= 256 GEMM columns, or G
// whole images per tile when one image needs less than half of them. Returns the fraction of useful columns.
static double pn_plan(int H, int W, int KH, int KW, int* Wt_out, int* R_out, int* G_out) {
  double best = 0.0;
  long long best_load = 0;
  int bw = 0, br = 0, bg = 1;
  for (int Wt = 1; Wt <= W && Wt + KW - 1 <= 256; ++Wt) {
    const int Wp = Wt + KW - 1;
    int R = PN_PIX / Wp;
    if (R > H) R = H;
    if (R < 1 || R + KH - 1 > 256) continue;
    const int hp = R + KH - 1;
    int G = 1;
    if (Wt == W && R == H) {
      G = 1 + (PN_PIX - R * Wp) / (hp * Wp);
      if (G > 64) G = 64;
    }
    const long long tiles = (long long)((W + Wt - 1) / Wt) * ((H + R - 1) / R);
    const double util = (double)G * H * W / (double)(tiles * PN_PIX);
    const long long load = tiles * hp * Wp;  // halo pixels fetched per image
    if (util > best + 1e-9 || (util > best - 1e-9 && load < best_load)) {
      best = util;
      best_load = load;
      bw = Wt;
      br = R;
      bg = G;
    }
  }
  *Wt_out = bw;
  *R_out = br;
  *G_out = bg;
  return best;
}

struct PnHost {
  Params p;
  PnParams pp;
  int smem;
  int halo_rows;
};

// Fills the kernel parameters; false when the geometry does not suit the pixels-as-N kernel.
static bool pn_setup(int n_img, int H, int W, int C, int KH, int KW, int n_out, int pad_h, int pad_w, int mirror,
                     PnHost* h, int n_phase = 0, const unsigned long long* phase_mask = nullptr) {
  if ((C & 63) || (n_out & 7) || KH * KW < 2 || KH * KW > 49) return false;
  Params& p = h->p;
  PnParams& pp = h->pp;
  p = Params{};
  pp = PnParams{};
  p.n_img = n_img; p.H = H; p.W = W; p.C = C; p.n_out = n_out; p.KH = KH; p.KW = KW;
  if (pn_plan(H, W, KH, KW, &p.Wt, &p.R, &pp.G) < 0.5) return false;
  p.Wp = p.Wt + KW - 1;
  p.tiles_x = (W + p.Wt - 1) / p.Wt;
  p.tiles_per_img = p.tiles_x * ((H + p.R - 1) / p.R);
  p.mirror = mirror;
  p.org_x = mirror ? -(KW - 1 - pad_w) : -pad_w;
  p.org_y = mirror ? -(KH - 1 - pad_h) : -pad_h;
  h->halo_rows = p.R + KH - 1;
  pp.img_rows = h->halo_rows * p.Wp;
  p.a_tx_bytes = pp.G * pp.img_rows * 128;
  const int span = (PN_PIX + (KH - 1) * p.Wp + (KW - 1)) * 128;
  p.a_stage_bytes = ((span > p.a_tx_bytes ? span : p.a_tx_bytes) + 1023) & ~1023;
  const int mw = n_out <= 64 ? 64 : 128;
  const int w_bytes = mw * 128;
  pp.tiles_co = (n_out + mw - 1) / mw;
  pp.groups = pp.G > 1 ? (n_img + pp.G - 1) / pp.G : n_img * p.tiles_per_img;
  const long long items = (long long)pp.groups * pp.tiles_co;
  if (items > 0x7fffffffLL || (long long)n_img * H * W > 0x7fffffffLL) return false;
  pp.total_items = (int)items;
  int total_w = KH * KW * (C >> 6);
  if (n_phase > 1 && n_phase <= 4 && phase_mask != nullptr) {
    // masks select by input-channel block (forward: C = n_phase x block) or by output tile (mirror: n_out = ...)
    const int blocks = mirror ? pp.tiles_co : (C >> 6);
    if (blocks % n_phase == 0) {
      pp.mask_by = mirror ? 2 : 1;
      pp.blk_per_phase = blocks / n_phase;
      for (int i = 0; i < n_phase; ++i) pp.phase_mask[i] = phase_mask[i];
      if (!mirror) {  // executed weight tiles per item (sizes the resident ring)
        total_w = 0;
        for (int i = 0; i < n_phase; ++i)
          total_w += pp.blk_per_phase * __builtin_popcountll(phase_mask[i] & ((1ull << (KH * KW)) - 1));
      }
    }
  }
  const bool masked_by_tile = pp.mask_by == 2;
  // two halo stages: a third one measured slower (it shrinks the weight ring, which matters more)
  for (pp.xst = 2; pp.xst >= 2; --pp.xst) {
    const int fixed = 1024 + pp.xst * p.a_stage_bytes + 1024 + PN_STAGE_BYTES + PN_TAB_BYTES;
    const int avail = 227 * 1024 - fixed;
    pp.resident = 0;
    if (!masked_by_tile && pp.tiles_co == 1 && total_w <= MAX_B_STAGES && total_w * w_bytes <= avail) {
      pp.resident = 1;
      pp.nb = total_w;
    } else {
      pp.nb = avail / w_bytes;
      if (pp.nb > 8) pp.nb = 8;
      if (pp.nb < (pp.xst == 3 ? 4 : 2)) continue;
    }
    h->smem = fixed + pp.nb * w_bytes;
    return true;
  }
  return false;
}

template <int MW, bool STATS>
static int launch_pn(const CUtensorMap& tx, const CUtensorMap& tw, const PnHost& h, cudaStream_t stream) {
  auto kern = conv_halo_pn_kernel<MW, STATS>;
  static int attr[64] = {};
  static int sms[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (attr[dev & 63] < h.smem) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, h.smem) != cudaSuccess)
      return DTF_ERR_CUDA;
    attr[dev & 63] = h.smem;
  }
  if (sms[dev & 63] == 0) cudaDeviceGetAttribute(&sms[dev & 63], cudaDevAttrMultiProcessorCount, dev);
  const int grid = h.pp.total_items < sms[dev & 63] ? h.pp.total_items : sms[dev & 63];
  kern<<<grid, NUM_THREADS, h.smem, stream>>>(tx, tw, h.p, h.pp);
  return DTF_CHECK_LAUNCH();
}

}  // namespace halo
}  // namespace dtf

using namespace dtf;

// Choose the tile shape (Wt output columns x R rows, (Wt+KW-1)*R <= 128 GEMM rows) with the best utilisation.
static double halo_plan(int H, int W, int KH, int KW, int* Wt_out, int* R_out) {
  double best = 0.0;
  int bw = 0, br = 0;
  for (int Wt = 1; Wt <= W && Wt + KW - 1 <= 128; ++Wt) {
    const int Wp = Wt + KW - 1;
    int R = 128 / Wp;
    if (R > H) R = H;
    if (R < 1 || R + KH - 1 > 256) continue;
    const long long tiles = (long long)((W + Wt - 1) / Wt) * ((H + R - 1) / R);
    const double util = (double)H * W / (double)(tiles * 128);
    if (util > best + 1e-9) {
      best = util;
      bw = Wt;
      br = R;
    }
  }
  *Wt_out = bw;
  *R_out = br;
  return best;
}

// 1 if dtf_conv_halo_bf16 supports the geometry (stride 1, C % 64 == 0, tile utilisation >= 60 %).
static int legacy_halo_supported(int H, int W, int C, int KH, int KW, int n_out) {
  if ((C & 63) || (n_out & 7) || KH * KW < 2 || KH > 7 || KW > 7) return 0;
  int Wt, R;
  return halo_plan(H, W, KH, KW, &Wt, &R) >= 0.60 ? 1 : 0;
}
extern "C" int dtf_conv_halo_supported(int H, int W, int C, int KH, int KW, int n_out) {
  if (KH > 7 || KW > 7) return 0;
  if (dtf::halo::g_halo_persistent >= 2) {
    dtf::halo::PnHost h;
    if (dtf::halo::pn_setup(1 << 10, H, W, C, KH, KW, n_out, 0, 0, 0, &h)) return 1;
  }
  return legacy_halo_supported(H, W, C, KH, KW, n_out);
}

// src: x (fprop) or dy (mirror = 1, data gradient) [n_img, H, W, C]; W2d: [KH*KW*C, n_out] row-major bf16 (HWIO for
// fprop, per-tap transposed filters for the gradient); out [n_img, H, W, n_out] bf16. SAME padding, stride 1.
// flags: bit0 relu, bit8 use the descriptor base-offset field for the shifted tap views.
extern "C" int dtf_conv_halo_bf16_ex(const void* src, const void* W2d, void* out, const float* bias, float* stats,
                                     int n_img, int H, int W, int C, int KH, int KW, int n_out, int pad_h, int pad_w,
                                     int mirror, int flags, int n_phase, const unsigned long long* phase_mask,
                                     void* stream_);
extern "C" int dtf_conv_halo_bf16(const void* src, const void* W2d, void* out, const float* bias, int n_img, int H,
                                  int W, int C, int KH, int KW, int n_out, int pad_h, int pad_w, int mirror, int flags,
                                  void* stream_) {
  return dtf_conv_halo_bf16_ex(src, W2d, out, bias, nullptr, n_img, H, W, C, KH, KW, n_out, pad_h, pad_w, mirror, flags,
                               0, nullptr, stream_);
}
// Rows of the per-tile statistics buffer [rows][2][n_out] that dtf_conv_halo_bf16_ex fills (0: this geometry does
// not run on the pixels-as-N kernel and cannot emit statistics).
extern "C" int dtf_conv_halo_stats_rows(int n_img, int H, int W, int C, int KH, int KW, int n_out) {
  if (KH > 7 || KW > 7 || dtf::halo::g_halo_persistent < 2) return 0;
  dtf::halo::PnHost h;
  if (!dtf::halo::pn_setup(n_img, H, W, C, KH, KW, n_out, 0, 0, 0, &h)) return 0;
  return h.pp.groups;
}
// stats (optional): per pixel tile, the column sums and sums of squares of the fp32 results (bias included, before
// the ReLU) -- the batch-norm statistics of the layer's output without another pass over it.
// n_phase / phase_mask (host array, optional): the filter matrix has structurally zero (tap, channel-phase) blocks --
// space-to-depth packed filters: channels (forward) or output channels (mirror) are n_phase equal groups and bit t of
// phase_mask[g] says whether tap t has weights for group g; absent tiles are neither loaded nor multiplied.
extern "C" int dtf_conv_halo_bf16_ex(const void* src, const void* W2d, void* out, const float* bias, float* stats,
                                     int n_img, int H, int W, int C, int KH, int KW, int n_out, int pad_h, int pad_w,
                                     int mirror, int flags, int n_phase, const unsigned long long* phase_mask,
                                     void* stream_) {
  using namespace dtf::halo;
  if (KH > 7 || KW > 7) return DTF_ERR_UNSUPPORTED;
  EncodeTiledFn fn = encode_fn();
  if (!fn) return DTF_ERR_DRIVER;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PnHost h;
  if (g_halo_persistent >= 2 && pn_setup(n_img, H, W, C, KH, KW, n_out, pad_h, pad_w, mirror, &h, n_phase, phase_mask)) {
    h.p.out = out; h.p.bias = bias; h.p.relu = flags & 1; h.p.stats = stats;
    CUtensorMap tx, tw;
    cuuint32_t es4[4] = {1, 1, 1, 1};
    cuuint64_t xd[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)n_img};
    cuuint64_t xs[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
    cuuint32_t xb[4] = {64, (cuuint32_t)h.p.Wp, (cuuint32_t)h.halo_rows, (cuuint32_t)h.pp.G};
    cuuint64_t wd[2] = {(cuuint64_t)n_out, (cuuint64_t)KH * KW * C};
    cuuint64_t ws[1] = {(cuuint64_t)n_out * 2};
    cuuint32_t wb[2] = {64, 64};
    if (fn(&tx, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(src), xd, xs, xb, es4,
           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS ||
        fn(&tw, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(W2d), wd, ws, wb, es4,
           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return DTF_ERR_TENSORMAP;
    if (stats != nullptr) return n_out <= 64 ? launch_pn<64, true>(tx, tw, h, stream) : launch_pn<128, true>(tx, tw, h, stream);
    return n_out <= 64 ? launch_pn<64, false>(tx, tw, h, stream) : launch_pn<128, false>(tx, tw, h, stream);
  }
  if (stats != nullptr || !legacy_halo_supported(H, W, C, KH, KW, n_out)) return DTF_ERR_UNSUPPORTED;
  Params p{};
  p.n_img = n_img; p.H = H; p.W = W; p.C = C; p.n_out = n_out; p.KH = KH; p.KW = KW;
  halo_plan(H, W, KH, KW, &p.Wt, &p.R);
  p.Wp = p.Wt + KW - 1;
  p.tiles_x = (W + p.Wt - 1) / p.Wt;
  p.tiles_per_img = p.tiles_x * ((H + p.R - 1) / p.R);
  p.mirror = mirror;
  p.org_x = mirror ? -(KW - 1 - pad_w) : -pad_w;
  p.org_y = mirror ? -(KH - 1 - pad_h) : -pad_h;
  const int halo_rows = p.R + KH - 1;
  p.a_tx_bytes = halo_rows * p.Wp * 128;
  const int span = (128 + (KH - 1) * p.Wp + (KW - 1)) * 128;
  p.a_stage_bytes = ((span > p.a_tx_bytes ? span : p.a_tx_bytes) + 1023) & ~1023;
  p.base_off_mode = (flags >> 8) & 0xff;
  p.out = out; p.bias = bias; p.relu = flags & 1;
  CUtensorMap tx, tw;
  {
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)n_img};
    cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
    cuuint32_t box[4] = {64, (cuuint32_t)p.Wp, (cuuint32_t)halo_rows, 1};
    cuuint32_t es[4] = {1, 1, 1, 1};
    if (fn(&tx, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(src), dims, strides, box, es,
           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return DTF_ERR_TENSORMAP;
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)n_out, (cuuint64_t)KH * KW * C};
    cuuint64_t strides[1] = {(cuuint64_t)n_out * 2};
    cuuint32_t box[2] = {64, 64};
    cuuint32_t es[2] = {1, 1};
    if (fn(&tw, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(W2d), dims, strides, box, es,
           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return DTF_ERR_TENSORMAP;
  }
  if (g_halo_persistent) {
    const int rc = (n_out <= 64) ? launch_persistent<64>(tx, tw, p, stream) : launch_persistent<128>(tx, tw, p, stream);
    if (rc != DTF_ERR_UNSUPPORTED) return rc;
  }
  if (n_out <= 64) return launch<64, 4>(tx, tw, p, stream);
  return launch<128, 4>(tx, tw, p, stream);
}

// 2 (default): pixels-as-N persistent kernel; 1: persistent channels-as-N kernel; 0: one tile per CTA.
// Returns the previous setting.
extern "C" int dtf_set_persistent_halo(int mode) {
  const int prev = dtf::halo::g_halo_persistent;
  dtf::halo::g_halo_persistent = mode < 0 ? 0 : (mode > 2 ? 2 : mode);
  return prev;
}

// 1 if dtf_conv_wgrad_halo_bf16 supports the geometry (stride 1, C % 64 == 0, n_out % 64 == 0, <= 16 taps).
extern "C" int dtf_conv_wgrad_halo_supported(int H, int W, int C, int KH, int KW, int n_out) {
  if ((C & 63) || (n_out & 63) || KH * KW < 2 || KH * KW > 16) return 0;
  int Wt, R;
  return halo_plan(H, W, KH, KW, &Wt, &R) >= 0.60 ? 1 : 0;
}

// dW[KH*KW*C, n_out] (fp32, += with red.add) from x[n_img,H,W,C] and dy[n_img,H,W,n_out]; SAME padding, stride 1.
extern "C" int dtf_conv_wgrad_halo_bf16(const void* x, const void* dy, float* dw, int n_img, int H, int W, int C,
                                        int KH, int KW, int n_out, int pad_h, int pad_w, void* stream_) {
  using namespace dtf::halo;
  if (!dtf_conv_wgrad_halo_supported(H, W, C, KH, KW, n_out)) return DTF_ERR_UNSUPPORTED;
  EncodeTiledFn fn = encode_fn();
  if (!fn) return DTF_ERR_DRIVER;
  WgradParams p{};
  p.n_img = n_img; p.H = H; p.W = W; p.C = C; p.n_out = n_out; p.KH = KH; p.KW = KW;
  halo_plan(H, W, KH, KW, &p.Wt, &p.R);
  p.Wp = p.Wt + KW - 1;
  p.tiles_x = (W + p.Wt - 1) / p.Wt;
  p.tiles_per_img = p.tiles_x * ((H + p.R - 1) / p.R);
  p.org_x = -pad_w;
  p.org_y = -pad_h;
  const int halo_rows = p.R + KH - 1;
  p.a_tx_bytes = halo_rows * p.Wp * 128;
  const int span = (128 + (KH - 1) * p.Wp + (KW - 1) + 1) * 128;
  p.a_stage_bytes = ((span > p.a_tx_bytes ? span : p.a_tx_bytes) + 1023) & ~1023;
  p.b_tx_bytes = p.R * p.Wt * 128;
  p.total_tiles = n_img * p.tiles_per_img;
  const int blocks_xy = (n_out / 64) * (C / 64);
  int splits = 148 / blocks_xy;  // one wave of (1 CTA per SM): never spill a few CTAs into a second wave
  if (splits > p.total_tiles) splits = p.total_tiles;
  if (splits < 1) splits = 1;
  p.tiles_per_split = (p.total_tiles + splits - 1) / splits;
  splits = (p.total_tiles + p.tiles_per_split - 1) / p.tiles_per_split;
  p.dw = dw;
  CUtensorMap tx, tdy;
  {
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)n_img};
    cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
    cuuint32_t box[4] = {64, (cuuint32_t)p.Wp, (cuuint32_t)halo_rows, 1};
    cuuint32_t es[4] = {1, 1, 1, 1};
    if (fn(&tx, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(x), dims, strides, box, es,
           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return DTF_ERR_TENSORMAP;
  }
  {
    cuuint64_t dims[4] = {(cuuint64_t)n_out, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)n_img};
    cuuint64_t strides[3] = {(cuuint64_t)n_out * 2, (cuuint64_t)W * n_out * 2, (cuuint64_t)H * W * n_out * 2};
    cuuint32_t box[4] = {64, (cuuint32_t)p.Wt, 1, 1};
    cuuint32_t es[4] = {1, 1, 1, 1};
    if (fn(&tdy, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(dy), dims, strides, box, es,
           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return DTF_ERR_TENSORMAP;
  }
  const int smem = WG_STAGES * (p.a_stage_bytes + WG_B_BYTES) + 256 + 1024;
  if (smem > 227 * 1024) return DTF_ERR_TOO_LARGE;
  static int attr[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (attr[dev & 63] < smem) {
    if (cudaFuncSetAttribute(wgrad_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess)
      return DTF_ERR_CUDA;
    attr[dev & 63] = smem;
  }
  dim3 grid(n_out / 64, C / 64, splits);
  wgrad_halo_kernel<<<grid, NUM_THREADS, smem, reinterpret_cast<cudaStream_t>(stream_)>>>(tx, tdy, p);
  return DTF_CHECK_LAUNCH();
}
