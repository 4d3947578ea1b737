// K-GEMM / K-CONV-FPROP / K-CONV-DGRAD / K-CONV-WGRAD for sm_100a.
//
// One warp-specialised kernel family: D[M,N] (+)= A[M,K] * B[K,N], bf16 inputs, fp32 accumulation in TMEM.
//   * B is always an "MN-major" operand: a row-major [K, N] matrix in global memory (TF's HWIO filters and
//     [in,out] FC weights are exactly that), streamed by TMA into 128B-swizzled tiles.
//   * A is one of
//       A_TMA_K     dense row-major [M, K]              (FC forward, FC dX)           -> TMA, K-major tiles
//       A_TMA_MN    dense row-major [K, M] (i.e. A^T)   (FC dW: X^T dY)               -> TMA, MN-major tiles
//       A_GATHER_K  implicit im2col of an NHWC tensor   (conv fprop, conv dgrad/deconv) -> cp.async gather
//       A_GATHER_MN implicit im2col^T of an NHWC tensor (conv wgrad)                  -> cp.async gather
//   * tcgen05.mma (cta_group::1, M=128, N=BLOCK_N, K=16) issued by one thread; accumulator lives in TMEM and is
//     read back with tcgen05.ld for the fused epilogue (bias, ReLU, bf16/fp32 store, or split-K fp32 red.add).
//
// Reference behaviour being replaced: tf.nn.conv2d / conv2d_transpose / matmul calls of
// /root/reference/distribute_tools.py:69-143,190-208 (library cuDNN/cuBLAS kernels in TF1).
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include "ptx.cuh"
#include "dtf_internal.h"

namespace dtf {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;  // 64 bf16 = 128 B = one swizzle row
constexpr int UMMA_K = 16;
constexpr int NUM_THREADS = 192;  // warps 0-3: gather producers + epilogue; warp 4: MMA issuer; warp 5: TMA
constexpr int GATHER_LAG = 2;     // cp.async groups kept in flight per producer thread

enum AMode { A_TMA_K = 0, A_TMA_MN = 1, A_GATHER_K = 2, A_GATHER_MN = 3 };
enum Geo { GEO_FPROP = 0, GEO_DGRAD = 1 };

struct IgemmParams {
  int M, N, K;
  // gather source (NHWC bf16) and geometry
  const __nv_bfloat16* src;
  int H, W, C;      // source tensor spatial dims / channels
  int PH, PW;       // pixel grid indexed by the GEMM row (K-mode) or by the reduction index (MN-mode)
  int KH, KW;
  int stride, pad_h, pad_w;
  int geo;
  // epilogue
  void* out;
  int ldc;
  const float* bias;
  int relu;
  int out_fp32;
  int atomic;
  int kb_per_split;
  // optional bf16 [M, N] (row stride ldc) added to the result before the ReLU: residual inputs, and the second
  // gradient of a tensor with two consumers (so autograd never has to sum them in a separate pass)
  const __nv_bfloat16* addend;
  // optional batch-norm statistics of the (bf16-rounded) output: per-CTA partial rows [grid][2][N] of column sums
  // and sums of squares, accumulated in shared memory across the tiles of a persistent CTA
  float* stats;
};

template <int BN, int STAGES>
struct SmemLayout {
  static constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;  // 16 KB
  static constexpr int B_BYTES = BN * BLOCK_K * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int BAR_OFFSET = STAGES * STAGE_BYTES;
  static constexpr int TOTAL = BAR_OFFSET + 256 + 1024;  // barriers + slack for 1024 B alignment
};

// Decompose k (im2col column) -> tap, channel offset. C is a multiple of 8 so a 16 B chunk never straddles taps.
struct PixelCoord {
  int n, y, x;
};

// Fused epilogue of one 128 x BN accumulator tile: TMEM -> registers -> (bias, ReLU) -> bf16 / fp32 store or
// split-K red.add. Called by the four epilogue warps (warp index == TMEM lane quarter).
// `warp` = TMEM lane quarter of the calling warp (warp index % 4); with two warps per quarter (persistent kernel, wide
// tiles) each takes every second 32-column group (cc0 / cstep).
template <int BN>
__device__ __forceinline__ void igemm_epilogue(const IgemmParams& p, uint32_t tmem_base, int tile_m, int tile_n,
                                               int warp, int cc0 = 0, int cstep = 1) {
  const int r = warp * 32 + (threadIdx.x & 31);  // accumulator row == TMEM lane
  const int m = tile_m * BLOCK_M + r;
  const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
#pragma unroll 1
  for (int cc = cc0; cc < BN / 32; cc += cstep) {
    uint32_t v[32];
    tmem_ld32(lane_addr + cc * 32, v);
    tmem_ld_wait();
    const int n0 = tile_n * BN + cc * 32;
    if (m < p.M && n0 < p.N) {
      float f[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
      if (p.bias != nullptr) {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (n0 + j < p.N) f[j] += __ldg(p.bias + n0 + j);
      }
      if (p.addend != nullptr) {
        const __nv_bfloat16* ad = p.addend + (size_t)m * p.ldc + n0;
        for (int j = 0; j < 32; ++j)
          if (n0 + j < p.N) f[j] += __bfloat162float(ad[j]);
      }
      if (p.relu) {
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
      }
      const bool full = (n0 + 32 <= p.N);
      if (p.atomic) {
        float* o = reinterpret_cast<float*>(p.out) + (size_t)m * p.ldc + n0;
        if (full && ((p.ldc & 3) == 0)) {
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};\n" ::"l"(o + j), "f"(f[j]), "f"(f[j + 1]),
                         "f"(f[j + 2]), "f"(f[j + 3])
                         : "memory");
        } else {
          for (int j = 0; j < 32; ++j)
            if (n0 + j < p.N) red_add_f32(o + j, f[j]);
        }
      } else if (p.out_fp32) {
        float* o = reinterpret_cast<float*>(p.out) + (size_t)m * p.ldc + n0;
        if (full && ((p.ldc & 3) == 0)) {
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            *reinterpret_cast<float4*>(o + j) = make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
        } else {
          for (int j = 0; j < 32; ++j)
            if (n0 + j < p.N) o[j] = f[j];
        }
      } else {
        __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + (size_t)m * p.ldc + n0;
        if (full && ((p.ldc & 7) == 0)) {
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
            if (n0 + j < p.N) o[j] = __float2bfloat16(f[j]);
        }
      }
    }
  }
}


// Coalescing epilogue for bf16 outputs: a warp's 32 rows x COLS columns go through a shared-memory staging tile
// (16-byte chunks XOR-swizzled by the row) so that every global store instruction writes complete 32-byte sectors of
// few rows (COLS = 64: four full 128-byte lines per instruction; COLS = 32: eight 64-byte row segments) instead of
// 32 scattered 16-byte pieces. `stage` is this warp's private 32 * COLS * 2 bytes.
template <int BN, int COLS, bool STATS = false>
__device__ __forceinline__ void igemm_epilogue_staged(const IgemmParams& p, uint32_t tmem_base, int tile_m,
                                                      int tile_n, int warp, uint32_t stage, uint32_t acc = 0,
                                                      int cg0 = 0, int cgstep = 1) {
  // `stage` / `acc`: 32-bit shared-memory addresses (explicit st.shared / ld.shared, see ptx.cuh)
  constexpr int ROWB = COLS * 2;        // bytes per staged row
  constexpr int CPR = ROWB / 16;        // 16-byte chunks per row (8 or 4)
  constexpr int RPI = 32 / CPR;         // rows written per store instruction
  constexpr int NIT = 32 / RPI;         // store instructions per staged group
  const int lane = threadIdx.x & 31;
  const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
  const int row0 = tile_m * BLOCK_M + warp * 32;
  __nv_bfloat16* outp = reinterpret_cast<__nv_bfloat16*>(p.out);
  const int chunk = lane & (CPR - 1);
  const bool has_add = p.addend != nullptr;
#pragma unroll 1
  for (int cg = cg0; cg < BN / COLS; cg += cgstep) {
    const int n0 = tile_n * BN + cg * COLS;
    const int n = n0 + chunk * 8;
    // the addend pieces of this group are requested first, all at once, so their latency overlaps the TMEM load
    // and the staging below (a load issued between the stores could not be hoisted: out and addend may alias)
    uint4 av[NIT];
    if (has_add) {
#pragma unroll
      for (int i = 0; i < NIT; ++i) {
        const int m = row0 + i * RPI + lane / CPR;
        av[i] = (m < p.M && n + 8 <= p.N) ? __ldg(reinterpret_cast<const uint4*>(p.addend + (size_t)m * p.ldc + n))
                                          : make_uint4(0, 0, 0, 0);
      }
    }
#pragma unroll
    for (int h = 0; h < COLS / 32; ++h) {
      uint32_t v[32];
      tmem_ld32(lane_addr + cg * COLS + h * 32, v);
      tmem_ld_wait();
      float f[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
      if (p.bias != nullptr) {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (n0 + h * 32 + j < p.N) f[j] += __ldg(p.bias + n0 + h * 32 + j);
      }
      if (p.relu && !has_add) {  // with an addend the ReLU follows the add, in the write-out below
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
      }
#pragma unroll
      for (int c = 0; c < 4; ++c) {  // four 16-byte chunks of this 32-column half
        uint4 q;
        q.x = pack_bf16x2(f[c * 8 + 0], f[c * 8 + 1]);
        q.y = pack_bf16x2(f[c * 8 + 2], f[c * 8 + 3]);
        q.z = pack_bf16x2(f[c * 8 + 4], f[c * 8 + 5]);
        q.w = pack_bf16x2(f[c * 8 + 6], f[c * 8 + 7]);
        const int ch = h * 4 + c;
        st_shared_v4(stage + lane * ROWB + ((ch ^ (lane & (CPR - 1))) << 4), q);
      }
    }
    __syncwarp();
    float cs[8], cq[8];  // STATS: this lane's column sums / sums of squares over its NIT rows of the group
    if (STATS) {
#pragma unroll
      for (int j = 0; j < 8; ++j) cs[j] = cq[j] = 0.f;
    }
#pragma unroll
    for (int i = 0; i < NIT; ++i) {
      const int r = i * RPI + lane / CPR;
      uint4 q = ld_shared_v4(stage + r * ROWB + ((chunk ^ (r & (CPR - 1))) << 4));
      const int m = row0 + r;
      if (m < p.M && n + 8 <= p.N) {
        if (STATS) {  // statistics of exactly the values that are stored (bf16-rounded)
          const __nv_bfloat162* qs = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 x = __bfloat1622float2(qs[e]);
            cs[2 * e] += x.x;
            cs[2 * e + 1] += x.y;
            cq[2 * e] = fmaf(x.x, x.x, cq[2 * e]);
            cq[2 * e + 1] = fmaf(x.y, x.y, cq[2 * e + 1]);
          }
        }
        if (has_add) {  // add in fp32, ReLU, round once more
          const __nv_bfloat162* qa = reinterpret_cast<const __nv_bfloat162*>(&q);
          const __nv_bfloat162* aa = reinterpret_cast<const __nv_bfloat162*>(&av[i]);
          uint32_t o[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 x = __bfloat1622float2(qa[e]), y = __bfloat1622float2(aa[e]);
            float s0 = x.x + y.x, s1 = x.y + y.y;
            if (p.relu) {
              s0 = fmaxf(s0, 0.f);
              s1 = fmaxf(s1, 0.f);
            }
            o[e] = pack_bf16x2(s0, s1);
          }
          q = make_uint4(o[0], o[1], o[2], o[3]);
        }
        *reinterpret_cast<uint4*>(outp + (size_t)m * p.ldc + n) = q;
      }
    }
    if (STATS) {
      // lanes chunk, chunk + CPR, ... hold the same 8 columns: fold them, then the first CPR lanes add the group's
      // sums to this warp's private accumulators (fixed order: bitwise reproducible)
#pragma unroll
      for (int off = CPR; off < 32; off <<= 1) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          cs[j] += __shfl_xor_sync(0xffffffffu, cs[j], off);
          cq[j] += __shfl_xor_sync(0xffffffffu, cq[j], off);
        }
      }
      if (lane < CPR && n + 8 <= p.N) {
        const uint32_t a0 = acc + (uint32_t)n * 4u, a1 = acc + (uint32_t)(p.N + n) * 4u;
        float4 t = ld_shared_f4(a0);
        st_shared_f4(a0, make_float4(t.x + cs[0], t.y + cs[1], t.z + cs[2], t.w + cs[3]));
        t = ld_shared_f4(a0 + 16);
        st_shared_f4(a0 + 16, make_float4(t.x + cs[4], t.y + cs[5], t.z + cs[6], t.w + cs[7]));
        t = ld_shared_f4(a1);
        st_shared_f4(a1, make_float4(t.x + cq[0], t.y + cq[1], t.z + cq[2], t.w + cq[3]));
        t = ld_shared_f4(a1 + 16);
        st_shared_f4(a1 + 16, make_float4(t.x + cq[4], t.y + cq[5], t.z + cq[6], t.w + cq[7]));
      }
    }
    __syncwarp();
  }
}

// Pull the addend rows of a tile towards L2 ahead of the epilogue that will read them (one 128-byte line per call).
template <int BN>
__device__ __forceinline__ void prefetch_addend_tile(const IgemmParams& p, int tile_m, int tile_n, int tile_row) {
  const int m = tile_m * BLOCK_M + tile_row;  // one epilogue thread per tile row
  if (m >= p.M) return;
  const char* row = reinterpret_cast<const char*>(p.addend + (size_t)m * p.ldc + (size_t)tile_n * BN);
  const int bytes = min(BN, p.N - tile_n * BN) * 2;
  for (int b = 0; b < bytes; b += 128) asm volatile("prefetch.global.L2 [%0];\n" ::"l"(row + b));
}

// K-AG-GEMM support: the B operand (weights) is sharded over the ranks of a parameter-server group and is NOT
// gathered beforehand: the TMA producer picks, per k-block (mode 1: B sharded by rows / K) or per n-tile (mode 2:
// B sharded by columns / N), the tensor map that points into the OWNER's symmetric heap, so weight tiles cross
// NVLink straight into shared memory while tcgen05.mma consumes the previous stages.
struct PeerMaps {
  CUtensorMap m[8];
};
struct AgSel {
  int mode;          // 0: plain (single map), 1: owner by k-block, 2: owner by n-tile
  int world;
  long long unit;    // arena elements covered by one k-block (mode 1) or by one n-tile (mode 2)
  long long b_off;   // arena element offset of the weight matrix
  long long bounds[9];
};
__device__ __forceinline__ int ag_owner(const AgSel& ag, long long idx) {
  const long long e = ag.b_off + idx * ag.unit;
  int o = 0;
#pragma unroll
  for (int r = 1; r < 8; ++r)
    if (r < ag.world && e >= ag.bounds[r]) o = r;
  return o;
}

template <int BN, int STAGES, int AMODE>
__device__ __forceinline__ void igemm_body(const CUtensorMap& tma_a, const CUtensorMap* tma_b_list, const AgSel& ag,
                                           const IgemmParams& p) {
  const CUtensorMap& tma_b = tma_b_list[0];
  using L = SmemLayout<BN, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::BAR_OFFSET);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // tiles are linearised in grid.x (grid.y is limited to 65535 blocks: 8.4 M rows): consecutive blocks share an A tile
  const int tiles_n_grid = (p.N + BN - 1) / BN;
  const int tile_n = blockIdx.x % tiles_n_grid;
  const int tile_m = blockIdx.x / tiles_n_grid;
  const int total_kb = (p.K + BLOCK_K - 1) / BLOCK_K;
  const int kb0 = blockIdx.z * p.kb_per_split;
  const int kb1 = min(total_kb, kb0 + p.kb_per_split);
  constexpr bool kGather = (AMODE == A_GATHER_K || AMODE == A_GATHER_MN);
  constexpr int LAG = (GATHER_LAG < STAGES - 1) ? GATHER_LAG : STAGES - 1;  // must stay below the ring depth
  constexpr bool kAMn = (AMODE == A_TMA_MN || AMODE == A_GATHER_MN);

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], kGather ? 129 : 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full_bar, 1);
    mbar_fence_init();
  }
  if (warp == 4) {
    tmem_alloc<BN>(tmem_slot);
  }
  if (warp == 5 && lane == 0) {
    tma_prefetch_desc(&tma_b);
    if (!kGather) tma_prefetch_desc(&tma_a);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const uint32_t smem_base = smem_u32(smem);
  auto a_stage = [&](int s) { return smem_base + s * L::STAGE_BYTES; };
  auto b_stage = [&](int s) { return smem_base + s * L::STAGE_BYTES + L::A_BYTES; };

  if (warp < 4) {
    // ===================================================== A producer (gather modes), then epilogue
    if constexpr (AMODE == A_GATHER_K) {
      const int r = threadIdx.x;  // tile row owned by this thread
      const int m = tile_m * BLOCK_M + r;
      const bool row_ok = m < p.M;
      int n_img = 0, py = 0, px = 0;
      if (row_ok) {
        n_img = m / (p.PH * p.PW);
        int rem = m - n_img * (p.PH * p.PW);
        py = rem / p.PW;
        px = rem - py * p.PW;
      }
      const __nv_bfloat16* img = p.src + (size_t)n_img * p.H * p.W * p.C;
      const uint32_t row_off = r * 128;
      const uint32_t sw = r & 7;
      int it = 0;
      // Fast path: with C a multiple of 64 every k-block (64 channels = 128 B) lies inside ONE filter tap, so the
      // tap is tracked with counters (no divisions) and the bounds test is done once per k-block.
      const bool one_tap_per_kb = (p.C & 63) == 0;
      const int cpb = p.C >> 6;  // k-blocks per tap
      int f_tap = 0, f_cblk = 0, f_kh = 0, f_kw = 0;
      if (one_tap_per_kb) {
        f_tap = kb0 / cpb;
        f_cblk = kb0 - f_tap * cpb;
        f_kh = f_tap / p.KW;
        f_kw = f_tap - f_kh * p.KW;
      }
      for (int kb = kb0; kb < kb1; ++kb, ++it) {
        const int s = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1;
        mbar_wait(&empty_bar[s], ph ^ 1);
        const uint32_t dst_row = a_stage(s) + row_off;
        if (one_tap_per_kb) {
          const char* srcp = reinterpret_cast<const char*>(p.src);
          uint32_t nbytes = 0;
          if (row_ok) {
            int iy, ix;
            bool ok;
            if (p.geo == GEO_FPROP) {
              iy = py * p.stride - p.pad_h + f_kh;
              ix = px * p.stride - p.pad_w + f_kw;
              ok = (iy >= 0) && (iy < p.H) && (ix >= 0) && (ix < p.W);
            } else {
              const int ty = py + p.pad_h - f_kh;
              const int tx = px + p.pad_w - f_kw;
              if (p.stride == 1) {
                iy = ty;
                ix = tx;
                ok = (ty >= 0) && (tx >= 0) && (iy < p.H) && (ix < p.W);
              } else {
                iy = ty / p.stride;
                ix = tx / p.stride;
                ok = (ty >= 0) && (tx >= 0) && (iy * p.stride == ty) && (ix * p.stride == tx) && (iy < p.H) &&
                     (ix < p.W);
              }
            }
            if (ok) {
              srcp = reinterpret_cast<const char*>(img + ((size_t)iy * p.W + ix) * p.C + (f_cblk << 6));
              nbytes = 16;
            }
          }
#pragma unroll
          for (int c = 0; c < 8; ++c) cp_async16(dst_row + ((c ^ sw) << 4), srcp + (nbytes ? c * 16 : 0), nbytes);
          if (++f_cblk == cpb) {
            f_cblk = 0;
            if (++f_kw == p.KW) {
              f_kw = 0;
              ++f_kh;
            }
          }
        } else {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const int k = kb * BLOCK_K + c * 8;
          const void* srcp = p.src;
          uint32_t nbytes = 0;
          if (row_ok && k < p.K) {
            const int tap = k / p.C;
            const int ci = k - tap * p.C;
            const int kh = tap / p.KW;
            const int kw = tap - kh * p.KW;
            int iy, ix;
            bool ok;
            if (p.geo == GEO_FPROP) {
              iy = py * p.stride - p.pad_h + kh;
              ix = px * p.stride - p.pad_w + kw;
              ok = (iy >= 0) && (iy < p.H) && (ix >= 0) && (ix < p.W);
            } else {
              const int ty = py + p.pad_h - kh;
              const int tx = px + p.pad_w - kw;
              iy = ty / p.stride;
              ix = tx / p.stride;
              ok = (ty >= 0) && (tx >= 0) && (iy * p.stride == ty) && (ix * p.stride == tx) && (iy < p.H) &&
                   (ix < p.W);
            }
            if (ok) {
              srcp = img + ((size_t)iy * p.W + ix) * p.C + ci;
              nbytes = 16;
            }
          }
          cp_async16(dst_row + ((c ^ sw) << 4), srcp, nbytes);
        }
        }
        cp_async_commit();
        if (it >= LAG) {
          cp_async_wait<LAG>();
          fence_proxy_async_smem();
          mbar_arrive(&full_bar[(it - LAG) % STAGES]);
        }
      }
      cp_async_wait<0>();
      fence_proxy_async_smem();
      for (int j = max(0, it - LAG); j < it; ++j) mbar_arrive(&full_bar[j % STAGES]);
    } else if constexpr (AMODE == A_GATHER_MN) {
      // A^T tile: [2 atoms][64 reduction rows (pixels)][64 M' elements = 128 B]
      const int krow = threadIdx.x & 63;
      const int atom = threadIdx.x >> 6;
      const uint32_t sw = krow & 7;
      const uint32_t dst_off = atom * 8192 + krow * 128;
      // M' chunks owned by this thread are fixed for the whole tile: decode (tap, ci) once.
      int tap_kh[8], tap_kw[8], tap_ci[8];
      bool tap_ok[8];
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const int mp = tile_m * BLOCK_M + atom * 64 + c * 8;
        tap_ok[c] = mp < p.M;
        const int tap = mp / p.C;
        tap_ci[c] = mp - tap * p.C;
        tap_kh[c] = tap / p.KW;
        tap_kw[c] = tap - tap_kh[c] * p.KW;
      }
      int it = 0;
      for (int kb = kb0; kb < kb1; ++kb, ++it) {
        const int s = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1;
        const int pix = kb * BLOCK_K + krow;
        const bool pix_ok = pix < p.K;
        int n_img = 0, py = 0, px = 0;
        if (pix_ok) {
          n_img = pix / (p.PH * p.PW);
          int rem = pix - n_img * (p.PH * p.PW);
          py = rem / p.PW;
          px = rem - py * p.PW;
        }
        const __nv_bfloat16* img = p.src + (size_t)n_img * p.H * p.W * p.C;
        const int by = py * p.stride - p.pad_h;
        const int bx = px * p.stride - p.pad_w;
        mbar_wait(&empty_bar[s], ph ^ 1);
        const uint32_t dst_row = a_stage(s) + dst_off;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const int iy = by + tap_kh[c];
          const int ix = bx + tap_kw[c];
          const bool ok = pix_ok && tap_ok[c] && (iy >= 0) && (iy < p.H) && (ix >= 0) && (ix < p.W);
          const void* srcp = ok ? (const void*)(img + ((size_t)iy * p.W + ix) * p.C + tap_ci[c]) : (const void*)p.src;
          cp_async16(dst_row + ((c ^ sw) << 4), srcp, ok ? 16u : 0u);
        }
        cp_async_commit();
        if (it >= LAG) {
          cp_async_wait<LAG>();
          fence_proxy_async_smem();
          mbar_arrive(&full_bar[(it - LAG) % STAGES]);
        }
      }
      cp_async_wait<0>();
      fence_proxy_async_smem();
      for (int j = max(0, it - LAG); j < it; ++j) mbar_arrive(&full_bar[j % STAGES]);
    }

    // ----------------------------------------------------- epilogue: TMEM -> registers -> global
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    igemm_epilogue<BN>(p, tmem_base, tile_m, tile_n, warp);
  } else if (warp == 4) {
    // ===================================================== MMA issuer
    // The warp walks the loop with warp-uniform state (descriptors stay in uniform registers, ring index and phase
    // are counters); one elected lane issues the MMAs and the commits.
    constexpr uint32_t idesc = make_idesc_bf16(BLOCK_M, BN, kAMn, /*b_mn=*/true);
    const uint64_t adesc0 = kAMn ? make_smem_desc(a_stage(0), /*lbo=*/8192, /*sbo=*/1024) : make_smem_desc(a_stage(0), 16, 1024);
    const uint64_t bdesc0 = make_smem_desc(b_stage(0), /*lbo=*/8192, /*sbo=*/1024);
    constexpr uint32_t kAStep = kAMn ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
    constexpr uint32_t kBStep = (UMMA_K * 128) >> 4;
    constexpr uint32_t kStage16 = L::STAGE_BYTES >> 4;
    int s = 0;
    uint32_t ph = 0, accum = 0;
    for (int kb = kb0; kb < kb1; ++kb) {
      mbar_wait(&full_bar[s], ph);
      tc_fence_after();
      if (elect_one()) {
        const uint64_t adesc = adesc0 + (uint32_t)s * kStage16;
        const uint64_t bdesc = bdesc0 + (uint32_t)s * kStage16;
#pragma unroll
        for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
          umma_bf16(tmem_base, adesc + k * kAStep, bdesc + k * kBStep, idesc, (k > 0) ? 1u : accum);
        umma_commit(&empty_bar[s]);
        if (kb == kb1 - 1) umma_commit(tmem_full_bar);
      }
      __syncwarp();
      accum = 1;
      if (++s == STAGES) {
        s = 0;
        ph ^= 1;
      }
    }
    if (kb1 <= kb0 && elect_one()) umma_commit(tmem_full_bar);
    __syncwarp();
  } else {
    // ===================================================== TMA producer (B always; A in dense modes)
    if (lane == 0) {
      int it = 0;
      for (int kb = kb0; kb < kb1; ++kb, ++it) {
        const int s = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1;
        mbar_wait(&empty_bar[s], ph ^ 1);
        mbar_arrive_expect_tx(&full_bar[s], kGather ? L::B_BYTES : L::STAGE_BYTES);
        const CUtensorMap* bmap = tma_b_list;
        if (ag.mode == 1) bmap = tma_b_list + ag_owner(ag, kb);
        else if (ag.mode == 2) bmap = tma_b_list + ag_owner(ag, tile_n);
        if constexpr (AMODE == A_TMA_K) {
          tma_load_2d(a_stage(s), &tma_a, &full_bar[s], kb * BLOCK_K, tile_m * BLOCK_M);
        } else if constexpr (AMODE == A_TMA_MN) {
          tma_load_2d(a_stage(s), &tma_a, &full_bar[s], tile_m * BLOCK_M, kb * BLOCK_K);
          tma_load_2d(a_stage(s) + 8192, &tma_a, &full_bar[s], tile_m * BLOCK_M + 64, kb * BLOCK_K);
        }
#pragma unroll
        for (int j = 0; j < BN / 64; ++j)
          tma_load_2d(b_stage(s) + j * 8192, bmap, &full_bar[s], tile_n * BN + j * 64, kb * BLOCK_K);
      }
    }
    __syncwarp();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tmem_dealloc<BN>(tmem_base);
  }
}


// Persistent variant for the TMA-fed (dense) modes: one CTA per SM slot walks tiles t = blockIdx.x, +gridDim.x, ...
// The TMA producer keeps the smem ring full ACROSS tiles, the accumulator is double buffered in TMEM (2 x BN
// columns), so the epilogue of tile i (tcgen05.ld + stores) overlaps the loads and MMAs of tile i+1, and the
// per-tile costs of the one-shot kernel (launch, TMEM alloc, barrier init, exposed first-load latency) are paid once.
// Wide tiles (BN = 256) run one CTA per SM; four epilogue warps then cannot keep up with the bandwidth-bound 1x1 layers
// (each warp walks ~1500 dependent instructions per tile: measured, every instruction added to the epilogue showed up
// 1:1 in the kernel time). Those instances therefore get EIGHT epilogue warps -- warps 6..9 join warps 0..3, two warps
// per TMEM lane quarter, each taking every second column group.
template <int BN>
struct PersistCfg {
  static constexpr int EW = (BN >= 256) ? 8 : 4;
  static constexpr int THREADS = NUM_THREADS + (EW - 4) * 32;
  static constexpr int STAGE_COLS = (BN >= 256) ? 64 : 32;
  static constexpr int STAGING_BYTES = EW * 32 * STAGE_COLS * 2;
};

template <int BN, int STAGES, int AMODE>
__global__ void __launch_bounds__(PersistCfg<BN>::THREADS)
igemm_persistent_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                        const IgemmParams p, int tiles_m, int tiles_n, int splits) {
  static_assert(AMODE == A_TMA_K || AMODE == A_TMA_MN, "persistent kernel is for the TMA-fed modes");
  using L = SmemLayout<BN, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::BAR_OFFSET);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* acc_full = empty_bar + STAGES;   // [2]
  uint64_t* acc_empty = acc_full + 2;        // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  constexpr int kStageCols = PersistCfg<BN>::STAGE_COLS;  // wide tiles run one CTA per SM: 4 KB per warp
  constexpr int EW = PersistCfg<BN>::EW;
  uint8_t* stage_base = smem + L::BAR_OFFSET + 256;  // EW warps x (32 rows x kStageCols x 2 B) epilogue staging
  float* stats_acc = reinterpret_cast<float*>(stage_base + PersistCfg<BN>::STAGING_BYTES);  // [4 quarters][2][N]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr bool kAMn = (AMODE == A_TMA_MN);
  // bf16 output with 16-byte aligned rows and whole 8-column groups: use the coalescing epilogue
  const bool staged = !p.atomic && !p.out_fp32 && (p.ldc % 8 == 0) && (p.N % 8 == 0) &&
                      ((reinterpret_cast<uintptr_t>(p.out) & 15) == 0) &&
                      ((reinterpret_cast<uintptr_t>(p.addend) & 15) == 0);
  const int total_kb = (p.K + BLOCK_K - 1) / BLOCK_K;
  const int total_tiles = tiles_m * tiles_n * splits;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&acc_full[b], 1);
      mbar_init(&acc_empty[b], EW);  // one arrival per epilogue warp
    }
    mbar_fence_init();
  }
  if (warp == 4) tmem_alloc<2 * BN>(tmem_slot);
  if (warp == 5 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t smem_base = smem_u32(smem);
  auto a_stage = [&](int s) { return smem_base + s * L::STAGE_BYTES; };
  auto b_stage = [&](int s) { return smem_base + s * L::STAGE_BYTES + L::A_BYTES; };
  auto decode = [&](int t, int& tile_m, int& tile_n, int& kb0, int& kb1) {
    tile_n = t % tiles_n;
    const int r = t / tiles_n;
    tile_m = r % tiles_m;
    const int z = r / tiles_m;
    kb0 = z * p.kb_per_split;
    kb1 = min(total_kb, kb0 + p.kb_per_split);
  };

  if (warp == 5) {
    if (lane == 0) {
      int it = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        int tile_m, tile_n, kb0, kb1;
        decode(t, tile_m, tile_n, kb0, kb1);
        for (int kb = kb0; kb < kb1; ++kb, ++it) {
          const int s = it % STAGES;
          mbar_wait(&empty_bar[s], ((it / STAGES) & 1) ^ 1);
          mbar_arrive_expect_tx(&full_bar[s], L::STAGE_BYTES);
          if constexpr (AMODE == A_TMA_K) {
            tma_load_2d(a_stage(s), &tma_a, &full_bar[s], kb * BLOCK_K, tile_m * BLOCK_M);
          } else {
            tma_load_2d(a_stage(s), &tma_a, &full_bar[s], tile_m * BLOCK_M, kb * BLOCK_K);
            tma_load_2d(a_stage(s) + 8192, &tma_a, &full_bar[s], tile_m * BLOCK_M + 64, kb * BLOCK_K);
          }
#pragma unroll
          for (int j = 0; j < BN / 64; ++j)
            tma_load_2d(b_stage(s) + j * 8192, &tma_b, &full_bar[s], tile_n * BN + j * 64, kb * BLOCK_K);
        }
      }
    }
    __syncwarp();
  } else if (warp == 4) {
    // warp-uniform loop state (descriptors in uniform registers); one elected lane issues MMAs and commits
    constexpr uint32_t idesc = make_idesc_bf16(BLOCK_M, BN, kAMn, true);
    const uint64_t adesc0 = kAMn ? make_smem_desc(a_stage(0), 8192, 1024) : make_smem_desc(a_stage(0), 16, 1024);
    const uint64_t bdesc0 = make_smem_desc(b_stage(0), 8192, 1024);
    constexpr uint32_t kAStep = kAMn ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
    constexpr uint32_t kBStep = (UMMA_K * 128) >> 4;
    constexpr uint32_t kStage16 = L::STAGE_BYTES >> 4;
    int s = 0, buf = 0;
    uint32_t ph = 0, acc_par = 1;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      int tile_m, tile_n, kb0, kb1;
      decode(t, tile_m, tile_n, kb0, kb1);
      mbar_wait(&acc_empty[buf], acc_par);  // epilogue has drained this accumulator
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + buf * BN;
      uint32_t accum = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&full_bar[s], ph);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t adesc = adesc0 + (uint32_t)s * kStage16;
          const uint64_t bdesc = bdesc0 + (uint32_t)s * kStage16;
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
            umma_bf16(d_tmem, adesc + k * kAStep, bdesc + k * kBStep, idesc, (k > 0) ? 1u : accum);
          umma_commit(&empty_bar[s]);
          if (kb == kb1 - 1) umma_commit(&acc_full[buf]);
        }
        __syncwarp();
        accum = 1;
        if (++s == STAGES) {
          s = 0;
          ph ^= 1;
        }
      }
      if ((buf ^= 1) == 0) acc_par ^= 1;
    }
  } else {
    const int q = warp & 3;                // TMEM lane quarter this warp may read
    const int half = (warp < 4) ? 0 : 1;   // which column groups of a tile (second epilogue warp set: warps 6..9)
    constexpr int kSplit = EW / 4;
    const uint32_t my_stage = smem_u32(stage_base) + (half * 4 + q) * (32 * kStageCols * 2);
    const bool want_stats = staged && p.stats != nullptr;
    const uint32_t acc_u32 = smem_u32(stats_acc);
    const uint32_t my_acc = acc_u32 + (uint32_t)(q * 2 * p.N) * 4u;  // shared by the two warps of a quarter
    if (want_stats) {
      for (int i = (half * 4 + q) * 32 + lane; i < 4 * 2 * p.N; i += EW * 32) st_shared_f32(acc_u32 + i * 4, 0.f);
      asm volatile("bar.sync 1, %0;\n" ::"r"(EW * 32) : "memory");
    }
    int lt = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++lt) {
      int tile_m, tile_n, kb0, kb1;
      decode(t, tile_m, tile_n, kb0, kb1);
      const int buf = lt & 1;
      if (p.addend != nullptr && half == 0) {  // this tile's rows (first tile only) and the next tile's, a tile ahead
        if (lt == 0) prefetch_addend_tile<BN>(p, tile_m, tile_n, q * 32 + lane);
        const int tn = t + gridDim.x;
        if (tn < total_tiles) {
          int m2, n2, k0, k1;
          decode(tn, m2, n2, k0, k1);
          prefetch_addend_tile<BN>(p, m2, n2, q * 32 + lane);
        }
      }
      mbar_wait(&acc_full[buf], (lt >> 1) & 1);
      tc_fence_after();
      if (want_stats)
        igemm_epilogue_staged<BN, kStageCols, true>(p, tmem_base + buf * BN, tile_m, tile_n, q, my_stage, my_acc, half,
                                                    kSplit);
      else if (staged)
        igemm_epilogue_staged<BN, kStageCols>(p, tmem_base + buf * BN, tile_m, tile_n, q, my_stage, 0u, half,
                                              kSplit);
      else igemm_epilogue<BN>(p, tmem_base + buf * BN, tile_m, tile_n, q, half, kSplit);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[buf]);
    }
    if (want_stats) {  // one partial row per CTA: the four quarters are added in a fixed order (reproducible)
      asm volatile("bar.sync 1, %0;\n" ::"r"(EW * 32) : "memory");
      float* o = p.stats + (size_t)blockIdx.x * 2 * p.N;
      const uint32_t qs = (uint32_t)(2 * p.N) * 4u;
      for (int i = (half * 4 + q) * 32 + lane; i < 2 * p.N; i += EW * 32) {
        const uint32_t a = acc_u32 + i * 4;
        o[i] = (ld_shared_f32(a) + ld_shared_f32(a + qs)) + (ld_shared_f32(a + 2 * qs) + ld_shared_f32(a + 3 * qs));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tmem_dealloc<2 * BN>(tmem_base);
  }
}

constexpr int kMaxSmemOptin = 227 * 1024;
// Shared memory and grid of the persistent kernel for a problem (the statistics accumulators are 32 N bytes).
template <int BN, int STAGES>
static void persistent_plan(int M, int N, int splits, bool with_stats, int* smem_out, int* grid_out) {
  using L = SmemLayout<BN, STAGES>;
  constexpr int kStage = PersistCfg<BN>::STAGING_BYTES;  // epilogue staging: 32 KB (wide tiles) or 8 KB
  const int smem = L::TOTAL + kStage + (with_stats ? 4 * 2 * N * 4 : 0);
  const int tiles_n = (N + BN - 1) / BN, tiles_m = (M + BLOCK_M - 1) / BLOCK_M;
  const int total = tiles_n * tiles_m * splits;
  const int per_sm = (smem <= 112 * 1024) ? 2 : 1;
  *smem_out = smem;
  *grid_out = total < 148 * per_sm ? total : 148 * per_sm;
}

template <int BN, int STAGES, int AMODE>
static int launch_persistent(const CUtensorMap& ta, const CUtensorMap& tb, const IgemmParams& p, int splits,
                             cudaStream_t stream) {
  using L = SmemLayout<BN, STAGES>;
  constexpr int kStage = PersistCfg<BN>::STAGING_BYTES;
  static bool attr_set[64] = {};
  auto kern = igemm_persistent_kernel<BN, STAGES, AMODE>;
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    const int want = L::TOTAL + kStage + 4 * 2 * 2048 * 4;  // room for the statistics accumulators of N <= 2048
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             want < kMaxSmemOptin ? want : kMaxSmemOptin) != cudaSuccess)
      return DTF_ERR_CUDA;
    attr_set[dev & 63] = true;
  }
  int smem, grid;
  persistent_plan<BN, STAGES>(p.M, p.N, splits, p.stats != nullptr, &smem, &grid);
  if (smem > kMaxSmemOptin) return DTF_ERR_TOO_LARGE;
  const int tiles_n = (p.N + BN - 1) / BN, tiles_m = (p.M + BLOCK_M - 1) / BLOCK_M;
  kern<<<grid, PersistCfg<BN>::THREADS, smem, stream>>>(ta, tb, p, tiles_m, tiles_n, splits);
  return cudaGetLastError() == cudaSuccess ? DTF_OK : DTF_ERR_CUDA;
}

template <int BN, int STAGES, int AMODE>
__global__ void __launch_bounds__(NUM_THREADS)
igemm_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
             const IgemmParams p) {
  AgSel ag;
  ag.mode = 0;
  igemm_body<BN, STAGES, AMODE>(tma_a, &tma_b, ag, p);
}

template <int BN, int STAGES>
__global__ void __launch_bounds__(NUM_THREADS)
igemm_ag_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ PeerMaps pm,
                const __grid_constant__ AgSel ag, const IgemmParams p) {
  igemm_body<BN, STAGES, A_TMA_K>(tma_a, pm.m, ag, p);
}

// --------------------------------------------------------------------------- SIMT fallback (tiny / unaligned)
// C[M,N] = act(A*B + bias); a_t: A stored [K,M]; b_t: B stored [N,K]. One thread per output element.
__global__ void gemm_simt_kernel(const __nv_bfloat16* A, const __nv_bfloat16* B, void* C, const float* bias, int M,
                                 int N, int K, int lda, int ldb, int ldc, int a_t, int b_t, int relu, int out_fp32,
                                 int accumulate, const __nv_bfloat16* addend = nullptr) {
  const int n = blockIdx.y * blockDim.x + threadIdx.x;  // rows on grid.x (no 65535 limit), column blocks on grid.y
  const int m = blockIdx.x * blockDim.y + threadIdx.y;
  if (m >= M || n >= N) return;
  float acc = 0.f;
  for (int k = 0; k < K; ++k) {
    const float a = __bfloat162float(a_t ? A[(size_t)k * lda + m] : A[(size_t)m * lda + k]);
    const float b = __bfloat162float(b_t ? B[(size_t)n * ldb + k] : B[(size_t)k * ldb + n]);
    acc = fmaf(a, b, acc);
  }
  if (bias) acc += bias[n];
  if (addend) acc += __bfloat162float(addend[(size_t)m * ldc + n]);
  if (relu) acc = fmaxf(acc, 0.f);
  if (out_fp32) {
    float* o = reinterpret_cast<float*>(C) + (size_t)m * ldc + n;
    *o = accumulate ? (*o + acc) : acc;
  } else {
    reinterpret_cast<__nv_bfloat16*>(C)[(size_t)m * ldc + n] = __float2bfloat16(acc);
  }
}

// --------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || sym == nullptr) return nullptr;
  fn = reinterpret_cast<EncodeTiledFn>(sym);
  return fn;
}

// 2D bf16 row-major [rows, cols] with leading dimension ld (elements); box = [box_rows, 64 cols], 128B swizzle.
static int make_map_2d(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld,
                       uint32_t box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return DTF_ERR_DRIVER;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {64, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? DTF_OK : DTF_ERR_TENSORMAP;
}

template <int BN, int STAGES, int AMODE>
static int launch_igemm(const CUtensorMap& ta, const CUtensorMap& tb, const IgemmParams& p, int splits,
                        cudaStream_t stream) {
  using L = SmemLayout<BN, STAGES>;
  static bool attr_set[64] = {};  // per (template instance, device)
  auto kern = igemm_kernel<BN, STAGES, AMODE>;
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL);
    if (e != cudaSuccess) return DTF_ERR_CUDA;
    attr_set[dev & 63] = true;
  }
  const long long tiles = (long long)((p.N + BN - 1) / BN) * ((p.M + BLOCK_M - 1) / BLOCK_M);
  if (tiles > 0x7fffffffLL) return DTF_ERR_TOO_LARGE;
  dim3 grid((unsigned)tiles, 1, splits);
  kern<<<grid, NUM_THREADS, L::TOTAL, stream>>>(ta, tb, p);
  return cudaGetLastError() == cudaSuccess ? DTF_OK : DTF_ERR_CUDA;
}

static int g_use_persistent = 1;
static int g_wide_tiles = 1;

template <int AMODE>
static int dispatch_bn(const CUtensorMap& ta, const CUtensorMap& tb, const IgemmParams& p, int splits,
                       cudaStream_t stream) {
  // few k-blocks per tile: a deep ring is useless, so use 2 stages and let more CTAs share an SM (their
  // prologues / epilogues overlap with each other's main loops)
  const bool shallow = p.kb_per_split <= 2;
  if constexpr (AMODE == A_TMA_K || AMODE == A_TMA_MN) {
    if (g_use_persistent) {
      if (p.N <= 64) return launch_persistent<64, 4, AMODE>(ta, tb, p, splits, stream);
      // 128 x 256 tiles halve the A re-reads and make the tile MMA-bound instead of smem-bound; use them when
      // there are still enough tiles to fill the machine
      const long long tiles256 = (long long)((p.M + BLOCK_M - 1) / BLOCK_M) * ((p.N + 255) / 256) * splits;
      if (g_wide_tiles && p.N >= 256 && tiles256 >= 148) return launch_persistent<256, 3, AMODE>(ta, tb, p, splits, stream);
      return launch_persistent<128, 3, AMODE>(ta, tb, p, splits, stream);
    }
  }
  if (p.N <= 64) {
    if (shallow) return launch_igemm<64, 2, AMODE>(ta, tb, p, splits, stream);
    return launch_igemm<64, 4, AMODE>(ta, tb, p, splits, stream);
  }
  if (shallow) return launch_igemm<128, 2, AMODE>(ta, tb, p, splits, stream);
  return launch_igemm<128, 3, AMODE>(ta, tb, p, splits, stream);
}

static int pick_splits(int M, int N, int total_kb, int want_split) {
  if (!want_split) return 1;
  const int tiles = ((M + BLOCK_M - 1) / BLOCK_M) * ((N + 127) / 128);
  int s = (2 * 148 + tiles - 1) / tiles;  // aim for ~2 CTAs per SM
  s = max(1, min(s, total_kb));
  return s;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace dtf

using namespace dtf;

// 1 (default): dense TMA-fed GEMMs use the persistent, TMEM-double-buffered kernel; 0: one tile per CTA.
extern "C" int dtf_set_persistent_gemm(int on) {
  const int prev = g_use_persistent | (g_wide_tiles << 1);
  g_use_persistent = on & 1;
  g_wide_tiles = (on >> 1) & 1;
  return prev;
}

// flags
//   bit0 relu, bit1 out_fp32, bit2 atomic accumulate into fp32 out (split-K allowed), bit3 A is stored [K,M]
extern "C" int dtf_gemm_bf16_ex(const void* A, const void* B, void* C, const float* bias, const void* addend,
                                float* stats, int M, int N, int K, int lda, int ldb, int ldc, int flags, void* stream_);
extern "C" int dtf_gemm_bf16(const void* A, const void* B, void* C, const float* bias, int M, int N, int K, int lda,
                             int ldb, int ldc, int flags, void* stream_) {
  return dtf_gemm_bf16_ex(A, B, C, bias, nullptr, nullptr, M, N, K, lda, ldb, ldc, flags, stream_);
}
// Which persistent instantiation dispatch_bn picks for a dense problem (mirrors it): 64, 128 or 256.
static int persistent_bn(int M, int N, int splits) {
  if (N <= 64) return 64;
  const long long tiles256 = (long long)((M + BLOCK_M - 1) / BLOCK_M) * ((N + 255) / 256) * splits;
  return (g_wide_tiles && N >= 256 && tiles256 >= 148) ? 256 : 128;
}
// Rows of the batch-norm statistics buffer [rows][2][N] that dtf_gemm_bf16_ex fills for this problem when given a
// `stats` pointer, or 0 when the epilogue cannot produce them (then run dtf_bn_stats on the output). The persistent
// kernel keeps per-warp column sums / sums of squares of the stored bf16 values in shared memory across its tiles and
// writes one partial row per CTA (`grid` rows) for dtf_bn_finalize_rows.
extern "C" int dtf_gemm_stats_rows(const void* A, const void* B, const void* C, int M, int N, int K, int lda, int ldb,
                                   int ldc) {
  (void)K;
  if (!g_use_persistent || M <= 0 || N <= 0 || N > 2048 || (N % 8) || (ldc % 8)) return 0;
  if (!aligned16(A) || !aligned16(B) || !aligned16(C) || (lda % 8) || (ldb % 8) || (M * (long long)N < 4096)) return 0;
  int smem = 0, grid = 0;
  switch (persistent_bn(M, N, 1)) {
    case 64: persistent_plan<64, 4>(M, N, 1, true, &smem, &grid); break;
    case 128: persistent_plan<128, 3>(M, N, 1, true, &smem, &grid); break;
    default: persistent_plan<256, 3>(M, N, 1, true, &smem, &grid); break;
  }
  return smem <= kMaxSmemOptin ? grid : 0;
}
// Same with `addend` (bf16 [M, N], row stride ldc, may be NULL): C = act(A*B + bias + addend); bf16 outputs only.
// `stats`: NULL, or a buffer of dtf_gemm_stats_rows(...) x 2 x N floats that receives per-warp partial column sums and
// sums of squares of the output (plain bf16 outputs without addend / ReLU / split-K only).
extern "C" int dtf_gemm_bf16_ex(const void* A, const void* B, void* C, const float* bias, const void* addend,
                                float* stats, int M, int N, int K, int lda, int ldb, int ldc, int flags, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  const int relu = flags & 1, out_fp32 = (flags >> 1) & 1, atomic = (flags >> 2) & 1, a_t = (flags >> 3) & 1;
  if (M <= 0 || N <= 0 || K <= 0) return DTF_ERR_ARG;
  if (stats != nullptr && (relu || out_fp32 || atomic || a_t || addend != nullptr ||
                           dtf_gemm_stats_rows(A, B, C, M, N, K, lda, ldb, ldc) == 0))
    return DTF_ERR_UNSUPPORTED;
  if (addend != nullptr && (out_fp32 || atomic)) return DTF_ERR_UNSUPPORTED;
  const bool tc_ok = aligned16(A) && aligned16(B) && (lda % 8 == 0) && (ldb % 8 == 0) && (M * (long long)N >= 4096);
  if (!tc_ok) {
    dim3 block(32, 8), grid((M + 7) / 8, (N + 31) / 32);
    gemm_simt_kernel<<<grid, block, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(A),
                                                 reinterpret_cast<const __nv_bfloat16*>(B), C, bias, M, N, K, lda, ldb,
                                                 ldc, a_t, 0, relu, out_fp32 | atomic, atomic,
                                                 reinterpret_cast<const __nv_bfloat16*>(addend));
    return cudaGetLastError() == cudaSuccess ? DTF_OK : DTF_ERR_CUDA;
  }
  CUtensorMap ta, tb;
  int rc;
  if (a_t) rc = make_map_2d(&ta, A, K, M, lda, 64);
  else rc = make_map_2d(&ta, A, M, K, lda, 128);
  if (rc) return rc;
  rc = make_map_2d(&tb, B, K, N, ldb, 64);
  if (rc) return rc;
  IgemmParams p{};
  p.M = M; p.N = N; p.K = K;
  p.out = C; p.ldc = ldc; p.bias = bias; p.relu = relu; p.out_fp32 = out_fp32; p.atomic = atomic;
  p.addend = reinterpret_cast<const __nv_bfloat16*>(addend);
  p.stats = stats;
  const int total_kb = (K + BLOCK_K - 1) / BLOCK_K;
  const int splits = pick_splits(M, N, total_kb, atomic);
  p.kb_per_split = (total_kb + splits - 1) / splits;
  const int eff_splits = (total_kb + p.kb_per_split - 1) / p.kb_per_split;
  if (a_t) return dispatch_bn<A_TMA_MN>(ta, tb, p, eff_splits, stream);
  return dispatch_bn<A_TMA_K>(ta, tb, p, eff_splits, stream);
}

// Implicit-GEMM convolution family on NHWC bf16 tensors. `geo`: 0 = fprop (gather from x), 1 = dgrad / transposed
// conv (gather from dy). W2d is the row-major [KH*KW*Csrc, Nout] bf16 filter matrix (HWIO for fprop; per-tap
// transposed HWOI for dgrad). Output [Nimg*PH*PW, Nout] row-major (= NHWC).
extern "C" int dtf_conv_igemm_bf16(const void* src, const void* W2d, void* out, const float* bias, int n_img, int H,
                                   int W, int C, int PH, int PW, int KH, int KW, int n_out, int stride, int pad_h,
                                   int pad_w, int geo, int flags, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  if ((C % 8) != 0 || (n_out % 8) != 0 || !aligned16(src) || !aligned16(W2d)) return DTF_ERR_ALIGN;
  CUtensorMap ta{}, tb;
  IgemmParams p{};
  p.M = n_img * PH * PW; p.N = n_out; p.K = KH * KW * C;
  int rc = make_map_2d(&tb, W2d, p.K, n_out, n_out, 64);
  if (rc) return rc;
  p.src = reinterpret_cast<const __nv_bfloat16*>(src);
  p.H = H; p.W = W; p.C = C; p.PH = PH; p.PW = PW; p.KH = KH; p.KW = KW;
  p.stride = stride; p.pad_h = pad_h; p.pad_w = pad_w; p.geo = geo;
  p.out = out; p.ldc = n_out; p.bias = bias; p.relu = flags & 1; p.out_fp32 = (flags >> 1) & 1; p.atomic = 0;
  p.kb_per_split = (p.K + BLOCK_K - 1) / BLOCK_K;
  return dispatch_bn<A_GATHER_K>(ta, tb, p, 1, stream);
}

// Filter gradient: dW[KH*KW*C, n_out] (fp32, accumulated with red.add, split over pixels) =
//   im2col(x)^T [KH*KW*C, P] * dy [P, n_out],  P = n_img*PH*PW output pixels (fprop geometry).
extern "C" int dtf_conv_wgrad_bf16(const void* x, const void* dy, float* dw, int n_img, int H, int W, int C, int PH,
                                   int PW, int KH, int KW, int n_out, int stride, int pad_h, int pad_w,
                                   void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  if ((C % 8) != 0 || (n_out % 8) != 0 || !aligned16(x) || !aligned16(dy)) return DTF_ERR_ALIGN;
  CUtensorMap ta{}, tb;
  IgemmParams p{};
  p.M = KH * KW * C; p.N = n_out; p.K = n_img * PH * PW;
  int rc = make_map_2d(&tb, dy, p.K, n_out, n_out, 64);
  if (rc) return rc;
  p.src = reinterpret_cast<const __nv_bfloat16*>(x);
  p.H = H; p.W = W; p.C = C; p.PH = PH; p.PW = PW; p.KH = KH; p.KW = KW;
  p.stride = stride; p.pad_h = pad_h; p.pad_w = pad_w; p.geo = GEO_FPROP;
  p.out = dw; p.ldc = n_out; p.bias = nullptr; p.relu = 0; p.out_fp32 = 1; p.atomic = 1;
  const int total_kb = (p.K + BLOCK_K - 1) / BLOCK_K;
  const int splits = pick_splits(p.M, p.N, total_kb, 1);
  p.kb_per_split = (total_kb + splits - 1) / splits;
  const int eff_splits = (total_kb + p.kb_per_split - 1) / p.kb_per_split;
  return dispatch_bn<A_GATHER_MN>(ta, tb, p, eff_splits, stream);
}

// K-AG-GEMM: C[M,N] = act(A[M,K] * B[K,N] + bias) where B lives SHARDED in the symmetric heaps of a rank group:
// peer_b[r] is rank r's mapping of the matrix (same offset in every heap), `bounds` (world+1 arena element offsets)
// gives each rank's owned range and b_arena_off the arena offset of B's first element. shard_by_n = 0: the matrix is
// the row-major [K,N] weight itself (owner changes along K, boundaries multiples of 64 rows); shard_by_n = 1: the
// matrix is a transposed shadow [K,N] = W^T of a weight W[N,K] whose owners are row ranges of W, i.e. column ranges
// here (boundaries multiples of the N tile). No gather pass, no local copy: tiles come over NVLink by TMA.
extern "C" int dtf_gemm_bf16_ag(const void* A, const void* const* peer_b, void* C, const float* bias, int M, int N,
                                int K, int lda, int ldb, int ldc, int flags, int world, const long long* bounds,
                                long long b_arena_off, int shard_by_n, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  if (world < 1 || world > 8 || M <= 0 || N <= 0 || K <= 0) return DTF_ERR_ARG;
  if (!aligned16(A) || (lda % 8) || (ldb % 8)) return DTF_ERR_ALIGN;
  CUtensorMap ta;
  PeerMaps pm;
  int rc = make_map_2d(&ta, A, M, K, lda, 128);
  if (rc) return rc;
  for (int r = 0; r < world; ++r) {
    if (!aligned16(peer_b[r])) return DTF_ERR_ALIGN;
    rc = make_map_2d(&pm.m[r], peer_b[r], K, N, ldb, 64);
    if (rc) return rc;
  }
  for (int r = world; r < 8; ++r) pm.m[r] = pm.m[0];
  const bool small_n = N <= 64;
  AgSel ag;
  ag.mode = shard_by_n ? 2 : 1;
  ag.world = world;
  ag.b_off = b_arena_off;
  // mode 1: one k-block = 64 rows of B = 64*ldb elements of W. mode 2: one n-tile = BN columns of W^T = BN rows of W,
  // and a row of W has K elements (W is [N, K] row-major with leading dimension K).
  ag.unit = shard_by_n ? (long long)(small_n ? 64 : 128) * K : (long long)BLOCK_K * ldb;
  for (int r = 0; r <= 8; ++r) ag.bounds[r] = bounds[r <= world ? r : world];
  IgemmParams p{};
  p.M = M; p.N = N; p.K = K;
  p.out = C; p.ldc = ldc; p.bias = bias; p.relu = flags & 1; p.out_fp32 = (flags >> 1) & 1; p.atomic = 0;
  p.kb_per_split = (K + BLOCK_K - 1) / BLOCK_K;
  int dev = 0;
  cudaGetDevice(&dev);
  static bool attr64[64] = {}, attr128[64] = {};
  if (small_n) {
    using L = SmemLayout<64, 4>;
    auto kern = igemm_ag_kernel<64, 4>;
    if (!attr64[dev & 63]) {
      if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL) != cudaSuccess) return DTF_ERR_CUDA;
      attr64[dev & 63] = true;
    }
    // N <= 64: one column tile; tiles are linearised in grid.x (see igemm_body)
    kern<<<dim3(((N + 63) / 64) * ((M + BLOCK_M - 1) / BLOCK_M), 1, 1), NUM_THREADS, L::TOTAL, stream>>>(ta, pm, ag, p);
  } else {
    using L = SmemLayout<128, 3>;
    auto kern = igemm_ag_kernel<128, 3>;
    if (!attr128[dev & 63]) {
      if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL) != cudaSuccess) return DTF_ERR_CUDA;
      attr128[dev & 63] = true;
    }
    kern<<<dim3(((N + 127) / 128) * ((M + BLOCK_M - 1) / BLOCK_M), 1, 1), NUM_THREADS, L::TOTAL, stream>>>(ta, pm, ag, p);
  }
  return cudaGetLastError() == cudaSuccess ? DTF_OK : DTF_ERR_CUDA;
}
