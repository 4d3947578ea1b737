// Peer-memory collectives over NVLink 5 / NVSwitch, fused with the optimizer (no NCCL on these paths).
//
// Every rank owns one "symmetric heap" (same layout on every rank, peers mapped with CUDA IPC). Kernels address a
// peer's copy of a buffer as peer[r] + offset. Cross-rank ordering uses monotonically increasing flags written with
// st.release.sys into the peer's flag page and polled with ld.acquire.sys (bounded spins: a lost peer traps instead
// of hanging the box).
//
//   K-AR      dtf_comm_allreduce_optim   one-shot: every rank reads all peers' gradient bucket, sums in rank order
//                                        (bitwise identical on all ranks), x1/N, optimizer update, bf16 shadow.
//   K-RS-OPT  dtf_comm_rs_optim_ag       two-shot / parameter-server: each owner reduces its slice from all peers,
//                                        applies the optimizer to its fp32 master shard and pushes the updated
//                                        bf16 weights (and optionally fp32 master) into every peer (all-gather by stores).
//   K-ASYNC   dtf_comm_async_push/apply  workers store gradients into the owner's mailbox with a version stamp; the
//                                        owner applies whatever has arrived (stale pushes allowed, no waiting).
//   barrier   dtf_comm_barrier           device-side token barrier (sync token queue / done queue / init handshake).
//
// Reference behaviour being replaced: Tower.average_gradients (/root/reference/distribute_tower.py:78-114),
// SyncReplicasOptimizer push/apply + token queue (/root/reference/distribute_train.py:146-160), replica_device_setter
// variable placement (:109-110), done queue (:43-46,86-90,202-204).
#include <cuda_runtime.h>
#include <stdlib.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <string.h>
#include "dtf_internal.h"
#include "optim.cuh"

namespace dtf {

constexpr int MAX_RANKS = 8;
constexpr int MAX_COMM_BLOCKS = 128;
// flag page layout (uint32 words), at heap offset 0 on every rank:
//   [0, MAX_COMM_BLOCKS*MAX_RANKS)            sig[block][src_rank]      written by peers
//   [SIG_WORDS, SIG_WORDS+MAX_COMM_BLOCKS)    ctr[block]                local barrier epoch per block
//   [MAIL_FLAG_OFF, +MAX_RANKS)               mail_flag[src_rank]       async push versions (written by peers)
//   [APPLIED_OFF, +MAX_RANKS)                 applied[src_rank]         local
//   [DONE_CTR_OFF]                            local block-completion counter
//   [HEARTBEAT_OFF, +MAX_RANKS)               heartbeat[src_rank]       written by peers
constexpr int SIG_WORDS = MAX_COMM_BLOCKS * MAX_RANKS;
constexpr int CTR_OFF = SIG_WORDS;
constexpr int MAIL_FLAG_OFF = CTR_OFF + MAX_COMM_BLOCKS;
constexpr int APPLIED_OFF = MAIL_FLAG_OFF + MAX_RANKS;
constexpr int DONE_CTR_OFF = APPLIED_OFF + MAX_RANKS;
constexpr int HEARTBEAT_OFF = DONE_CTR_OFF + 8;
constexpr int FLAG_PAGE_BYTES = 8192;
static_assert((HEARTBEAT_OFF + MAX_RANKS) * 4 <= FLAG_PAGE_BYTES, "flag page too small");

struct CommCtx {
  int rank, world;
  uint8_t* peer[MAX_RANKS];  // heap base of every rank as mapped in this process (peer[rank] is local)
  uint8_t* mc;               // NVLS multicast mapping of the heap (nullptr when not available)
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_cg_f4(const float4* p) {
  float4 v;
  asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];\n" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

#ifndef DTF_COMM_SPIN_LIMIT
#define DTF_COMM_SPIN_LIMIT (1ull << 31)
#endif

// Barrier among the same-index blocks of all ranks. Also orders memory: everything written (by this GPU) before the
// barrier is visible to peers after it.
__device__ void block_barrier(const CommCtx& c) {
  __syncthreads();
  uint32_t* flags = reinterpret_cast<uint32_t*>(c.peer[c.rank]);
  __shared__ uint32_t epoch_s;
  if (threadIdx.x == 0) {
    __threadfence_system();
    const uint32_t e = flags[CTR_OFF + blockIdx.x] + 1;
    flags[CTR_OFF + blockIdx.x] = e;
    epoch_s = e;
  }
  __syncthreads();
  const uint32_t e = epoch_s;
  if (threadIdx.x < c.world) {
    const int r = threadIdx.x;
    uint32_t* remote = reinterpret_cast<uint32_t*>(c.peer[r]) + blockIdx.x * MAX_RANKS + c.rank;
    st_release_sys(remote, e);
    const uint32_t* mine = flags + blockIdx.x * MAX_RANKS + r;
    unsigned long long spins = 0;
    while ((int32_t)(ld_acquire_sys(mine) - e) < 0) {
      if (++spins > DTF_COMM_SPIN_LIMIT) __trap();
    }
  }
  __syncthreads();
}

__device__ __forceinline__ uint2 pack4_bf16(float4 v) {
  __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
  uint2 u;
  u.x = *reinterpret_cast<uint32_t*>(&a);
  u.y = *reinterpret_cast<uint32_t*>(&b);
  return u;
}

__device__ __forceinline__ float4 reduce_peers(const CommCtx& c, size_t grad_off, long long i4) {
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int r = 0; r < MAX_RANKS; ++r) {
    if (r < c.world) {
      const float4 v = ld_cg_f4(reinterpret_cast<const float4*>(c.peer[r] + grad_off) + i4);
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
  }
  return s;
}

// Remote loads cost ~2 us each: keep U independent 16-byte loads per peer in flight per thread. q0 is the first
// float4 index of this thread, `stride` the distance between its U elements, n4 the range length (indices relative
// to base4). Sums are formed in rank order, so every rank computes bit-identical results.
constexpr int COMM_U = 4;
__device__ __forceinline__ void reduce_peers_u(const CommCtx& c, size_t grad_off, long long base4, long long q0,
                                               long long stride, long long n4, float4 (&s)[COMM_U]) {
#pragma unroll
  for (int u = 0; u < COMM_U; ++u) s[u] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int r = 0; r < MAX_RANKS; ++r) {
    if (r < c.world) {
      const float4* g = reinterpret_cast<const float4*>(c.peer[r] + grad_off) + base4;
      float4 v[COMM_U];
#pragma unroll
      for (int u = 0; u < COMM_U; ++u) {
        const long long q = q0 + u * stride;
        v[u] = (q < n4) ? ld_cg_f4(g + q) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < COMM_U; ++u) {
        s[u].x += v[u].x; s[u].y += v[u].y; s[u].z += v[u].z; s[u].w += v[u].w;
      }
    }
  }
}

__device__ __forceinline__ float4 optim4(const OptimArgs& o, float* master, float* m, float* v, long long i,
                                         float4 g) {
  float4 p = *reinterpret_cast<float4*>(master + i);
  p.x = optim_update(o.kind, p.x, g.x * o.gscale, m, v, i + 0, o.lr, o.b1, o.b2, o.eps, o.wd, o.lr_t);
  p.y = optim_update(o.kind, p.y, g.y * o.gscale, m, v, i + 1, o.lr, o.b1, o.b2, o.eps, o.wd, o.lr_t);
  p.z = optim_update(o.kind, p.z, g.z * o.gscale, m, v, i + 2, o.lr, o.b1, o.b2, o.eps, o.wd, o.lr_t);
  p.w = optim_update(o.kind, p.w, g.w * o.gscale, m, v, i + 3, o.lr, o.b1, o.b2, o.eps, o.wd, o.lr_t);
  *reinterpret_cast<float4*>(master + i) = p;
  return p;
}

// ---- K-AR: one-shot allreduce (+ optimizer). [start, start+count) in elements, multiples of 4.
// If master == nullptr the averaged gradient is written to avg_out (plain allreduce-mean).
__global__ void __launch_bounds__(512)
allreduce_optim_kernel(CommCtx c, size_t grad_off, long long start, long long count, OptimArgs o, float* master,
                       float* m, float* v, __nv_bfloat16* shadow, float* avg_out) {
  block_barrier(c);
  const long long n4 = count >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long q0 = blockIdx.x * (long long)blockDim.x + threadIdx.x; q0 < n4; q0 += stride * COMM_U) {
    float4 gs[COMM_U];
    reduce_peers_u(c, grad_off, start >> 2, q0, stride, n4, gs);
#pragma unroll
    for (int u = 0; u < COMM_U; ++u) {
      const long long q = q0 + u * stride;
      if (q >= n4) continue;
      const long long i = start + (q << 2);
      const float4 g = gs[u];
      if (master != nullptr) {
        const float4 p = optim4(o, master, m, v, i, g);
        if (shadow) *reinterpret_cast<uint2*>(shadow + i) = pack4_bf16(p);
      } else {
        *reinterpret_cast<float4*>(avg_out + i) =
            make_float4(g.x * o.gscale, g.y * o.gscale, g.z * o.gscale, g.w * o.gscale);
      }
    }
  }
  block_barrier(c);  // nobody may overwrite its gradient bucket while a peer still reads it
}

// ---- K-RS-OPT (+ all-gather by peer stores). bounds[r]..bounds[r+1] is rank r's owned element range (multiples
// of 4). master/m/v are local (owner-side state); shadow_off / master_off are heap offsets of the replicated copies
// that every peer receives (master_off may be SIZE_MAX to skip the fp32 broadcast).
struct Bounds {
  long long b[MAX_RANKS + 1];
};
// element ranges whose bf16 weights are NOT pushed to the peers (consumers pull them with K-AG-GEMM instead)
struct NoPush {
  int n;
  long long lo[8], hi[8];
};
__device__ __forceinline__ bool in_nopush(const NoPush& np, long long i) {
  bool hit = false;
#pragma unroll
  for (int k = 0; k < 8; ++k)
    if (k < np.n && i >= np.lo[k] && i < np.hi[k]) hit = true;
  return hit;
}
__global__ void __launch_bounds__(512)
rs_optim_ag_kernel(CommCtx c, size_t grad_off, Bounds bounds, OptimArgs o, float* master, float* m, float* v,
                   size_t shadow_off, size_t master_off, NoPush np) {
  block_barrier(c);
  const long long start = bounds.b[c.rank];
  const long long n4 = (bounds.b[c.rank + 1] - start) >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long q0 = blockIdx.x * (long long)blockDim.x + threadIdx.x; q0 < n4; q0 += stride * COMM_U) {
    float4 gs[COMM_U];
    reduce_peers_u(c, grad_off, start >> 2, q0, stride, n4, gs);
#pragma unroll
    for (int u = 0; u < COMM_U; ++u) {
      const long long q = q0 + u * stride;
      if (q >= n4) continue;
      const long long i = start + (q << 2);
      const float4 p = optim4(o, master, m, v, i, gs[u]);
      const uint2 pb = pack4_bf16(p);
      const bool local_only = in_nopush(np, i);
#pragma unroll
      for (int r = 0; r < MAX_RANKS; ++r) {
        if (r < c.world && (!local_only || r == c.rank)) {
          *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(c.peer[r] + shadow_off) + i) = pb;
          if (master_off != (size_t)-1 && r != c.rank)
            *reinterpret_cast<float4*>(reinterpret_cast<float*>(c.peer[r] + master_off) + i) = p;
        }
      }
    }
  }
  block_barrier(c);  // all shards have landed everywhere; gradients may be overwritten
}

// ---- K-RS-OPT over NVLS: the owner reduces its slice with ONE multimem.ld_reduce per 16 bytes (the NVSwitch pulls the
// eight copies and adds them in the switch), applies the optimizer, and publishes the new bf16 weights with ONE
// multimem.st (the switch replicates the store into every GPU's shadow arena). Wire bytes per GPU drop from
// 2 (N-1)/N S to about S/N in each direction.
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const void* mc_addr) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];\n"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc_addr)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_b64(void* mc_addr, uint2 v) {
  asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};\n" ::"l"(mc_addr), "f"(__uint_as_float(v.x)),
               "f"(__uint_as_float(v.y))
               : "memory");
}
__global__ void __launch_bounds__(512)
rs_optim_ag_nvls_kernel(CommCtx c, size_t grad_off, Bounds bounds, OptimArgs o, float* master, float* m, float* v,
                        size_t shadow_off, NoPush np) {
  block_barrier(c);
  const long long start = bounds.b[c.rank];
  const long long n4 = (bounds.b[c.rank + 1] - start) >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const float4* gmc = reinterpret_cast<const float4*>(c.mc + grad_off) + (start >> 2);
  for (long long q0 = blockIdx.x * (long long)blockDim.x + threadIdx.x; q0 < n4; q0 += stride * COMM_U) {
    float4 gs[COMM_U];
#pragma unroll
    for (int u = 0; u < COMM_U; ++u) {
      const long long q = q0 + u * stride;
      gs[u] = (q < n4) ? multimem_ld_reduce_f4(gmc + q) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < COMM_U; ++u) {
      const long long q = q0 + u * stride;
      if (q >= n4) continue;
      const long long i = start + (q << 2);
      const float4 p = optim4(o, master, m, v, i, gs[u]);
      const uint2 pb = pack4_bf16(p);
      if (in_nopush(np, i))
        *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(c.peer[c.rank] + shadow_off) + i) = pb;
      else
        multimem_st_b64(reinterpret_cast<__nv_bfloat16*>(c.mc + shadow_off) + i, pb);
    }
  }
  block_barrier(c);
}

// ---- all-gather of owned slices by peer stores (16-byte units): used for checkpoint gathers and weight pulls.
__global__ void __launch_bounds__(512)
allgather_slices_kernel(CommCtx c, size_t buf_off, Bounds bounds16) {
  block_barrier(c);
  const long long start = bounds16.b[c.rank];
  const long long n = bounds16.b[c.rank + 1] - start;
  const uint4* src = reinterpret_cast<const uint4*>(c.peer[c.rank] + buf_off) + start;
  for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < n; q += (long long)gridDim.x * blockDim.x) {
    const uint4 val = src[q];
#pragma unroll
    for (int r = 0; r < MAX_RANKS; ++r)
      if (r < c.world && r != c.rank) reinterpret_cast<uint4*>(c.peer[r] + buf_off)[start + q] = val;
  }
  block_barrier(c);
}

// ---- K-ASYNC-PUSH: copy my gradient slice for every owner into that owner's mailbox[me], then publish `version`.
// mailbox layout on each rank: [world][slice_cap] floats at mailbox_off.
__global__ void __launch_bounds__(512)
async_push_kernel(CommCtx c, size_t grad_off, size_t mailbox_off, long long slice_cap, Bounds bounds,
                  uint32_t version) {
  const float4* grad = reinterpret_cast<const float4*>(c.peer[c.rank] + grad_off);
  for (int o = 0; o < c.world; ++o) {
    const long long s4 = bounds.b[o] >> 2, n4 = (bounds.b[o + 1] - bounds.b[o]) >> 2;
    float4* dst = reinterpret_cast<float4*>(c.peer[o] + mailbox_off) + (((long long)c.rank * slice_cap) >> 2);
    for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < n4; q += (long long)gridDim.x * blockDim.x)
      dst[q] = grad[s4 + q];
  }
  // publish after every block has finished its stores
  __threadfence_system();
  __syncthreads();
  __shared__ bool last;
  uint32_t* flags = reinterpret_cast<uint32_t*>(c.peer[c.rank]);
  if (threadIdx.x == 0) {
    const uint32_t prev = atomicAdd(flags + DONE_CTR_OFF, 1u);
    last = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (last) {
    __threadfence_system();
    if (threadIdx.x < c.world)
      st_release_sys(reinterpret_cast<uint32_t*>(c.peer[threadIdx.x]) + MAIL_FLAG_OFF + c.rank, version);
    if (threadIdx.x == 0) flags[DONE_CTR_OFF] = 0;
  }
}
// ---- owner-side apply: for every worker whose mailbox version is newer than the last applied one, apply its
// (possibly stale) gradient to the owned shard and push the new bf16 weights to every rank. Never waits.
// snap[w] (8 words, caller-provided) receives the version applied for worker w (0 = nothing new).
// A worker may overwrite its mailbox while the owner reads it: like use_locking=False in the reference, a torn
// (mixed-version) gradient is accepted.
__global__ void async_snapshot_kernel(CommCtx c, uint32_t* snap) {
  // one consistent view of "what has arrived" for the whole apply grid
  uint32_t* flags = reinterpret_cast<uint32_t*>(c.peer[c.rank]);
  const int w = threadIdx.x;
  if (w < MAX_RANKS) {
    uint32_t s = 0;
    if (w < c.world) {
      const uint32_t ver = ld_acquire_sys(flags + MAIL_FLAG_OFF + w);
      if (ver != flags[APPLIED_OFF + w]) s = ver;
    }
    snap[w] = s;
  }
}
__global__ void __launch_bounds__(512)
async_apply_kernel(CommCtx c, size_t mailbox_off, long long slice_cap, Bounds bounds, OptimArgs o, float* master,
                   float* m, float* v, size_t shadow_off, const uint32_t* snap) {
  const long long start = bounds.b[c.rank];
  const long long n4 = (bounds.b[c.rank + 1] - start) >> 2;
  for (int w = 0; w < c.world; ++w) {
    if (snap[w] == 0) continue;
    const float4* mail = reinterpret_cast<const float4*>(c.peer[c.rank] + mailbox_off) + (((long long)w * slice_cap) >> 2);
    for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < n4; q += (long long)gridDim.x * blockDim.x) {
      const long long i = start + (q << 2);
      const float4 g = ld_cg_f4(mail + q);
      const float4 p = optim4(o, master, m, v, i, g);
      const uint2 pb = pack4_bf16(p);
#pragma unroll
      for (int r = 0; r < MAX_RANKS; ++r)
        if (r < c.world) *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(c.peer[r] + shadow_off) + i) = pb;
    }
  }
}
__global__ void async_commit_kernel(CommCtx c, const uint32_t* snap) {
  uint32_t* flags = reinterpret_cast<uint32_t*>(c.peer[c.rank]);
  const int w = threadIdx.x;
  if (w < c.world && snap[w] != 0) flags[APPLIED_OFF + w] = snap[w];
}

__global__ void barrier_kernel(CommCtx c) { block_barrier(c); }

// heartbeat: bump my counter in every peer's flag page; read what peers wrote into mine
__global__ void heartbeat_kernel(CommCtx c, uint32_t beat, uint32_t* seen_out) {
  const int r = threadIdx.x;
  if (r < c.world) {
    st_release_sys(reinterpret_cast<uint32_t*>(c.peer[r]) + HEARTBEAT_OFF + c.rank, beat);
    seen_out[r] = ld_acquire_sys(reinterpret_cast<uint32_t*>(c.peer[c.rank]) + HEARTBEAT_OFF + r);
  }
}

}  // namespace dtf

using namespace dtf;

struct DtfComm {
  CommCtx ctx;
  size_t heap_bytes;
  void* opened[MAX_RANKS];
};

extern "C" int dtf_comm_flag_page_bytes(void) { return FLAG_PAGE_BYTES; }
extern "C" int dtf_comm_max_blocks(void) { return MAX_COMM_BLOCKS; }

// Allocate the local symmetric heap (zeroed) and return its IPC handle (64 bytes).
extern "C" int dtf_comm_heap_alloc(size_t bytes, void** ptr_out, void* handle_out64) {
  void* p = nullptr;
  if (cudaMalloc(&p, bytes) != cudaSuccess) return DTF_ERR_CUDA;
  if (cudaMemset(p, 0, bytes) != cudaSuccess) return DTF_ERR_CUDA;
  cudaIpcMemHandle_t h;
  if (cudaIpcGetMemHandle(&h, p) != cudaSuccess) return DTF_ERR_CUDA;
  memcpy(handle_out64, &h, sizeof(h));
  *ptr_out = p;
  return cudaDeviceSynchronize() == cudaSuccess ? DTF_OK : DTF_ERR_CUDA;
}
extern "C" int dtf_comm_heap_free(void* ptr) { return cudaFree(ptr) == cudaSuccess ? DTF_OK : DTF_ERR_CUDA; }

// handles: world x 64 bytes (all ranks' IPC handles, rank order). Returns an opaque communicator.
extern "C" int dtf_comm_create(int rank, int world, void* local_heap, size_t heap_bytes, const void* handles,
                               void** comm_out) {
  if (world < 1 || world > MAX_RANKS || rank < 0 || rank >= world) return DTF_ERR_ARG;
  DtfComm* c = new DtfComm();
  memset(c, 0, sizeof(*c));
  c->ctx.rank = rank;
  c->ctx.world = world;
  c->heap_bytes = heap_bytes;
  for (int r = 0; r < world; ++r) {
    if (r == rank) {
      c->ctx.peer[r] = reinterpret_cast<uint8_t*>(local_heap);
      continue;
    }
    cudaIpcMemHandle_t h;
    memcpy(&h, reinterpret_cast<const uint8_t*>(handles) + (size_t)r * 64, sizeof(h));
    void* p = nullptr;
    if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
      delete c;
      return DTF_ERR_CUDA;
    }
    c->opened[r] = p;
    c->ctx.peer[r] = reinterpret_cast<uint8_t*>(p);
  }
  *comm_out = c;
  return DTF_OK;
}
// Single-process variant (tests / in-process multi-tower): peers are plain device pointers on peer-enabled GPUs.
extern "C" int dtf_comm_create_local(int rank, int world, void* const* heaps, size_t heap_bytes, void** comm_out) {
  if (world < 1 || world > MAX_RANKS || rank < 0 || rank >= world) return DTF_ERR_ARG;
  DtfComm* c = new DtfComm();
  memset(c, 0, sizeof(*c));
  c->ctx.rank = rank;
  c->ctx.world = world;
  c->heap_bytes = heap_bytes;
  for (int r = 0; r < world; ++r) c->ctx.peer[r] = reinterpret_cast<uint8_t*>(heaps[r]);
  *comm_out = c;
  return DTF_OK;
}
extern "C" int dtf_comm_destroy(void* comm) {
  DtfComm* c = reinterpret_cast<DtfComm*>(comm);
  for (int r = 0; r < MAX_RANKS; ++r)
    if (c->opened[r]) cudaIpcCloseMemHandle(c->opened[r]);
  delete c;
  return DTF_OK;
}
// Attach the NVLS multicast mapping of the heap (from cuMulticast* / torch symmetric memory); 0 detaches it.
extern "C" int dtf_comm_set_multicast(void* comm, void* mc_base) {
  reinterpret_cast<DtfComm*>(comm)->ctx.mc = reinterpret_cast<uint8_t*>(mc_base);
  return DTF_OK;
}
extern "C" void* dtf_comm_peer_ptr(void* comm, int r) { return reinterpret_cast<DtfComm*>(comm)->ctx.peer[r]; }

static OptimArgs make_optim(int kind, float lr, float b1, float b2, float eps, float wd, float gscale,
                            long long step) {
  OptimArgs o;
  o.kind = kind; o.lr = lr; o.b1 = b1; o.b2 = b2; o.eps = eps; o.wd = wd; o.gscale = gscale;
  o.lr_t = lr;
  if (kind == 2) o.lr_t = lr * sqrtf(1.f - powf(b2, (float)step)) / (1.f - powf(b1, (float)step));
  return o;
}
// Threads per collective CTA (DTF_COMM_THREADS, default 512). Smaller CTAs (256 / 128 threads, 64 registers) were tried
// so that collective CTAs could sit next to the persistent compute CTAs; the 2-GPU step time did not move (26.0 ms in
// every variant: the multi-GPU step is paced by the slowest GPU of the group, not by the collectives), and the
// stand-alone collective got slower, so the full-size CTAs stay.
static int comm_threads() {
  static int t = 0;
  if (t == 0) {
    const char* e = getenv("DTF_COMM_THREADS");
    t = e ? atoi(e) : 512;
    if (t < 32 || t > 512 || (t & 31)) t = 512;
  }
  return t;
}
static int comm_grid(long long n4, int want) {
  long long g = (n4 + comm_threads() - 1) / comm_threads();
  if (g < 1) g = 1;
  if (g > want) g = want;
  if (g > MAX_COMM_BLOCKS) g = MAX_COMM_BLOCKS;
  return (int)g;
}

// The grid must be identical on every rank for a given call, so it is derived from (count, blocks) only.
extern "C" int dtf_comm_allreduce_optim(void* comm, size_t grad_off, long long start, long long count, int kind,
                                        float lr, float b1, float b2, float eps, float wd, float gscale,
                                        long long step, float* master, float* m, float* v, void* shadow,
                                        float* avg_out, int blocks, void* stream) {
  DtfComm* c = reinterpret_cast<DtfComm*>(comm);
  if ((start & 3) || (count & 3) || (grad_off & 15)) return DTF_ERR_ALIGN;
  if (master == nullptr && avg_out == nullptr) return DTF_ERR_ARG;
  const OptimArgs o = make_optim(kind, lr, b1, b2, eps, wd, gscale, step);
  const int grid = comm_grid(count >> 2, blocks);
  allreduce_optim_kernel<<<grid, comm_threads(), 0, dtf_stream(stream)>>>(c->ctx, grad_off, start, count, o, master, m, v,
                                                               (__nv_bfloat16*)shadow, avg_out);
  return DTF_CHECK_LAUNCH();
}
extern "C" int dtf_comm_rs_optim_ag(void* comm, size_t grad_off, const long long* bounds, int kind, float lr,
                                    float b1, float b2, float eps, float wd, float gscale, long long step,
                                    float* master, float* m, float* v, size_t shadow_off, size_t master_off,
                                    int n_nopush, const long long* nopush_ranges, int blocks, void* stream) {
  DtfComm* c = reinterpret_cast<DtfComm*>(comm);
  Bounds b;
  long long maxn = 0;
  for (int r = 0; r <= c->ctx.world; ++r) {
    b.b[r] = bounds[r];
    if (bounds[r] & 3) return DTF_ERR_ALIGN;
    if (r > 0 && bounds[r] - bounds[r - 1] > maxn) maxn = bounds[r] - bounds[r - 1];
  }
  if ((grad_off & 15) || (shadow_off & 15)) return DTF_ERR_ALIGN;
  const OptimArgs o = make_optim(kind, lr, b1, b2, eps, wd, gscale, step);
  const int grid = comm_grid(maxn >> 2, blocks);
  NoPush np;
  np.n = 0;
  if (n_nopush < 0 || n_nopush > 8) return DTF_ERR_ARG;
  for (int k = 0; k < n_nopush; ++k) {
    np.lo[k] = nopush_ranges[2 * k];
    np.hi[k] = nopush_ranges[2 * k + 1];
    np.n = k + 1;
  }
  if (c->ctx.mc != nullptr && master_off == (size_t)-1)
    rs_optim_ag_nvls_kernel<<<grid, comm_threads(), 0, dtf_stream(stream)>>>(c->ctx, grad_off, b, o, master, m, v, shadow_off,
                                                                  np);
  else
    rs_optim_ag_kernel<<<grid, comm_threads(), 0, dtf_stream(stream)>>>(c->ctx, grad_off, b, o, master, m, v, shadow_off,
                                                             master_off, np);
  return DTF_CHECK_LAUNCH();
}
// bounds16 in units of 16 bytes relative to buf_off
extern "C" int dtf_comm_allgather_slices(void* comm, size_t buf_off, const long long* bounds16, int blocks,
                                         void* stream) {
  DtfComm* c = reinterpret_cast<DtfComm*>(comm);
  Bounds b;
  long long maxn = 0;
  for (int r = 0; r <= c->ctx.world; ++r) {
    b.b[r] = bounds16[r];
    if (r > 0 && bounds16[r] - bounds16[r - 1] > maxn) maxn = bounds16[r] - bounds16[r - 1];
  }
  if (buf_off & 15) return DTF_ERR_ALIGN;
  allgather_slices_kernel<<<comm_grid(maxn, blocks), comm_threads(), 0, dtf_stream(stream)>>>(c->ctx, buf_off, b);
  return DTF_CHECK_LAUNCH();
}
extern "C" int dtf_comm_async_push(void* comm, size_t grad_off, size_t mailbox_off, long long slice_cap,
                                   const long long* bounds, unsigned int version, int blocks, void* stream) {
  DtfComm* c = reinterpret_cast<DtfComm*>(comm);
  Bounds b;
  for (int r = 0; r <= c->ctx.world; ++r) {
    b.b[r] = bounds[r];
    if (bounds[r] & 3) return DTF_ERR_ALIGN;
    if (r > 0 && bounds[r] - bounds[r - 1] > slice_cap) return DTF_ERR_TOO_LARGE;
  }
  if ((slice_cap & 3) || (grad_off & 15) || (mailbox_off & 15)) return DTF_ERR_ALIGN;
  int grid = blocks < 1 ? 1 : (blocks > MAX_COMM_BLOCKS ? MAX_COMM_BLOCKS : blocks);
  async_push_kernel<<<grid, comm_threads(), 0, dtf_stream(stream)>>>(c->ctx, grad_off, mailbox_off, slice_cap, b, version);
  return DTF_CHECK_LAUNCH();
}
extern "C" int dtf_comm_async_apply(void* comm, size_t mailbox_off, long long slice_cap, const long long* bounds,
                                    int kind, float lr, float b1, float b2, float eps, float wd, float gscale,
                                    long long step, float* master, float* m, float* v, size_t shadow_off,
                                    unsigned int* applied_out_dev, int blocks, void* stream) {
  DtfComm* c = reinterpret_cast<DtfComm*>(comm);
  Bounds b;
  for (int r = 0; r <= c->ctx.world; ++r) b.b[r] = bounds[r];
  const OptimArgs o = make_optim(kind, lr, b1, b2, eps, wd, gscale, step);
  int grid = blocks < 1 ? 1 : (blocks > MAX_COMM_BLOCKS ? MAX_COMM_BLOCKS : blocks);
  cudaStream_t st = dtf_stream(stream);
  async_snapshot_kernel<<<1, 32, 0, st>>>(c->ctx, applied_out_dev);
  async_apply_kernel<<<grid, comm_threads(), 0, st>>>(c->ctx, mailbox_off, slice_cap, b, o, master, m, v, shadow_off,
                                           applied_out_dev);
  async_commit_kernel<<<1, 32, 0, st>>>(c->ctx, applied_out_dev);
  return DTF_CHECK_LAUNCH();
}
extern "C" int dtf_comm_barrier(void* comm, void* stream) {
  DtfComm* c = reinterpret_cast<DtfComm*>(comm);
  barrier_kernel<<<1, 32, 0, dtf_stream(stream)>>>(c->ctx);
  return DTF_CHECK_LAUNCH();
}
extern "C" int dtf_comm_heartbeat(void* comm, unsigned int beat, unsigned int* seen_out_dev, void* stream) {
  DtfComm* c = reinterpret_cast<DtfComm*>(comm);
  heartbeat_kernel<<<1, 32, 0, dtf_stream(stream)>>>(c->ctx, beat, seen_out_dev);
  return DTF_CHECK_LAUNCH();
}
