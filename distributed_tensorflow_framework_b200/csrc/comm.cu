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
//   [TICKET_OFF]                              arrival ticket counter of this owner (peers atom.add.sys into it)
//   [ARRIVE_OFF, +MAX_RANKS)                  arrive[src_rank] = ticket + 1 of that rank's latest arrival (peers)
//   [BK_CALL_OFF]                             local call counter of backup-worker collectives (+1 by the last block)
//   [BK_CALL_OFF + 1]                         blocks of the running backup-worker collective that have finished
//   [PUSH_VER_OFF]                            local async push version (bumped in-kernel: survives graph replay)
//   [OWNER_STEP_OFF, +MAX_RANKS)              owner_step[o]: number of updates owner o has applied (written by o)
//   [BASE_STEP_OFF, +MAX_RANKS)               local snapshot of owner_step[] taken when this step's forward began
//   [MAIL_BASE_OFF, +MAX_RANKS)               mail_base[src_rank]: owner step the pushed gradient was computed at
//   [DROPPED_OFF]                             local count of async pushes dropped as too stale
//   [ERR_OFF, +4)                             local error word: code, block, peer waited for, epoch
constexpr int TICKET_OFF = HEARTBEAT_OFF + MAX_RANKS;
constexpr int ARRIVE_OFF = TICKET_OFF + 8;
constexpr int BK_CALL_OFF = ARRIVE_OFF + MAX_RANKS;
constexpr int PUSH_VER_OFF = BK_CALL_OFF + MAX_COMM_BLOCKS;
constexpr int OWNER_STEP_OFF = PUSH_VER_OFF + 8;
constexpr int BASE_STEP_OFF = OWNER_STEP_OFF + MAX_RANKS;
constexpr int MAIL_BASE_OFF = BASE_STEP_OFF + MAX_RANKS;
constexpr int DROPPED_OFF = MAIL_BASE_OFF + MAX_RANKS;
constexpr int ERR_OFF = DROPPED_OFF + 8;
//   [KIND_OFF, +MAX_COMM_BLOCKS*MAX_RANKS)    kind[block][src_rank]: which collective the peer entered with the epoch
//                                             it published in sig[block][src_rank] (protocol check, see block_barrier)
constexpr int KIND_OFF = ERR_OFF + 8;
constexpr int FLAG_PAGE_BYTES = 16384;
static_assert((KIND_OFF + SIG_WORDS) * 4 <= FLAG_PAGE_BYTES, "flag page too small");

constexpr uint32_t ERR_BARRIER_TIMEOUT = 1;  // a peer did not reach the barrier within the time bound
constexpr uint32_t ERR_ARRIVAL_TIMEOUT = 2;  // fewer than replicas_to_aggregate gradients arrived within the bound
constexpr uint32_t ERR_KIND_MISMATCH = 3;    // a peer entered a DIFFERENT collective at this barrier epoch
// collective kinds carried through the barriers
constexpr uint32_t KIND_AR = 1, KIND_RS = 2, KIND_AG = 3, KIND_BARRIER = 4;

struct CommCtx {
  int rank, world;
  uint8_t* peer[MAX_RANKS];  // heap base of every rank as mapped in this process (peer[rank] is local)
  uint8_t* mc;               // NVLS multicast mapping of the heap (nullptr when not available)
  uint32_t* err_host;        // 4 words of mapped pinned host memory: the host watchdog reads them without a sync
  unsigned long long timeout_ns;  // bound of every device-side wait (0 = wait forever)
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

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;\n" : "=l"(t));
  return t;
}

// A wait that outlives the time bound records what it was waiting for -- in the local flag page (so that the
// kernels queued behind this one drain immediately) and in mapped host memory (so that the host watchdog sees it
// without synchronising) -- and makes the kernel return. No __trap(): the context stays usable, the host raises
// WorkerLostError and the session reloads the last checkpoint.
__device__ __noinline__ void report_timeout(const CommCtx& c, uint32_t code, uint32_t peer, uint32_t epoch) {
  uint32_t* flags = reinterpret_cast<uint32_t*>(c.peer[c.rank]);
  if (atomicCAS(flags + ERR_OFF, 0u, code) == 0u) {
    flags[ERR_OFF + 1] = blockIdx.x;
    flags[ERR_OFF + 2] = peer;
    flags[ERR_OFF + 3] = epoch;
    if (c.err_host != nullptr) {
      c.err_host[1] = blockIdx.x;
      c.err_host[2] = peer;
      c.err_host[3] = epoch;
      __threadfence_system();
      st_release_sys(c.err_host, code);
    }
  }
}
__device__ __forceinline__ bool comm_failed(const CommCtx& c) {
  return reinterpret_cast<volatile uint32_t*>(c.peer[c.rank])[ERR_OFF] != 0u;
}

// Barrier among the same-index blocks of all ranks. Also orders memory: everything written (by this GPU) before the
// barrier is visible to peers after it. Returns false (for every thread of the block) when a peer did not show up
// within c.timeout_ns or an earlier wait on this rank already failed; the caller must then return at once.
// `kind` names the collective that is calling: every rank publishes it next to its epoch, and a rank that finds a
// peer at the SAME epoch with a different kind records ERR_KIND_MISMATCH -- the ranks did not issue the same sequence
// of collectives (for example a checkpoint gather launched by the chief alone pairs with the peers' next gradient
// collective: the failure mode is silent data corruption otherwise). SURVEY 5.2: epoch-stamped barriers that make
// peer-memory protocol faults detectable.
__device__ bool block_barrier(const CommCtx& c, uint32_t kind) {
  __syncthreads();
  uint32_t* flags = reinterpret_cast<uint32_t*>(c.peer[c.rank]);
  __shared__ uint32_t epoch_s;
  __shared__ int ok_s;
  if (threadIdx.x == 0) {
    __threadfence_system();
    const uint32_t e = flags[CTR_OFF + blockIdx.x] + 1;
    flags[CTR_OFF + blockIdx.x] = e;
    epoch_s = e;
    ok_s = comm_failed(c) ? 0 : 1;
  }
  __syncthreads();
  const uint32_t e = epoch_s;
  const bool healthy = ok_s != 0;
  if (threadIdx.x < c.world) {
    const int r = threadIdx.x;
    uint32_t* remote = reinterpret_cast<uint32_t*>(c.peer[r]) + blockIdx.x * MAX_RANKS + c.rank;
    remote[KIND_OFF] = kind;  // ordered before the epoch by the release store below
    st_release_sys(remote, e);
    if (healthy) {
      const uint32_t* mine = flags + blockIdx.x * MAX_RANKS + r;
      unsigned long long t0 = 0;
      uint32_t spins = 0, seen;
      while ((int32_t)((seen = ld_acquire_sys(mine)) - e) < 0) {
        if ((++spins & 1023u) == 0u && c.timeout_ns != 0ull) {
          const unsigned long long now = globaltimer_ns();
          if (t0 == 0) t0 = now;
          if (now - t0 > c.timeout_ns || comm_failed(c)) {
            report_timeout(c, ERR_BARRIER_TIMEOUT, (uint32_t)r, e);
            ok_s = 0;
            break;
          }
        }
      }
      // the peer is at exactly this epoch (not already one ahead): its kind word belongs to this barrier
      if (ok_s && seen == e && ld_acquire_sys(mine + KIND_OFF) != kind) {
        report_timeout(c, ERR_KIND_MISMATCH, (uint32_t)r, e);
        ok_s = 0;
      }
    }
  }
  __syncthreads();
  return ok_s != 0;
}

__device__ __forceinline__ uint2 pack4_bf16(float4 v) {
  __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
  uint2 u;
  u.x = *reinterpret_cast<uint32_t*>(&a);
  u.y = *reinterpret_cast<uint32_t*>(&b);
  return u;
}

// Remote loads cost ~2 us each: keep U independent 16-byte loads per peer in flight per thread. q0 is the first
// float4 index of this thread, `stride` the distance between its U elements, n4 the range length (indices relative
// to base4). Sums are formed in rank order, so every rank computes bit-identical results. `mask` selects the ranks
// whose gradients enter the sum (all ranks, or the first replicas_to_aggregate arrivals in backup-worker mode).
constexpr int COMM_U = 4;
__device__ __forceinline__ void reduce_peers_u(const CommCtx& c, size_t grad_off, long long base4, long long q0,
                                               long long stride, long long n4, uint32_t mask, float4 (&s)[COMM_U]) {
#pragma unroll
  for (int u = 0; u < COMM_U; ++u) s[u] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int r = 0; r < MAX_RANKS; ++r) {
    if (r < c.world && ((mask >> r) & 1u)) {
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
allreduce_optim_kernel(CommCtx c, size_t grad_off, long long start, long long count, OptimArgs o,
                       const float* __restrict__ hyper, float* master, float* m, float* v, __nv_bfloat16* shadow,
                       float* avg_out) {
  optim_load_hyper(o, hyper);
  if (!block_barrier(c, KIND_AR)) return;
  const long long n4 = count >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long q0 = blockIdx.x * (long long)blockDim.x + threadIdx.x; q0 < n4; q0 += stride * COMM_U) {
    float4 gs[COMM_U];
    reduce_peers_u(c, grad_off, start >> 2, q0, stride, n4, 0xffffffffu, gs);
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
  block_barrier(c, KIND_AR);  // nobody may overwrite its gradient bucket while a peer still reads it
}

// ---- K-RS-OPT (+ all-gather by peer stores). bounds[r]..bounds[r+1] is rank r's owned element range (multiples
// of 4). master/m/v are local (owner-side state); shadow_off / master_off are heap offsets of the replicated copies
// that every peer receives. The fp32 master is broadcast only inside [f32_lo, f32_hi): the region of the arena that
// holds the variables the forward / backward kernels read in fp32 (biases, batch-norm gamma / beta) -- without it a
// non-owner would keep computing with the initial values of the parameters it does not own.
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
struct PushF32 {
  size_t master_off;
  long long lo, hi;
};

// Backup workers (SyncReplicasOptimizer(replicas_to_aggregate = R < N), /root/reference/distribute_train.py:146-155):
// every rank takes a ticket from each owner's arrival counter (a remote atomic in the owner's flag page -- served by
// the owner's memory system, not its SMs) and publishes it; the owner aggregates the ranks whose ticket of this call
// is among the first R, drops the rest (their gradient is stale by TF's rule) and the caller scales by 1/R. Each
// owner decides for itself, exactly like TF's per-variable accumulators. Tickets of call k are k*N .. k*N+N-1: no
// rank can arrive for call k+1 before every rank has arrived for call k (the exit barrier of call k needs them all).
// The call index k is one counter per rank, read by every block when the kernel starts and advanced by the block
// that finishes last (backup_finish): kernels of one communicator are stream-ordered, so every block of call k reads
// k whatever the grid sizes of the calls were. Returns the bit mask of aggregated ranks, or 0 after a time-out.
__device__ void backup_finish(const CommCtx& c) {
  uint32_t* flags = reinterpret_cast<uint32_t*>(c.peer[c.rank]);
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(flags + BK_CALL_OFF + 1, 1u) == gridDim.x - 1) {
      flags[BK_CALL_OFF + 1] = 0;
      __threadfence();
      atomicAdd(flags + BK_CALL_OFF, 1u);
    }
  }
}
__device__ uint32_t backup_arrive_and_select(const CommCtx& c, int R) {
  uint32_t* flags = reinterpret_cast<uint32_t*>(c.peer[c.rank]);
  __shared__ uint32_t mask_s;
  __shared__ uint32_t call_s;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    call_s = *reinterpret_cast<volatile uint32_t*>(flags + BK_CALL_OFF);
    mask_s = 0;
  }
  __syncthreads();
  const uint32_t k = call_s;
  if (blockIdx.x == 0 && threadIdx.x < c.world) {
    uint32_t* owner = reinterpret_cast<uint32_t*>(c.peer[threadIdx.x]);
    const uint32_t t = atomicAdd_system(owner + TICKET_OFF, 1u);
    st_release_sys(owner + ARRIVE_OFF + c.rank, t + 1u);
  }
  if (threadIdx.x == 0) {
    const uint32_t lo = k * (uint32_t)c.world;
    unsigned long long t0 = 0;
    uint32_t spins = 0, mask = 0;
    bool failed = comm_failed(c);
    while (!failed) {
      mask = 0;
      int cnt = 0;
      for (int r = 0; r < c.world; ++r) {
        const uint32_t a = ld_acquire_sys(flags + ARRIVE_OFF + r);
        const uint32_t ord = a - 1u - lo;  // wraps to a huge value for tickets of earlier calls
        if (a != 0u && ord < (uint32_t)R) {
          mask |= 1u << r;
          ++cnt;
        }
      }
      if (cnt >= R) break;
      if ((++spins & 255u) == 0u && c.timeout_ns != 0ull) {
        const unsigned long long now = globaltimer_ns();
        if (t0 == 0) t0 = now;
        if (now - t0 > c.timeout_ns || comm_failed(c)) {
          report_timeout(c, ERR_ARRIVAL_TIMEOUT, (uint32_t)cnt, k);
          failed = true;
        }
      }
    }
    mask_s = failed ? 0u : mask;
  }
  __syncthreads();
  return mask_s;
}

__global__ void __launch_bounds__(512)
rs_optim_ag_kernel(CommCtx c, size_t grad_off, Bounds bounds, OptimArgs o, const float* __restrict__ hyper,
                   float* master, float* m, float* v, size_t shadow_off, PushF32 f32, NoPush np, int backup_r) {
  optim_load_hyper(o, hyper);
  uint32_t mask = 0xffffffffu;
  const bool backup = backup_r > 0 && backup_r < c.world;
  if (backup) {
    mask = backup_arrive_and_select(c, backup_r);
    if (mask == 0u) {
      backup_finish(c);
      return;
    }
  } else if (!block_barrier(c, KIND_RS)) {
    return;
  }
  const long long start = bounds.b[c.rank];
  const long long n4 = (bounds.b[c.rank + 1] - start) >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long q0 = blockIdx.x * (long long)blockDim.x + threadIdx.x; q0 < n4; q0 += stride * COMM_U) {
    float4 gs[COMM_U];
    reduce_peers_u(c, grad_off, start >> 2, q0, stride, n4, mask, gs);
#pragma unroll
    for (int u = 0; u < COMM_U; ++u) {
      const long long q = q0 + u * stride;
      if (q >= n4) continue;
      const long long i = start + (q << 2);
      const float4 p = optim4(o, master, m, v, i, gs[u]);
      const uint2 pb = pack4_bf16(p);
      const bool local_only = in_nopush(np, i);
      const bool push_f32 = i >= f32.lo && i < f32.hi;
#pragma unroll
      for (int r = 0; r < MAX_RANKS; ++r) {
        if (r < c.world && (!local_only || r == c.rank)) {
          *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(c.peer[r] + shadow_off) + i) = pb;
          if (push_f32 && r != c.rank)
            *reinterpret_cast<float4*>(reinterpret_cast<float*>(c.peer[r] + f32.master_off) + i) = p;
        }
      }
    }
  }
  block_barrier(c, KIND_RS);  // all shards have landed everywhere; gradients may be overwritten
  if (backup) backup_finish(c);
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
__device__ __forceinline__ void multimem_st_f4(void* mc_addr, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};\n" ::"l"(mc_addr), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}
__global__ void __launch_bounds__(512)
rs_optim_ag_nvls_kernel(CommCtx c, size_t grad_off, Bounds bounds, OptimArgs o, const float* __restrict__ hyper,
                        float* master, float* m, float* v, size_t shadow_off, PushF32 f32, NoPush np) {
  optim_load_hyper(o, hyper);
  if (!block_barrier(c, KIND_RS)) return;
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
      // the owner's own master was written by optim4; the multicast store rewrites it with the same value
      if (i >= f32.lo && i < f32.hi) multimem_st_f4(reinterpret_cast<float*>(c.mc + f32.master_off) + i, p);
    }
  }
  block_barrier(c, KIND_RS);
}

// ---- all-gather of owned slices by peer stores (16-byte units): used for checkpoint gathers and weight pulls.
__global__ void __launch_bounds__(512)
allgather_slices_kernel(CommCtx c, size_t buf_off, Bounds bounds16) {
  if (!block_barrier(c, KIND_AG)) return;
  const long long start = bounds16.b[c.rank];
  const long long n = bounds16.b[c.rank + 1] - start;
  const uint4* src = reinterpret_cast<const uint4*>(c.peer[c.rank] + buf_off) + start;
  for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < n; q += (long long)gridDim.x * blockDim.x) {
    const uint4 val = src[q];
#pragma unroll
    for (int r = 0; r < MAX_RANKS; ++r)
      if (r < c.world && r != c.rank) reinterpret_cast<uint4*>(c.peer[r] + buf_off)[start + q] = val;
  }
  block_barrier(c, KIND_AG);
}

// ---- K-ASYNC. The version stamps and step counters live in the flag page and are advanced by the kernels, so a
// captured CUDA graph that replays push + apply keeps publishing NEW versions (a by-value version argument would be
// frozen at capture and the owner would apply once and never again).
// begin: remember which update of every owner this step's weights correspond to (the push is stamped with it).
__global__ void async_begin_kernel(CommCtx c) {
  uint32_t* flags = reinterpret_cast<uint32_t*>(c.peer[c.rank]);
  const int o = threadIdx.x;
  if (o < c.world) flags[BASE_STEP_OFF + o] = ld_acquire_sys(flags + OWNER_STEP_OFF + o);
}
// push: copy my gradient slice for every owner into that owner's mailbox[me], then publish the next version together
// with the owner step the gradient was computed at. mailbox layout on each rank: [world][slice_cap] floats.
__global__ void __launch_bounds__(512)
async_push_kernel(CommCtx c, size_t grad_off, size_t mailbox_off, long long slice_cap, Bounds bounds) {
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
  __shared__ uint32_t ver_s;
  uint32_t* flags = reinterpret_cast<uint32_t*>(c.peer[c.rank]);
  if (threadIdx.x == 0) {
    const uint32_t prev = atomicAdd(flags + DONE_CTR_OFF, 1u);
    last = (prev == gridDim.x - 1);
    if (last) {
      ver_s = flags[PUSH_VER_OFF] + 1u;
      flags[PUSH_VER_OFF] = ver_s;
    }
  }
  __syncthreads();
  if (last) {
    __threadfence_system();
    if (threadIdx.x < c.world) {
      uint32_t* owner = reinterpret_cast<uint32_t*>(c.peer[threadIdx.x]);
      owner[MAIL_BASE_OFF + c.rank] = flags[BASE_STEP_OFF + threadIdx.x];
      st_release_sys(owner + MAIL_FLAG_OFF + c.rank, ver_s);
    }
    if (threadIdx.x == 0) flags[DONE_CTR_OFF] = 0;
  }
}
// ---- owner-side apply: for every worker whose mailbox version is newer than the last applied one, apply its
// (possibly stale) gradient to the owned shard and push the new bf16 weights to every rank. Never waits.
// snap[w] (8 words, caller-provided) receives the version applied for worker w (0 = nothing new or dropped).
// A push computed more than max_staleness owner updates ago is dropped (max_staleness < 0: no bound).
// A worker may overwrite its mailbox while the owner reads it: like use_locking=False in the reference, a torn
// (mixed-version) gradient is accepted.
__global__ void async_snapshot_kernel(CommCtx c, uint32_t* snap, int max_staleness) {
  // one consistent view of "what has arrived" for the whole apply grid
  uint32_t* flags = reinterpret_cast<uint32_t*>(c.peer[c.rank]);
  const int w = threadIdx.x;
  if (w < MAX_RANKS) {
    uint32_t s = 0;
    if (w < c.world) {
      const uint32_t ver = ld_acquire_sys(flags + MAIL_FLAG_OFF + w);
      if (ver != flags[APPLIED_OFF + w]) {
        s = ver;
        const uint32_t age = flags[OWNER_STEP_OFF + c.rank] - flags[MAIL_BASE_OFF + w];
        if (max_staleness >= 0 && (int32_t)age > max_staleness) {
          s = 0;
          flags[APPLIED_OFF + w] = ver;  // consumed without applying
          atomicAdd(flags + DROPPED_OFF, 1u);
        }
      }
    }
    snap[w] = s;
  }
}
__global__ void __launch_bounds__(512)
async_apply_kernel(CommCtx c, size_t mailbox_off, long long slice_cap, Bounds bounds, OptimArgs o,
                   const float* __restrict__ hyper, float* master, float* m, float* v, size_t shadow_off, PushF32 f32,
                   const uint32_t* snap) {
  optim_load_hyper(o, hyper);
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
      const bool push_f32 = i >= f32.lo && i < f32.hi;
#pragma unroll
      for (int r = 0; r < MAX_RANKS; ++r)
        if (r < c.world) {
          *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(c.peer[r] + shadow_off) + i) = pb;
          if (push_f32 && r != c.rank)
            *reinterpret_cast<float4*>(reinterpret_cast<float*>(c.peer[r] + f32.master_off) + i) = p;
        }
    }
  }
}
__global__ void async_commit_kernel(CommCtx c, const uint32_t* snap) {
  uint32_t* flags = reinterpret_cast<uint32_t*>(c.peer[c.rank]);
  const int w = threadIdx.x;
  __shared__ uint32_t napplied, base;
  if (w == 0) {
    napplied = 0;
    base = flags[OWNER_STEP_OFF + c.rank];
  }
  __syncthreads();
  if (w < c.world && snap[w] != 0) {
    flags[APPLIED_OFF + w] = snap[w];
    atomicAdd(&napplied, 1u);
  }
  __syncthreads();
  // publish how many updates this owner has applied so far (the staleness clock of its shard) to every rank
  if (napplied != 0 && w < c.world)
    st_release_sys(reinterpret_cast<uint32_t*>(c.peer[w]) + OWNER_STEP_OFF + c.rank, base + napplied);
}

__global__ void barrier_kernel(CommCtx c) { block_barrier(c, KIND_BARRIER); }

// heartbeat: bump my counter in every peer's flag page; read what peers wrote into mine
__global__ void heartbeat_kernel(CommCtx c, uint32_t beat, uint32_t* seen_out) {
  const int r = threadIdx.x;
  if (r < c.world) {
    st_release_sys(reinterpret_cast<uint32_t*>(c.peer[r]) + HEARTBEAT_OFF + c.rank, beat);
    seen_out[r] = ld_acquire_sys(reinterpret_cast<uint32_t*>(c.peer[c.rank]) + HEARTBEAT_OFF + r);
  }
}

// test / fault-injection helper: keep the stream busy for `ns` nanoseconds (an injected straggler)
__global__ void sleep_kernel(unsigned long long ns) {
  const unsigned long long t0 = globaltimer_ns();
  while (globaltimer_ns() - t0 < ns) __nanosleep(1000);
}

}  // namespace dtf

using namespace dtf;

struct DtfComm {
  CommCtx ctx;
  size_t heap_bytes;
  void* opened[MAX_RANKS];
  uint32_t* err_host;  // mapped pinned host memory (4 words)
};

extern "C" int dtf_comm_flag_page_bytes(void) { return FLAG_PAGE_BYTES; }
extern "C" int dtf_comm_max_blocks(void) { return MAX_COMM_BLOCKS; }

// Allocate the local symmetric heap (zeroed) and return its IPC handle (64 bytes).
extern "C" int dtf_comm_heap_alloc(size_t bytes, void** ptr_out, void* handle_out64) {
  void* p = nullptr;
  if (cudaMalloc(&p, bytes) != cudaSuccess) return DTF_ERR_CUDA;
  if (cudaMemset(p, 0, bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) {
    cudaFree(p);
    return DTF_ERR_CUDA;
  }
  cudaIpcMemHandle_t h;
  if (cudaIpcGetMemHandle(&h, p) != cudaSuccess) {
    cudaFree(p);
    return DTF_ERR_CUDA;
  }
  memcpy(handle_out64, &h, sizeof(h));
  *ptr_out = p;
  return DTF_OK;
}
extern "C" int dtf_comm_heap_free(void* ptr) { return cudaFree(ptr) == cudaSuccess ? DTF_OK : DTF_ERR_CUDA; }

static unsigned long long default_timeout_ns() {
  const char* e = getenv("DTF_COMM_TIMEOUT_MS");
  const long long ms = e ? atoll(e) : 30000;
  return ms <= 0 ? 0ull : (unsigned long long)ms * 1000000ull;
}
static void comm_finish_create(DtfComm* c) {
  c->ctx.timeout_ns = default_timeout_ns();
  c->err_host = nullptr;
  void* h = nullptr;
  if (cudaHostAlloc(&h, 64, cudaHostAllocMapped) == cudaSuccess) {
    memset(h, 0, 64);
    void* d = nullptr;
    if (cudaHostGetDevicePointer(&d, h, 0) == cudaSuccess) {
      c->err_host = reinterpret_cast<uint32_t*>(h);
      c->ctx.err_host = reinterpret_cast<uint32_t*>(d);
    } else {
      cudaFreeHost(h);
    }
  }
  (void)cudaGetLastError();
}
static void comm_release(DtfComm* c) {
  for (int r = 0; r < MAX_RANKS; ++r)
    if (c->opened[r]) cudaIpcCloseMemHandle(c->opened[r]);
  if (c->err_host) cudaFreeHost(c->err_host);
  delete c;
}

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
      (void)cudaGetLastError();
      comm_release(c);  // closes the mappings opened so far
      return DTF_ERR_CUDA;
    }
    c->opened[r] = p;
    c->ctx.peer[r] = reinterpret_cast<uint8_t*>(p);
  }
  comm_finish_create(c);
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
  comm_finish_create(c);
  *comm_out = c;
  return DTF_OK;
}
extern "C" int dtf_comm_destroy(void* comm) {
  comm_release(reinterpret_cast<DtfComm*>(comm));
  return DTF_OK;
}
// Attach the NVLS multicast mapping of the heap (from cuMulticast* / torch symmetric memory); 0 detaches it.
extern "C" int dtf_comm_set_multicast(void* comm, void* mc_base) {
  reinterpret_cast<DtfComm*>(comm)->ctx.mc = reinterpret_cast<uint8_t*>(mc_base);
  return DTF_OK;
}
extern "C" void* dtf_comm_peer_ptr(void* comm, int r) { return reinterpret_cast<DtfComm*>(comm)->ctx.peer[r]; }
// Bound of every device-side wait in milliseconds (0 = wait forever). Default 30 s, or DTF_COMM_TIMEOUT_MS.
extern "C" int dtf_comm_set_timeout_ms(void* comm, long long ms) {
  reinterpret_cast<DtfComm*>(comm)->ctx.timeout_ns = ms <= 0 ? 0ull : (unsigned long long)ms * 1000000ull;
  return DTF_OK;
}
// Host watchdog: copies the 4-word error record (code, block, peer / arrivals, epoch) out of mapped host memory.
// No CUDA call, no synchronisation: safe to poll from any thread while kernels run. Returns the code (0 = healthy).
extern "C" int dtf_comm_error(void* comm, unsigned int* out4) {
  DtfComm* c = reinterpret_cast<DtfComm*>(comm);
  if (c->err_host == nullptr) return 0;
  const volatile uint32_t* e = c->err_host;
  const uint32_t code = e[0];
  if (out4 != nullptr) {
    out4[0] = code; out4[1] = e[1]; out4[2] = e[2]; out4[3] = e[3];
  }
  return (int)code;
}
// After a recovery (every rank reloaded the checkpoint and re-synchronised on the control plane) clear the record.
extern "C" int dtf_comm_clear_error(void* comm, void* stream) {
  DtfComm* c = reinterpret_cast<DtfComm*>(comm);
  if (c->err_host) memset(c->err_host, 0, 16);
  uint32_t* flags = reinterpret_cast<uint32_t*>(c->ctx.peer[c->ctx.rank]);
  return cudaMemsetAsync(flags + ERR_OFF, 0, 16, dtf_stream(stream)) == cudaSuccess ? DTF_OK : DTF_ERR_CUDA;
}

static OptimArgs make_optim(int kind, float lr, float b1, float b2, float eps, float wd, float gscale,
                            long long step) {
  OptimArgs o;
  o.kind = kind; o.lr = lr; o.b1 = b1; o.b2 = b2; o.eps = eps; o.wd = wd; o.gscale = gscale;
  o.lr_t = lr;
  if (kind == 2) o.lr_t = lr * sqrtf(1.f - powf(b2, (float)step)) / (1.f - powf(b1, (float)step));
  return o;
}
// Threads per collective CTA (DTF_COMM_THREADS, default 512). Smaller CTAs (256 / 128 threads, 64 registers) were tried
// so that collective CTAs could sit next to the persistent compute CTAs; the 2-GPU step time did not move, and the
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
static int make_push_f32(PushF32* f32, size_t master_off, long long f32_lo, long long f32_hi) {
  f32->master_off = master_off;
  f32->lo = f32->hi = 0;
  if (master_off != (size_t)-1 && f32_hi > f32_lo) {
    if ((master_off & 15) || (f32_lo & 3) || (f32_hi & 3)) return DTF_ERR_ALIGN;
    f32->lo = f32_lo;
    f32->hi = f32_hi;
  }
  return DTF_OK;
}

// The grid must be identical on every rank for a given call, so it is derived from (count, blocks) only.
extern "C" int dtf_comm_allreduce_optim(void* comm, size_t grad_off, long long start, long long count, int kind,
                                        float lr, float b1, float b2, float eps, float wd, float gscale,
                                        long long step, const float* hyper, float* master, float* m, float* v,
                                        void* shadow, float* avg_out, int blocks, void* stream) {
  DtfComm* c = reinterpret_cast<DtfComm*>(comm);
  if ((start & 3) || (count & 3) || (grad_off & 15)) return DTF_ERR_ALIGN;
  if (master == nullptr && avg_out == nullptr) return DTF_ERR_ARG;
  const OptimArgs o = make_optim(kind, lr, b1, b2, eps, wd, gscale, step);
  const int grid = comm_grid(count >> 2, blocks);
  allreduce_optim_kernel<<<grid, comm_threads(), 0, dtf_stream(stream)>>>(c->ctx, grad_off, start, count, o, hyper,
                                                                          master, m, v, (__nv_bfloat16*)shadow, avg_out);
  return DTF_CHECK_LAUNCH();
}
extern "C" int dtf_comm_rs_optim_ag(void* comm, size_t grad_off, const long long* bounds, int kind, float lr,
                                    float b1, float b2, float eps, float wd, float gscale, long long step,
                                    const float* hyper, float* master, float* m, float* v, size_t shadow_off,
                                    size_t master_off, long long f32_lo, long long f32_hi, int n_nopush,
                                    const long long* nopush_ranges, int backup_r, int blocks, void* stream) {
  DtfComm* c = reinterpret_cast<DtfComm*>(comm);
  Bounds b;
  long long maxn = 0;
  for (int r = 0; r <= c->ctx.world; ++r) {
    b.b[r] = bounds[r];
    if (bounds[r] & 3) return DTF_ERR_ALIGN;
    if (r > 0 && bounds[r] - bounds[r - 1] > maxn) maxn = bounds[r] - bounds[r - 1];
  }
  if ((grad_off & 15) || (shadow_off & 15)) return DTF_ERR_ALIGN;
  PushF32 f32;
  const int rc = make_push_f32(&f32, master_off, f32_lo, f32_hi);
  if (rc != DTF_OK) return rc;
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
  if (backup_r < 0 || backup_r > c->ctx.world) return DTF_ERR_ARG;
  const bool backup = backup_r > 0 && backup_r < c->ctx.world;
  // the in-switch reduction adds ALL ranks: backup-worker mode (a subset of the ranks) needs the peer-load kernel
  if (c->ctx.mc != nullptr && !backup)
    rs_optim_ag_nvls_kernel<<<grid, comm_threads(), 0, dtf_stream(stream)>>>(c->ctx, grad_off, b, o, hyper, master, m, v,
                                                                             shadow_off, f32, np);
  else
    rs_optim_ag_kernel<<<grid, comm_threads(), 0, dtf_stream(stream)>>>(c->ctx, grad_off, b, o, hyper, master, m, v,
                                                                        shadow_off, f32, np, backup_r);
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
extern "C" int dtf_comm_async_begin(void* comm, void* stream) {
  DtfComm* c = reinterpret_cast<DtfComm*>(comm);
  async_begin_kernel<<<1, 32, 0, dtf_stream(stream)>>>(c->ctx);
  return DTF_CHECK_LAUNCH();
}
extern "C" int dtf_comm_async_push(void* comm, size_t grad_off, size_t mailbox_off, long long slice_cap,
                                   const long long* bounds, int blocks, void* stream) {
  DtfComm* c = reinterpret_cast<DtfComm*>(comm);
  Bounds b;
  for (int r = 0; r <= c->ctx.world; ++r) {
    b.b[r] = bounds[r];
    if (bounds[r] & 3) return DTF_ERR_ALIGN;
    if (r > 0 && bounds[r] - bounds[r - 1] > slice_cap) return DTF_ERR_TOO_LARGE;
  }
  if ((slice_cap & 3) || (grad_off & 15) || (mailbox_off & 15)) return DTF_ERR_ALIGN;
  int grid = blocks < 1 ? 1 : (blocks > MAX_COMM_BLOCKS ? MAX_COMM_BLOCKS : blocks);
  async_push_kernel<<<grid, comm_threads(), 0, dtf_stream(stream)>>>(c->ctx, grad_off, mailbox_off, slice_cap, b);
  return DTF_CHECK_LAUNCH();
}
extern "C" int dtf_comm_async_apply(void* comm, size_t mailbox_off, long long slice_cap, const long long* bounds,
                                    int kind, float lr, float b1, float b2, float eps, float wd, float gscale,
                                    long long step, const float* hyper, float* master, float* m, float* v,
                                    size_t shadow_off, size_t master_off, long long f32_lo, long long f32_hi,
                                    int max_staleness, unsigned int* applied_out_dev, int blocks, void* stream) {
  DtfComm* c = reinterpret_cast<DtfComm*>(comm);
  Bounds b;
  for (int r = 0; r <= c->ctx.world; ++r) b.b[r] = bounds[r];
  PushF32 f32;
  const int rc = make_push_f32(&f32, master_off, f32_lo, f32_hi);
  if (rc != DTF_OK) return rc;
  const OptimArgs o = make_optim(kind, lr, b1, b2, eps, wd, gscale, step);
  int grid = blocks < 1 ? 1 : (blocks > MAX_COMM_BLOCKS ? MAX_COMM_BLOCKS : blocks);
  cudaStream_t st = dtf_stream(stream);
  async_snapshot_kernel<<<1, 32, 0, st>>>(c->ctx, applied_out_dev, max_staleness);
  async_apply_kernel<<<grid, comm_threads(), 0, st>>>(c->ctx, mailbox_off, slice_cap, b, o, hyper, master, m, v,
                                                      shadow_off, f32, applied_out_dev);
  async_commit_kernel<<<1, 32, 0, st>>>(c->ctx, applied_out_dev);
  return DTF_CHECK_LAUNCH();
}
// counters of the asynchronous protocol: out[0] = pushes dropped as too stale, out[1] = updates applied by this owner,
// out[2] = versions pushed by this rank (synchronises the stream)
extern "C" int dtf_comm_async_stats(void* comm, unsigned int* out3, void* stream) {
  DtfComm* c = reinterpret_cast<DtfComm*>(comm);
  const uint32_t* flags = reinterpret_cast<const uint32_t*>(c->ctx.peer[c->ctx.rank]);
  cudaStream_t st = dtf_stream(stream);
  if (cudaMemcpyAsync(out3 + 0, flags + DROPPED_OFF, 4, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
      cudaMemcpyAsync(out3 + 1, flags + OWNER_STEP_OFF + c->ctx.rank, 4, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
      cudaMemcpyAsync(out3 + 2, flags + PUSH_VER_OFF, 4, cudaMemcpyDeviceToHost, st) != cudaSuccess)
    return DTF_ERR_CUDA;
  return cudaStreamSynchronize(st) == cudaSuccess ? DTF_OK : DTF_ERR_CUDA;
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
extern "C" int dtf_debug_sleep(long long ns, void* stream) {
  if (ns < 0) return DTF_ERR_ARG;
  sleep_kernel<<<1, 1, 0, dtf_stream(stream)>>>((unsigned long long)ns);
  return DTF_CHECK_LAUNCH();
}
