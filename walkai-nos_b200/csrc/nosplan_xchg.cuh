// nosplan_xchg.cuh — the one exchange step of the sharded scorer, done by our own kernels
// over NVLink peer memory instead of a library all-reduce.
//
// Every rank owns a contiguous range of nodes and scores all demand rows against it; "first node wins"
// over the whole list (tryRepartition, mig_controller.go:146-166) is then the minimum global node index
// over ranks.  Each rank keeps an exchange buffer that all peers have mapped (CUDA IPC):
//
//     keys[2][max_keys]   int32, this rank's OWN winners of a step (-1 = none; as unsigned, above every index)
//     pushed[world]       pushed[r] = last step whose keys rank r has completed in ITS buffer
//     read[world]         read[r]   = last step whose keys rank r has finished reading from THIS buffer
//
// Step s uses keys[s & 1].  A rank writes only its own buffer: the scorers fold their winners into it with
// local atomics (the scorer CTA that finishes last stores s into the rank's slot of every rank's pushed[]),
// and the finishing kernel PULLS: it waits until all of pushed[] reached s, reads every rank's keys[s & 1]
// with coalesced 128-byte loads over NVLink and keeps the unsigned minimum, then stores s into its slot of
// every rank's read[].  (A first version pushed 4-byte atomicMin's into every peer from the scorers'
// epilogue: rows x tiles x ranks scattered NVLink transactions, 1.6 M per pass at 8 ranks, which cost more
// than the scoring of a shard.  Pulling moves the same information in (ranks - 1) x rows x 4 contiguous bytes.)
// Re-use of keys[s & 1] at step s waits for read[] >= s - 2 (in the kernel that resets it), so two halves
// are enough.  Nothing here needs a peer's SMs, so a rank that spins cannot block the ranks it waits for; a
// spin that lasts longer than `timeout_ns` gives up and raises *err.
#pragma once
#include "nosplan_kernels.cuh"

namespace nsp {

struct XchgPeers {
  int32_t* keys[kMaxWorld];     // each rank's keys[2][max_keys]
  uint32_t* pushed[kMaxWorld];  // each rank's pushed[world]
  uint32_t* read[kMaxWorld];    // each rank's read[world]
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// a flag store that follows ONE __threadfence_system(): the fence orders everything before it system-wide, the stores
// themselves are then posted back to back (a release store per peer would drain the write path once per peer: 8 peers
// cost 17 us at the end of the last scorer CTA)
__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_relaxed_sys_v4(const uint4* p) {  // peer memory: never from a stale L1 line
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// threads 0 .. world-1 of the CTA wait until flags[t] has reached `need`; the CTA meets at a barrier
__device__ __forceinline__ void wait_flags(const uint32_t* flags, int world, uint32_t need, unsigned long long timeout_ns, int* err) {
  if ((int)threadIdx.x < world) {
    const unsigned long long t0 = globaltimer_ns();
    while ((int32_t)(ld_acquire_sys(flags + threadIdx.x) - need) < 0) {
      if (globaltimer_ns() - t0 > timeout_ns) {
        atomicExch(err, 1 + (int)threadIdx.x);
        break;
      }
      __nanosleep(100);
    }
  }
  __syncthreads();
}

// g = global node index of the CTA's winner for `row`: folded into the rank's own buffer (or the plain first_node[])
__device__ __forceinline__ void push_row(const PushArgs&, int32_t* first_node, int row, int g) {
  atomicMin(reinterpret_cast<unsigned*>(first_node) + row, (unsigned)g);
}
// every thread of the CTA calls this once, after its last push_row (CTAs without work included): the CTA that
// finishes last over all scorer launches of the pass tells every rank that this rank's keys of the step are complete
__device__ __forceinline__ void push_finish(const PushArgs& P) {
  if (P.world == 0) return;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(P.ticket, 1u);
    if (prev == P.total - 1u) {  // every CTA's winners are ordered before this point
      *P.ticket = 0;
      __threadfence_system();
      for (int r = 0; r < P.world; ++r) st_relaxed_sys(P.flags[r] + P.rank, P.step);
    }
  }
}

// in[d]: global node index of this rank's winner for row d, or -1: copied into the rank's own keys[step & 1]
// (after every rank has read what the half held two steps ago), then published like the scorers do.
__global__ void xchg_publish_kernel(XchgPeers P, int world, int rank, int max_keys, uint32_t step, int n,
                                    const int32_t* __restrict__ in, unsigned int* __restrict__ done,
                                    unsigned long long timeout_ns, int* __restrict__ err) {
  if (step > 2u) wait_flags(P.read[rank], world, step - 2u, timeout_ns, err);
  int32_t* k = P.keys[rank] + (size_t)(step & 1u) * (size_t)max_keys;
  for (int d = blockIdx.x * blockDim.x + threadIdx.x; d < n; d += gridDim.x * blockDim.x) k[d] = in[d];
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(done, 1u);
    if (prev == gridDim.x - 1) {
      *done = 0;
      __threadfence_system();
      for (int r = 0; r < world; ++r) st_relaxed_sys(P.pushed[r] + rank, step);
    }
  }
}

// out[d] = lowest non-negative key of row d over all ranks, or -1
__global__ void xchg_finish_kernel(XchgPeers P, int world, int rank, int max_keys, uint32_t step, int n,
                                   int32_t* __restrict__ out, unsigned int* __restrict__ done,
                                   unsigned long long timeout_ns, int* __restrict__ err) {
  wait_flags(P.pushed[rank], world, step, timeout_ns, err);
  if (*reinterpret_cast<volatile int*>(err) != 0) {
    // a peer never completed a step (this one or an earlier one): no minimum exists.  Every key reads -2, nothing is
    // signalled; the host sees the error at its next status call and refuses further passes until the exchange is rebuilt.
    for (int d = blockIdx.x * blockDim.x + threadIdx.x; d < n; d += gridDim.x * blockDim.x) out[d] = -2;
    return;
  }
  const size_t half = (size_t)(step & 1u) * (size_t)max_keys;  // max_keys is a multiple of 4: halves stay 16-byte aligned
  const int n4 = n >> 2;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += gridDim.x * blockDim.x) {
    // all ranks' loads are issued before the first one is used: one NVLink round trip per thread, not one per rank
    uint4 v[kMaxWorld];
#pragma unroll
    for (int r = 0; r < kMaxWorld; ++r)
      if (r < world) v[r] = ld_relaxed_sys_v4(reinterpret_cast<const uint4*>(P.keys[r] + half) + i);
    uint4 m = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
#pragma unroll
    for (int r = 0; r < kMaxWorld; ++r)
      if (r < world) {
        m.x = min(m.x, v[r].x);
        m.y = min(m.y, v[r].y);
        m.z = min(m.z, v[r].z);
        m.w = min(m.w, v[r].w);
      }
    reinterpret_cast<uint4*>(out)[i] = m;  // 0xffffffff == -1: no rank had a node
  }
  for (int d = (n4 << 2) + blockIdx.x * blockDim.x + threadIdx.x; d < n; d += gridDim.x * blockDim.x) {
    uint32_t m = 0xffffffffu;
    for (int r = 0; r < world; ++r) m = min(m, ld_relaxed_sys(reinterpret_cast<const uint32_t*>(P.keys[r] + half) + d));
    out[d] = (int32_t)m;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(done + 1, 1u);
    if (prev == gridDim.x - 1) {  // every block has read: the peers may re-use this half two steps from now
      done[1] = 0;
      __threadfence_system();
      for (int r = 0; r < world; ++r) st_relaxed_sys(P.read[r] + rank, step);
    }
  }
}

}  // namespace nsp
