// nosplan_xchg.cuh — the one exchange step of the sharded scorer, done by our own kernels
// over NVLink peer memory instead of a library all-reduce.
//
// Every rank owns a contiguous range of nodes and scores all demand rows against it; "first node wins"
// over the whole list (tryRepartition, mig_controller.go:146-166) is then the minimum global node index
// over ranks.  Each rank keeps an exchange buffer that all peers have mapped (CUDA IPC):
//
//     keys[2][max_keys]   int32, rests at kNoNode; peers fold their winners in with atomicMin
//     flags[world]        flags[r] = number of the last step rank r has finished pushing here
//
// Step s uses keys[s & 1].  push: every rank atomicMin's its winners into ALL ranks' keys[s & 1]
// (red.min over NVLink, fire and forget), fences, and the last block of the grid stores s into its
// slot of every rank's flags.  finish: wait until all of flags[] reached s, read the local keys[s & 1]
// (kNoNode -> -1) and put it back to rest for step s + 2.  Two buffers are enough: a peer can only
// push step s + 2 after it saw our flag of step s + 1, and our push of s + 1 is stream-ordered after
// our finish of s.  Nothing here needs the peers' SMs, so a rank that spins in finish cannot block
// the ranks it waits for; a spin that lasts longer than `timeout_ns` gives up and raises *err.
#pragma once
#include "nosplan_kernels.cuh"

namespace nsp {

constexpr int kMaxWorld = 16;

struct XchgPeers {
  int32_t* keys[kMaxWorld];    // each rank's keys[2][max_keys]
  uint32_t* flags[kMaxWorld];  // each rank's flags[world]
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// in[d]: global node index of this rank's winner for row d, or -1
__global__ void xchg_push_kernel(XchgPeers P, int world, int rank, int max_keys, uint32_t step, int n,
                                 const int32_t* __restrict__ in, unsigned int* __restrict__ done) {
  const size_t half = (size_t)(step & 1u) * (size_t)max_keys;
  for (int d = blockIdx.x * blockDim.x + threadIdx.x; d < n; d += gridDim.x * blockDim.x) {
    const int k = in[d];
    if (k < 0) continue;
    for (int r = 0; r < world; ++r) atomicMin_system(P.keys[r] + half + d, k);
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(done, 1u);
    if (prev == gridDim.x - 1) {  // every block's pushes are ordered before this point
      *done = 0;
      __threadfence_system();
      for (int r = 0; r < world; ++r) st_release_sys(P.flags[r] + rank, step);
    }
  }
}

__global__ void xchg_finish_kernel(int32_t* __restrict__ keys, const uint32_t* __restrict__ flags, int world, int max_keys,
                                   uint32_t step, int n, int32_t* __restrict__ out, unsigned long long timeout_ns,
                                   int* __restrict__ err) {
  if (threadIdx.x < world) {
    const unsigned long long t0 = globaltimer_ns();
    while ((int32_t)(ld_acquire_sys(flags + threadIdx.x) - step) < 0) {
      if (globaltimer_ns() - t0 > timeout_ns) {
        atomicExch(err, 1 + (int)threadIdx.x);
        break;
      }
      __nanosleep(200);
    }
  }
  __syncthreads();
  volatile int32_t* k = keys + (size_t)(step & 1u) * (size_t)max_keys;
  for (int d = blockIdx.x * blockDim.x + threadIdx.x; d < n; d += gridDim.x * blockDim.x) {
    const int v = k[d];
    out[d] = v == kNoNode ? -1 : v;
    k[d] = kNoNode;
  }
}

}  // namespace nsp
