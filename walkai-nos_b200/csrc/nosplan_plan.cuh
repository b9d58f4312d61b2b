// nosplan_plan.cuh — the planner proper: Plan(snapshot, pendingPods) as ONE persistent CTA per
// partitioning kind (MIG and MPS chains touch disjoint nodes and profiles, so they run
// concurrently).  The pod order is a true dependency chain (each commit changes what the next pod
// sees), so one "sequencer" warp walks the pods; the other warps of the CTA are parked on a named
// barrier and are woken only for the data-parallel step — the first-node-wins search over nodes.
//
// Reference semantics replayed per pod:
//   gate        Controller.profileAlreadyPresent   mig_controller.go:121-144
//   repartition Controller.tryRepartition          mig_controller.go:146-166 (first node that updates)
//   schedule    Node.AddPod first-fit              mig/node.go:201-210, mig/gpu.go:289-303
//                                                  slicing/node.go:196-205, slicing/gpu.go:142-156
#pragma once
#include "nosplan_kernels.cuh"

namespace nsp {

constexpr int kPlanThreads = 1024;

struct PlanArgs {
  // pods of this kind, in planner order
  int M;
  const int32_t* pod_sig;  // [M] index into the demand table
  // demand table (deduplicated request vectors), CSR
  int D;
  const int32_t* dem_off;
  const int32_t* dem_key;
  const int32_t* dem_qty;
  int flags;
  // cluster
  int N, tg, n_keys;  // n_keys = P (MIG) or S (MPS)
  const int32_t* gpu_off;
  const uint8_t* valid;
  // MIG state (working copy)
  const MigTables* T;
  uint4* m_used;
  uint4* m_free;
  const uint8_t* m_model;
  // MPS state (working copy)
  const int32_t* sizes;
  MpsGpu* p_gpu;
  uint4* p_used_lo;
  uint4* p_used_hi;
  // scratch (global): first-fit bitmaps [n_keys+1][words] (last = GPUs of valid nodes), bounds [2][D]
  uint32_t* bitmaps;
  int words;
  int32_t* bounds;
  // outputs [M]
  uint8_t* gate_hit;
  int32_t* repart_node;
  int32_t* sched_node;
  int32_t* sched_gpu;
  const int32_t* skip_if;  // device flag: non-zero = the chain fast path owns this plan (may be null)
};

// Out-of-line copies of the node walks for the planner kernel.  Inlining the MPS walk into the
// sequencer let the compiler overlay the walk's scratch arrays with the caller's demand arrays
// (observed on nvcc 12.9: the caller's key[] read back as 0 after the call); a real call gives the
// walk its own frame.
__device__ __noinline__ bool plan_mps_walk(const int* sz, int S, MpsGpu* gp, int g0, int cnt, int k, const int* sid,
                                           const int* qty, bool commit) {
  return mps_walk_node<true>(sz, S, gp, g0, cnt, k, sid, qty, commit, nullptr, nullptr);
}
__device__ __noinline__ bool plan_mig_walk(const MigTables* T, const uint4* used, uint4* fr, const uint8_t* model, int g0,
                                           int cnt, int k, const int* pid, const int* qty, bool commit) {
  return mig_walk_node<true>(T, used, fr, model, g0, cnt, k, pid, qty, commit, nullptr);
}

__device__ __forceinline__ void bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

template <bool MPS>
__device__ __forceinline__ uint32_t plan_free_of(const PlanArgs& a, int g, int key) {
  if constexpr (MPS) {
    const MpsGpu* G = a.p_gpu + g;
    return mps_byte(__ldcg(&G->f0), __ldcg(&G->f1), key);
  } else {
    return byte_of(__ldcg(a.m_free + g), key);
  }
}
template <bool MPS>
__device__ __forceinline__ uint32_t plan_used_of(const PlanArgs& a, int g, int key) {
  if constexpr (MPS) {
    return mps_byte(__ldcg(a.p_used_lo + g), __ldcg(a.p_used_hi + g), key);
  } else {
    return byte_of(__ldcg(a.m_used + g), key);
  }
}

// add delta to byte `key` of a packed 16-byte map in global memory
__device__ __forceinline__ void bump_byte(uint4* v, int key, int delta) {
  uint32_t* w = reinterpret_cast<uint32_t*>(v) + (key >> 2);
  const int sh = (key & 3) * 8;
  const uint32_t old = __ldcg(w);
  const uint32_t b = (((old >> sh) & 0xffu) + (uint32_t)delta) & 0xffu;
  __stcg(w, (old & ~(0xffu << sh)) | (b << sh));
}

// GPU.AddPod on GPU g (mig/gpu.go:289-303, slicing/gpu.go:142-156): keys in ascending order; stops at
// the first key that does not fit, KEEPING the deductions already made.  Returns true if all fit.
// Maintains the first-fit bitmaps.
template <bool MPS>
__device__ inline bool plan_gpu_add_pod(const PlanArgs& a, int g, int k, const int* key, const int* qty) {
  for (int j = 0; j < k; ++j) {
    const int q = qty[j];
    const int f = (int)plan_free_of<MPS>(a, g, key[j]);
    if (f < q) return false;
    if constexpr (MPS) {
      MpsGpu* G = a.p_gpu + g;
      if (key[j] < 16) bump_byte(&G->f0, key[j], -q);
      else bump_byte(&G->f1, key[j] - 16, -q);
      if (key[j] < 16) bump_byte(a.p_used_lo + g, key[j], q);
      else bump_byte(a.p_used_hi + g, key[j] - 16, q);
      __stcg(&G->fmask, __ldcg(&G->fmask) | (1u << key[j]));  // FreeProfiles[r] -= q creates the key
      __stcg(&G->cap, __ldcg(&G->cap) - (uint32_t)(q * a.sizes[key[j]]));
    } else {
      bump_byte(a.m_free + g, key[j], -q);
      bump_byte(a.m_used + g, key[j], q);
    }
    if (f - q == 0 && q > 0) a.bitmaps[(size_t)key[j] * a.words + (g >> 5)] &= ~(1u << (g & 31));
  }
  return true;
}

template <bool MPS>
__global__ void __launch_bounds__(kPlanThreads, 1) plan_kernel(const PlanArgs a) {
  __shared__ int sh_present[kMaxSizes];  // Node.Geometry()[key] summed over valid nodes (gate)
  __shared__ int sh_sz[kMaxSizes];
  __shared__ int sh_cmd, sh_sig, sh_from, sh_best;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int NT = kPlanThreads;
  if (a.skip_if && *a.skip_if) return;

  // ---- init: presence counters, first-fit bitmaps, search bounds -------------------------------------
  if (tid < kMaxSizes) {
    sh_present[tid] = 0;
    sh_sz[tid] = (MPS && tid < a.n_keys) ? a.sizes[tid] : 1;
  }
  __syncthreads();
  for (int w = tid; w < a.words * (a.n_keys + 1); w += NT) a.bitmaps[w] = 0;
  for (int i = tid; i < 2 * a.D; i += NT) a.bounds[i] = 0;
  __syncthreads();
  for (int n = tid; n < a.N; n += NT) {
    if (!a.valid[n]) continue;
    for (int g = a.gpu_off[n]; g < a.gpu_off[n + 1]; ++g) {
      atomicOr(&a.bitmaps[(size_t)a.n_keys * a.words + (g >> 5)], 1u << (g & 31));
      for (int key = 0; key < a.n_keys; ++key) {
        const int f = (int)plan_free_of<MPS>(a, g, key), u = (int)plan_used_of<MPS>(a, g, key);
        if (f + u) atomicAdd(&sh_present[key], f + u);
        if (f > 0) atomicOr(&a.bitmaps[(size_t)key * a.words + (g >> 5)], 1u << (g & 31));
      }
    }
  }
  __threadfence_block();
  __syncthreads();

  int32_t* lb_rep = a.bounds;         // [D] no node below it can update for this demand
  int32_t* lb_fit = a.bounds + a.D;   // [D] no GPU below it can take this demand

  // the node search executed by every thread of the CTA when the sequencer asks for it
  auto scan_part = [&]() {
    const int sig = sh_sig;
    const int e0 = __ldg(a.dem_off + sig), k = __ldg(a.dem_off + sig + 1) - e0;
    int key[kMaxKeys], qty[kMaxKeys];
    for (int j = 0; j < kMaxKeys; ++j) {
      key[j] = j < k ? __ldg(a.dem_key + e0 + j) : 0;
      qty[j] = j < k ? __ldg(a.dem_qty + e0 + j) : 0;
    }
    for (int n = sh_from + tid; n < a.N; n += NT) {
      if (n > *(volatile int*)&sh_best) break;
      if (!a.valid[n]) continue;
      const int g0 = a.gpu_off[n], cnt = a.gpu_off[n + 1] - g0;
      bool u;
      if constexpr (MPS) u = plan_mps_walk(sh_sz, a.n_keys, a.p_gpu, g0, cnt, k, key, qty, false);
      else u = plan_mig_walk(a.T, a.m_used, a.m_free, a.m_model, g0, cnt, k, key, qty, false);
      if (u) {
        atomicMin(&sh_best, n);
        break;
      }
    }
  };

  if (warp != 0) {
    // helpers: parked until the sequencer needs a node search
    for (;;) {
      bar_sync(1, NT);
      if (sh_cmd == 0) break;
      scan_part();
      bar_sync(2, NT);
    }
    return;
  }

  // ---- sequencer warp ------------------------------------------------------------------------------------
  for (int i = 0; i < a.M; ++i) {
    const int sig = __ldg(a.pod_sig + i);
    const int e0 = __ldg(a.dem_off + sig), k = __ldg(a.dem_off + sig + 1) - e0;
    int key[kMaxKeys], qty[kMaxKeys];
    for (int j = 0; j < kMaxKeys; ++j) {
      key[j] = j < k ? __ldg(a.dem_key + e0 + j) : 0;
      qty[j] = j < k ? __ldg(a.dem_qty + e0 + j) : 0;
    }
    int out_gate = 0, out_rep = -1, out_node = -1, out_gpu = -1;
    if (k > 0 && a.N > 0) {  // mig_controller.go:68-71, :78-81
      // gate: every requested profile has used+free > 0 on some valid node
      bool present = false;
      if (a.flags & NOSPLAN_PLAN_GATE) {
        present = true;
        for (int j = 0; j < k; ++j) present &= sh_present[key[j]] > 0;
      }
      out_gate = present;
      if (!present) {
        int from = lb_rep[sig];
        int best = kNoNode;
        if (from < a.N) {
          if (lane == 0) {
            sh_cmd = 1;
            sh_sig = sig;
            sh_from = from;
            sh_best = kNoNode;
          }
          __syncwarp();
          bar_sync(1, NT);
          scan_part();
          bar_sync(2, NT);
          best = sh_best;
          __syncwarp();
        }
        if (best != kNoNode) {
          // commit on node `best`: lanes = key for the presence delta, lane 0 applies
          const int g0 = a.gpu_off[best], cnt = a.gpu_off[best + 1] - g0;
          int before = 0;
          if (lane < a.n_keys)
            for (int s = 0; s < cnt; ++s) before += (int)plan_free_of<MPS>(a, g0 + s, lane);
          __syncwarp();
          if (lane == 0) {
            if constexpr (MPS) plan_mps_walk(sh_sz, a.n_keys, a.p_gpu, g0, cnt, k, key, qty, true);
            else plan_mig_walk(a.T, a.m_used, a.m_free, a.m_model, g0, cnt, k, key, qty, true);
            __threadfence_block();
          }
          __syncwarp();
          if (lane < a.n_keys) {
            int after = 0;
            for (int s = 0; s < cnt; ++s) {
              const int g = g0 + s;
              const int f = (int)plan_free_of<MPS>(a, g, lane);
              after += f;
              uint32_t* w = &a.bitmaps[(size_t)lane * a.words + (g >> 5)];
              if (f > 0) *w |= 1u << (g & 31);
              else *w &= ~(1u << (g & 31));
            }
            sh_present[lane] += after - before;
          }
          __syncwarp();
          out_rep = best;
          // every cached bound at or above this node is stale
          for (int d = lane; d < a.D; d += 32) {
            lb_rep[d] = min(lb_rep[d], best);
            lb_fit[d] = min(lb_fit[d], g0);
          }
          __syncwarp();
          lb_rep[sig] = best;
        } else if (lane == 0) {
          lb_rep[sig] = a.N;
        }
        __syncwarp();
      }
      if (a.flags & NOSPLAN_PLAN_SCHED) {
        // first GPU (global GPU order == node order, then slot order) where every key fits
        int found = -1;
        for (int wbase = lb_fit[sig] >> 5; wbase < a.words && found < 0; wbase += 32) {
          const int w = wbase + lane;
          uint32_t bits = 0;
          if (w < a.words) {
            bits = a.bitmaps[(size_t)a.n_keys * a.words + w];
            for (int j = 0; j < k; ++j)
              if (qty[j] > 0) bits &= a.bitmaps[(size_t)key[j] * a.words + w];
          }
          int mine = -1;
          while (bits) {
            const int g = (w << 5) + __ffs(bits) - 1;
            bits &= bits - 1;
            bool fits = true;
            for (int j = 0; j < k; ++j)
              if (qty[j] > 1) fits &= (int)plan_free_of<MPS>(a, g, key[j]) >= qty[j];
            if (fits) {
              mine = g;
              break;
            }
          }
          const uint32_t hit = __ballot_sync(0xffffffffu, mine >= 0);
          if (hit) found = __shfl_sync(0xffffffffu, mine, __ffs(hit) - 1);
        }
        if (found >= 0) {
          // node of the winning GPU: binary search in gpu_off
          int lo = 0, hi = a.N - 1;
          while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (a.gpu_off[mid] <= found) lo = mid;
            else hi = mid - 1;
          }
          // skip empty nodes that share the offset
          while (a.gpu_off[lo + 1] <= found) ++lo;
          const int g0 = a.gpu_off[lo];
          if (lane == 0) {
            // Node.AddPod tries the node's GPUs in order: the ones before the winner fail, keeping
            // their partial deductions (multi-key pods only)
            if (k > 1)
              for (int g = g0; g < found; ++g) plan_gpu_add_pod<MPS>(a, g, k, key, qty);
            plan_gpu_add_pod<MPS>(a, found, k, key, qty);
            __threadfence_block();
          }
          __syncwarp();
          out_node = lo;
          out_gpu = found - g0;
          if (lane == 0) lb_fit[sig] = found;
          for (int d = lane; d < a.D; d += 32) lb_rep[d] = min(lb_rep[d], lo);
        } else if (lane == 0) {
          lb_fit[sig] = a.tg;
        }
        __syncwarp();
      }
    }
    if (lane == 0) {
      if (a.gate_hit) a.gate_hit[i] = (uint8_t)out_gate;
      if (a.repart_node) a.repart_node[i] = out_rep;
      if (a.sched_node) a.sched_node[i] = out_node;
      if (a.sched_gpu) a.sched_gpu[i] = out_gpu;
    }
  }
  if (lane == 0) sh_cmd = 0;
  __syncwarp();
  bar_sync(1, NT);
}



// =================================================================================================================
// Fast path of the planner for the reference's own behaviour (gate on, scheduling on) when every pod
// requests exactly one profile with quantity >= 1.
//
// If every requested profile already exists somewhere in the cluster (used+free > 0), the gate
// (mig_controller.go:121-144) stops every pod before tryRepartition, and it keeps doing so: AddPod
// moves devices from free to used and never changes used+free.  What is left is Node.AddPod
// first-fit, and a single-profile pod only reads and writes its own profile's counters — so the pods
// of different profiles are independent chains.  One warp walks each chain in pod order.  Within a
// chain "first GPU with free >= q" is monotone (free only decreases), so the warp keeps one cursor
// per quantity with the 32 GPUs under it cached in registers (lane = GPU), and each pod costs a
// ballot plus two fire-and-forget atomics.  Chains update different bytes of the same packed words,
// hence atomics.  plan_presence_kernel decides eligibility on the device; the general kernel runs
// (and this one does nothing) when a requested profile is absent.
struct ChainArgs {
  int n_chains;
  const int32_t* chain_key;   // [n_chains] profile / size id
  const int32_t* chain_off;   // [n_chains+1] into chain_pod / chain_q
  const int32_t* chain_pod;   // kind-local pod index, planner order
  const int32_t* chain_q;     // requested quantity (>= 1)
  int N, tg, n_keys;
  const int32_t* gpu_off;
  const int32_t* gpu_node;    // [tg] node of each GPU, -1 if the node is invalid
  uint4* m_used;
  uint4* m_free;
  const int32_t* sizes;
  MpsGpu* p_gpu;
  uint4* p_used_lo;
  uint4* p_used_hi;
  int32_t* present;           // [n_keys] sum of used+free over valid nodes (plan_presence_kernel)
  int32_t* eligible;          // [1] 1 = every chain's key is present -> this kernel owns the plan
  uint8_t* gate_hit;
  int32_t* repart_node;
  int32_t* sched_node;
  int32_t* sched_gpu;
};

template <bool MPS>
__global__ void plan_presence_kernel(ChainArgs a) {
  __shared__ int sh[kMaxSizes];
  if (threadIdx.x < kMaxSizes) sh[threadIdx.x] = 0;
  __syncthreads();
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g < a.tg && a.gpu_node[g] >= 0) {
    for (int key = 0; key < a.n_keys; ++key) {
      int v;
      if constexpr (MPS) v = (int)mps_byte(a.p_gpu[g].f0, a.p_gpu[g].f1, key) + (int)mps_byte(a.p_used_lo[g], a.p_used_hi[g], key);
      else v = (int)byte_of(a.m_free[g], key) + (int)byte_of(a.m_used[g], key);
      if (v) atomicAdd(&sh[key], v);
    }
  }
  __syncthreads();
  if (threadIdx.x < a.n_keys && sh[threadIdx.x]) atomicAdd(&a.present[threadIdx.x], sh[threadIdx.x]);
}

__global__ void plan_eligible_kernel(ChainArgs a) {
  // one warp: eligible iff every chain's key is present
  bool ok = true;
  for (int c = threadIdx.x; c < a.n_chains; c += 32) ok &= a.present[a.chain_key[c]] > 0;
  ok = __all_sync(0xffffffffu, ok);
  if (threadIdx.x == 0) *a.eligible = ok ? 1 : 0;
}

constexpr int kChainQ = 8;  // quantities 1..kChainQ each get a cursor; the host keeps larger ones off this path

template <bool MPS>
__global__ void __launch_bounds__(32) plan_chain_kernel(ChainArgs a) {
  if (*a.eligible == 0) return;
  const int c = blockIdx.x, lane = threadIdx.x;
  const int key = a.chain_key[c];
  const int j0 = a.chain_off[c], j1 = a.chain_off[c + 1];
  const int sz = MPS ? a.sizes[key] : 0;
  int base[kChainQ], cache[kChainQ];  // window start (multiple of 32) and this lane's GPU free count
  // read through an atomic so the value is ordered after this warp's earlier deductions at L2
  auto load_free = [&](int g) -> int {
    if (g >= a.tg || a.gpu_node[g] < 0) return 0;
    uint32_t* w;
    if constexpr (MPS) w = reinterpret_cast<uint32_t*>(&a.p_gpu[g].f0) + (key >> 2);
    else w = reinterpret_cast<uint32_t*>(a.m_free + g) + (key >> 2);
    return (int)((atomicOr(w, 0u) >> ((key & 3) * 8)) & 0xffu);
  };
#pragma unroll
  for (int qi = 0; qi < kChainQ; ++qi) {
    base[qi] = 0;
    cache[qi] = load_free(lane);
  }
  for (int j = j0; j < j1; ++j) {
    const int q = __ldg(a.chain_q + j);
    const int pod = __ldg(a.chain_pod + j);
    const int qi = q - 1;
    int found = -1;
#pragma unroll
    for (int t = 0; t < kChainQ; ++t) {
      if (t != qi) continue;
      while (base[t] < a.tg) {
        const uint32_t hit = __ballot_sync(0xffffffffu, cache[t] >= q);
        if (hit) {
          found = base[t] + __ffs(hit) - 1;
          break;
        }
        base[t] += 32;
        cache[t] = load_free(base[t] + lane);
      }
    }
    int node = -1, slot = -1;
    if (found >= 0) {
      node = a.gpu_node[found];
      slot = found - a.gpu_off[node];
      // every cursor whose window covers the GPU sees the deduction
#pragma unroll
      for (int t = 0; t < kChainQ; ++t)
        if (found - base[t] == lane) cache[t] -= q;
      if (lane == 0) {
        const int sh = (key & 3) * 8;
        if constexpr (MPS) {
          uint32_t* fw = reinterpret_cast<uint32_t*>(&a.p_gpu[found].f0) + (key >> 2);  // f0,f1 are contiguous
          uint32_t* uw = (key < 16 ? reinterpret_cast<uint32_t*>(a.p_used_lo + found) + (key >> 2)
                                   : reinterpret_cast<uint32_t*>(a.p_used_hi + found) + ((key - 16) >> 2));
          atomicSub(fw, (uint32_t)q << sh);
          atomicAdd(uw, (uint32_t)q << sh);
          atomicOr(&a.p_gpu[found].fmask, 1u << key);
          atomicSub(&a.p_gpu[found].cap, (uint32_t)(q * sz));
        } else {
          atomicSub(reinterpret_cast<uint32_t*>(a.m_free + found) + (key >> 2), (uint32_t)q << sh);
          atomicAdd(reinterpret_cast<uint32_t*>(a.m_used + found) + (key >> 2), (uint32_t)q << sh);
        }
      }
    }
    if (lane == 0) {
      a.gate_hit[pod] = 1;
      a.repart_node[pod] = -1;
      a.sched_node[pod] = node;
      a.sched_gpu[pod] = slot;
    }
  }
}

}  // namespace nsp
