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

}  // namespace nsp
