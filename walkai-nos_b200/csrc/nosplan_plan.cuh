// nosplan_plan.cuh — the planner: argument carrier shared by both planner paths and the per-profile
// chain fast path.  MIG and MPS pods touch disjoint nodes and profiles, so the two kinds are planned
// concurrently on two streams.  The general path is in nosplan_seq.cuh.
//
// Reference semantics replayed per pod:
//   gate        Controller.profileAlreadyPresent   mig_controller.go:121-144
//   repartition Controller.tryRepartition          mig_controller.go:146-166 (first node that updates)
//   schedule    Node.AddPod first-fit              mig/node.go:201-210, mig/gpu.go:289-303
//                                                  slicing/node.go:196-205, slicing/gpu.go:142-156
#pragma once
#include "nosplan_kernels.cuh"

namespace nsp {

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
  const int32_t* gpu_slot;    // [tg] position of the GPU inside its node
  uint4* m_used;
  uint4* m_free;
  const int32_t* sizes;
  MpsGpu* p_gpu;
  uint4* p_used_lo;
  uint4* p_used_hi;
  int32_t* present;           // [n_keys] sum of used+free over valid nodes (plan_presence_kernel)
  int32_t* eligible;          // [1] 1 = every chain's key is present -> this kernel owns the plan
  uint8_t* winmax;            // [n_keys][n_win] largest free count of the key among the 32 GPUs of a window
  int n_win;                  //                 (plan_winmax_kernel; an upper bound while the chains run)
  int sched;                  // 0: gate only (NOSPLAN_PLAN_SCHED off): every pod is stopped by the gate, nothing else
  // The chains only own the pods the gate stops.  range[0..1] (device) = the kind-local pod indices [lo, hi) this
  // launch may place: hi is the first pod whose profile is absent (plan_find_miss_kernel); that pod goes through
  // tryRepartition on the sequencer, after which the chains continue behind it.  dead[key] != 0: the profile is
  // absent and its last tryRepartition found no node (nothing can change that before the next commit): its pods
  // get "not stopped by the gate, no node, not scheduled" without another search.
  const int32_t* range;
  const int32_t* dead;
  const int32_t* pod_sig;     // [M] request row of every pod of the kind (plan_find_miss_kernel)
  const int32_t* dem_off;
  const int32_t* dem_key;
  int M;
  int32_t* limit_out;         // [2] device: range of the next chain launch; plan_find_miss_kernel writes [1]
  int ignore_gate;            // batch-demand mode: plain first-fit for every pod of the range, whatever the gate would say
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

// One warp per window of 32 GPUs (the windows of plan_chain_kernel).  First-fit only ever takes free
// devices away, so the maxima stay upper bounds while the chains run: a window whose maximum is below the
// requested quantity can be skipped without loading it.
template <bool MPS>
__global__ void __launch_bounds__(32) plan_winmax_kernel(ChainArgs a) {
  const int w = blockIdx.x, lane = threadIdx.x;
  const int g = w * 32 + lane;
  const bool live = g < a.tg && a.gpu_node[g] >= 0;
  uint4 f0 = make_uint4(0, 0, 0, 0), f1 = f0;
  if (live) {
    if constexpr (MPS) {
      f0 = a.p_gpu[g].f0;
      f1 = a.p_gpu[g].f1;
    } else {
      f0 = a.m_free[g];
    }
  }
  for (int key = 0; key < a.n_keys; ++key) {
    const unsigned v = MPS ? mps_byte(f0, f1, key) : byte_of(f0, key);
    const unsigned m = __reduce_max_sync(0xffffffffu, v);
    if (lane == 0) a.winmax[(size_t)key * a.n_win + w] = (uint8_t)m;
  }
}

// limit_out[1] = the first pod at or after limit_out[0] that the gate does not stop and whose profile is not known to
// be out of reach (dead), M if there is none.  One block; limit_out[1] must hold M on entry.
__global__ void plan_find_miss_kernel(ChainArgs a) {
  const int lo = a.limit_out[0];
  int mine = a.M;
  for (int j = lo + threadIdx.x; j < a.M; j += blockDim.x) {
    const int key = __ldg(a.dem_key + __ldg(a.dem_off + __ldg(a.pod_sig + j)));
    if (a.present[key] <= 0 && !a.dead[key]) {
      mine = j;
      break;  // this thread's later pods are further down the order
    }
  }
  if (mine < a.M) atomicMin(a.limit_out + 1, mine);
}

constexpr int kChainQ = 8;  // quantities 1..kChainQ each get a cursor; the host keeps larger ones off this path

// NQ: cursors kept (quantities 1..NQ occur in the chains; 2 for every BASELINE config): the per-step
// bookkeeping over all cursors is what a single latency-bound warp spends its time on
template <bool MPS, int NQ>
__global__ void __launch_bounds__(32) plan_chain_kernel(ChainArgs a) {
  const int c = blockIdx.x, lane = threadIdx.x;
  const int key = a.chain_key[c];
  NSP_CHECK(key >= 0 && key < a.n_keys);
  int j0 = a.chain_off[c], j1 = a.chain_off[c + 1];
  {  // the chain's pods inside [range[0], range[1]): chain_pod is ascending
    const int lo = a.range[0], hi = a.range[1];
    int x0 = j0, x1 = j1;
    while (x0 < x1) {
      const int mid = (x0 + x1) >> 1;
      if (__ldg(a.chain_pod + mid) < lo) x0 = mid + 1;
      else x1 = mid;
    }
    int y0 = x0, y1 = j1;
    while (y0 < y1) {
      const int mid = (y0 + y1) >> 1;
      if (__ldg(a.chain_pod + mid) < hi) y0 = mid + 1;
      else y1 = mid;
    }
    j0 = x0;
    j1 = y0;
  }
  if (j0 >= j1) return;
  if (!a.ignore_gate && a.present[key] <= 0) {
    // only reachable for a dead profile (plan_find_miss_kernel stops the range at any other absent one): the gate
    // does not stop the pod (mig_controller.go:83-86), tryRepartition finds no node, nothing of it is free
    for (int j = j0 + lane; j < j1; j += 32) {
      const int pod = a.chain_pod[j];
      a.gate_hit[pod] = 0;
      a.repart_node[pod] = -1;
      a.sched_node[pod] = -1;
      a.sched_gpu[pod] = -1;
    }
    return;
  }
  if (!a.sched) {
    for (int j = j0 + lane; j < j1; j += 32) {
      const int pod = a.chain_pod[j];
      a.gate_hit[pod] = 1;
      a.repart_node[pod] = -1;
      a.sched_node[pod] = -1;
      a.sched_gpu[pod] = -1;
    }
    return;
  }
  const int sz = MPS ? a.sizes[key] : 0;
  int base[NQ], cache[NQ];  // window start (multiple of 32) and this lane's GPU free count
  int wnode[NQ], wslot[NQ];  // node and slot of this lane's GPU in each window (-1: not a valid GPU)
  // read through an atomic so the value is ordered after this warp's earlier deductions at L2
  auto load_win = [&](int g, int* node, int* slot) -> int {
    *node = -1;
    *slot = -1;
    if (g >= a.tg) return 0;
    uint32_t* w;
    if constexpr (MPS) w = reinterpret_cast<uint32_t*>(&a.p_gpu[g].f0) + (key >> 2);
    else w = reinterpret_cast<uint32_t*>(a.m_free + g) + (key >> 2);
    // three independent loads (one L2 latency per window); rows of invalid nodes are zero
    const uint32_t word = atomicOr(w, 0u);
    const int n = __ldg(a.gpu_node + g);
    const int sl = __ldg(a.gpu_slot + g);
    if (n < 0) return 0;
    *node = n;
    *slot = sl;
    return (int)((word >> ((key & 3) * 8)) & 0xffu);
  };
#pragma unroll
  for (int qi = 0; qi < NQ; ++qi) {
    base[qi] = 0;
    cache[qi] = load_win(lane, &wnode[qi], &wslot[qi]);
  }
  for (int jb = j0; jb < j1; jb += 32) {
    // the next 32 pods of the chain: one coalesced load, then a shuffle per pod
    const int jl = jb + lane;
    const int q32 = jl < j1 ? __ldg(a.chain_q + jl) : 1;
    const int pod32 = jl < j1 ? __ldg(a.chain_pod + jl) : 0;
    const int nj = min(32, j1 - jb);
    // Runs of pods asking for the same quantity are placed together: GPU l of the window can take
    // floor(free/q) of them, an inclusive warp prefix sum of those capacities tells every pod of the
    // run (one per lane) which GPU it lands on — first-fit in pod order, 32 pods per step.
    int t32 = 0;
    while (t32 < nj) {
      const int q = __shfl_sync(0xffffffffu, q32, t32);
      const int qi = q - 1;
      const uint32_t same = __ballot_sync(0xffffffffu, lane >= t32 && lane < nj && q32 == q);
      const uint32_t rest = ~(same >> t32);
      const int r = rest ? __ffs(rest) - 1 : 32 - t32;  // run length (>= 1)
#pragma unroll
      for (int t = 0; t < NQ; ++t) {
        if (t != qi) continue;
        int served = 0;
        while (served < r) {
          const int jl0 = t32 + served;  // lane of the first pod still to place
          if (base[t] >= a.tg) {         // no GPU left for this quantity: the rest of the run stays pending
            if (lane >= jl0 && lane < t32 + r) {
              a.gate_hit[pod32] = 1;
              a.repart_node[pod32] = -1;
              a.sched_node[pod32] = -1;
              a.sched_gpu[pod32] = -1;
            }
            break;
          }
          const int capl = q == 1 ? cache[t] : cache[t] / q;
          int P = capl;  // inclusive prefix sum over the window's GPUs
#pragma unroll
          for (int off = 1; off < 32; off <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, P, off);
            if (lane >= off) P += v;
          }
          const int tot = __shfl_sync(0xffffffffu, P, 31);
          if (tot == 0) {  // window exhausted for this quantity: every GPU before the next one has free < q
            // next window that may still hold free >= q: 32 window maxima per load (1024 GPUs per L2 round trip
            // instead of 32); the maxima are upper bounds, the window itself is then loaded and decides
            int w = base[t] / 32 + 1;
            const uint8_t* wm = a.winmax + (size_t)key * a.n_win;
            while (w < a.n_win) {
              const int mx = (w + lane < a.n_win) ? (int)__ldg(wm + w + lane) : 0;
              const uint32_t hit = __ballot_sync(0xffffffffu, mx >= q);
              if (hit) {
                w += __ffs(hit) - 1;
                break;
              }
              w += 32;
            }
            base[t] = min(w, a.n_win) * 32;  // n_win * 32 >= tg: nothing left for this quantity
            cache[t] = load_win(base[t] + lane, &wnode[t], &wslot[t]);
            continue;
          }
          const int m = min(r - served, tot);
          const int j = lane - jl0;  // this lane's pod is the j-th of the run still to place
          const bool mine = j >= 0 && j < m;
          int lo = 0;  // number of GPUs whose inclusive prefix is <= j  ==  index of the GPU that takes pod j
#pragma unroll
          for (int step = 16; step >= 1; step >>= 1) {
            const int pv = __shfl_sync(0xffffffffu, P, (lo + step - 1) & 31);
            if (pv <= j) lo += step;
          }
          lo = min(lo, 31);
          NSP_CHECK(!mine || (pod32 >= 0 && pod32 < a.M));
          const int node = __shfl_sync(0xffffffffu, wnode[t], lo);
          const int slot = __shfl_sync(0xffffffffu, wslot[t], lo);
          if (mine) {
            a.gate_hit[pod32] = 1;
            a.repart_node[pod32] = -1;
            a.sched_node[pod32] = node;
            a.sched_gpu[pod32] = slot;
          }
          // what each GPU of the window gives up
          const int take = min(max(m - (P - capl), 0), capl);
          const int dq = take * q;
          const int wb = base[t];
#pragma unroll
          for (int t2 = 0; t2 < NQ; ++t2)
            if (base[t2] == wb) cache[t2] -= dq;  // windows are 32-aligned: same base == same GPUs
          if (dq > 0) {
            const int g = wb + lane;
            const int sh = (key & 3) * 8;
            if constexpr (MPS) {
              uint32_t* fw = reinterpret_cast<uint32_t*>(&a.p_gpu[g].f0) + (key >> 2);  // f0,f1 are contiguous
              uint32_t* uw = (key < 16 ? reinterpret_cast<uint32_t*>(a.p_used_lo + g) + (key >> 2)
                                       : reinterpret_cast<uint32_t*>(a.p_used_hi + g) + ((key - 16) >> 2));
              atomicSub(fw, (uint32_t)dq << sh);
              atomicAdd(uw, (uint32_t)dq << sh);
              atomicOr(&a.p_gpu[g].fmask, 1u << key);
              atomicSub(&a.p_gpu[g].cap, (uint32_t)(dq * sz));
            } else {
              atomicSub(reinterpret_cast<uint32_t*>(a.m_free + g) + (key >> 2), (uint32_t)dq << sh);
              atomicAdd(reinterpret_cast<uint32_t*>(a.m_used + g) + (key >> 2), (uint32_t)dq << sh);
            }
          }
          served += m;
        }
      }
      t32 += r;
    }
  }
}

// ---- batch-demand mode (NEW BEHAVIOUR, oracle/batch_plan.py): commit step of one round -------------------------------
// The batched scorer has scored the rows {key: lacking[key]} against every node of the round's starting state
// (updated[D][N] and, per GPU, the geometry it would apply / the slice layout it would leave).  One warp now walks the
// rows in the round's order and, per row, the nodes in list order: a node whose trial update succeeded and that no row
// took yet in this round is committed while the row's demand is not covered by the free devices of the key on the nodes
// committed before it.  32 nodes per step: the running demand is an exclusive prefix sum over the candidates' gains.
struct BatchCommitArgs {
  int D, N, tg, n_keys;
  const int32_t* row_key;      // [D] in the round's order
  const int32_t* row_index;    // [D] row of the scorer outputs
  const int32_t* row_lacking;  // [D]
  const uint8_t* updated;      // [D][N]
  const int8_t* choice;        // MIG [D][tg]
  const uint8_t* out_free;     // MPS [D][tg][S]
  const uint32_t* out_mask;    // MPS [D][tg]
  const int32_t* gpu_off;
  const MigTables* T;
  const uint8_t* m_model;
  uint4* m_used;
  uint4* m_free;
  MpsGpu* p_gpu;
  uint8_t* taken;              // [N] zero on entry
  uint8_t* repartitioned;      // [N] accumulated over the rounds
};

template <bool MPS>
__global__ void __launch_bounds__(32) batch_commit_kernel(const BatchCommitArgs a) {
  const int lane = threadIdx.x;
  for (int r = 0; r < a.D; ++r) {
    const int key = a.row_key[r], row = a.row_index[r];
    int rem = a.row_lacking[r];
    for (int base = 0; base < a.N && rem > 0; base += 32) {
      const int n = base + lane;
      bool cand = false;
      int gain = 0, g0 = 0, g1 = 0;
      if (n < a.N) {
        cand = a.updated[(size_t)row * a.N + n] != 0 && !a.taken[n];
        if (cand) {
          g0 = a.gpu_off[n];
          g1 = a.gpu_off[n + 1];
          for (int g = g0; g < g1; ++g) {  // free devices of the key on the node after the trial update
            if constexpr (MPS) {
              gain += (int)a.out_free[((size_t)row * a.tg + g) * a.n_keys + key];
            } else {
              const int c = a.choice[(size_t)row * a.tg + g];
              if (c >= 0) {
                const int m = a.m_model[g];
                gain += (int)byte_of(a.T->cand[a.T->geom_off[m] + c], key) - (int)byte_of(a.m_used[g], key);
              } else {
                gain += (int)byte_of(a.m_free[g], key);
              }
            }
          }
        }
      }
      int incl = cand ? gain : 0;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      const int excl = incl - (cand ? gain : 0);
      const bool commit = cand && rem - excl > 0;
      if (commit) {
        for (int g = g0; g < g1; ++g) {
          if constexpr (MPS) {
            const uint8_t* f = a.out_free + ((size_t)row * a.tg + g) * a.n_keys;
            uint32_t w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            for (int s = 0; s < a.n_keys; ++s) w[s >> 2] |= (uint32_t)f[s] << ((s & 3) * 8);
            a.p_gpu[g].f0 = make_uint4(w[0], w[1], w[2], w[3]);
            a.p_gpu[g].f1 = make_uint4(w[4], w[5], w[6], w[7]);
            a.p_gpu[g].fmask = a.out_mask[(size_t)row * a.tg + g];
          } else {
            const int c = a.choice[(size_t)row * a.tg + g];
            if (c >= 0) a.m_free[g] = sub4(a.T->cand[a.T->geom_off[a.m_model[g]] + c], a.m_used[g]);  // ApplyGeometry
          }
        }
        a.taken[n] = 1;
        a.repartitioned[n] = 1;
      }
      // what the committed nodes of this step cover: the gains of the candidates up to the last committed one
      const uint32_t cm = __ballot_sync(0xffffffffu, commit);
      if (cm) rem -= __shfl_sync(0xffffffffu, incl, 31 - __clz(cm));
      __syncwarp();
    }
  }
}

}  // namespace nsp
