// nosplan_seq.cuh — the general (sequential) planner: Plan(snapshot, pendingPods) for any flags and any
// request vectors.  The pod order is a dependency chain, so one warp (the "sequencer") replays
// Controller.Reconcile pod by pod; what makes it fast is that it never walks nodes to answer the two
// searches of a reconcile — it keeps small derived structures and only re-derives the node a commit
// touched:
//
//   per GPU   gm[g]   mask of keys this GPU could still provide more of:
//                     MIG  some applicable geometry has cand[p] > used[p]+free[p]   (then
//                          countProvidedProfiles > 0 for any positive need, gpu.go:198-223)
//                     MPS  cap = MemoryGB - sum(size*used) >= size[s]  (createSlices can succeed after the
//                          free slices are deleted, slicing/gpu.go:188-212)
//             f1[g]   mask of keys with free >= 1
//   per node  NU1     exact "UpdateGeometryFor({key: 1}) returns true": some GPU can grow the key and no
//                     GPU up to and including it holds a free one (residual rule, node.go:161-170)
//             NB      superset for any quantity / several keys: some GPU can grow the key
//   transposed, in shared memory: NU1t/NBt[key] (bit per node), FBt[key] (bit per GPU, first-fit).
//
// tryRepartition (mig_controller.go:146-166) = first set bit of NU1t (quantity 1) or first NBt candidate
// that passes the residual walk; Node.AddPod first-fit (mig/node.go:201-210) = first set bit of the AND
// of FBt rows.  A commit loads the node into shared memory, mutates it there (warp-cooperative
// candidate evaluation for MIG: lanes = allowed geometries), writes it back and re-derives its masks.
#pragma once
#include "nosplan_plan.cuh"

namespace nsp {

constexpr int kSeqThreads = 256;

// The MPS per-GPU update as the sequencer calls it.  Round 1 kept this out of line (__noinline__) because an inlined build
// had read the caller's request vector back as zeros after the call and blamed the compiler.  Round 2 looked for the cause
// instead: with every local and shared array index asserted on the device (-DNOSPLAN_CHECKS) no access leaves its array in
// any parity, limit, stress or full-size test, and the inlined build — with and without those assertions — passes the same
// suites bit for bit (profiles/r2_sanitizer.md), so the call is inlined again.  -DNOSPLAN_SEQ_NOINLINE restores the call.
#ifdef NOSPLAN_SEQ_NOINLINE
__device__ __noinline__
#else
__device__ __forceinline__
#endif
bool seq_mps_gpu_update(const int* sz, uint4 f0, uint4 f1, uint32_t origmask, uint32_t cap, int k,
                                                const int* sid, const int* resid, int* created, uint32_t* restored) {
  return mps_gpu_update(sz, f0, f1, origmask, cap, k, sid, resid, created, restored);
}
constexpr int kNodeWords = 18;  // shared-memory image of one GPU: MIG used[4] free[4]; MPS free[8] used[8] fmask cap

struct SeqArgs {
  int M;
  const int32_t* pod_sig;
  int D;
  const int32_t* dem_off;
  const int32_t* dem_key;
  const int32_t* dem_qty;
  int flags;
  int N, tg, n_keys;
  const int32_t* gpu_off;
  const uint8_t* valid;
  const int32_t* gpu_node;
  const MigTables* T;
  uint4* m_used;
  uint4* m_free;
  const uint8_t* m_model;
  const int32_t* sizes;
  MpsGpu* p_gpu;
  uint4* p_used_lo;
  uint4* p_used_hi;
  uint32_t* gm;       // [tg]
  uint32_t* f1;       // [tg]
  uint32_t* nbt;      // [n_keys][nw]   (global copy, built by the prep kernels)
  uint32_t* nu1t;     // [n_keys][nw]
  uint32_t* fbt;      // [n_keys+1][gw] last row: GPUs of valid nodes
  int32_t* present;   // [n_keys]
  int nw, gw;
  int use_smem;
  uint8_t* gate_hit;
  int32_t* repart_node;
  int32_t* sched_node;
  int32_t* sched_gpu;
  const int32_t* skip_if;
  // hybrid planning (the chains own the pods the gate stops, see ChainArgs): the profile counts after these pods, and
  // per key "absent and no node could be repartitioned for it" (cleared by any commit, which may create any profile)
  int32_t* present_out;
  int32_t* dead;
};

// ---- derived masks of one GPU ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t nonzero_bytes16(const uint4& v) {
  uint32_t m = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) m |= (byte_of(v, i) != 0u) ? (1u << i) : 0u;
  return m;
}
// keys p with a[p] > b[p]
__device__ __forceinline__ uint32_t gt_bytes16(const uint4& a, const uint4& b) {
  uint32_t m = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) m |= (byte_of(a, i) > byte_of(b, i)) ? (1u << i) : 0u;
  return m;
}
// MIG, one thread: mask of profiles some applicable geometry would add to
__device__ inline uint32_t mig_grow_mask(const MigTables* __restrict__ T, int model, const uint4& u, const uint4& f) {
  const uint4 cur = add4(u, f);
  uint32_t m = 0;
  for (int c = T->geom_off[model]; c < T->geom_off[model + 1]; ++c) {
    const uint4 cd = T->cand[c];
    if (all_ge4(cd, u)) m |= gt_bytes16(cd, cur);
  }
  return m;
}
// MPS: sizes are ascending, so the keys that fit in cap are a prefix
__device__ __forceinline__ uint32_t mps_cap_mask(const int* sz, int S, uint32_t cap) {
  int c = 0;
  while (c < S && (uint32_t)sz[c] <= cap) ++c;
  return c >= 32 ? 0xffffffffu : ((1u << c) - 1u);
}
__device__ __forceinline__ uint32_t mps_free1_mask(const uint4& f0, const uint4& f1) {
  return nonzero_bytes16(f0) | (nonzero_bytes16(f1) << 16);
}

// ---- prep kernels (whole grid): derive everything from the working state ---------------------------------------
template <bool MPS>
__global__ void seq_prep_gpu_kernel(SeqArgs a) {
  __shared__ int sh[kMaxSizes];
  __shared__ int s_sz[kMaxSizes];
  if (a.skip_if && *a.skip_if) return;
  if (threadIdx.x < kMaxSizes) {
    sh[threadIdx.x] = 0;
    s_sz[threadIdx.x] = (MPS && threadIdx.x < a.n_keys) ? a.sizes[threadIdx.x] : 1;
  }
  __syncthreads();
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g < a.tg) {
    uint32_t gm = 0, f1 = 0;
    if (a.gpu_node[g] >= 0) {
      if constexpr (MPS) {
        const MpsGpu G = a.p_gpu[g];
        gm = mps_cap_mask(s_sz, a.n_keys, G.cap);
        f1 = mps_free1_mask(G.f0, G.f1);
        for (int key = 0; key < a.n_keys; ++key) {
          const int v = (int)mps_byte(G.f0, G.f1, key) + (int)mps_byte(a.p_used_lo[g], a.p_used_hi[g], key);
          if (v) atomicAdd(&sh[key], v);
        }
      } else {
        const uint4 u = a.m_used[g], f = a.m_free[g];
        gm = mig_grow_mask(a.T, a.m_model[g], u, f);
        f1 = nonzero_bytes16(f);
        for (int key = 0; key < a.n_keys; ++key) {
          const int v = (int)byte_of(u, key) + (int)byte_of(f, key);
          if (v) atomicAdd(&sh[key], v);
        }
      }
      atomicOr(&a.fbt[(size_t)a.n_keys * a.gw + (g >> 5)], 1u << (g & 31));
      for (uint32_t mk = f1; mk; mk &= mk - 1) atomicOr(&a.fbt[(size_t)(__ffs(mk) - 1) * a.gw + (g >> 5)], 1u << (g & 31));
    }
    a.gm[g] = gm;
    a.f1[g] = f1;
  }
  __syncthreads();
  if (threadIdx.x < a.n_keys && sh[threadIdx.x]) atomicAdd(&a.present[threadIdx.x], sh[threadIdx.x]);
}

// node masks from the per-GPU masks (gm/f1 of the node's GPUs, in GPU order)
__device__ __forceinline__ void node_masks(const uint32_t* gm, const uint32_t* f1, int cnt, uint32_t* nb, uint32_t* nu1) {
  uint32_t any = 0, exact = 0, blocked = 0;
  for (int s = 0; s < cnt; ++s) {
    any |= gm[s];
    blocked |= f1[s];            // a free device on this GPU (or an earlier one) covers a quantity-1 request
    exact |= gm[s] & ~blocked;
  }
  *nb = any;
  *nu1 = exact;
}

__global__ void seq_prep_node_kernel(SeqArgs a) {
  if (a.skip_if && *a.skip_if) return;
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= a.N || !a.valid[n]) return;
  const int g0 = a.gpu_off[n], cnt = a.gpu_off[n + 1] - g0;
  uint32_t nb, nu1;
  node_masks(a.gm + g0, a.f1 + g0, cnt, &nb, &nu1);
  for (uint32_t mk = nb; mk; mk &= mk - 1) atomicOr(&a.nbt[(size_t)(__ffs(mk) - 1) * a.nw + (n >> 5)], 1u << (n & 31));
  for (uint32_t mk = nu1; mk; mk &= mk - 1) atomicOr(&a.nu1t[(size_t)(__ffs(mk) - 1) * a.nw + (n >> 5)], 1u << (n & 31));
}

// ---- the sequencer ---------------------------------------------------------------------------------------------
template <bool MPS>
__global__ void __launch_bounds__(kSeqThreads, 1) plan_seq_kernel(const SeqArgs a) {
  extern __shared__ __align__(16) uint32_t dyn[];
  __shared__ int sh_present[kMaxSizes];
  __shared__ int sh_sz[kMaxSizes];
  __shared__ int sh_hint[kMaxSizes];                                   // first FBt word that may hold a bit, per key
  __shared__ uint32_t sh_node[NOSPLAN_MAX_GPUS_PER_NODE][kNodeWords];  // image of the node being mutated
  __shared__ uint32_t sh_gm[NOSPLAN_MAX_GPUS_PER_NODE], sh_f1[NOSPLAN_MAX_GPUS_PER_NODE];
  __shared__ uint8_t sh_model[NOSPLAN_MAX_GPUS_PER_NODE];
  __shared__ int key[kMaxKeys], qty[kMaxKeys];  // request vector of the pod being replayed (uniform)
  if (a.skip_if && *a.skip_if) return;
  const int tid = threadIdx.x, lane = tid & 31;
  const int K = a.n_keys;
  uint32_t* nbt = a.use_smem ? dyn : a.nbt;
  uint32_t* nu1t = a.use_smem ? dyn + (size_t)K * a.nw : a.nu1t;
  uint32_t* fbt = a.use_smem ? dyn + (size_t)2 * K * a.nw : a.fbt;
  if (a.use_smem) {
    for (int i = tid; i < K * a.nw; i += kSeqThreads) {
      nbt[i] = a.nbt[i];
      nu1t[i] = a.nu1t[i];
    }
    for (int i = tid; i < (K + 1) * a.gw; i += kSeqThreads) fbt[i] = a.fbt[i];
  }
  if (tid < kMaxSizes) {
    sh_present[tid] = tid < K ? a.present[tid] : 0;
    sh_sz[tid] = (MPS && tid < K) ? a.sizes[tid] : 1;
    sh_hint[tid] = 0;
  }
  __syncthreads();
  if (tid >= 32) return;  // one warp replays the pods

  // -- node image helpers (all 32 lanes call them) ---------------------------------------------------------------
  auto load_node = [&](int g0, int cnt) {
    if (lane < cnt) {
      const int g = g0 + lane;
      if constexpr (MPS) {
        const uint4 f0 = __ldcg(&a.p_gpu[g].f0), f1 = __ldcg(&a.p_gpu[g].f1);
        const uint4 u0 = __ldcg(a.p_used_lo + g), u1 = __ldcg(a.p_used_hi + g);
        uint32_t* r = sh_node[lane];
        r[0] = f0.x; r[1] = f0.y; r[2] = f0.z; r[3] = f0.w; r[4] = f1.x; r[5] = f1.y; r[6] = f1.z; r[7] = f1.w;
        r[8] = u0.x; r[9] = u0.y; r[10] = u0.z; r[11] = u0.w; r[12] = u1.x; r[13] = u1.y; r[14] = u1.z; r[15] = u1.w;
        r[16] = __ldcg(&a.p_gpu[g].fmask);
        r[17] = __ldcg(&a.p_gpu[g].cap);
      } else {
        const uint4 u = __ldcg(a.m_used + g), f = __ldcg(a.m_free + g);
        uint32_t* r = sh_node[lane];
        r[0] = u.x; r[1] = u.y; r[2] = u.z; r[3] = u.w; r[4] = f.x; r[5] = f.y; r[6] = f.z; r[7] = f.w;
        sh_model[lane] = __ldg(a.m_model + g);
      }
      sh_gm[lane] = a.gm[g];
      sh_f1[lane] = a.f1[g];
    }
    __syncwarp();
  };
  auto store_node = [&](int g0, int cnt, uint32_t changed) {
    if (lane < cnt && ((changed >> lane) & 1u)) {
      const int g = g0 + lane;
      const uint32_t* r = sh_node[lane];
      if constexpr (MPS) {
        __stcg(&a.p_gpu[g].f0, make_uint4(r[0], r[1], r[2], r[3]));
        __stcg(&a.p_gpu[g].f1, make_uint4(r[4], r[5], r[6], r[7]));
        __stcg(a.p_used_lo + g, make_uint4(r[8], r[9], r[10], r[11]));
        __stcg(a.p_used_hi + g, make_uint4(r[12], r[13], r[14], r[15]));
        __stcg(&a.p_gpu[g].fmask, r[16]);
        __stcg(&a.p_gpu[g].cap, r[17]);
      } else {
        __stcg(a.m_used + g, make_uint4(r[0], r[1], r[2], r[3]));
        __stcg(a.m_free + g, make_uint4(r[4], r[5], r[6], r[7]));
      }
    }
    __syncwarp();
  };
  // byte `key` of the image's free / used map of GPU slot s
  auto img_free = [&](int s, int key) -> int {
    return (int)reinterpret_cast<const uint8_t*>(sh_node[s])[(MPS ? 0 : 16) + key];
  };
  auto img_used = [&](int s, int key) -> int {
    return (int)reinterpret_cast<const uint8_t*>(sh_node[s])[(MPS ? 32 : 0) + key];
  };
  auto img_add = [&](int s, int base, int key, int delta) {  // one lane only
    uint32_t& w = sh_node[s][base + (key >> 2)];
    const int sh = (key & 3) * 8;
    const uint32_t b = (((w >> sh) & 0xffu) + (uint32_t)delta) & 0xffu;
    w = (w & ~(0xffu << sh)) | (b << sh);
  };
  // lane = key: sum over the node image of used+free (Node.Geometry()[key])
  auto node_sum = [&](int cnt) -> int {
    int sum = 0;
    if (lane < K)
      for (int s = 0; s < cnt; ++s) sum += img_free(s, lane) + img_used(s, lane);
    return sum;
  };
  // re-derive the masks and bitmap bits of node n from its image
  auto refresh_node = [&](int n, int g0, int cnt, uint32_t changed) {
    if constexpr (MPS) {
      if (lane < cnt && ((changed >> lane) & 1u)) {
        const uint32_t* r = sh_node[lane];
        sh_gm[lane] = mps_cap_mask(sh_sz, K, r[17]);
        sh_f1[lane] = mps_free1_mask(make_uint4(r[0], r[1], r[2], r[3]), make_uint4(r[4], r[5], r[6], r[7]));
      }
    } else {
      // lanes = allowed geometries of the GPU's model
      for (uint32_t mk = changed; mk; mk &= mk - 1) {
        const int s = __ffs(mk) - 1;
        const uint32_t* r = sh_node[s];
        const uint4 u = make_uint4(r[0], r[1], r[2], r[3]), f = make_uint4(r[4], r[5], r[6], r[7]);
        const int m = sh_model[s];
        const int c0 = a.T->geom_off[m], C = a.T->geom_off[m + 1] - c0;
        uint32_t mine = 0;
        if (lane < C) {
          const uint4 cd = a.T->cand[c0 + lane];
          if (all_ge4(cd, u)) mine = gt_bytes16(cd, add4(u, f));
        }
        const uint32_t gmask = __reduce_or_sync(0xffffffffu, mine);
        if (lane == 0) {
          sh_gm[s] = gmask;
          sh_f1[s] = nonzero_bytes16(f);
        }
      }
    }
    __syncwarp();
    if (lane < cnt && ((changed >> lane) & 1u)) {
      a.gm[g0 + lane] = sh_gm[lane];
      a.f1[g0 + lane] = sh_f1[lane];
    }
    uint32_t nb, nu1;
    node_masks(sh_gm, sh_f1, cnt, &nb, &nu1);
    if (lane < K) {
      const uint32_t bit = 1u << (n & 31);
      uint32_t* w = &nbt[(size_t)lane * a.nw + (n >> 5)];
      *w = ((nb >> lane) & 1u) ? (*w | bit) : (*w & ~bit);
      w = &nu1t[(size_t)lane * a.nw + (n >> 5)];
      *w = ((nu1 >> lane) & 1u) ? (*w | bit) : (*w & ~bit);
      for (uint32_t mk = changed; mk; mk &= mk - 1) {
        const int s = __ffs(mk) - 1;
        const int g = g0 + s;
        const uint32_t gb = 1u << (g & 31);
        uint32_t* fw = &fbt[(size_t)lane * a.gw + (g >> 5)];
        if ((sh_f1[s] >> lane) & 1u) {
          *fw |= gb;
          sh_hint[lane] = min(sh_hint[lane], g >> 5);
        } else {
          *fw &= ~gb;
        }
      }
    }
    __syncwarp();
  };
  // the residual walk of Node.UpdateGeometryFor reduced to "does any GPU update": one lane, node from global
  auto node_would_update = [&](int n, int k, const int* key, const int* qty) -> bool {
    const int g0 = a.gpu_off[n], cnt = a.gpu_off[n + 1] - g0;
    int resid[kMaxKeys];
    for (int j = 0; j < kMaxKeys; ++j) resid[j] = j < k ? qty[j] : 0;
    for (int s = 0; s < cnt; ++s) {
      const int g = g0 + s;
      const uint32_t gmask = a.gm[g];
      uint4 fa, fb;
      if constexpr (MPS) {
        fa = __ldcg(&a.p_gpu[g].f0);
        fb = __ldcg(&a.p_gpu[g].f1);
      } else {
        fa = __ldcg(a.m_free + g);
        fb = fa;
      }
      bool left = false;
      for (int j = 0; j < k; ++j) {
        const int fj = MPS ? (int)mps_byte(fa, fb, key[j]) : (int)byte_of(fa, key[j]);
        if (resid[j] - fj > 0) {
          // MIG: any missing key the GPU can grow.  MPS: the smallest missing size fits in cap; the keys
          // are ascending sizes and gm is a prefix mask, so testing the first missing key is the same.
          if ((gmask >> key[j]) & 1u) return true;
          if (MPS) break;
        }
      }
      for (int j = 0; j < k; ++j) {
        const int fj = MPS ? (int)mps_byte(fa, fb, key[j]) : (int)byte_of(fa, key[j]);
        resid[j] = max(resid[j] - fj, 0);
        left |= resid[j] > 0;
      }
      if (!left) return false;
    }
    return false;
  };

#ifdef NOSPLAN_SEQ_PROFILE
  long long tp[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  long long t0 = clock64();
#define SEQ_TICK(i) { long long t1 = clock64(); tp[i] += t1 - t0; t0 = t1; }
#else
#define SEQ_TICK(i)
#endif
  int e0_32 = 0, k_32 = 0;
  for (int i = 0; i < a.M; ++i) {
    if ((i & 31) == 0) {  // the next 32 pods: one coalesced load, then a shuffle per pod
      const int il = i + lane;
      const int sg = il < a.M ? __ldg(a.pod_sig + il) : 0;
      e0_32 = __ldg(a.dem_off + sg);
      k_32 = __ldg(a.dem_off + sg + 1) - e0_32;
    }
    const int e0 = __shfl_sync(0xffffffffu, e0_32, i & 31), k = __shfl_sync(0xffffffffu, k_32, i & 31);
    __syncwarp();
    NSP_CHECK(k >= 0 && k <= kMaxKeys);
    if (lane < kMaxKeys) {
      key[lane] = lane < k ? __ldg(a.dem_key + e0 + lane) : 0;
      qty[lane] = lane < k ? __ldg(a.dem_qty + e0 + lane) : 0;
      NSP_CHECK(key[lane] >= 0 && key[lane] < K);
    }
    __syncwarp();
    uint32_t reqmask = 0;   // keys with a positive quantity
    for (int j = 0; j < k; ++j)
      if (qty[j] > 0) reqmask |= 1u << key[j];
    int out_gate = 0, out_rep = -1, out_node = -1, out_gpu = -1;
    if (k > 0 && a.N > 0) {  // mig_controller.go:68-71, :78-81
      bool present = false;
      if (a.flags & NOSPLAN_PLAN_GATE) {  // profileAlreadyPresent, :121-144
        present = true;
        for (int j = 0; j < k; ++j) present &= sh_present[key[j]] > 0;
      }
      out_gate = present;
      if (!present && reqmask) {
        // ---- tryRepartition: first node whose UpdateGeometryFor returns true --------------------------
        const bool exact = (k == 1 && qty[0] == 1);
        int best = -1;
        for (int wbase = 0; wbase < a.nw && best < 0; wbase += 32) {
          const int w = wbase + lane;
          uint32_t bits = 0;
          if (w < a.nw) {
            if (exact) {
              bits = nu1t[(size_t)key[0] * a.nw + w];
            } else {
              for (int j = 0; j < k; ++j)
                if (qty[j] > 0) bits |= nbt[(size_t)key[j] * a.nw + w];
            }
          }
          uint32_t words = __ballot_sync(0xffffffffu, bits != 0u);
          while (words && best < 0) {
            const int src = __ffs(words) - 1;
            words &= words - 1;
            const uint32_t wb = __shfl_sync(0xffffffffu, bits, src);
            const int n = ((wbase + src) << 5) + lane;
            bool upd = (wb >> lane) & 1u;
            if (upd && !exact) upd = node_would_update(n, k, key, qty);
            const uint32_t hit = __ballot_sync(0xffffffffu, upd);
            if (hit) best = ((wbase + src) << 5) + __ffs(hit) - 1;
          }
        }
        SEQ_TICK(0)
        if (best >= 0) {
          NSP_CHECK(best < a.N);
          const int g0 = __ldg(a.gpu_off + best), cnt = __ldg(a.gpu_off + best + 1) - g0;
          NSP_CHECK(cnt >= 0 && cnt <= NOSPLAN_MAX_GPUS_PER_NODE && g0 >= 0 && g0 + cnt <= a.tg);
          load_node(g0, cnt);
          const int before = node_sum(cnt);
          SEQ_TICK(1)
          uint32_t changed = 0;
          if constexpr (MPS) {
            // slicing.Node.UpdateGeometryFor on the image: lane 0 runs the closed-form walk GPU by GPU
            if (lane == 0 && k == 1) {
              // one requested size: slicing/gpu.go:162-220 in closed form (same algebra as mps_score_kernel)
              const int sid = key[0], m = sh_sz[sid];
              int resid = qty[0];
              for (int s = 0; s < cnt && resid > 0; ++s) {
                const int fj = img_free(s, sid);
                const int miss = resid - fj;
                if (miss <= 0) break;  // covered: the residual drops to 0
                if (!((sh_gm[s] >> sid) & 1u)) {  // the size does not fit in cap: this GPU cannot update
                  resid -= fj;
                  continue;
                }
                uint32_t* r = sh_node[s];
                uint8_t* fb = reinterpret_cast<uint8_t*>(r);
                const uint32_t omask = r[16];
                int otot = 0;
                for (uint32_t mk = omask; mk; mk &= mk - 1) otot += sh_sz[__ffs(mk) - 1] * (int)fb[__ffs(mk) - 1];
                int spare = (int)r[17] - otot;
                const bool in_orig = (omask >> sid) & 1u;
                const int k1 = min(miss, spare / m);                       // (a) spare capacity first
                spare = spare - k1 * m + otot + (in_orig ? k1 * m : 0);    // (b) every original free key deleted
                int created = in_orig ? 0 : k1;
                const int k2 = min(miss - k1, spare / m);                  // (c) create in the space made
                created += k2;
                spare -= k2 * m;
                uint32_t restored = 0;                                     // (d) restore, ascending size
                for (uint32_t mk = omask; mk; mk &= mk - 1) {
                  const int kk = __ffs(mk) - 1;
                  const int need = sh_sz[kk] * (int)fb[kk];
                  if (spare >= need) {
                    restored |= 1u << kk;
                    spare -= need;
                  } else {
                    fb[kk] = 0;
                  }
                }
                const int nf = (((restored >> sid) & 1u) ? fj : 0) + created;
                fb[sid] = (uint8_t)nf;
                r[16] = restored | (created > 0 ? (1u << sid) : 0u);
                changed |= 1u << s;
                resid = max(resid - nf, 0);
              }
            } else if (lane == 0) {
              int resid[kMaxKeys];
              for (int j = 0; j < kMaxKeys; ++j) resid[j] = j < k ? qty[j] : 0;
              for (int s = 0; s < cnt; ++s) {
                uint32_t* r = sh_node[s];
                // this GPU updates only if the smallest missing size fits in its cap (gm is that prefix mask)
                bool left = false, can = false;
                for (int j = 0; j < k; ++j) {
                  left |= resid[j] > 0;
                  if (resid[j] - img_free(s, key[j]) > 0) {
                    can = (sh_gm[s] >> key[j]) & 1u;
                    break;
                  }
                }
                if (!left) break;  // nothing required any more
                if (!can) {
                  for (int j = 0; j < k; ++j) resid[j] = max(resid[j] - img_free(s, key[j]), 0);
                  continue;
                }
                const uint4 f0 = make_uint4(r[0], r[1], r[2], r[3]), f1 = make_uint4(r[4], r[5], r[6], r[7]);
                int created[kMaxKeys];
                uint32_t restored = r[16];
                const bool upd = seq_mps_gpu_update(sh_sz, f0, f1, r[16], r[17], k, key, resid, created, &restored);
                for (int j = 0; j < k; ++j) NSP_CHECK(key[j] >= 0 && key[j] < K && created[j] >= 0);  // the request vector is intact
                if (upd) {
                  uint32_t w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                  uint32_t nmask = restored;
                  for (uint32_t mk = restored; mk; mk &= mk - 1) {
                    const int kk = __ffs(mk) - 1;
                    w[kk >> 2] |= mps_byte(f0, f1, kk) << ((kk & 3) * 8);
                  }
                  for (int j = 0; j < k; ++j)
                    if (created[j] > 0) {
                      const int kk = key[j];
                      const uint32_t old = (w[kk >> 2] >> ((kk & 3) * 8)) & 0xffu;
                      w[kk >> 2] = (w[kk >> 2] & ~(0xffu << ((kk & 3) * 8))) | ((old + (uint32_t)created[j]) << ((kk & 3) * 8));
                      nmask |= 1u << kk;
                    }
                  for (int t = 0; t < 8; ++t) r[t] = w[t];
                  r[16] = nmask;
                  changed |= 1u << s;
                }
                for (int j = 0; j < k; ++j) resid[j] = max(resid[j] - img_free(s, key[j]), 0);
              }
            }
            changed = __shfl_sync(0xffffffffu, changed, 0);
          } else {
            // mig.Node.UpdateGeometryFor on the image: per GPU, lanes evaluate the allowed geometries
            int resid[kMaxKeys];
            for (int j = 0; j < kMaxKeys; ++j) resid[j] = j < k ? qty[j] : 0;
            for (int s = 0; s < cnt; ++s) {
              uint32_t* r = sh_node[s];
              // a GPU can only update if it can grow a key it is still asked for (gm, countProvidedProfiles > 0)
              uint32_t need = 0;
              bool left = false;
              for (int j = 0; j < k; ++j) {
                left |= resid[j] > 0;
                if (resid[j] - img_free(s, key[j]) > 0) need |= 1u << key[j];
              }
              if (!left) break;  // nothing is required any more: no later GPU can update (node.go:164-169)
              if ((sh_gm[s] & need) == 0u) {
                for (int j = 0; j < k; ++j) resid[j] = max(resid[j] - img_free(s, key[j]), 0);
                continue;
              }
              const uint4 u = make_uint4(r[0], r[1], r[2], r[3]), f = make_uint4(r[4], r[5], r[6], r[7]);
              const uint4 cur = add4(u, f);
              const int m = sh_model[s];
              const int c0 = a.T->geom_off[m], C = a.T->geom_off[m + 1] - c0;
              uint32_t prov = 0, k0 = 0;
              uint4 cd = make_uint4(0, 0, 0, 0);
              if (lane < C) {
                cd = a.T->cand[c0 + lane];
                if (all_ge4(cd, u)) {  // CanApplyGeometry
                  for (int j = 0; j < k; ++j) {
                    const int needed = resid[j] - (int)byte_of(f, key[j]);
                    const int add = (int)byte_of(cd, key[j]) - (int)byte_of(cur, key[j]);
                    if (needed > 0 && add > 0) prov += (uint32_t)min(add, needed);
                  }
                  if (prov) k0 = key0_of(a.T->total[c0 + lane], sad16(cd, cur), a.T->rank[c0 + lane]);
                }
              }
              const uint32_t pmax = __reduce_max_sync(0xffffffffu, prov);
              uint4 nf = f;
              if (pmax) {
                const uint32_t kmine = (prov == pmax) ? k0 : 0u;
                const uint32_t kmax = __reduce_max_sync(0xffffffffu, kmine);
                const int win = __ffs(__ballot_sync(0xffffffffu, prov == pmax && k0 == kmax)) - 1;  // lowest index on ties
                uint4 wc;
                wc.x = __shfl_sync(0xffffffffu, cd.x, win);
                wc.y = __shfl_sync(0xffffffffu, cd.y, win);
                wc.z = __shfl_sync(0xffffffffu, cd.z, win);
                wc.w = __shfl_sync(0xffffffffu, cd.w, win);
                nf = sub4(wc, u);  // ApplyGeometry
                if (lane == 0) {
                  r[4] = nf.x; r[5] = nf.y; r[6] = nf.z; r[7] = nf.w;
                }
                changed |= 1u << s;
              }
              for (int j = 0; j < k; ++j) resid[j] = max(resid[j] - (int)byte_of(nf, key[j]), 0);
              __syncwarp();
            }
          }
          __syncwarp();
          SEQ_TICK(2)
          store_node(g0, cnt, changed);
          const int after = node_sum(cnt);
          if (lane < K) sh_present[lane] += after - before;
          refresh_node(best, g0, cnt, changed);
          SEQ_TICK(3)
          out_rep = best;
        }
      }
      if (a.flags & NOSPLAN_PLAN_SCHED) {
        // ---- Node.AddPod first-fit: first GPU (node order, then slot order) where every key fits --------
        int start = 0;
        for (int j = 0; j < k; ++j)
          if (qty[j] > 0) start = max(start, sh_hint[key[j]]);
        int found = -1;
        for (int wbase = start; wbase < a.gw && found < 0; wbase += 32) {
          const int w = wbase + lane;
          uint32_t bits = 0;
          if (w < a.gw) {
            bits = fbt[(size_t)K * a.gw + w];
            for (int j = 0; j < k; ++j)
              if (qty[j] > 0) bits &= fbt[(size_t)key[j] * a.gw + w];
          }
          int mine = -1;
          while (bits) {
            const int g = (w << 5) + __ffs(bits) - 1;
            bits &= bits - 1;
            bool fits = true;
            for (int j = 0; j < k; ++j)
              if (qty[j] > 1) {
                int fj;
                if constexpr (MPS) fj = (int)mps_byte(__ldcg(&a.p_gpu[g].f0), __ldcg(&a.p_gpu[g].f1), key[j]);
                else fj = (int)byte_of(__ldcg(a.m_free + g), key[j]);
                fits &= fj >= qty[j];
              }
            if (fits) {
              mine = g;
              break;
            }
          }
          const uint32_t hit = __ballot_sync(0xffffffffu, mine >= 0);
          if (hit) found = __shfl_sync(0xffffffffu, mine, __ffs(hit) - 1);
        }
        SEQ_TICK(4)
        if (found >= 0) {
          NSP_CHECK(found < a.tg);
          const int node = __ldg(a.gpu_node + found);
          NSP_CHECK(node >= 0 && node < a.N);
          const int g0 = __ldg(a.gpu_off + node), cnt = __ldg(a.gpu_off + node + 1) - g0;
          NSP_CHECK(cnt > 0 && cnt <= NOSPLAN_MAX_GPUS_PER_NODE && found - g0 < cnt);
          // single-key quantity-1 pods: everything before `found` for this key is exhausted
          if (k == 1 && qty[0] == 1 && lane == 0) sh_hint[key[0]] = found >> 5;
          load_node(g0, cnt);
          uint32_t changed = 1u << (found - g0);
          if (k > 1) changed |= changed - 1u;  // earlier GPUs of the node may keep partial deductions
          if (lane == 0) {
            // Node.AddPod tries the node's GPUs in order; the ones before the winner fail on some key and
            // keep the deductions made before it (gpu.go:289-303); single-key pods never deduct there
            for (int s = 0; s <= found - g0; ++s) {
              if (s < found - g0 && k == 1) continue;
              for (int j = 0; j < k; ++j) {
                const int q = qty[j];
                if (img_free(s, key[j]) < q) break;
                img_add(s, MPS ? 0 : 4, key[j], -q);
                img_add(s, MPS ? 8 : 0, key[j], q);
                if constexpr (MPS) {
                  sh_node[s][16] |= 1u << key[j];  // FreeProfiles[r] -= q creates the key
                  sh_node[s][17] -= (uint32_t)(q * sh_sz[key[j]]);
                }
              }
            }
          }
          __syncwarp();
          SEQ_TICK(5)
          store_node(g0, cnt, changed);
          refresh_node(node, g0, cnt, changed);
          SEQ_TICK(6)
          out_node = node;
          out_gpu = found - g0;
        } else if (k == 1 && qty[0] == 1 && lane == 0) {
          sh_hint[key[0]] = a.gw;  // no GPU holds a free one: nothing to scan until a commit sets a bit again
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
    if (a.dead) {
      if (out_rep >= 0) {
        if (lane < K) a.dead[lane] = 0;  // the new geometry may hold any profile
      } else if (lane == 0 && k == 1 && !out_gate && sh_present[key[0]] <= 0) {
        a.dead[key[0]] = 1;
      }
    }
    __syncwarp();
    SEQ_TICK(7)
  }
  if (a.present_out && lane < K) a.present_out[lane] = sh_present[lane];
#ifdef NOSPLAN_SEQ_PROFILE
  if (lane == 0)
    printf("seq profile (Mcycles) mps=%d search %lld load %lld commit %lld refresh %lld fit %lld addpod %lld refresh2 %lld rest %lld\n",
           (int)MPS, tp[0] >> 20, tp[1] >> 20, tp[2] >> 20, tp[3] >> 20, tp[4] >> 20, tp[5] >> 20, tp[6] >> 20, tp[7] >> 20);
#endif
}

}  // namespace nsp
