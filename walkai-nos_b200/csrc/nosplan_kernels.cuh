// nosplan_kernels.cuh — device data layout and kernels of the planner hot path (sm_100a).
//
// Integer scoring, no tensor cores: the work is byte/small-int compares over SoA state, bounded by
// HBM/L2 bandwidth for the streaming form and by the INT32 pipes once state is tiled in shared
// memory (DESIGN.md "Kernels").  Reference semantics: SURVEY §A, cited per function below.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "nosplan.h"

// -DNOSPLAN_CHECKS: every index into a fixed-size local or shared array, and every table index read from memory, is
// asserted on the device (compute-sanitizer is closed on the pool this was developed on: tools/run_checked.sh runs the
// parity suites on such a build instead; a failed assertion aborts the kernel and the C ABI reports the CUDA error).
#ifdef NOSPLAN_CHECKS
#include <cassert>
#define NSP_CHECK(c) assert(c)
#else
#define NSP_CHECK(c) ((void)0)
#endif

namespace nsp {

constexpr int kMaxProf = NOSPLAN_MAX_MIG_PROFILES;               // 16 bytes = one uint4 per map
constexpr int kMaxModels = NOSPLAN_MAX_MODELS;
constexpr int kMaxGeoms = NOSPLAN_MAX_GEOMS;
constexpr int kMaxTotalGeoms = kMaxModels * kMaxGeoms;
constexpr int kMaxSizes = NOSPLAN_MAX_MPS_SIZES;                  // 32 bytes = two uint4 per map
constexpr int kMaxKeys = NOSPLAN_MAX_DEMAND_KEYS;
constexpr int kNoNode = 0x7fffffff;
constexpr int kMaxWorld = 16;  // ranks of the first-node exchange (nosplan_xchg.cuh)

// Where a scorer CTA's pass ends (nosplan_score_pass_dev).  The per-row winners always go into first_node[] with
// a local unsigned atomicMin (-1 = no node = above every index).  world == 0: that is all.  world >= 1:
// first_node[] is the rank's own half of the exchange buffer, and the CTA that finishes last over all scorer
// launches of the pass tells every rank that this rank's keys of the step are complete (nosplan_xchg.cuh).
struct PushArgs {
  int world;
  int rank;
  uint32_t step;
  unsigned int total;          // CTAs of all scorer launches of this pass (each takes one ticket)
  unsigned int* ticket;        // zero on entry, left at zero
  uint32_t* flags[kMaxWorld];  // every rank's pushed[world]
};

// ---- MIG geometry tables (global memory, one per handle) ---------------------------------------------
struct MigTables {
  int n_models;
  int n_prof;
  int n_geoms;                        // total over models
  int geom_off[kMaxModels + 1];       // CSR over models
  int init_choice[kMaxModels];        // GetFewestSlicesGeometry index (partitioning.go:67-79), -1 if empty
  uint4 cand[kMaxTotalGeoms];         // 16 x u8 quantity per profile id
  uint16_t total[kMaxTotalGeoms];     // totalSliceCount (gpu.go:262-268)
  uint8_t rank[kMaxTotalGeoms];       // rank of Geometry.Id() among the model's ids, bytewise ascending
};

__device__ __forceinline__ uint32_t byte_of(const uint4& v, int i) {
  NSP_CHECK(i >= 0 && i < 16);
  uint32_t w = (i < 8) ? ((i < 4) ? v.x : v.y) : ((i < 12) ? v.z : v.w);
  return (w >> ((i & 3) * 8)) & 0xffu;
}
__device__ __forceinline__ uint32_t word_of(const uint4& v, int i) {
  return (i < 2) ? ((i == 0) ? v.x : v.y) : ((i == 2) ? v.z : v.w);
}
__device__ __forceinline__ uint4 add4(const uint4& a, const uint4& b) {  // bytewise, no carries (<=255 checked on host)
  return make_uint4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
}
__device__ __forceinline__ uint4 sub4(const uint4& a, const uint4& b) {  // bytewise a-b where a>=b bytewise
  return make_uint4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w);
}
// every byte of a >= the matching byte of b  (CanApplyGeometry's used-device check, gpu.go:105-109)
__device__ __forceinline__ bool all_ge4(const uint4& a, const uint4& b) {
  return (__vcmpgeu4(a.x, b.x) & __vcmpgeu4(a.y, b.y) & __vcmpgeu4(a.z, b.z) & __vcmpgeu4(a.w, b.w)) == 0xffffffffu;
}
// sum over bytes |a-b|  (geometryDistance, gpu.go:245-260; missing keys are 0 on both sides)
__device__ __forceinline__ uint32_t sad16(const uint4& a, const uint4& b) {
  return __vsadu4(a.x, b.x) + __vsadu4(a.y, b.y) + __vsadu4(a.z, b.z) + __vsadu4(a.w, b.w);
}
// demand-independent part of geometrySelectionScore (gpu.go:225-243): total up, distance down, id down.
// 0 is reserved for "cannot apply"; a feasible key is never 0 because 255-rank >= 223.
__device__ __forceinline__ uint32_t key0_of(uint32_t total, uint32_t dist, uint32_t rank) {
  return (total << 20) | ((4095u - dist) << 8) | (255u - rank);
}

// ---- generic (unpacked) node walk on AoS state ----------------------------------------------------------
// mig.Node.UpdateGeometryFor (node.go:145-177) over GPUs [g0, g0+cnt) with demand (pid[], qty[]) of k
// keys.  If commit, free[] is rewritten (ApplyGeometry, gpu.go:134-154); choice (may be null) gets
// the applied geometry index per GPU or -1.  Returns anyGpuUpdated.
template <bool CG>
__device__ __forceinline__ uint4 ld_state(const uint4* p) {
  if constexpr (CG) return __ldcg(p);
  else return *p;
}
template <bool CG>
__device__ __forceinline__ void st_state(uint4* p, const uint4& v) {
  if constexpr (CG) __stcg(p, v);
  else *p = v;
}

// CG = state loads/stores bypass L1 (the planner kernel mutates the state it reads within one launch).
template <bool CG = false>
__device__ inline bool mig_walk_node(const MigTables* __restrict__ T, const uint4* used, uint4* fr,
                                     const uint8_t* __restrict__ model, int g0, int cnt, int k,
                                     const int* __restrict__ pid, const int* __restrict__ qty, bool commit,
                                     int8_t* choice) {
  NSP_CHECK(k >= 0 && k <= kMaxKeys);
  NSP_CHECK(cnt >= 0 && (choice == nullptr || cnt <= NOSPLAN_MAX_GPUS_PER_NODE));
  int resid[kMaxKeys];
#pragma unroll
  for (int j = 0; j < kMaxKeys; ++j) resid[j] = (j < k) ? max(qty[j], 0) : 0;
  bool any = false;
  if (k == 0) cnt = 0;  // node.go:150-152
  for (int s = 0; s < cnt; ++s) {
    const int g = g0 + s;
    const uint4 u = ld_state<CG>(used + g);
    const uint4 f = ld_state<CG>(fr + g);
    const uint4 cur = add4(u, f);  // GetGeometry, gpu.go:84-95
    const int m = model[g];
    NSP_CHECK(m >= 0 && m < T->n_models);
    const int c0 = T->geom_off[m], c1 = T->geom_off[m + 1];
    int best = -1;
    uint32_t best_prov = 0, best_k0 = 0;
    for (int c = c0; c < c1; ++c) {
      const uint4 cd = T->cand[c];
      if (!all_ge4(cd, u)) continue;  // CanApplyGeometry
      int prov = 0;                   // countProvidedProfiles, gpu.go:198-223
      for (int j = 0; j < k; ++j) {
        const int fj = (int)byte_of(f, pid[j]);
        const int needed = resid[j] - fj;
        if (needed <= 0) continue;
        const int add = (int)byte_of(cd, pid[j]) - (int)byte_of(cur, pid[j]);
        if (add <= 0) continue;
        prov += min(add, needed);
      }
      if (prov <= 0) continue;
      const uint32_t k0 = key0_of(T->total[c], sad16(cd, cur), T->rank[c]);
      if (best < 0 || (uint32_t)prov > best_prov || ((uint32_t)prov == best_prov && k0 > best_k0)) {
        best = c;
        best_prov = prov;
        best_k0 = k0;
      }
    }
    uint4 nf = f;
    if (best >= 0) {
      any = true;
      nf = sub4(T->cand[best], u);  // free[p] = cand[p]-used[p]; keys outside cand vanish (cand[p]=0=used[p])
      if (commit) st_state<CG>(fr + g, nf);
    }
    if (choice) choice[s] = (int8_t)(best >= 0 ? best - c0 : -1);
    for (int j = 0; j < k; ++j) resid[j] = max(resid[j] - (int)byte_of(nf, pid[j]), 0);  // node.go:164-169
  }
  return any;
}

// ---- pack kernels: int32 rows from the host -> byte-packed device state ------------------------------------
// The range checks of the snapshot run here as well, so the host never walks the state arrays:
// status[0] = lowest failing code (atomicMin; ~0 = clean), status[1] = evaluations per demand row.
// MIG code = gpu * 64 + (0: unknown model | 1 + profile: count out of range); the host turns the
// winning code back into the message of the first bad entry in GPU order.
//
// MIG: used/free [tg][P] int32 -> uint4 per GPU (AoS, CSR order) and slot-major copies for the scorer:
// sused/sfree[slot][npad] (uint4), smodel[slot][npad] (0xff = no such GPU); gnode/gslot per GPU.
// Rows of invalid nodes are not interpreted (the controller never builds those nodes): zeros.
__global__ void mig_pack_kernel(int N, int P, int gs, int npad, int max_count, const MigTables* __restrict__ T,
                                const int32_t* __restrict__ gpu_off, const uint8_t* __restrict__ valid,
                                const int32_t* __restrict__ model, const int32_t* __restrict__ used,
                                const int32_t* __restrict__ fr, uint4* __restrict__ a_used, uint4* __restrict__ a_free,
                                uint8_t* __restrict__ a_model, uint4* __restrict__ s_used, uint4* __restrict__ s_free,
                                uint8_t* __restrict__ s_model, uint8_t* __restrict__ s_cnt, int32_t* __restrict__ gnode,
                                int32_t* __restrict__ gslot, unsigned long long* __restrict__ status) {
  // one thread per (node, slot); the grid covers npad * gs of them in whole warps
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int n = i / gs, s = i - n * gs;
  unsigned long long bad = ~0ull;
  unsigned int evals = 0;
  if (n < npad) {
    const int g0 = n < N ? gpu_off[n] : 0;
    const int cnt = n < N ? gpu_off[n + 1] - g0 : 0;
    const bool ok = n < N && valid[n];
    if (s == 0) s_cnt[n] = ok ? (uint8_t)cnt : 0;  // an invalid node behaves like a node the controller skips
    uint4 u = make_uint4(0, 0, 0, 0), f = u;
    uint8_t m = 0xff;
    if (s < cnt) {
      const int g = g0 + s;
      uint32_t uw[4] = {0, 0, 0, 0}, fw[4] = {0, 0, 0, 0};
      int md = 0;
      if (ok) {
        md = model[g];
        if (md < 0 || md >= T->n_models) {
          bad = (unsigned long long)g * 64ull;
          md = 0;
        } else {
          evals = (unsigned int)(T->geom_off[md + 1] - T->geom_off[md]);
        }
        for (int p = 0; p < P; ++p) {
          const int uv = used[(size_t)g * P + p], fv = fr[(size_t)g * P + p];
          if (uv < 0 || fv < 0 || uv > max_count || fv > max_count || uv + fv > max_count)
            bad = min(bad, (unsigned long long)g * 64ull + 1ull + (unsigned long long)p);
          uw[p >> 2] |= (uint32_t)(uv & 0xff) << ((p & 3) * 8);
          fw[p >> 2] |= (uint32_t)(fv & 0xff) << ((p & 3) * 8);
        }
      }
      u = make_uint4(uw[0], uw[1], uw[2], uw[3]);
      f = make_uint4(fw[0], fw[1], fw[2], fw[3]);
      m = (uint8_t)md;
      a_used[g] = u;
      a_free[g] = f;
      a_model[g] = m;
      gnode[g] = ok ? n : -1;
      gslot[g] = s;
    }
    s_used[(size_t)s * npad + n] = u;
    s_free[(size_t)s * npad + n] = f;
    s_model[(size_t)s * npad + n] = ok ? m : (uint8_t)0xff;
  }
  // one atomic per warp for each of the two status words
  evals = __reduce_add_sync(0xffffffffu, evals);
  const bool anybad = __any_sync(0xffffffffu, bad != ~0ull);
  if (anybad && bad != ~0ull) atomicMin(&status[0], bad);
  if ((threadIdx.x & 31) == 0 && evals) atomicAdd(&status[1], (unsigned long long)evals);
}

__global__ void mig_unpack_kernel(int tg, int P, const uint4* __restrict__ a_used, const uint4* __restrict__ a_free,
                                  int32_t* __restrict__ used, int32_t* __restrict__ fr) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= tg) return;
  const uint4 u = a_used[g], f = a_free[g];
  for (int p = 0; p < P; ++p) {
    used[(size_t)g * P + p] = (int32_t)byte_of(u, p);
    fr[(size_t)g * P + p] = (int32_t)byte_of(f, p);
  }
}

// ---- MIG batched scorer ----------------------------------------------------------------------------------------
// One thread per node of a tile, looping over a chunk of demand rows.  The demand-independent work
// of gpu.go:168-188 (CanApplyGeometry, geometryDistance, totalSliceCount, the id rank) is hoisted
// out of the demand loop: the prologue builds key0[slot][cand][node] in shared memory once per tile
// and every (demand, node, GPU, candidate) evaluation then costs one conflict-free LDS, the
// provided-profiles count for the demand's keys, and a compare.
struct MigScoreArgs {
  const MigTables* T;
  const uint4* s_used;
  const uint4* s_free;
  const uint8_t* s_model;
  const uint8_t* s_cnt;
  const int32_t* gpu_off;
  int N, npad, gs, cmax, tg;
  int k16;  // tables allow the 16-bit key path: <= 16 geometries per model, every quantity <= 63
  int D, d_chunk;
  const int32_t* dem_off;
  const int32_t* dem_key;
  const int32_t* dem_qty;
  int32_t* first_node;  // unsigned atomicMin target, pre-filled with -1 (0xffffffff = no node)
  int node_base;        // global index of node 0 (0 unless the node list is sharded)
  uint8_t* updated;     // [D][N] or null
  int8_t* choice;       // [D][tg] or null
  // optional: score only the rows row_list[row_meta[1] .. + row_meta[2]) (device-side list and counts)
  const int32_t* row_list;
  const int32_t* row_meta;
  const uint8_t* row_ok;  // optional: rows_check_kernel's verdict per row (0 = breaks an ABI limit: skipped)
};

// Shared-memory layout of one tile (TN nodes, gs GPU slots, cw = ceil(cmax/4) words of 4 candidates):
//   candT [n_models][16][cw] u32   candidate quantities transposed: byte cc of word i = cand[4i+cc][p]
//   candT2 [n_models][16][2cw] u32 the same, one candidate per 16-bit lane (K16 path)
//   rank2 [gs][2cw][TN]      u32   per (node, slot): half-word h of word i describes candidate c = 2i+h:
//                                  0 if it cannot be applied, else rank << 8 | (255 - c), rank = 1 + the
//                                  position of its demand-independent score key0 among the GPU's
//                                  candidates (equal key0 -> equal rank; the index bits break that tie).
//                                  K16 path: rank << 4 | (15 - c), or 0x8000 if it cannot be applied, so
//                                  that a whole key provided << 9 | that fits one SIGNED 16-bit lane
//                                  (provided <= 63) and two candidates are scored per DPX instruction
//                                  (VIADDMNMX.S16x2.RELU, VIMNMX3.S16x2); an inapplicable lane is negative
//                                  whatever it provides and never wins against the running maximum
//   uf [gs][16][TN]          u16   per (node, slot, profile): free | (used + free) << 8
//   model [gs][TN]           u8
__host__ __device__ inline size_t mig_score_smem_bytes(int n_models, int gs, int cmax, int tn) {
  const size_t cw = (size_t)(cmax + 3) / 4;
  return (size_t)n_models * 16 * cw * 4 * 3 + (size_t)gs * 2 * cw * tn * 4 + (size_t)gs * 16 * tn * 2 + (size_t)gs * tn + 16;
}

// Shared-memory tile of the MIG scorers (layout above).
struct MigTile {
  uint32_t* candT;
  uint32_t* candT2;
  uint32_t* rank2;
  uint16_t* uf;
  uint8_t* model;
};
template <int TN>
__device__ __forceinline__ MigTile mig_tile_carve(unsigned char* smem, int n_models, int gs, int cw) {
  MigTile t;
  t.candT = reinterpret_cast<uint32_t*>(smem);
  t.candT2 = t.candT + (size_t)n_models * 16 * cw;
  t.rank2 = t.candT2 + (size_t)n_models * 16 * cw * 2;
  t.uf = reinterpret_cast<uint16_t*>(t.rank2 + (size_t)gs * 2 * cw * TN);
  t.model = reinterpret_cast<uint8_t*>(t.uf + (size_t)gs * 16 * TN);
  return t;
}

// Fills the tile: the transposed candidate tables and, per (node, slot), feasibility + the
// demand-independent key of every candidate reduced to a rank half-word.  Ends with __syncthreads().
template <int TN, int DS, bool K16>
__device__ __forceinline__ void mig_tile_prologue(const MigScoreArgs& a, int cw, const MigTile& t, int tid, int dsub, int n,
                                                  int cnt) {
  const MigTables* __restrict__ T = a.T;
  const int n_models = T->n_models;
  uint32_t* s_candT = t.candT;
  uint32_t* s_candT2 = t.candT2;
  uint32_t* s_rank2 = t.rank2;
  uint16_t* s_uf = t.uf;
  uint8_t* s_model = t.model;
  // transposed candidate table (demand- and node-independent)
  for (int i = threadIdx.x; i < n_models * 16 * cw; i += TN * DS) {
    const int w = i % cw, p = (i / cw) & 15, m = i / (cw * 16);
    const int c0 = T->geom_off[m], cc = T->geom_off[m + 1] - c0;
    uint32_t v = 0;
    for (int b = 0; b < 4; ++b) {
      const int c = w * 4 + b;
      if (c < cc) v |= byte_of(T->cand[c0 + c], p) << (8 * b);
    }
    s_candT[i] = v;
    if (K16) {
      s_candT2[2 * i] = (v & 0xffu) | ((v & 0xff00u) << 8);
      s_candT2[2 * i + 1] = ((v >> 16) & 0xffu) | ((v >> 24) << 16);
    }
  }
  // per (node, slot): feasibility + demand-independent key of every candidate, reduced to a rank byte
  for (int s = dsub; s < cnt; s += DS) {  // the DS threads of a node split its GPUs
    const uint4 u = a.s_used[(size_t)s * a.npad + n];
    const uint4 f = a.s_free[(size_t)s * a.npad + n];
    const int m = a.s_model[(size_t)s * a.npad + n];
    const uint4 cur = add4(u, f);
    NSP_CHECK(m >= 0 && m < T->n_models);
    const int c0 = T->geom_off[m], cc = T->geom_off[m + 1] - c0;
    NSP_CHECK(cc >= 0 && cc <= kMaxGeoms && cc <= 4 * cw);
    s_model[s * TN + tid] = (uint8_t)m;
#pragma unroll
    for (int q = 0; q < 16; ++q) s_uf[(s * 16 + q) * TN + tid] = (uint16_t)(byte_of(f, q) | (byte_of(cur, q) << 8));
    uint32_t k0[kMaxGeoms];
    for (int c = 0; c < cc; ++c) {
      const uint4 cd = T->cand[c0 + c];
      k0[c] = all_ge4(cd, u) ? key0_of(T->total[c0 + c], sad16(cd, cur), T->rank[c0 + c]) : 0u;
    }
    for (int w = 0; w < 2 * cw; ++w) {
      uint32_t r2 = 0;
      for (int b = 0; b < 2; ++b) {
        const int c = w * 2 + b;
        if (c < cc && k0[c] != 0u) {
          uint32_t r = 1;
          for (int c2 = 0; c2 < cc; ++c2) r += (k0[c2] != 0u && k0[c2] < k0[c]) ? 1u : 0u;
          r2 |= (K16 ? ((r << 4) | (uint32_t)(15 - c)) : ((r << 8) | (uint32_t)(255 - c))) << (16 * b);
        } else if (K16) {
          r2 |= 0x8000u << (16 * b);  // not applicable (or no such candidate): the lane's key stays negative
        }
      }
      s_rank2[((size_t)s * 2 * cw + w) * TN + tid] = r2;
    }
  }
  __syncthreads();
}

// One demand row with any number of keys against one node (the general form of gpu.go:160-223 +
// node.go:161-170); returns "the node would be updated".
template <int TN, bool K16>
__device__ __forceinline__ bool mig_row_generic(const MigScoreArgs& a, int cw, const MigTile& t, int tid, int cnt, int g0, int d,
                                                int k, int e0) {
  const uint32_t* s_candT = t.candT;
  const uint32_t* s_rank2 = t.rank2;
  const uint16_t* s_uf = t.uf;
  const uint8_t* s_model = t.model;
  bool any = false;
  NSP_CHECK(k >= 0 && k <= kMaxKeys);
  int pid[kMaxKeys], resid[kMaxKeys];
  for (int j = 0; j < kMaxKeys; ++j) {
    pid[j] = (j < k) ? __ldg(a.dem_key + e0 + j) : 0;
    NSP_CHECK(pid[j] >= 0 && pid[j] < 16);
    resid[j] = (j < k) ? max(__ldg(a.dem_qty + e0 + j), 0) : 0;
  }
  for (int s = 0; s < cnt; ++s) {
    int ch = -1;
    if (k > 0) {
      const uint32_t* ctm = s_candT + (size_t)s_model[s * TN + tid] * 16 * cw;
      int uj[kMaxKeys], fj[kMaxKeys];
      for (int j = 0; j < k; ++j) {
        const uint32_t x = s_uf[(s * 16 + pid[j]) * TN + tid];
        fj[j] = (int)(x & 0xffu);
        uj[j] = (int)(x >> 8) - fj[j];
      }
      uint32_t best = 0;
      for (int c = 0; c < a.cmax; ++c) {
        const uint32_t rk = (s_rank2[((size_t)s * 2 * cw + (c >> 1)) * TN + tid] >> (16 * (c & 1))) & 0xffffu;
        if (K16 ? (rk & 0x8000u) != 0u : rk == 0u) continue;  // not applicable; else ordered by (rank, lowest index)
        int prov = 0;
        for (int j = 0; j < k; ++j) {
          const int needed = resid[j] - fj[j];
          const int add = (int)((ctm[pid[j] * cw + (c >> 2)] >> (8 * (c & 3))) & 0xffu) - (uj[j] + fj[j]);
          if (needed > 0 && add > 0) prov += min(add, needed);
        }
        if (prov > 0) best = max(best, ((uint32_t)prov << 16) | rk);
      }
      if (best != 0u) {
        any = true;
        ch = K16 ? 15 - (int)(best & 15u) : 255 - (int)(best & 0xffu);
      }
      for (int j = 0; j < k; ++j) {
        const int nf = (ch >= 0) ? (int)((ctm[pid[j] * cw + (ch >> 2)] >> (8 * (ch & 3))) & 0xffu) - uj[j] : fj[j];
        resid[j] = max(resid[j] - nf, 0);
      }
    }
    if (a.choice) a.choice[(size_t)d * a.tg + g0 + s] = (int8_t)ch;
  }
      return any;
}

// TN nodes per tile; DS threads share one node's tile (thread j of them takes every DS-th demand row of
// the chunk), which raises occupancy without more shared memory; CW = words of 4 candidates per GPU
// (0 = runtime), a template parameter so the candidate loop unrolls completely.
template <int TN, int DS, int CW, bool K16>
__global__ void __launch_bounds__(TN * DS) mig_score_kernel(const MigScoreArgs a) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int cw = CW ? CW : ((a.cmax + 3) >> 2);
  const MigTile tile = mig_tile_carve<TN>(smem, a.T->n_models, a.gs, cw);
  uint32_t* const s_candT = tile.candT;
  uint32_t* const s_candT2 = tile.candT2;
  uint32_t* const s_rank2 = tile.rank2;
  uint16_t* const s_uf = tile.uf;
  uint8_t* const s_model = tile.model;

  const int tid = threadIdx.x % TN;   // node of the tile
  const int dsub = threadIdx.x / TN;  // which share of the demand chunk
  const int d_begin = blockIdx.y * a.d_chunk;
  const int32_t* rl = nullptr;
  int n_rows = a.D;
  if (a.row_list) {
    rl = a.row_list + a.row_meta[1];
    n_rows = a.row_meta[2];
    if (d_begin >= n_rows) return;  // uniform over the CTA, before any barrier
  }
  // grid.x may be smaller than the number of tiles (a launch that usually finds no rows should not start one CTA per
  // 128 nodes of a million-node list just to leave): the CTA then takes every gridDim.x-th tile
  const int n_tiles = (a.N + TN - 1) / TN;
  for (int tile_i = blockIdx.x; tile_i < n_tiles; tile_i += gridDim.x) {
  if (tile_i != (int)blockIdx.x) __syncthreads();  // the previous tile's shared memory is still being read
  const int n = tile_i * TN + tid;
  const int cnt = (n < a.N) ? a.s_cnt[n] : 0;
  const int g0 = (n < a.N) ? a.gpu_off[n] : 0;
  mig_tile_prologue<TN, DS, K16>(a, cw, tile, tid, dsub, n, cnt);

  const int d_end = min(n_rows, d_begin + a.d_chunk);
  const int lane = threadIdx.x & 31;
  uint32_t k16, k8;  // opaque 2^16 and 2^8 (see the inner loop)
  asm("mov.u32 %0, 65536;" : "=r"(k16));
  asm("mov.u32 %0, 256;" : "=r"(k8));
  uint32_t k512;
  asm("mov.u32 %0, 512;" : "=r"(k512));
  for (int di = d_begin + dsub; di < d_end; di += DS) {
    const int d = rl ? rl[di] : di;
    if (a.row_ok && !a.row_ok[d]) continue;  // uniform over the CTA
    const int e0 = __ldg(a.dem_off + d), k = __ldg(a.dem_off + d + 1) - e0;  // uniform over the CTA
    bool any = false;
    if (k == 1) {
      // single requested profile (every BASELINE config): all per-demand state in registers.
      // Every candidate of every GPU is evaluated: provided = clamp(cand[p]-cur[p], 0, needed), then a
      // (provided, rank, lowest index) compare folded into one integer max.
      const int p = __ldg(a.dem_key + e0);
      NSP_CHECK(p >= 0 && p < 16);
      int resid = max(__ldg(a.dem_qty + e0), 0);
      // running pointers: one add per slot instead of re-deriving every shared-memory address
      const uint16_t* uf = s_uf + (size_t)p * TN + tid;
      const uint32_t* rk = s_rank2 + tid;
      const uint8_t* md = s_model + tid;
      uint32_t anyacc = 0;
      for (int s = 0; s < cnt; ++s, uf += 16 * TN, rk += 2 * cw * TN, md += TN) {
        // request covered: every candidate of the remaining GPUs provides 0 (gpu.go:172-175), nothing can change
        if (resid == 0 && !a.choice) break;
        const uint32_t x = *uf;
        const int fj = (int)(x & 0xffu), curj = (int)(x >> 8);
        const int needed = resid - fj;
        // Every candidate of the GPU is evaluated when something is still needed:
        // provided = clamp(cand[p] - cur[p], 0, needed), forced to 0 for a candidate that cannot be applied
        // (its rank half-word is 0), then (provided, rank, lowest index) folded into one integer max.
        uint32_t best = 0, prov;
        if constexpr (K16) {
          // two candidates per instruction: key = provided << 9 | rank << 4 | (15 - index) per signed 16-bit
          // lane; a lane that cannot be applied carries 0x8000 and stays below the running maximum (>= 0)
          uint32_t best2 = 0;
          if (needed > 0) {
            const uint32_t* ct2 = s_candT2 + ((size_t)(*md) * 16 + p) * (2 * cw);
            const uint32_t nd2 = (uint32_t)min(needed, 32767) * 0x10001u;
            const uint32_t nc2 = ((uint32_t)(-curj) & 0xffffu) * 0x10001u;
#pragma unroll
            for (int w = 0; w < cw; ++w) {
              const uint32_t ra = rk[(2 * w) * TN];
              const uint32_t rb = rk[(2 * w + 1) * TN];
              const uint32_t pa = __viaddmin_s16x2_relu(ct2[2 * w], nc2, nd2);
              const uint32_t pb = __viaddmin_s16x2_relu(ct2[2 * w + 1], nc2, nd2);
              best2 = __vimax3_s16x2(best2, pa * k512 + ra, pb * k512 + rb);  // IMAD: the FMA pipe, not the ALU
            }
          }
          best = max(best2 & 0xffffu, best2 >> 16);
          prov = best >> 9;  // 0 unless some applicable candidate provides something
        } else {
          // key = provided << 16 | rank << 8 | (255 - index)
          if (needed > 0) {
            const uint32_t* ct = s_candT + ((size_t)(*md) * 16 + p) * cw;
            const int ncur = -curj;
#pragma unroll
            for (int w = 0; w < cw; ++w) {
              const uint32_t ra = rk[(2 * w) * TN];
              const uint32_t rb = rk[(2 * w + 1) * TN];
              const uint32_t cb4 = ct[w];
              // Field extraction with multiplies (IMAD / IMAD.HI run on the FMA pipe, the ALU pipe is the
              // busy one): x >> 16 == mulhi(x, 2^16), x & 0xffff == x - (x >> 16) * 2^16, x >> 24 ==
              // mulhi(x, 2^8).  The constants come from asm so the compiler cannot turn them back into shifts.
              const uint32_t ah = __umulhi(ra, k16), bh = __umulhi(rb, k16);
              const uint32_t rk4[4] = {ra - ah * k16, ah, rb - bh * k16, bh};
              const int cb3 = (int)__umulhi(cb4, k8);
              const int cb2 = (int)__umulhi(cb4 * k8, k8);  // (x << 8) >> 24
              uint32_t key[4];
#pragma unroll
              for (int b = 0; b < 4; ++b) {
                const int cb = (b == 3) ? cb3 : (b == 2) ? cb2 : (int)__byte_perm(cb4, 0u, 0x4440u | (uint32_t)b);
                const uint32_t r = rk4[b];  // rank << 8 | 255-index, 0 if not applicable
                const int pv = min(__viaddmin_s32_relu(cb, ncur, needed), (int)r);
                key[b] = (uint32_t)pv * k16 + r;
              }
              best = __vimax3_u32(best, key[0], key[1]);
              best = __vimax3_u32(best, key[2], key[3]);
            }
          }
          prov = best >> 16;
        }
        anyacc |= prov;
        if (a.choice) {
          const int ch = prov == 0u ? -1 : (K16 ? 15 - (int)(best & 15u) : 255 - (int)(best & 0xffu));
          a.choice[(size_t)d * a.tg + g0 + s] = (int8_t)ch;
        }
        // node.go:164-169 with the winner applied (gpu.go:134-154): the GPU's free count of the profile
        // becomes cand - used = free + (cand - cur) >= free + provided, with equality unless the request is
        // covered; either way what is still missing afterwards is needed - provided (0 if nothing was)
        resid = max(needed, 0) - (int)prov;
      }
      any = anyacc != 0u;
    } else {
      any = mig_row_generic<TN, K16>(a, cw, tile, tid, cnt, g0, d, k, e0);
    }
    if (a.updated && n < a.N) a.updated[(size_t)d * a.N + n] = any ? 1 : 0;
    // -1 read as unsigned is above every node index: the minimum needs no final pass
    const unsigned mine = any ? (unsigned)(n + a.node_base) : 0xffffffffu;
    const unsigned wmin = __reduce_min_sync(0xffffffffu, mine);
    if (lane == 0 && wmin != 0xffffffffu) atomicMin(reinterpret_cast<unsigned*>(a.first_node) + d, wmin);
  }
  }  // tiles of this CTA
}

// re-walks the winning node of every row to emit its per-GPU choices
__global__ void mig_finalize_kernel(const MigTables* __restrict__ T, const uint4* __restrict__ a_used,
                                    const uint4* __restrict__ a_free, const uint8_t* __restrict__ a_model,
                                    const int32_t* __restrict__ gpu_off, int D, int gs, const int32_t* __restrict__ dem_off,
                                    const int32_t* __restrict__ dem_key, const int32_t* __restrict__ dem_qty,
                                    int32_t* __restrict__ first_node, int8_t* __restrict__ first_choice, int node_base, int N) {
  const int d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= D) return;
  const int fn = first_node[d] < 0 ? -1 : first_node[d] - node_base;  // (see mig_finalize16_kernel)
  if (!first_choice) return;
  int8_t ch[NOSPLAN_MAX_GPUS_PER_NODE];
  for (int s = 0; s < NOSPLAN_MAX_GPUS_PER_NODE; ++s) ch[s] = -1;
  if (fn >= 0 && fn < N) {
    const int e0 = dem_off[d], k = dem_off[d + 1] - e0;
    int pid[kMaxKeys], qty[kMaxKeys];
    for (int j = 0; j < kMaxKeys; ++j) {
      pid[j] = j < k ? dem_key[e0 + j] : 0;
      qty[j] = j < k ? dem_qty[e0 + j] : 0;
    }
    const int g0 = gpu_off[fn], cnt = gpu_off[fn + 1] - g0;
    mig_walk_node(T, a_used, const_cast<uint4*>(a_free), a_model, g0, cnt, k, pid, qty, false, ch);
  }
  for (int s = 0; s < gs; ++s) first_choice[(size_t)d * gs + s] = ch[s];
}

__global__ void fill_i32_kernel(int32_t* p, int n, int32_t v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// GPU.InitGeometry over every GPU (gpu.go:120-129): apply the model's fewest-kinds geometry unless a
// used device would be deleted.
__global__ void mig_init_geometry_kernel(const MigTables* __restrict__ T, int tg, int P, const uint4* __restrict__ a_used,
                                         const uint4* __restrict__ a_free, const uint8_t* __restrict__ a_model,
                                         uint8_t* __restrict__ err, int32_t* __restrict__ out_free) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= tg) return;
  const uint4 u = a_used[g];
  uint4 f = a_free[g];
  const int m = a_model[g];
  const int ic = T->init_choice[m];
  bool e = true;
  if (ic >= 0) {
    const uint4 cd = T->cand[T->geom_off[m] + ic];
    if (all_ge4(cd, u)) {
      f = sub4(cd, u);
      e = false;
    }
  } else {
    // empty geometry list: nil geometry is not allowed by AllowsGeometry -> error
    e = true;
  }
  err[g] = e ? 1 : 0;
  for (int p = 0; p < P; ++p) out_free[(size_t)g * P + p] = (int32_t)byte_of(f, p);
}

// =================================================================================================================
// MPS ("slicing")
// Packed per-GPU state: free bytes [32] as two uint4 (key = size id), fmask = keys present in
// FreeProfiles, cap = MemoryGB - sum(size*used) (what free slices may occupy), used bytes kept apart
// (only AddPod and read-back need them).
struct MpsGpu {
  uint4 f0, f1;
  uint32_t fmask;
  uint32_t cap;
};

__device__ __forceinline__ uint32_t mps_byte(const uint4& lo, const uint4& hi, int i) {
  NSP_CHECK(i >= 0 && i < 32);
  return i < 16 ? byte_of(lo, i) : byte_of(hi, i - 16);
}

// slicing.GPU.UpdateGeometryFor (slicing/gpu.go:162-220) in closed form on one GPU.
//   sz[]: size in GB per key; demand keys sid[0..k) ascending (= ascending size, the sort of :174-182),
//   resid[]: remaining required quantity per key (0 = key absent).
// Outputs: created[j] slices created for demand key j that survive, restored = orig keys whose
// original count was put back (:214-216).  Returns updated.  FreeProfiles afterwards:
//   free[key] = (restored bit ? orig : 0) + created(key);  present iff restored bit or created > 0.
__device__ inline bool mps_gpu_update(const int* __restrict__ sz, const uint4& f0, const uint4& f1, uint32_t origmask,
                                      uint32_t cap, int k, const int* sid, const int* resid, int* created,
                                      uint32_t* restored_out) {
  NSP_CHECK(k >= 0 && k <= kMaxKeys);
  int miss[kMaxKeys];
  bool any_missing = false;
  for (int j = 0; j < k; ++j) {
    NSP_CHECK(sid[j] >= 0 && sid[j] < kMaxSizes);
    created[j] = 0;
    const int fj = (int)mps_byte(f0, f1, sid[j]);  // a present key holds its byte, an absent one 0
    miss[j] = max(resid[j] - fj, 0);               // getMissingSlices, :222-232
    any_missing |= miss[j] > 0;
  }
  *restored_out = origmask;
  if (!any_missing) return false;  // :166-168
  int origtot = 0;
  for (uint32_t mk = origmask; mk; mk &= mk - 1) {
    const int key = __ffs(mk) - 1;
    origtot += sz[key] * (int)mps_byte(f0, f1, key);
  }
  int spare = (int)cap - origtot;  // MemoryGB - getTotSlicesMemory (:256-265)
  uint32_t restored = origmask;
  bool updated = false;
  for (int j = 0; j < k; ++j) {
    if (miss[j] <= 0) continue;
    const int m = sz[sid[j]];
    // (a) spare capacity first, :188-197 — createSlice until it fails == min(q, spare / m)
    int k1 = min(miss[j], spare / m);
    created[j] += k1;
    spare -= k1 * m;
    miss[j] -= k1;
    updated |= k1 > 0;
    // (b) delete every key of the ORIGINAL free map, :199-202 (whole entries: also what was just created under them)
    for (uint32_t mk = restored; mk; mk &= mk - 1) {
      const int key = __ffs(mk) - 1;
      spare += sz[key] * (int)mps_byte(f0, f1, key);
    }
    restored = 0;
    for (int i = 0; i < k; ++i)
      if ((origmask >> sid[i]) & 1u) {
        spare += created[i] * sz[sid[i]];
        created[i] = 0;
      }
    // (c) create in the space made, :203-212
    int k2 = min(miss[j], spare / m);
    created[j] += k2;
    spare -= k2 * m;
    miss[j] -= k2;
    updated |= k2 > 0;
    // (d) restore the original free slices, ascending size, all-or-nothing per key, :214-216
    for (uint32_t mk = origmask; mk; mk &= mk - 1) {
      const int key = __ffs(mk) - 1;
      const int need = sz[key] * (int)mps_byte(f0, f1, key);
      if (spare >= need) {
        restored |= 1u << key;
        spare -= need;
      }
    }
  }
  *restored_out = restored;
  return updated;
}

// slicing.Node.UpdateGeometryFor (slicing/node.go:120-152) over GPUs g[0..cnt).  If commit, the
// packed state is rewritten.  out_free/out_mask (may be null): FreeProfiles after the walk.
template <bool CG = false>
__device__ inline bool mps_walk_node(const int* __restrict__ sz, int S, MpsGpu* gp, int g0, int cnt, int k,
                                     const int* __restrict__ sid, const int* __restrict__ qty, bool commit,
                                     uint8_t* out_free, uint32_t* out_mask) {
  NSP_CHECK(k >= 0 && k <= kMaxKeys && S >= 0 && S <= kMaxSizes);
  int resid[kMaxKeys];
  for (int j = 0; j < kMaxKeys; ++j) resid[j] = (j < k) ? max(qty[j], 0) : 0;
  bool any = false;
  for (int s = 0; s < cnt; ++s) {
    MpsGpu G;
    G.f0 = ld_state<CG>(&gp[g0 + s].f0);
    G.f1 = ld_state<CG>(&gp[g0 + s].f1);
    if constexpr (CG) {
      G.fmask = __ldcg(&gp[g0 + s].fmask);
      G.cap = __ldcg(&gp[g0 + s].cap);
    } else {
      G.fmask = gp[g0 + s].fmask;
      G.cap = gp[g0 + s].cap;
    }
    int created[kMaxKeys];
    uint32_t restored = G.fmask;
    bool upd = false;
    if (k > 0) upd = mps_gpu_update(sz, G.f0, G.f1, G.fmask, G.cap, k, sid, resid, created, &restored);
    any |= upd;
    if (!upd) {
      // untouched (when nothing was created every original key is restored: state identical)
      for (int j = 0; j < k; ++j) resid[j] = max(resid[j] - (int)mps_byte(G.f0, G.f1, sid[j]), 0);
      if (out_free) {
        for (int key = 0; key < S; ++key) out_free[(size_t)s * S + key] = (uint8_t)mps_byte(G.f0, G.f1, key);
        out_mask[s] = G.fmask;
      }
      continue;
    }
    uint32_t w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    uint32_t nmask = restored;
    for (uint32_t mk = restored; mk; mk &= mk - 1) {
      const int key = __ffs(mk) - 1;
      w[key >> 2] |= mps_byte(G.f0, G.f1, key) << ((key & 3) * 8);
    }
    for (int j = 0; j < k; ++j) {
      if (created[j] > 0) {
        const int key = sid[j];
        const uint32_t old = (w[key >> 2] >> ((key & 3) * 8)) & 0xffu;
        w[key >> 2] = (w[key >> 2] & ~(0xffu << ((key & 3) * 8))) | ((old + (uint32_t)created[j]) << ((key & 3) * 8));
        nmask |= 1u << key;
      }
    }
    const uint4 n0 = make_uint4(w[0], w[1], w[2], w[3]), n1 = make_uint4(w[4], w[5], w[6], w[7]);
    for (int j = 0; j < k; ++j) resid[j] = max(resid[j] - (int)mps_byte(n0, n1, sid[j]), 0);
    if (commit) {
      st_state<CG>(&gp[g0 + s].f0, n0);
      st_state<CG>(&gp[g0 + s].f1, n1);
      if constexpr (CG) __stcg(&gp[g0 + s].fmask, nmask);
      else gp[g0 + s].fmask = nmask;
    }
    if (out_free) {
      for (int key = 0; key < S; ++key) out_free[(size_t)s * S + key] = (uint8_t)mps_byte(n0, n1, key);
      out_mask[s] = nmask;
    }
  }
  return any;
}

// MPS: one thread per node walks its GPUs.  code = gpu * 128 + (0: memory out of range | 1 + 2s: count
// of size s out of range | 2 + 2s: free_mask misses a non-zero free count | 127: slices exceed memory,
// the GPU slicing.NewGPU rejects, slicing/gpu.go:92-95).
__global__ void mps_pack_kernel(int N, int S, int max_count, int max_mem, const int32_t* __restrict__ sizes,
                                const int32_t* __restrict__ gpu_off, const uint8_t* __restrict__ valid,
                                const int32_t* __restrict__ mem, const int32_t* __restrict__ used, const int32_t* __restrict__ fr,
                                const uint32_t* __restrict__ fmask, MpsGpu* __restrict__ gp, uint4* __restrict__ used_lo,
                                uint4* __restrict__ used_hi, int32_t* __restrict__ gnode, int32_t* __restrict__ gslot,
                                unsigned long long* __restrict__ status) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const bool ok = valid[n] != 0;
  unsigned long long bad = ~0ull;
  const int g0 = gpu_off[n], g1 = gpu_off[n + 1];
  for (int g = g0; g < g1; ++g) {
    uint32_t uw[8] = {0, 0, 0, 0, 0, 0, 0, 0}, fw[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    long long tot = 0;
    int usedtot = 0;
    uint32_t fm = 0;
    const int mg = ok ? mem[g] : 0;
    const unsigned long long base = (unsigned long long)g * 128ull;
    if (mg < 0 || mg > max_mem) bad = min(bad, base);
    for (int s = 0; s < S; ++s) {
      const int u = ok ? used[(size_t)g * S + s] : 0, f = ok ? fr[(size_t)g * S + s] : 0;
      const bool bit = fmask ? ((fmask[g] >> s) & 1u) : (f > 0);
      if (u < 0 || f < 0 || u > max_count || f > max_count || u + f > max_count) bad = min(bad, base + 1ull + 2ull * s);
      else if (f > 0 && !bit) bad = min(bad, base + 2ull + 2ull * s);
      uw[s >> 2] |= (uint32_t)(u & 0xff) << ((s & 3) * 8);
      fw[s >> 2] |= (uint32_t)(f & 0xff) << ((s & 3) * 8);
      usedtot += sizes[s] * u;
      tot += (long long)sizes[s] * ((long long)u + f);
      if (bit) fm |= 1u << s;
    }
    if (ok && tot > mg) bad = min(bad, base + 127ull);
    MpsGpu G;
    G.f0 = make_uint4(fw[0], fw[1], fw[2], fw[3]);
    G.f1 = make_uint4(fw[4], fw[5], fw[6], fw[7]);
    G.fmask = fm;
    G.cap = (uint32_t)(mg - usedtot);
    gp[g] = G;
    used_lo[g] = make_uint4(uw[0], uw[1], uw[2], uw[3]);
    used_hi[g] = make_uint4(uw[4], uw[5], uw[6], uw[7]);
    gnode[g] = ok ? n : -1;
    gslot[g] = g - g0;
  }
  if (bad != ~0ull) atomicMin(&status[0], bad);
  if (ok && g1 > g0) atomicAdd(&status[1], (unsigned long long)(g1 - g0));
}

__global__ void mps_unpack_kernel(int tg, int S, const MpsGpu* __restrict__ gp, const uint4* __restrict__ used_lo,
                                  const uint4* __restrict__ used_hi, int32_t* __restrict__ used, int32_t* __restrict__ fr,
                                  uint32_t* __restrict__ fmask) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= tg) return;
  const MpsGpu G = gp[g];
  const uint4 u0 = used_lo[g], u1 = used_hi[g];
  for (int s = 0; s < S; ++s) {
    used[(size_t)g * S + s] = (int32_t)mps_byte(u0, u1, s);
    fr[(size_t)g * S + s] = (int32_t)mps_byte(G.f0, G.f1, s);
  }
  if (fmask) fmask[g] = G.fmask;
}

// MPS batched scorer: one thread per (demand, node); lanes of a warp = consecutive nodes of one demand.
struct MpsScoreArgs {
  const int32_t* sizes;
  int S;
  const MpsGpu* gp;
  const int32_t* gpu_off;
  const uint8_t* valid;
  int N, tg, gs;
  int D, d_chunk;
  const int32_t* dem_off;
  const int32_t* dem_key;
  const int32_t* dem_qty;
  int32_t* first_node;  // unsigned atomicMin target, pre-filled with -1
  int node_base;
  uint8_t* updated;
  uint8_t* out_free;    // [D][tg][S] or null
  uint32_t* out_mask;   // [D][tg] or null
  // mode 0: rows [blockIdx.y * d_chunk, ...) of all D, tile = blockIdx.x (plain)
  // mode 1: work item blockIdx.x = (size id, row range of rows[]), tile blockIdx.y of perm[size id]: the
  //         nodes of a warp walk about equally far for that size (see nosplan_score16.cuh)
  // mode 2: rows[meta[1] .. + meta[2]) (the rows with several keys), plain tiles
  int mode;
  const int32_t* perm;  // [S][N]
  const int32_t* rows;
  const int32_t* work;
  const int32_t* meta;
  const uint8_t* row_ok;        // optional: rows_check_kernel's verdict per row (0 = breaks an ABI limit: skipped)
  PushArgs push;                // mps_score1_kernel: where the per-row winners go
  unsigned long long* counter;  // its CNT instantiation: (row, node, GPU) runs that reached the closed form
};

// Shared-memory tile: per (node, slot) the free bytes as nw = ceil(S/4) words, the key mask, and
// spare0 = cap - sum(size*free) (low 16 bits) | sum(size*free) (high 16 bits).
__host__ __device__ inline size_t mps_score_smem_bytes(int gs, int S, int tn) {
  const size_t nw = (size_t)(S + 3) / 4;
  return (size_t)gs * tn * 4 * (nw + 2);
}

// exact floor(n / m) for n, m < 2^16 with inv = (m == 1 ? 0 : 0xFFFFFFFF / m + 1)
__device__ __forceinline__ int div_small(int n, int m, uint32_t inv) { return m == 1 ? n : (int)__umulhi((uint32_t)n, inv); }

// TN nodes per tile, DS threads per node: thread j of them takes every DS-th demand row of the chunk
// (more resident warps over the same shared-memory tile; the per-row work is a dependent chain).
template <int TN, int DS>
__global__ void __launch_bounds__(TN * DS, 3) mps_score_kernel(const MpsScoreArgs a) {
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ int s_sz[kMaxSizes];
  __shared__ uint32_t s_inv[kMaxSizes];  // per size: multiplier of the exact multiply-high division
  const int tid = threadIdx.x % TN;
  const int dsub = threadIdx.x / TN;
  const int nw = (a.S + 3) >> 2;
  uint32_t* s_free = reinterpret_cast<uint32_t*>(smem);       // [gs][nw][TN]
  uint32_t* s_mask = s_free + (size_t)a.gs * nw * TN;          // [gs][TN]
  uint32_t* s_spare = s_mask + (size_t)a.gs * TN;              // [gs][TN]
  if (threadIdx.x < kMaxSizes) {
    const int m = threadIdx.x < a.S ? a.sizes[threadIdx.x] : 1;
    s_sz[threadIdx.x] = m;
    s_inv[threadIdx.x] = m == 1 ? 0u : 0xFFFFFFFFu / (uint32_t)m + 1u;
  }
  // which rows, which nodes (uniform over the CTA; a CTA without work leaves before any barrier)
  const int32_t* rl = nullptr;
  int r0 = 0, rlen = 0, tile = blockIdx.x, bucket = -1;
  if (a.mode == 1) {
    if ((int)blockIdx.x >= a.meta[0]) return;
    bucket = a.work[3 * blockIdx.x];
    rl = a.rows + a.work[3 * blockIdx.x + 1];
    rlen = a.work[3 * blockIdx.x + 2];
    tile = blockIdx.y;
  } else if (a.mode == 2) {
    r0 = blockIdx.y * a.d_chunk;
    if (r0 >= a.meta[2]) return;
    rl = a.rows + a.meta[1] + r0;
    rlen = min(a.d_chunk, a.meta[2] - r0);
  } else {
    r0 = blockIdx.y * a.d_chunk;
    rlen = min(a.d_chunk, a.D - r0);
  }
  __syncthreads();
  const int pos = tile * TN + tid;
  const int n = bucket >= 0 ? (pos < a.N ? a.perm[(size_t)bucket * a.N + pos] : a.N) : pos;
  const bool live = n < a.N && a.valid[n];
  const int g0 = n < a.N ? a.gpu_off[n] : 0;
  const int cnt = n < a.N ? a.gpu_off[n + 1] - g0 : 0;
  for (int s = dsub; s < cnt; s += DS) {  // the DS threads of a node split its GPUs
    const MpsGpu G = a.gp[g0 + s];
    int origtot = 0;
    for (uint32_t mk = G.fmask; mk; mk &= mk - 1) {
      const int key = __ffs(mk) - 1;
      origtot += s_sz[key] * (int)mps_byte(G.f0, G.f1, key);
    }
    const uint32_t w[8] = {G.f0.x, G.f0.y, G.f0.z, G.f0.w, G.f1.x, G.f1.y, G.f1.z, G.f1.w};
    for (int i = 0; i < nw; ++i) s_free[((size_t)s * nw + i) * TN + tid] = w[i];
    s_mask[s * TN + tid] = G.fmask;
    s_spare[s * TN + tid] = ((uint32_t)origtot << 16) | (uint32_t)((int)G.cap - origtot);
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  for (int i = dsub; i < rlen; i += DS) {
    const int d = rl ? rl[i] : r0 + i;
    if (a.row_ok && !a.row_ok[d]) continue;  // uniform over the CTA
    const int e0 = __ldg(a.dem_off + d), k = __ldg(a.dem_off + d + 1) - e0;
    bool any = false;
    if (k == 1 && !a.out_free) {
      // one requested size: slicing.GPU.UpdateGeometryFor in closed form, registers only
      const int sid = __ldg(a.dem_key + e0);
      int resid = max(__ldg(a.dem_qty + e0), 0);
      const int m = s_sz[sid];
      const uint32_t inv = s_inv[sid];
      const int wi = sid >> 2, sh = (sid & 3) * 8;
      if (live) {
        for (int s = 0; s < cnt && resid > 0; ++s) {
          const int fj = (s_free[((size_t)s * nw + wi) * TN + tid] >> sh) & 0xff;
          const int miss = resid - fj;  // getMissingSlices
          if (miss <= 0) break;         // this GPU already provides it; the residual drops to 0
          const uint32_t sp = s_spare[s * TN + tid];
          const uint32_t omask = s_mask[s * TN + tid];
          const int otot = (int)(sp >> 16);
          int spare = (int)(sp & 0xffffu);
          const bool in_orig = (omask >> sid) & 1u;
          // (a) spare capacity first
          const int k1 = min(miss, div_small(spare, m, inv));
          // (b) every original free key is deleted — with it whatever (a) created under such a key
          spare = spare - k1 * m + otot + (in_orig ? k1 * m : 0);
          int created = in_orig ? 0 : k1;
          // (c) create in the space made
          const int k2 = min(miss - k1, div_small(spare, m, inv));
          created += k2;
          spare -= k2 * m;
          const bool upd = (k1 | k2) > 0;
          // (d) restore the original keys, ascending size, each all-or-nothing
          bool restored_sid = in_orig;
          if (spare < otot) {
            restored_sid = false;
            for (uint32_t mk = omask; mk; mk &= mk - 1) {
              const int key = __ffs(mk) - 1;
              const int need = s_sz[key] * (int)((s_free[((size_t)s * nw + (key >> 2)) * TN + tid] >> ((key & 3) * 8)) & 0xff);
              if (spare >= need) {
                spare -= need;
                if (key == sid) restored_sid = true;
              }
            }
          }
          const int nf = upd ? created + (restored_sid ? fj : 0) : fj;
          any |= upd;
          resid = max(resid - nf, 0);
        }
      }
      if (a.updated && n < a.N) a.updated[(size_t)d * a.N + n] = any ? 1 : 0;
    } else {
      int sid[kMaxKeys], qty[kMaxKeys];
      for (int j = 0; j < kMaxKeys; ++j) {
        sid[j] = j < k ? __ldg(a.dem_key + e0 + j) : 0;
        qty[j] = j < k ? __ldg(a.dem_qty + e0 + j) : 0;
      }
      if (n < a.N) {
        uint8_t* of = a.out_free ? a.out_free + ((size_t)d * a.tg + g0) * a.S : nullptr;
        uint32_t* om = a.out_mask ? a.out_mask + (size_t)d * a.tg + g0 : nullptr;
        // an invalid node is never updated; its (untouched) layout is still reported
        any = mps_walk_node(s_sz, a.S, const_cast<MpsGpu*>(a.gp), g0, cnt, live ? k : 0, sid, qty, false, of, om);
        if (a.updated) a.updated[(size_t)d * a.N + n] = any ? 1 : 0;
      }
    }
    // -1 read as unsigned is above every node index: the minimum needs no final pass
    const unsigned mine = any ? (unsigned)(n + a.node_base) : 0xffffffffu;
    const unsigned wmin = __reduce_min_sync(0xffffffffu, mine);
    if (lane == 0 && wmin != 0xffffffffu) atomicMin(reinterpret_cast<unsigned*>(a.first_node) + d, wmin);
  }
}

}  // namespace nsp
