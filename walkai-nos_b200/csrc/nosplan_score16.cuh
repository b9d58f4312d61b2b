// nosplan_score16.cuh — the MIG batched scorer for the common tables (<= 16 geometries per model,
// quantities <= 63: 16-bit keys), organised so that the lanes of a warp stay together.
//
// What ncu showed on the plain tile kernel (profiles/r1_divergence.md): one warp = 32 consecutive
// nodes of one demand row, and a node leaves the GPU loop at the first GPU that covers the request
// (node.go:164-169), so ~14 of 32 lanes were alive in the candidate core.  How far a node walks depends
// on the requested profile much more than on the node.  Hence:
//   * at snapshot time every node gets, per profile, the cost of a quantity-1 walk (GPUs visited, GPUs
//     whose candidates get evaluated) and the nodes are counting-sorted by it, once per profile:
//     perm[p][position] = node.  This is scheduling data only; nothing of the result is precomputed.
//   * per scoring call the demand rows are bucketed by their profile (single-key rows; rows with
//     several keys go to the general kernel) and cut into work items (profile, row range).
//   * one CTA = (tile of 128 consecutive POSITIONS of perm[p], work item of profile p): its warps hold
//     nodes that walk about equally far for that profile.  The tile only needs profile p's column of
//     the state, so it is 33 KB instead of 60 KB.
//   * the demand-independent part of a candidate's score is evaluated when the snapshot arrives
//     (mig_rank16_kernel) instead of in every CTA's prologue.
// The arithmetic per (row, node, GPU, candidate) is the K16 path of mig_score_kernel, unchanged.
#pragma once
#include "nosplan_seq.cuh"
#include "nosplan_xchg.cuh"

namespace nsp {

constexpr int kRowCap = 512;     // rows per work item (their ids and quantities are staged in shared memory)
constexpr int kCostBuckets = 96;  // 2 * visited + 3 * evaluated <= 5 * 16

// ---- snapshot time ------------------------------------------------------------------------------------
// cost[p][n]: relative cost of scoring a quantity-1 request of profile p on node n.
__global__ void mig_cost_kernel(int N, int npad, int gs, int P, const uint8_t* __restrict__ s_cnt,
                                const uint32_t* __restrict__ maskg, uint8_t* __restrict__ cost) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= npad) return;
  const int cnt = n < N ? s_cnt[n] : 0;
  uint32_t open = (1u << P) - 1u;  // profiles whose walk is still going
  uint8_t c[16];
#pragma unroll
  for (int p = 0; p < 16; ++p) c[p] = 0;
  for (int s = 0; s < cnt && open; ++s) {
    const uint32_t mk = maskg[(size_t)n * gs + s];
    const uint32_t f1 = mk & 0xffffu, gm = mk >> 16;
#pragma unroll
    for (int p = 0; p < 16; ++p) {
      if (!((open >> p) & 1u)) continue;
      c[p] += 2;                         // visited
      if (!((f1 >> p) & 1u)) c[p] += 3;  // nothing free here: its candidates get evaluated
    }
    open &= ~(f1 | gm);  // a free device or a geometry that adds one covers a quantity-1 request
  }
  for (int p = 0; p < P; ++p) cost[(size_t)p * npad + n] = c[p];
}

// The demand-independent part of gpu.go:168-188 (CanApplyGeometry, geometryDistance, totalSliceCount, the
// id rank) depends only on the snapshot, so it is evaluated once when the snapshot arrives instead of in
// the prologue of every CTA of every scoring call: per (node, slot) the rank half-words of all
// candidates, and per (node, profile, slot) free | (used + free) << 8.
//   rankg [npad][gs][2cw] u32, ufg [16][npad][gs] u16 (profile-major: a profile's column is contiguous over nodes), maskg [npad][gs] u32 (profiles with a free device |
//   profiles some applicable geometry adds to, << 16)
__global__ void mig_rank16_kernel(const MigTables* __restrict__ T, int npad, int gs, int cw, const uint4* __restrict__ s_used,
                                  const uint4* __restrict__ s_free, const uint8_t* __restrict__ s_model,
                                  const uint8_t* __restrict__ s_cnt, uint32_t* __restrict__ rankg, uint16_t* __restrict__ ufg,
                                  uint32_t* __restrict__ maskg, uint8_t* __restrict__ modelg) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npad * gs) return;
  const int n = i / gs, s = i % gs;
  uint32_t* out = rankg + ((size_t)n * gs + s) * (2 * cw);
  modelg[(size_t)n * gs + s] = s < s_cnt[n] ? s_model[(size_t)s * npad + n] : 0;  // node-major copy (mig_stream_kernel)
  if (s >= s_cnt[n]) {
    for (int w = 0; w < 2 * cw; ++w) out[w] = 0x80008000u;
    for (int p = 0; p < 16; ++p) ufg[((size_t)p * npad + n) * gs + s] = 0;
    maskg[(size_t)n * gs + s] = 0;
    return;
  }
  const uint4 u = s_used[(size_t)s * npad + n];
  const uint4 f = s_free[(size_t)s * npad + n];
  const int m = s_model[(size_t)s * npad + n];
  const uint4 cur = add4(u, f);
  const int c0 = T->geom_off[m], cc = T->geom_off[m + 1] - c0;
  for (int p = 0; p < 16; ++p) ufg[((size_t)p * npad + n) * gs + s] = (uint16_t)(byte_of(f, p) | (byte_of(cur, p) << 8));
  uint32_t k0[16];
  uint32_t gm = 0;  // profiles some applicable geometry would add to
  for (int c = 0; c < 16; ++c) {
    k0[c] = 0u;
    if (c < cc) {
      const uint4 cd = T->cand[c0 + c];
      if (all_ge4(cd, u)) {
        k0[c] = key0_of(T->total[c0 + c], sad16(cd, cur), T->rank[c0 + c]);
        gm |= gt_bytes16(cd, cur);
      }
    }
  }
  maskg[(size_t)n * gs + s] = nonzero_bytes16(f) | (gm << 16);  // for mig_cost_kernel
  for (int w = 0; w < 2 * cw; ++w) {
    uint32_t r2 = 0;
    for (int b = 0; b < 2; ++b) {
      const int c = w * 2 + b;
      if (k0[c] != 0u) {
        uint32_t r = 1;
        for (int c2 = 0; c2 < 16; ++c2) r += (k0[c2] != 0u && k0[c2] < k0[c]) ? 1u : 0u;
        r2 |= ((r << 4) | (uint32_t)(15 - c)) << (16 * b);
      } else {
        r2 |= 0x8000u << (16 * b);  // cannot be applied (or no such candidate): the lane's key stays negative
      }
    }
    out[w] = r2;
  }
}

// perm[p][.]: nodes by descending cost (heavy tiles are scheduled first); one block per profile.
__global__ void mig_perm_kernel(int npad, const uint8_t* __restrict__ cost, int32_t* __restrict__ perm,
                                int32_t* __restrict__ prof_cost) {
  __shared__ int s_off[kCostBuckets];
  const int p = blockIdx.x;
  const uint8_t* c = cost + (size_t)p * npad;
  int32_t* out = perm + (size_t)p * npad;
  for (int i = threadIdx.x; i < kCostBuckets; i += blockDim.x) s_off[i] = 0;
  __syncthreads();
  for (int n = threadIdx.x; n < npad; n += blockDim.x) atomicAdd(&s_off[min((int)c[n], kCostBuckets - 1)], 1);
  __syncthreads();
  if (threadIdx.x == 0) {
    int run = 0;
    long long tot = 0;
    for (int b = kCostBuckets - 1; b >= 0; --b) {
      const int v = s_off[b];
      s_off[b] = run;
      run += v;
      tot += (long long)v * b;
    }
    prof_cost[p] = (int32_t)min(tot, 0x7fffffffLL);  // orders the work items of a scoring call
  }
  __syncthreads();
  for (int n = threadIdx.x; n < npad; n += blockDim.x) out[atomicAdd(&s_off[min((int)c[n], kCostBuckets - 1)], 1)] = n;
}

// MPS twin: cost[sid][n] of a quantity-1 request of size sid on node n (a GPU covers it if it has a free
// slice of the size, or room for one once its free slices are deleted, slicing/gpu.go:188-212).
__global__ void mps_cost_kernel(int N, int S, const int32_t* __restrict__ sizes, const int32_t* __restrict__ gpu_off,
                                const uint8_t* __restrict__ valid, const MpsGpu* __restrict__ gp, uint8_t* __restrict__ cost) {
  __shared__ int s_sz[kMaxSizes];
  if (threadIdx.x < kMaxSizes) s_sz[threadIdx.x] = threadIdx.x < S ? sizes[threadIdx.x] : 0x7fffffff;
  __syncthreads();
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const int g0 = gpu_off[n], cnt = valid[n] ? gpu_off[n + 1] - g0 : 0;
  uint32_t open = S >= 32 ? 0xffffffffu : ((1u << S) - 1u);
  uint8_t c[kMaxSizes];
#pragma unroll
  for (int k = 0; k < kMaxSizes; ++k) c[k] = 0;
  for (int s = 0; s < cnt && open; ++s) {
    const MpsGpu G = gp[g0 + s];
    const uint32_t f1 = mps_free1_mask(G.f0, G.f1);
    uint32_t room = 0;
#pragma unroll
    for (int k = 0; k < kMaxSizes; ++k) room |= ((uint32_t)s_sz[k] <= G.cap) ? (1u << k) : 0u;
#pragma unroll
    for (int k = 0; k < kMaxSizes; ++k) {
      if (!((open >> k) & 1u)) continue;
      c[k] = (uint8_t)min(c[k] + 2 + (((f1 >> k) & 1u) ? 0 : 3), 255);
    }
    open &= ~(f1 | room);
  }
  for (int k = 0; k < S; ++k) cost[(size_t)k * N + n] = c[k];
}

// ---- per call: rows by key -------------------------------------------------------------------------------------
// (MIG profiles and MPS sizes alike: "P" below is the number of keys of the kind, <= 32.)
// One launch.  Every row goes to the list of its bucket (bucket p < P: single-key rows of key p;
// bucket P: every other row); bucket b owns rows[b * D, (b + 1) * D), filled through an atomic cursor.
// The block that finishes last turns the counts into work items {key, start in rows[], length <= chunk},
// heaviest key first, and meta[0] = number of items, meta[1] = start of bucket P, meta[2] = its size; it
// leaves the cursors and the ticket at zero for the next call.  first_node[] is put to rest (-1) on the way.
constexpr int kRowBuckets = 33;  // up to 32 keys + "any other row"
constexpr int kNoRow = 63;

// The rows are validated here as well (the host does not walk them): a row that breaks a limit of the ABI
// joins no bucket, so no scorer ever sees it, and the lowest failing code row * 64 + sub of the call lands in *status
// (sub 0: dem_off descending, 1: too many keys, 2 + 3j / 3 + 3j / 4 + 3j: key j out of range / not
// ascending / quantity out of range); the host turns the code back into its message.
// state[0 .. kRowBuckets) = cursors, state[kRowBuckets] = ticket; all zero on entry and on exit
struct RowsGroupArgs {
  int D, P, chunk, max_keys, max_qty;
  int single_only;  // 1: a row that does not have exactly one key is an error (sub code kSubNotSingle)
  const int32_t* dem_off;
  const int32_t* dem_key;
  const int32_t* dem_qty;
  const int32_t* key_cost;
  int32_t* state;
  int32_t* rows;
  int32_t* work;
  int32_t* meta;
  int32_t* first_node;  // put to rest (-1) here; null = leave it alone
  unsigned long long* status;  // highest (call number << 32 | ~code) seen: never reset, the caller compares call numbers
  uint32_t status_call;
  // sharded pass: first_node is the rank's own half of the exchange buffer, which may only be overwritten once
  // every rank has read what it held two steps ago (nosplan_xchg.cuh); wait_world == 0: nothing to wait for
  const uint32_t* wait_read;
  int wait_world;
  uint32_t wait_need;
  int* wait_err;
};
constexpr int kSubNotSingle = 62;

// The ABI's limits on one demand row: -1 = fine, else the sub code (0: dem_off descending, 1: too many keys,
// 2 + 3j / 3 + 3j / 4 + 3j: key j out of range / not ascending / quantity out of range, kSubNotSingle).
__device__ __forceinline__ int row_check_sub(const RowsGroupArgs& g, int d, int* k_out, int* key0_out) {
  const int e0 = __ldg(g.dem_off + d), k = __ldg(g.dem_off + d + 1) - e0, nnz = __ldg(g.dem_off + g.D);
  *k_out = k;
  *key0_out = 0;
  if (k < 0 || e0 < 0 || e0 + k > nnz) return 0;  // entries outside [0, dem_off[D]) are never read
  if (k > g.max_keys) return 1;
  int prev = -1;
  for (int j = 0; j < k; ++j) {
    const int key = __ldg(g.dem_key + e0 + j), q = __ldg(g.dem_qty + e0 + j);
    if (key < 0 || key >= g.P) return 2 + 3 * j;
    if (j > 0 && key <= prev) return 3 + 3 * j;
    if (q < 0 || q > g.max_qty) return 4 + 3 * j;
    if (j == 0) *key0_out = key;
    prev = key;
  }
  if (g.single_only && k != 1) return kSubNotSingle;
  return -1;
}
__device__ __forceinline__ void row_report(const RowsGroupArgs& g, int d, int sub) {
  // the lowest code of this call wins: codes are stored inverted under the call number
  atomicMax(g.status, ((unsigned long long)g.status_call << 32) | (unsigned long long)(0xffffffffu - ((uint32_t)d * 64u + (uint32_t)sub)));
}

// The scorers that take the rows in input order (wide tables, NOSPLAN_PLAIN_ORDER) get the same check as a
// pre-pass: row_ok[d] = 1 for a row inside the limits (the kernels skip the others: their winner stays -1, nothing
// of them is indexed), the first bad row lands in *status, first_node[] is put to rest.
__global__ void rows_check_kernel(const RowsGroupArgs g, uint8_t* __restrict__ row_ok) {
  const int d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= g.D) return;
  if (g.first_node) g.first_node[d] = -1;
  int k, key0;
  const int sub = row_check_sub(g, d, &k, &key0);
  row_ok[d] = sub < 0 ? 1 : 0;
  if (sub >= 0) row_report(g, d, sub);
}

// block = index of this CTA among the n_blocks CTAs that group this kind's rows
__device__ __forceinline__ void rows_group_body(const RowsGroupArgs& g, int block, int n_blocks) {
  __shared__ int s_last;
  if (g.wait_world) wait_flags(g.wait_read, g.wait_world, g.wait_need, 2000000000ull, g.wait_err);
  const int D = g.D, P = g.P, chunk = g.chunk;
  const int d = block * blockDim.x + threadIdx.x;  // blockDim is a multiple of 32: whole warps reach the match
  const int lane = threadIdx.x & 31;
  int b = kNoRow;
  if (d < D) {
    if (g.first_node) g.first_node[d] = -1;  // "no node"; the scorers fold winners in with an unsigned atomicMin
    int k, key0;
    const int sub = row_check_sub(g, d, &k, &key0);
    if (sub >= 0) row_report(g, d, sub);
    else b = k == 1 ? key0 : P;
  }
  const uint32_t peers = __match_any_sync(0xffffffffu, b);  // one atomic per bucket per warp
  const int leader = __ffs(peers) - 1;
  int base = 0;
  if (b != kNoRow && lane == leader) base = atomicAdd(&g.state[b], __popc(peers));
  base = __shfl_sync(0xffffffffu, base, leader);
  NSP_CHECK(b == kNoRow || (b >= 0 && b <= P && base + __popc(peers & ((1u << lane) - 1u)) < D));
  if (b != kNoRow) g.rows[(size_t)b * D + base + __popc(peers & ((1u << lane) - 1u))] = d;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(&g.state[kRowBuckets], 1) == n_blocks - 1;
  __syncthreads();
  if (!s_last || threadIdx.x >= 32) return;
  __threadfence();
  // the last block, one warp: lane b holds bucket b
  const uint32_t full = 0xffffffffu;
  const int my_cnt = lane < P ? __ldcg(&g.state[lane]) : 0;
  const int my_cost = lane < P ? g.key_cost[lane] : -1;
  const int rest = __ldcg(&g.state[P]);
  // work items by descending key cost (ties: lower key first): the long CTAs start first
  int rank = 0;
  for (int j = 0; j < 32; ++j) {
    const int cj = __shfl_sync(full, my_cost, j);
    if (j < P && (cj > my_cost || (cj == my_cost && j < lane))) ++rank;
  }
  const int my_items = (my_cnt + chunk - 1) / chunk;
  int before = 0;  // items of the keys ranked before mine
  for (int j = 0; j < 32; ++j) {
    const int rj = __shfl_sync(full, rank, j), ij = __shfl_sync(full, my_items, j);
    if (j < P && rj < rank) before += ij;
  }
  if (lane < P)
    for (int it = 0; it < my_items; ++it) {
      g.work[3 * (before + it) + 0] = lane;
      g.work[3 * (before + it) + 1] = lane * D + it * chunk;
      g.work[3 * (before + it) + 2] = min(chunk, my_cnt - it * chunk);
    }
  int nw = lane < P ? my_items : 0;
  for (int o = 16; o; o >>= 1) nw += __shfl_xor_sync(full, nw, o);
  if (lane == 0) {
    g.meta[0] = nw;
    g.meta[1] = P * D;
    g.meta[2] = rest;
  }
  __syncwarp();
  for (int i = lane; i <= kRowBuckets; i += 32) g.state[i] = 0;
}

__global__ void rows_group_kernel(const RowsGroupArgs g) { rows_group_body(g, blockIdx.x, gridDim.x); }

// both kinds of a scoring pass in one launch: the first nb0 CTAs group kind 0's rows, the rest kind 1's
__global__ void rows_group2_kernel(const RowsGroupArgs g0, const RowsGroupArgs g1, int nb0) {
  if ((int)blockIdx.x < nb0) rows_group_body(g0, blockIdx.x, nb0);
  else rows_group_body(g1, blockIdx.x - nb0, gridDim.x - nb0);
}

// ---- the scorer -------------------------------------------------------------------------------------------
struct MigScore16Args {
  MigScoreArgs base;          // tables, slot-major state, demand CSR, outputs
  const int32_t* perm;        // [P][npad]
  const uint32_t* rankg;      // [npad][gs][2cw]  (mig_rank16_kernel)
  const uint16_t* ufg;        // [16][npad][gs]
  const int32_t* rows;        // grouped row ids
  const int32_t* work;        // work items
  const int32_t* meta;
  PushArgs push;                  // where the per-row winners go (plain first_node[] or the exchange buffers)
  unsigned long long* counter;    // CNT instantiation: candidate evaluations that reached the DPX core
};

// Shared memory of one CTA (all u32, addressed through 32-bit shared-window addresses):
//   ct2   [n_models][2cw]        profile p's quantity in every candidate, one candidate per 16-bit lane
//   slot  [gs][2cw + 2][TN]      per GPU slot, one column per node: 2cw words of rank half-words
//                                (rank << 4 | (15 - c), 0x8000 = cannot be applied), then
//                                uf = free | (-(used + free) & 0xffff) << 16 of profile p, then
//                                mo = shared address of the ct2 row of the GPU's model
//   row/qty/first [kRowCap]      the work item's rows, their quantities, their first updating node so far
// One pointer walks the slots (stride (2cw + 2) * TN words), every load has a constant offset from it.
__host__ __device__ inline size_t mig_score16_smem_bytes(int n_models, int gs, int cmax, int tn) {
  const size_t cw = (size_t)(cmax + 3) / 4;
  return ((size_t)n_models * 2 * cw * 4 + 15) / 16 * 16 + (size_t)gs * (2 * cw + 2) * tn * 4 + (size_t)kRowCap * 12;
}

__device__ __forceinline__ uint32_t lds_u32(uint32_t saddr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(saddr));
  return v;
}
__device__ __forceinline__ uint2 lds_v2(uint32_t saddr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(saddr));
  return v;
}

// OUT: the call wants updated[] / choice[] (the per-row stores stay out of the hot instantiation)
// RB: demand rows a thread scores per pass over its node's GPUs
// CNT: measurement instantiation — counts the (row, node, GPU, candidate) evaluations that really execute
//      (a (row, GPU) pair that enters the candidate core counts the model's candidates), see bench.py
// bx, by: the CTA's place in the (work item, tile) grid (the kernel's blockIdx, or derived from a linear index
// when both kinds share one launch, score_both_kernel)
template <int TN, int DS, int CW, bool OUT, int RB, bool CNT>
__device__ __forceinline__ void mig_score16_body(const MigScore16Args& A, const int bx, const int by) {
  const MigScoreArgs& a = A.base;
  if (bx >= A.meta[0]) {  // the grid is sized for the worst split of the rows
    push_finish(A.push);
    return;
  }
  unsigned long long executed = 0;
  extern __shared__ __align__(16) unsigned char smem[];
  constexpr int cw = CW;
  constexpr uint32_t kSlotStride = (2 * cw + 2) * TN * 4;  // bytes
  constexpr uint32_t kUfOff = 2 * cw * TN * 4;
  constexpr uint32_t kMoOff = (2 * cw + 1) * TN * 4;
  const MigTables* __restrict__ T = a.T;
  const int n_models = T->n_models;
  const size_t ct2_bytes = ((size_t)n_models * 2 * cw * 4 + 15) / 16 * 16;
  uint32_t* const s_ct2 = reinterpret_cast<uint32_t*>(smem);
  uint32_t* const s_slot = reinterpret_cast<uint32_t*>(smem + ct2_bytes);
  int32_t* const s_row = reinterpret_cast<int32_t*>(smem + ct2_bytes + (size_t)a.gs * kSlotStride);
  int32_t* const s_qty = s_row + kRowCap;
  int32_t* const s_first = s_qty + kRowCap;
  const uint32_t ct2_saddr = (uint32_t)__cvta_generic_to_shared(s_ct2);

  // grid: x = work item (heaviest profile first), y = tile of perm[p] (heaviest nodes first), so CTAs are
  // handed out roughly longest-first
  const int p = A.work[3 * bx + 0];
  const int r0 = A.work[3 * bx + 1];
  const int len = A.work[3 * bx + 2];
  NSP_CHECK(p >= 0 && p < 16 && len >= 0 && len <= kRowCap && r0 >= 0);
  const int tid = threadIdx.x % TN;
  const int dsub = threadIdx.x / TN;
  const int pos = by * TN + tid;
  const int n = pos < a.npad ? A.perm[(size_t)p * a.npad + pos] : a.N;  // positions past the nodes: nothing to do
  const int cnt = (n < a.N) ? a.s_cnt[n] : 0;
  const int g0 = (n < a.N) ? a.gpu_off[n] : 0;

  for (int i = threadIdx.x; i < len; i += TN * DS) {
    const int d = A.rows[r0 + i];
    s_row[i] = d;
    // beyond 16 GPUs x 255 devices a quantity can never be covered: larger values behave the same
    s_qty[i] = min(max(__ldg(a.dem_qty + __ldg(a.dem_off + d)), 0), 32767);
    s_first[i] = kNoNode;
  }
  for (int i = threadIdx.x; i < n_models * 2 * cw; i += TN * DS) {
    const int w = i % (2 * cw), m = i / (2 * cw);
    const int c0 = T->geom_off[m], cc = T->geom_off[m + 1] - c0;
    uint32_t v = 0;
    for (int b = 0; b < 2; ++b) {
      const int c = w * 2 + b;
      if (c < cc) v |= byte_of(T->cand[c0 + c], p) << (16 * b);
    }
    s_ct2[i] = v;
  }
  // per (node, slot): the rank half-words of the candidates and profile p's counts, made when the snapshot
  // arrived (mig_rank16_kernel); the DS threads of a node split its GPUs
  for (int s = dsub; s < cnt; s += DS) {
    uint32_t* col = s_slot + (size_t)s * (2 * cw + 2) * TN + tid;
    const uint32_t* src = A.rankg + ((size_t)n * a.gs + s) * (2 * cw);
#pragma unroll
    for (int w = 0; w < 2 * cw; ++w) col[w * TN] = __ldg(src + w);
    const uint32_t x = __ldg(A.ufg + ((size_t)p * a.npad + n) * a.gs + s);  // free | (used + free) << 8
    col[2 * cw * TN] = (x & 0xffu) | (((0u - (x >> 8)) & 0xffffu) << 16);
    col[(2 * cw + 1) * TN] = ct2_saddr + (uint32_t)a.s_model[(size_t)s * a.npad + n] * (2 * cw * 4);
  }
  __syncthreads();

  uint32_t k512;  // opaque 2^9: keeps "provided << 9 | rank" an IMAD (FMA pipe) instead of ALU shift + add
  asm("mov.u32 %0, 512;" : "=r"(k512));
  const uint32_t ps0 = (uint32_t)__cvta_generic_to_shared(s_slot) + (uint32_t)tid * 4u;
  // RB rows per thread and pass over the node's GPUs: the rank words and the candidate row of a GPU are
  // loaded once and scored against all of them (each keeps its own residual; a row that is already covered
  // has needed <= 0, which makes every provided count 0 without a branch).
  for (int i = dsub; i < len; i += RB * DS) {
    int res[RB];
    uint32_t any[RB];
#pragma unroll
    for (int r = 0; r < RB; ++r) {
      res[r] = (i + r * DS < len) ? s_qty[i + r * DS] : 0;  // uniform over the warp
      any[r] = 0;
    }
    uint32_t ps = ps0;
    for (int s = 0; s < cnt; ++s, ps += kSlotStride) {
      // every request covered: every candidate of the remaining GPUs provides 0 (gpu.go:172-175)
      int left = 0;
#pragma unroll
      for (int r = 0; r < RB; ++r) left |= res[r];
      if (left == 0) break;
      const uint32_t x = lds_u32(ps + kUfOff);
      const int fj = (int)(x & 0xffu);
      int need[RB], most = 0;
      uint32_t bst[RB];
#pragma unroll
      for (int r = 0; r < RB; ++r) {
        need[r] = res[r] - fj;
        most = max(most, need[r]);
        bst[r] = 0;
      }
      // two candidates per instruction: key = provided << 9 | rank << 4 | (15 - index) per signed 16-bit
      // lane; a lane that cannot be applied carries 0x8000 and stays below the running maximum (>= 0)
      if (most > 0) {
        if constexpr (CNT) {
          const int m = a.s_model[(size_t)s * a.npad + n];
          const int cc = T->geom_off[m + 1] - T->geom_off[m];
#pragma unroll
          for (int r = 0; r < RB; ++r) executed += (i + r * DS < len) ? cc : 0;
        }
        const uint32_t mo = lds_u32(ps + kMoOff);
        const uint32_t nc2 = __byte_perm(x, 0u, 0x3232u);  // -(used + free) in both lanes
        uint32_t nd[RB];
#pragma unroll
        for (int r = 0; r < RB; ++r) nd[r] = (uint32_t)need[r] * 0x10001u;  // a negative value stays negative in both lanes
#pragma unroll
        for (int w = 0; w < cw; ++w) {
          const uint32_t ra = lds_u32(ps + (2 * w) * TN * 4);
          const uint32_t rb = lds_u32(ps + (2 * w + 1) * TN * 4);
          const uint2 c2 = lds_v2(mo + 8 * w);
#pragma unroll
          for (int r = 0; r < RB; ++r)
            bst[r] = __vimax3_s16x2(bst[r], __viaddmin_s16x2_relu(c2.x, nc2, nd[r]) * k512 + ra,
                                    __viaddmin_s16x2_relu(c2.y, nc2, nd[r]) * k512 + rb);
        }
      }
#pragma unroll
      for (int r = 0; r < RB; ++r) {
        const uint32_t best = max(bst[r] & 0xffffu, bst[r] >> 16);
        const uint32_t prov = best >> 9;  // 0 unless some applicable candidate provides something
        any[r] |= prov;
        if constexpr (OUT) {
          if (a.choice && prov != 0u) a.choice[(size_t)s_row[i + r * DS] * a.tg + g0 + s] = (int8_t)(15 - (int)(best & 15u));
        }
        res[r] = max(need[r], 0) - (int)prov;  // what is still missing after this GPU (see mig_score_kernel)
      }
    }
#pragma unroll
    for (int r = 0; r < RB; ++r) {
      const int ir = i + r * DS;
      if (ir < len) {
        if constexpr (OUT) {
          if (a.updated && n < a.N) a.updated[(size_t)s_row[ir] * a.N + n] = any[r] ? 1 : 0;
        }
        // the row's first node: a checked atomicMin in shared memory, one global atomic per row at the end
        if (any[r] && n < s_first[ir]) atomicMin(&s_first[ir], n);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < len; i += TN * DS) {
    const int v = s_first[i];
    if (v != kNoNode) push_row(A.push, a.first_node, s_row[i], v + a.node_base);
  }
  push_finish(A.push);
  if constexpr (CNT) {
    executed = __reduce_add_sync(0xffffffffu, (unsigned)executed);  // < 2^32 per warp: <= 32 lanes x rows x GPUs x 16
    if ((threadIdx.x & 31) == 0 && executed) atomicAdd(A.counter, executed);
  }
}

template <int TN, int DS, int CW, bool OUT, int RB, bool CNT = false>
__global__ void __launch_bounds__(TN * DS, RB > 2 ? 2 : 3) mig_score16_kernel(const MigScore16Args A) {
  mig_score16_body<TN, DS, CW, OUT, RB, CNT>(A, blockIdx.x, blockIdx.y);
}

// The winner's per-GPU choices (nosplan_score_mig's first_choice): one thread per row re-walks the row's
// first node with the snapshot-time rank table; rows with several keys use the general walker.
__global__ void mig_finalize16_kernel(const MigTables* __restrict__ T, const uint32_t* __restrict__ rankg,
                                      const uint16_t* __restrict__ ufg, const uint4* __restrict__ a_used,
                                      const uint4* __restrict__ a_free, const uint8_t* __restrict__ a_model,
                                      const int32_t* __restrict__ gpu_off, int D, int gs, int cw,
                                      const int32_t* __restrict__ dem_off, const int32_t* __restrict__ dem_key,
                                      const int32_t* __restrict__ dem_qty, const int32_t* __restrict__ first_node,
                                      int8_t* __restrict__ first_choice, int node_base, int N, int npad) {
  const int d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= D) return;
  // first_node holds indices over the whole (possibly sharded) list: a winner on another rank has no choices here
  const int fn = first_node[d] < 0 ? -1 : first_node[d] - node_base;
  int8_t ch[NOSPLAN_MAX_GPUS_PER_NODE];
  for (int s = 0; s < NOSPLAN_MAX_GPUS_PER_NODE; ++s) ch[s] = -1;
  if (fn >= 0 && fn < N) {
    const int e0 = dem_off[d], k = dem_off[d + 1] - e0;
    const int g0 = gpu_off[fn], cnt = gpu_off[fn + 1] - g0;
    if (k == 1) {
      const int p = dem_key[e0];
      int resid = min(max(dem_qty[e0], 0), 32767);
      for (int s = 0; s < cnt && resid > 0; ++s) {
        const uint32_t x = ufg[((size_t)p * npad + fn) * gs + s];
        const int needed = resid - (int)(x & 0xffu);
        uint32_t best2 = 0;
        if (needed > 0) {
          const int m = a_model[g0 + s];
          const int c0 = T->geom_off[m], cc = T->geom_off[m + 1] - c0;
          const uint32_t nd2 = (uint32_t)needed * 0x10001u;
          const uint32_t nc2 = ((0u - (x >> 8)) & 0xffffu) * 0x10001u;
          const uint32_t* rk = rankg + ((size_t)fn * gs + s) * (2 * cw);
          for (int w = 0; w < 2 * cw; ++w) {
            uint32_t c2 = 0;
            if (2 * w < cc) c2 |= byte_of(T->cand[c0 + 2 * w], p);
            if (2 * w + 1 < cc) c2 |= byte_of(T->cand[c0 + 2 * w + 1], p) << 16;
            best2 = __vimax3_s16x2(best2, __viaddmin_s16x2_relu(c2, nc2, nd2) * 512u + rk[w], 0u);
          }
        }
        const uint32_t best = max(best2 & 0xffffu, best2 >> 16);
        const uint32_t prov = best >> 9;
        if (prov) ch[s] = (int8_t)(15 - (int)(best & 15u));
        resid = max(needed, 0) - (int)prov;
      }
    } else {
      int pid[kMaxKeys], qty[kMaxKeys];
      for (int j = 0; j < kMaxKeys; ++j) {
        pid[j] = j < k ? dem_key[e0 + j] : 0;
        qty[j] = j < k ? dem_qty[e0 + j] : 0;
      }
      mig_walk_node(T, a_used, const_cast<uint4*>(a_free), a_model, g0, cnt, k, pid, qty, false, ch);
    }
  }
  for (int s = 0; s < gs; ++s) first_choice[(size_t)d * gs + s] = ch[s];
}

// ---- MPS scorer for the rows of one size (mode 1 of mps_score_kernel, first_node / updated only) ---------------
// Same closed form of slicing.GPU.UpdateGeometryFor as mps_score_kernel's single-size path; the size is
// uniform over the CTA, so everything that depends on it is hoisted, the rows are staged in shared
// memory and one pointer walks the GPU slots.
// Shared memory (u32): per slot (3 + nw) * TN words: TN pairs {A = free[sid] | (sid in FreeProfiles) << 8,
// B = spare | origtot << 16 (spare = cap - origtot, origtot = GB held by free slices)}, then M[TN] = key
// mask, then the nw x TN words of free bytes (only the restore walk reads M and those); row/qty/first [kRowCap].
__host__ __device__ inline size_t mps_score1_smem_bytes(int gs, int S, int tn) {
  const size_t nw = (size_t)(S + 3) / 4;
  return (size_t)gs * (3 + nw) * tn * 4 + (size_t)kRowCap * 12;
}

template <int TN, int DS, bool OUT, bool CNT>
__device__ __forceinline__ void mps_score1_body(const MpsScoreArgs& a, const int bx, const int by) {
  if (bx >= a.meta[0]) {
    push_finish(a.push);
    return;
  }
  unsigned int executed = 0;
  extern __shared__ __align__(16) unsigned char smem[];
  const int nw = (a.S + 3) >> 2;
  const uint32_t slot_stride = (uint32_t)(3 + nw) * TN * 4;  // bytes
  uint32_t* const s_slot = reinterpret_cast<uint32_t*>(smem);
  int32_t* const s_row = reinterpret_cast<int32_t*>(smem + (size_t)a.gs * slot_stride);
  int32_t* const s_qty = s_row + kRowCap;
  int32_t* const s_first = s_qty + kRowCap;
  const int sid = a.work[3 * bx];
  const int32_t* rl = a.rows + a.work[3 * bx + 1];
  const int len = a.work[3 * bx + 2];
  NSP_CHECK(sid >= 0 && sid < a.S && len >= 0 && len <= kRowCap);
  const int tid = threadIdx.x % TN;
  const int dsub = threadIdx.x / TN;
  const int pos = by * TN + tid;
  const int n = pos < a.N ? a.perm[(size_t)sid * a.N + pos] : a.N;
  const bool live = n < a.N && a.valid[n];
  const int g0 = n < a.N ? a.gpu_off[n] : 0;
  const int cnt = live ? a.gpu_off[n + 1] - g0 : 0;  // an invalid node is never updated
  for (int i = threadIdx.x; i < len; i += TN * DS) {
    const int d = rl[i];
    s_row[i] = d;
    s_qty[i] = max(__ldg(a.dem_qty + __ldg(a.dem_off + d)), 0);
    s_first[i] = kNoNode;
  }
  // (sizes come straight from global memory: a handful of L1-resident words, one barrier less per CTA)
  for (int s = dsub; s < cnt; s += DS) {  // the DS threads of a node split its GPUs
    const MpsGpu G = a.gp[g0 + s];
    int origtot = 0;
    for (uint32_t mk = G.fmask; mk; mk &= mk - 1) {
      const int key = __ffs(mk) - 1;
      origtot += __ldg(a.sizes + key) * (int)mps_byte(G.f0, G.f1, key);
    }
    uint32_t* reg = s_slot + (size_t)s * (3 + nw) * TN;
    uint32_t* col = reg + tid;
    reg[2 * tid] = mps_byte(G.f0, G.f1, sid) | (((G.fmask >> sid) & 1u) << 8);      // A and B side by side:
    reg[2 * tid + 1] = ((uint32_t)origtot << 16) | (uint32_t)((int)G.cap - origtot);  // one 64-bit load per slot
    col[2 * TN] = G.fmask;
    const uint32_t w[8] = {G.f0.x, G.f0.y, G.f0.z, G.f0.w, G.f1.x, G.f1.y, G.f1.z, G.f1.w};
    for (int i = 0; i < nw; ++i) col[(3 + i) * TN] = w[i];
  }
  __syncthreads();

  const int m = __ldg(a.sizes + sid);
  const uint32_t inv = m == 1 ? 0u : 0xFFFFFFFFu / (uint32_t)m + 1u;  // exact multiply-high division (div_small)
  const uint32_t ps0 = (uint32_t)__cvta_generic_to_shared(s_slot);
  const uint32_t t4 = (uint32_t)tid * 4u;
  for (int i = dsub; i < len; i += DS) {
    int resid = s_qty[i];
    bool any = false;
    uint32_t ps = ps0;
    for (int s = 0; s < cnt && resid > 0; ++s, ps += slot_stride) {
      const uint2 ab = lds_v2(ps + 2 * t4);
      const uint32_t A = ab.x;
      const int fj = (int)(A & 0xffu);
      const int miss = resid - fj;  // getMissingSlices
      if (miss <= 0) break;         // this GPU already provides it; the residual drops to 0
      if constexpr (CNT) ++executed;
      const uint32_t sp = ab.y;
      const int otot = (int)(sp >> 16);
      int spare = (int)(sp & 0xffffu);
      const bool in_orig = (A >> 8) != 0u;
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
      // (d) restore the original keys, ascending size, each all-or-nothing.  Only whether key sid itself
      // comes back matters here, and only if it held something: walk the keys up to sid in that case.
      bool restored_sid = in_orig;
      if (spare < otot && fj > 0) {
        restored_sid = false;
        for (uint32_t mk = lds_u32(ps + 2 * TN * 4 + t4) & ((2u << sid) - 1u); mk; mk &= mk - 1) {
          const int key = __ffs(mk) - 1;
          const int need = __ldg(a.sizes + key) * (int)((lds_u32(ps + (3 + (key >> 2)) * TN * 4 + t4) >> ((key & 3) * 8)) & 0xffu);
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
    if constexpr (OUT) {
      if (a.updated && n < a.N) a.updated[(size_t)s_row[i] * a.N + n] = any ? 1 : 0;
    }
    // the row's first node: a checked atomicMin in shared memory, one global atomic per row at the end
    if (any && n < s_first[i]) atomicMin(&s_first[i], n);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < len; i += TN * DS) {
    const int v = s_first[i];
    if (v != kNoNode) push_row(a.push, a.first_node, s_row[i], v + a.node_base);
  }
  push_finish(a.push);
  if constexpr (CNT) {
    executed = __reduce_add_sync(0xffffffffu, executed);
    if ((threadIdx.x & 31) == 0 && executed) atomicAdd(a.counter, (unsigned long long)executed);
  }
}

template <int TN, int DS, bool OUT, bool CNT = false>
__global__ void __launch_bounds__(TN * DS, 4) mps_score1_kernel(const MpsScoreArgs a) {
  mps_score1_body<TN, DS, OUT, CNT>(a, blockIdx.x, blockIdx.y);
}

// Both kinds of a scoring pass in ONE launch (nosplan_score_pass_dev): the first n_mig CTAs are the MIG scorer's
// (work item, tile) grid in its own order (x fastest: heaviest profile first, heaviest nodes first), the rest the MPS
// scorer's.  One pool of CTAs drains over the SMs, so a small shard fills the GPU without two streams and the
// events that fork and join them.  TNM: nodes per MIG CTA; the MPS part keeps 128 nodes x 4 threads.
template <int TNM, int CW, bool CNT>
__global__ void __launch_bounds__(512, 3) score_both_kernel(const MigScore16Args A, const MpsScoreArgs a, int mig_x, int n_mig, int mps_x,
                                                            int n_mps_first) {
  // n_mps_first: how many of the MPS scorer's CTAs are handed out before the MIG scorer's (measurement knob
  // NOSPLAN_BOTH_ORDER: 0 = MIG first, the default; 1 = MPS first)
  int i = blockIdx.x;
  if (i < n_mps_first) {
    mps_score1_body<128, 4, false, CNT>(a, i % mps_x, i / mps_x);
    return;
  }
  i -= n_mps_first;
  if (i < n_mig) mig_score16_body<TNM, 512 / TNM, CW, false, 2, CNT>(A, i % mig_x, i / mig_x);
  else mps_score1_body<128, 4, false, CNT>(a, (i - n_mig + n_mps_first) % mps_x, (i - n_mig + n_mps_first) / mps_x);
}

}  // namespace nsp
