// nosplan_stream.cuh — the MIG scorer for the no-reuse regime: few demand rows against a node list far larger than
// the caches, where every byte of state is needed exactly once per launch and HBM bandwidth is the roof.
//
// The ordered scorer (nosplan_score16.cuh) is built around re-use: a CTA stages a tile of 128 nodes in shared
// memory behind a barrier, gathers the nodes through the per-profile order perm[p][.], and splits the rows of a
// work item over four threads per node.  With a handful of rows that is all overhead: three of a node's four threads
// have no row, the gather touches partial sectors (1.72x the ideal DRAM bytes) and every CTA waits at a barrier for
// 128 nodes' worth of loads (profiles/r1_stream_sweep.md: 20 % of HBM peak).  Here:
//   * one thread per node, nodes in STORAGE order, up to RS rows of one profile held in registers;
//   * a warp owns 32 consecutive nodes: their rank tables (snapshot-time, mig_rank16_kernel: gs * 2cw words per
//     node, 192 B for 8 GPUs x 12 geometries) are one contiguous block, fetched with fully coalesced 8-byte loads and
//     transposed through a warp-private slice of shared memory (odd 64-bit stride: conflict-free), so there is no
//     CTA-wide barrier anywhere, only __syncwarp;
//   * the profile's free|cur column (16 B per node), the models (8 B) and the GPU count go straight into
//     registers;
//   * the candidate arithmetic is the 16-bit DPX core of mig_score16_kernel, unchanged (same keys, same winners);
//   * a warp folds its winners with one checked atomicMin per row (the check keeps a million warps from queueing
//     on one address once an early node has won).
// Per node it reads gs * 2cw * 4 (ranks) + gs * 2 (the profile's column; ufg is profile-major so that 32 nodes' columns
// are one contiguous 512-byte block) + gs (models) + 1 bytes = 217 B for 8 GPUs x 12 geometries, all of it contiguous per warp.
#pragma once
#include "nosplan_score16.cuh"

namespace nsp {

constexpr int kStreamRows = 4;    // rows of one profile a thread scores per pass over its node
constexpr int kStreamWarps = 4;   // warps per CTA (each works alone)
constexpr int kStreamMaxGpus = 8; // GPU slots per node this kernel takes (wider snapshots stay with the ordered scorer)

struct MigStreamArgs {
  const MigTables* T;
  const uint32_t* rankg;     // [npad][gs][2cw]
  const uint16_t* ufg;       // [16][npad][gs]
  const uint8_t* modelg;     // [npad][gs] model of every slot (node-major copy made with the rank table)
  const uint8_t* s_cnt;      // [npad] GPUs of the node (0: invalid node or padding)
  int N, npad, gs;
  const int32_t* dem_off;
  const int32_t* dem_qty;
  const int32_t* rows;       // rows grouped by profile (rows_group_kernel, work items of <= kStreamRows rows)
  const int32_t* work;
  const int32_t* meta;
  int32_t* first_node;       // unsigned atomicMin target, at rest at -1
  int node_base;
};

__host__ __device__ inline size_t mig_stream_smem_bytes(int n_models, int gs, int cw) {
  const size_t w2 = (size_t)gs * cw;  // 64-bit words of rank table per node
  return (size_t)kStreamWarps * (32 * (w2 + 1) * 8 + (size_t)n_models * 2 * cw * 4 + 8);
}

__device__ __forceinline__ void cp_async8(uint32_t smem_addr, const void* gptr) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_addr), "l"(gptr) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// Persistent: the grid is a fixed number of CTAs; every warp walks the (work item, block of 32 nodes) pairs with a
// stride of the grid's warp count, so the number of work items (only known on the device) costs no empty CTAs.
// RS: rows of a work item (1, 2 or kStreamRows: a call with one row should not pay for four); GS: GPU slots per node as a
// compile-time constant (8, the usual node: the copy loop's index arithmetic folds away) or 0 = read it from the arguments
template <int CW, int RS, int GS>
__global__ void __launch_bounds__(kStreamWarps * 32) mig_stream_kernel(const MigStreamArgs A) {
  extern __shared__ __align__(16) unsigned char smem[];
  constexpr int cw = CW;
  constexpr int kStreamRows = RS;  // (shadows the namespace constant: the arrays below are sized for this instantiation)
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int gs = GS ? GS : A.gs;
  const int w2 = gs * cw;         // uint2 of rank table per node
  const int stride2 = w2 + 1;     // odd for the usual 8 x 3: 64-bit accesses of consecutive lanes hit distinct banks
  const MigTables* __restrict__ T = A.T;
  const int n_models = T->n_models;
  const size_t per_warp = (size_t)32 * (w2 + 1) * 8 + (size_t)n_models * 2 * cw * 4 + 8;
  uint2* const s_rank = reinterpret_cast<uint2*>(smem + per_warp * warp);
  uint32_t* const s_ct2 = reinterpret_cast<uint32_t*>(smem + per_warp * warp + (size_t)32 * (w2 + 1) * 8);
  const uint32_t s_rank_addr = (uint32_t)__cvta_generic_to_shared(s_rank);

  const int n_items = A.meta[0];
  const int n_blocks = A.npad >> 5;  // npad is a multiple of 32
  const long long total = (long long)n_items * n_blocks;
  const int warps = gridDim.x * kStreamWarps;
  uint32_t k512;
  asm("mov.u32 %0, 512;" : "=r"(k512));
  int cur_item = -1, p = 0, len = 0;
  int res0[kStreamRows], row[kStreamRows];
  for (long long q = (long long)blockIdx.x * kStreamWarps + warp; q < total; q += warps) {
    const int item = (int)(q / n_blocks);
    const int base = (int)(q - (long long)item * n_blocks) << 5;
    // 32 nodes' rank tables: one contiguous block, copied asynchronously (no registers in between), every node's row
    // to its padded place; the loads of the node's own column, models and count overlap it
    {
      const uint2* src = reinterpret_cast<const uint2*>(A.rankg) + (size_t)base * w2;
      const int tot = 32 * w2;
      if constexpr (GS != 0) {
#pragma unroll
        for (int k = 0; k < GS * CW; ++k) {  // 32 * w2 / 32 copies per lane, every index a constant plus the lane
          const int i = lane + 32 * k;
          const int node = i / (GS * CW), j = i - node * (GS * CW);
          cp_async8(s_rank_addr + (uint32_t)(node * (GS * CW + 1) + j) * 8u, src + i);
        }
      } else {
        for (int i = lane; i < tot; i += 32) {
          const int node = i / w2, j = i - node * w2;
          cp_async8(s_rank_addr + (uint32_t)(node * stride2 + j) * 8u, src + i);
        }
      }
    }
    if (item != cur_item) {  // the item's profile and rows (uniform over the warp)
      cur_item = item;
      p = A.work[3 * item + 0];
      const int r0 = A.work[3 * item + 1];
      len = A.work[3 * item + 2];  // <= kStreamRows
#pragma unroll
      for (int r = 0; r < kStreamRows; ++r) {
        row[r] = r < len ? A.rows[r0 + r] : -1;
        res0[r] = r < len ? min(max(__ldg(A.dem_qty + __ldg(A.dem_off + row[r])), 0), 32767) : 0;
      }
      __syncwarp();
      // the profile's quantity in every candidate of every model, one candidate per 16-bit lane (as in mig_score16_kernel)
      for (int i = lane; i < n_models * 2 * cw; i += 32) {
        const int w = i % (2 * cw), m = i / (2 * cw);
        const int c0 = T->geom_off[m], cc = T->geom_off[m + 1] - c0;
        uint32_t v = 0;
        for (int b = 0; b < 2; ++b) {
          const int c = w * 2 + b;
          if (c < cc) v |= byte_of(T->cand[c0 + c], p) << (16 * b);
        }
        s_ct2[i] = v;
      }
    }
    const int n = base + lane;
    NSP_CHECK(n < A.npad && p >= 0 && p < 16 && len >= 1 && len <= RS);
    const int cnt = A.s_cnt[n];
    NSP_CHECK(cnt <= gs && gs <= kStreamMaxGpus);
    // this node's column of the profile (gs x u16) and its models (gs x u8), at most kStreamMaxGpus = 8 slots:
    // two 64-bit registers and one, indexed with shifts (a register array indexed by the slot would go to local memory)
    unsigned long long uf0 = 0, uf1 = 0, md = 0;
    {
      const uint16_t* col = A.ufg + ((size_t)p * A.npad + n) * gs;
      const uint8_t* mo = A.modelg + (size_t)n * gs;
      if (gs == 8) {  // the usual node: one 16-byte and one 8-byte load
        const uint4 c = __ldg(reinterpret_cast<const uint4*>(col));
        uf0 = (unsigned long long)c.x | ((unsigned long long)c.y << 32);
        uf1 = (unsigned long long)c.z | ((unsigned long long)c.w << 32);
        const uint2 m = __ldg(reinterpret_cast<const uint2*>(mo));
        md = (unsigned long long)m.x | ((unsigned long long)m.y << 32);
      } else {
#pragma unroll
        for (int i = 0; i < kStreamMaxGpus; ++i)
          if (i < gs) {
            const unsigned long long v = __ldg(col + i);
            if (i < 4) uf0 |= v << (16 * i);
            else uf1 |= v << (16 * (i - 4));
            md |= (unsigned long long)__ldg(mo + i) << (8 * i);
          }
      }
    }
    int res[kStreamRows];
    uint32_t any[kStreamRows];
#pragma unroll
    for (int r = 0; r < kStreamRows; ++r) {
      res[r] = res0[r];
      any[r] = 0;
    }
    cp_async_wait_all();
    __syncwarp();

    const uint2* mine = s_rank + lane * stride2;
#pragma unroll 1
    for (int s = 0; s < cnt; ++s) {
      int left = 0;
#pragma unroll
      for (int r = 0; r < kStreamRows; ++r) left |= res[r];
      if (left == 0) break;  // every request covered: the remaining GPUs' candidates all provide 0 (gpu.go:172-175)
      const uint32_t x = (uint32_t)((s < 4 ? uf0 : uf1) >> (16 * (s & 3))) & 0xffffu;  // free | (used + free) << 8
      const int fj = (int)(x & 0xffu);
      int need[kStreamRows], most = 0;
      uint32_t bst[kStreamRows];
#pragma unroll
      for (int r = 0; r < kStreamRows; ++r) {
        need[r] = res[r] - fj;
        most = max(most, need[r]);
        bst[r] = 0;
      }
      if (most > 0) {
        const uint32_t m = (uint32_t)(md >> (8 * s)) & 0xffu;
        const uint32_t* ct = s_ct2 + m * (2 * cw);
        const uint32_t nc2 = ((0u - (x >> 8)) & 0xffffu) * 0x10001u;  // -(used + free) in both lanes
        uint32_t nd[kStreamRows];
#pragma unroll
        for (int r = 0; r < kStreamRows; ++r) nd[r] = (uint32_t)need[r] * 0x10001u;
#pragma unroll
        for (int w = 0; w < cw; ++w) {
          const uint2 rk = mine[s * cw + w];
          const uint32_t ca = ct[2 * w], cb = ct[2 * w + 1];
#pragma unroll
          for (int r = 0; r < kStreamRows; ++r)
            bst[r] = __vimax3_s16x2(bst[r], __viaddmin_s16x2_relu(ca, nc2, nd[r]) * k512 + rk.x,
                                    __viaddmin_s16x2_relu(cb, nc2, nd[r]) * k512 + rk.y);
        }
      }
#pragma unroll
      for (int r = 0; r < kStreamRows; ++r) {
        const uint32_t best = max(bst[r] & 0xffffu, bst[r] >> 16);
        const uint32_t prov = best >> 9;
        any[r] |= prov;
        res[r] = max(need[r], 0) - (int)prov;
      }
    }
#pragma unroll
    for (int r = 0; r < kStreamRows; ++r) {
      if (r >= len) break;
      const unsigned mine_n = (any[r] && n < A.N) ? (unsigned)(n + A.node_base) : 0xffffffffu;
      const unsigned wmin = __reduce_min_sync(0xffffffffu, mine_n);
      if (lane == 0 && wmin != 0xffffffffu) {
        unsigned* dst = reinterpret_cast<unsigned*>(A.first_node) + row[r];
        if (wmin < __ldcg(dst)) atomicMin(dst, wmin);
      }
    }
    __syncwarp();  // the block's rows in shared memory are free for the next copy
  }
}

}  // namespace nsp
