// nosplan.cu — host side of libnosplan.so: handle, validation, uploads, launches, the C ABI of
// include/nosplan.h.  No CPU fallback exists anywhere in this file: every compute entry point ends
// in a kernel launch on the handle's device or returns an error.
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <chrono>
#include <cstring>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <map>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include "nosplan.h"
#include "nosplan_kernels.cuh"
#include "nosplan_plan.cuh"
#include "nosplan_seq.cuh"
#include "nosplan_score16.cuh"
#include "nosplan_stream.cuh"
#include "nosplan_xchg.cuh"

using namespace nsp;

namespace {

// NVTX range around a phase of an ABI call (shows up in Nsight Systems / ncu --nvtx; costs nothing without a tool)
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};
#define NSP_RANGE(name) NvtxRange nsp_range_##__LINE__(name)

// grow-only device buffer
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t reserve(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    size_t want = std::max<size_t>(bytes, 256);
    cudaError_t e = cudaMalloc(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
  template <typename T>
  T* as() const { return reinterpret_cast<T*>(p); }
};

}  // namespace

struct nosplan {
  int device = 0;
  cudaStream_t stream = nullptr, stream2 = nullptr;
  std::string err;
  int sm_count = 148;

  // MIG tables
  bool have_tables = false;
  int P = 0;
  std::vector<std::string> names;
  MigTables h_tables;
  DevBuf d_tables;
  int cmax = 0;
  bool k16ok = false;  // <= 16 geometries per model, every quantity <= 63 (16-bit score keys)

  // MIG snapshot
  bool have_mig = false;
  int mig_n = 0, mig_tg = 0, mig_gs = 0, mig_npad = 0;
  int64_t mig_evals = 0;
  DevBuf d_mig_off, d_mig_valid, d_mig_model, d_mig_used0, d_mig_free0, d_mig_used, d_mig_free;
  DevBuf d_mig_sused, d_mig_sfree, d_mig_smodel, d_mig_scnt, d_mig_gnode, d_mig_gslot;
  bool mig_state_dirty = false;  // working copy differs from the snapshot
  DevBuf d_mig_cost, d_mig_perm;  // per-profile node order of the 16-bit-key scorer (nosplan_score16.cuh)
  DevBuf d_mig_rankg, d_mig_ufg, d_mig_maskg, d_mig_modelg;  // its snapshot-time tables
  DevBuf d_profcost;
  struct RowPlan {  // per scoring call and kind: demand rows grouped by key, work items (rows_*_kernel)
    DevBuf rows, work, meta, cnt;
  } rp[2];
  DevBuf d_mps_cost, d_mps_perm, d_mps_sizecost;  // per-size node order of the MPS scorer

  // MPS snapshot
  bool have_mps = false;
  int S = 0;
  int mps_n = 0, mps_tg = 0, mps_gs = 0;
  int64_t mps_evals = 0;
  DevBuf d_mps_sizes, d_mps_off, d_mps_valid, d_mps_gpu0, d_mps_ulo0, d_mps_uhi0, d_mps_gpu, d_mps_ulo, d_mps_uhi, d_mps_gnode, d_mps_gslot;
  bool mps_state_dirty = false;

  // scratch / staging
  DevBuf d_in[6];    // staging for int32 host inputs (MIG snapshot: 0-2; read-back: 3-5)
  DevBuf d_inp[4];   // staging of the MPS snapshot (its own, so both snapshots can be in flight)
  DevBuf d_dem[3];   // demand CSR
  DevBuf d_out[6];   // outputs
  DevBuf d_plan[32];
  DevBuf d_rowok[2];                       // per kind: rows_check_kernel's verdict per row (scorers in input order)
  DevBuf d_status, d_rowstatus;            // d_rowstatus: one word per kind: grouping call number << 32 | ~(lowest bad row code)
  uint32_t row_call[2] = {0, 0};           // number of the kind's last grouping call
  unsigned long long* h_status = nullptr;  // pinned: [0..1] MIG pack status, [2..3] MPS pack status, [4..5] row status, [6] exchange
  cudaEvent_t status_ev = nullptr, status_ev2 = nullptr;
  int32_t* h_limit = nullptr;              // pinned [2]: per kind, the first pod the chains could not take (nosplan_plan)
  int32_t (*h_range)[2] = nullptr;         // pinned [2][2]: per kind, the range handed to the next chain launch
  void* h_pin = nullptr;                   // pinned staging for results (nosplan_score_snapshot, nosplan_plan)
  size_t h_pin_cap = 0;
  // nosplan_plan's split of the pod list by kind, kept between calls: 50 000 pods are 0.8 MB of host arrays, and fresh
  // vectors cost more in page faults than the split itself
  struct PlanSide {
    std::vector<int32_t> pod_idx, pod_sig, off, key, qty;
    std::map<std::vector<int32_t>, int> sigs;           // general case
    std::vector<int> single;                             // [key * (kChainQ+1) + qty] -> row, single-key fast lookup
    int n_sigs = 0;
    bool chains_ok = true;                               // every pod: one key, 1 <= qty <= kChainQ
  } plan_side[2];

  // nosplan_score_pass_dev: fork/join of the two kinds, CUDA events around the dominant kernels, executed-work counters
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  static constexpr int kKevPool = 64;
  bool kernel_timing = false;
  cudaEvent_t kev[2][kKevPool][2] = {};
  int kev_n[2] = {0, 0};
  DevBuf d_counter;  // [2] evaluations executed by the CNT instantiations (MIG, MPS)
  DevBuf d_passbuf;  // local winners when the pass cannot push from the scorers
  int occ_cache[2][2][8] = {};  // [kind][out][cw] resident CTAs per SM of the ordered scorers (0 = not asked yet)

  // nosplan_score_snapshot: two private lanes (child handles), the MPS one driven by a worker thread
  std::vector<int32_t> tab_geom_off, tab_geom_counts;  // what nosplan_set_mig_tables was given (for the MIG lane)
  int tab_models = 0;
  uint64_t tab_version = 0;
  struct Lanes* lanes = nullptr;

  // sharded scoring (nosplan_set_node_base, nosplan_exchange_*)
  int mig_base = 0, mps_base = 0;
  struct {
    bool created = false, open = false;
    bool broken = false;               // a status call saw a time-out: the step counters of the ranks no longer agree
    int rank = 0, world = 1, max_keys = 0;
    int stride = 0;                    // max_keys rounded up to a multiple of 4: both halves stay 16-byte aligned
    void* local = nullptr;             // keys[2][stride] + pushed[kMaxWorld] + read[kMaxWorld]
    size_t flags_off = 0;
    void* mapped[kMaxWorld] = {};      // peers' buffers as mapped here (own slot = local)
    XchgPeers peers;
    uint32_t step = 0;
    DevBuf done, err;
  } xch;
};

void lanes_release(nosplan_t* h);  // defined with nosplan_score_snapshot

namespace {

int fail(nosplan_t* h, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (h) h->err = buf;
  return code;
}

#define CU(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e__ = (call);                                                                      \
    if (e__ != cudaSuccess)                                                                        \
      return fail(h, NOSPLAN_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
  } while (0)

int bind(nosplan_t* h) {
  CU(cudaSetDevice(h->device));
  return NOSPLAN_OK;
}

// MIG profile name must match \\d+g\\.\\d+gb somewhere (unanchored regexp.MatchString, profile.go:51-53)
bool mig_name_valid(const std::string& s) {
  const size_t n = s.size();
  for (size_t i = 0; i < n; ++i) {
    size_t j = i;
    while (j < n && isdigit((unsigned char)s[j])) ++j;
    if (j == i) continue;                                  // \d+
    if (j + 1 >= n || s[j] != 'g' || s[j + 1] != '.') continue;  // g\.
    size_t k = j + 2;
    while (k < n && isdigit((unsigned char)s[k])) ++k;
    if (k == j + 2) continue;                              // \d+
    if (k + 1 < n && s[k] == 'g' && s[k + 1] == 'b') return true;  // gb
  }
  return false;
}

int upload(nosplan_t* h, DevBuf& b, const void* src, size_t bytes, cudaStream_t st) {
  CU(b.reserve(bytes));
  if (bytes && src) CU(cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, st));
  return NOSPLAN_OK;
}

// status words of the pack kernels: [0] lowest failing code (~0 = clean), [1] evaluations per demand row
int status_reset(nosplan_t* h, int kind, cudaStream_t st) {
  CU(h->d_status.reserve(32));
  char* base = static_cast<char*>(h->d_status.p) + 16 * kind;
  CU(cudaMemsetAsync(base, 0xff, 8, st));
  CU(cudaMemsetAsync(base + 8, 0, 8, st));
  return NOSPLAN_OK;
}
unsigned long long* status_words(nosplan_t* h, int kind) { return h->d_status.as<unsigned long long>() + 2 * kind; }
// the read-back is enqueued right behind the pack kernel; work enqueued after it (tables that only the
// scorers need) is not waited for: the stream orders it before any later call
int status_enqueue(nosplan_t* h, int kind, cudaStream_t st) {
  CU(cudaMemcpyAsync(h->h_status + 2 * kind, status_words(h, kind), 16, cudaMemcpyDeviceToHost, st));
  CU(cudaEventRecord(kind ? h->status_ev2 : h->status_ev, st));
  return NOSPLAN_OK;
}
int status_wait(nosplan_t* h, int kind, unsigned long long out[2]) {
  CU(cudaEventSynchronize(kind ? h->status_ev2 : h->status_ev));
  out[0] = h->h_status[2 * kind];
  out[1] = h->h_status[2 * kind + 1];
  return NOSPLAN_OK;
}

int check_demands(nosplan_t* h, int D, const int32_t* off, const int32_t* key, const int32_t* qty, int n_keys) {
  if (D < 0) return fail(h, NOSPLAN_ERR_ARG, "negative demand count");
  if (D == 0) return NOSPLAN_OK;
  if (!off) return fail(h, NOSPLAN_ERR_ARG, "dem_off is null");
  if (off[0] != 0) return fail(h, NOSPLAN_ERR_ARG, "dem_off[0] must be 0");
  for (int d = 0; d < D; ++d) {
    int k = off[d + 1] - off[d];
    if (k < 0) return fail(h, NOSPLAN_ERR_ARG, "dem_off not ascending at row %d", d);
    if (k > NOSPLAN_MAX_DEMAND_KEYS)
      return fail(h, NOSPLAN_ERR_RANGE, "demand row %d has %d keys (max %d)", d, k, NOSPLAN_MAX_DEMAND_KEYS);
    for (int e = off[d]; e < off[d + 1]; ++e) {
      if (key[e] < 0 || key[e] >= n_keys) return fail(h, NOSPLAN_ERR_RANGE, "demand row %d: key %d out of range", d, key[e]);
      if (e > off[d] && key[e] <= key[e - 1]) return fail(h, NOSPLAN_ERR_ARG, "demand row %d: keys must be strictly ascending", d);
      if (qty[e] < 0 || qty[e] > NOSPLAN_MAX_DEMAND_QTY)
        return fail(h, NOSPLAN_ERR_RANGE, "demand row %d: quantity %d out of range", d, qty[e]);
    }
  }
  return NOSPLAN_OK;
}

// chunking of the demand axis so the grid is a small multiple of the SM count
void pick_chunks(int sm_count, int tiles, int D, int min_chunk, int* d_chunk, int* gy, int ctas_per_sm = 4) {
  int want_ctas = sm_count * ctas_per_sm;
  int y = std::max(1, want_ctas / std::max(1, tiles));
  int chunk = std::max(min_chunk, (D + y - 1) / y);
  chunk = std::min(chunk, std::max(D, 1));
  *d_chunk = chunk;
  *gy = (D + chunk - 1) / chunk;
  if (*gy < 1) *gy = 1;
}

bool wide_keys_forced() {  // measurement knob: NOSPLAN_KEY_BITS=32 forces the wide-key scorer
  static const bool f = [] {
    const char* e = std::getenv("NOSPLAN_KEY_BITS");
    return e && std::atoi(e) == 32;
  }();
  return f;
}

bool trace_on() {  // NOSPLAN_TRACE=1: host-side phase times of nosplan_plan on stderr
  static const bool f = [] {
    const char* e = std::getenv("NOSPLAN_TRACE");
    return e && std::atoi(e) == 1;
  }();
  return f;
}

bool plain_order_forced() {  // measurement knob: NOSPLAN_PLAIN_ORDER=1 scores rows and nodes in input order
  static const bool f = [] {
    const char* e = std::getenv("NOSPLAN_PLAIN_ORDER");
    return e && std::atoi(e) == 1;
  }();
  return f;
}

int min_chunk() {  // measurement knob: fewest rows per work item of 128 nodes
  static const int w = [] {
    const char* e = std::getenv("NOSPLAN_MIN_CHUNK");
    const int v = e ? std::atoi(e) : 64;
    return v >= 8 && v <= 512 ? v : 64;
  }();
  return w;
}

// Waves of CTAs a scorer launch is split into.  Measured (profiles/r2_bench_cfg5.md, "work-item size"): a kind scored on
// its own wants 4; the launch that carries both kinds wants 2 — its MIG and MPS CTAs fill each other's tails, so the
// larger work items win (fewer tile loads).  NOSPLAN_SCORE_WAVES=n overrides both.
int score_waves(bool both_kinds = false) {
  static const int w = [] {
    const char* e = std::getenv("NOSPLAN_SCORE_WAVES");
    const int v = e ? std::atoi(e) : 0;
    return v >= 1 && v <= 16 ? v : 0;
  }();
  return w ? w : both_kinds ? 2 : 4;
}

int mps_min_chunk() {  // measurement knob: fewest rows per MPS work item (128 nodes) in a scoring pass
  static const int w = [] {
    const char* e = std::getenv("NOSPLAN_MPS_MIN_CHUNK");
    const int v = e ? std::atoi(e) : 128;
    return v >= 8 && v <= 512 ? v : 128;
  }();
  return w;
}

#ifndef NOSPLAN_RB
#define NOSPLAN_RB 2
#endif
constexpr unsigned long long kXchgTimeoutNs = 2000000000ull;  // a rank that waits this long for a peer gives up
constexpr int kMig16Tile = 64;  // default nodes per CTA of the ordered MIG scorer (measured: profiles/r2_sweeps.md)
constexpr int kScoreTN = 128;
constexpr int kScoreDS = 4;  // threads per node in the MIG scorer

size_t mig_score_smem(const nosplan_t* h, int tn) {
  return mig_score_smem_bytes(h->h_tables.n_models, h->mig_gs, h->cmax, tn);
}

// picks the instantiation for the table's candidate count (CW words of 4 candidates; 0 = any) and key width
template <typename Kern>
int launch_mig_score_grid(nosplan_t* h, Kern kern, const MigScoreArgs& a, int tiles, size_t smem, int chunk_cap, cudaStream_t st) {
  CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // the demand axis is cut so that the grid is a whole number of waves of what really fits on an SM
  // (registers and shared memory decide); two waves: the rows' costs differ, smaller CTAs even them out
  int occ = 0;
  CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kScoreTN * kScoreDS, smem));
  MigScoreArgs b = a;
  int gy = 1;
  pick_chunks(h->sm_count, tiles, a.D, 16, &b.d_chunk, &gy, std::max(occ, 1) * score_waves());
  if (chunk_cap > 0 && b.d_chunk > chunk_cap) {
    b.d_chunk = chunk_cap;
    gy = (a.D + chunk_cap - 1) / chunk_cap;
  }
  kern<<<dim3(tiles, gy), kScoreTN * kScoreDS, smem, st>>>(b);
  return NOSPLAN_OK;
}
int launch_mig_score_kernel(nosplan_t* h, const MigScoreArgs& a, dim3 grid, size_t smem, cudaStream_t st) {
  const int cw = (a.cmax + 3) / 4;
  const int tiles = (int)grid.x;
  if (a.k16 && cw <= 4 && !wide_keys_forced()) {
    // <= 16 geometries per model and quantities <= 63: 16-bit keys, two candidates per DPX instruction
    switch (cw) {
      case 1: return launch_mig_score_grid(h, mig_score_kernel<kScoreTN, kScoreDS, 1, true>, a, tiles, smem, 0, st);
      case 2: return launch_mig_score_grid(h, mig_score_kernel<kScoreTN, kScoreDS, 2, true>, a, tiles, smem, 0, st);
      case 3: return launch_mig_score_grid(h, mig_score_kernel<kScoreTN, kScoreDS, 3, true>, a, tiles, smem, 0, st);
      default: return launch_mig_score_grid(h, mig_score_kernel<kScoreTN, kScoreDS, 4, true>, a, tiles, smem, 0, st);
    }
  }
  switch (cw) {
    case 1: return launch_mig_score_grid(h, mig_score_kernel<kScoreTN, kScoreDS, 1, false>, a, tiles, smem, 0, st);
    case 2: return launch_mig_score_grid(h, mig_score_kernel<kScoreTN, kScoreDS, 2, false>, a, tiles, smem, 0, st);
    case 3: return launch_mig_score_grid(h, mig_score_kernel<kScoreTN, kScoreDS, 3, false>, a, tiles, smem, 0, st);
    case 4: return launch_mig_score_grid(h, mig_score_kernel<kScoreTN, kScoreDS, 4, false>, a, tiles, smem, 0, st);
    default: return launch_mig_score_grid(h, mig_score_kernel<kScoreTN, kScoreDS, 0, false>, a, tiles, smem, 0, st);
  }
}

// ---- rows by key (rows_group_kernel / rows_group2_kernel) -----------------------------------------------------------
// Scratch of one kind's row grouping for D rows cut into items of `chunk` rows; max_work = the grid extent that
// covers any split of the rows.  Allocates only when a buffer has to grow.
int reserve_rows(nosplan_t* h, int kind, int D, int n_keys, int chunk, cudaStream_t st, int* max_work) {
  auto& rp = h->rp[kind];
  *max_work = (D + chunk - 1) / chunk + n_keys;  // every key may end with a short item
  CU(rp.rows.reserve(sizeof(int32_t) * (size_t)D * (n_keys + 1)));
  CU(rp.work.reserve(sizeof(int32_t) * 3 * (size_t)*max_work));
  CU(rp.meta.reserve(sizeof(int32_t) * 4));
  if (!rp.cnt.p) {  // cursors + ticket; the kernel leaves them at zero for the next call
    CU(rp.cnt.reserve(sizeof(int32_t) * (kRowBuckets + 1)));
    CU(cudaMemsetAsync(rp.cnt.p, 0, sizeof(int32_t) * (kRowBuckets + 1), st));
  }
  if (!h->d_rowstatus.p) {
    CU(h->d_rowstatus.reserve(16));
    CU(cudaMemsetAsync(h->d_rowstatus.p, 0, 16, st));
  }
  return NOSPLAN_OK;
}
// the row code a status word holds for the kind's last grouping call, ~0 = every row was fine
unsigned long long row_status_code(const nosplan_t* h, int kind, unsigned long long word) {
  if (h->row_call[kind] == 0 || (uint32_t)(word >> 32) != h->row_call[kind]) return ~0ull;  // call 0 = this kind was never grouped
  return (unsigned long long)(0xffffffffu - (uint32_t)word);
}
RowsGroupArgs group_args(nosplan_t* h, int kind, int D, int n_keys, int chunk, const int32_t* dem_off, const int32_t* dem_key,
                         const int32_t* dem_qty, const int32_t* key_cost, int32_t* first_node, bool single_only) {
  auto& rp = h->rp[kind];
  RowsGroupArgs g{};
  g.D = D;
  g.P = n_keys;
  g.chunk = chunk;
  g.max_keys = NOSPLAN_MAX_DEMAND_KEYS;
  g.max_qty = NOSPLAN_MAX_DEMAND_QTY;
  g.single_only = single_only ? 1 : 0;
  g.dem_off = dem_off;
  g.dem_key = dem_key;
  g.dem_qty = dem_qty;
  g.key_cost = key_cost;
  g.state = rp.cnt.as<int32_t>();
  g.rows = rp.rows.as<int32_t>();
  g.work = rp.work.as<int32_t>();
  g.meta = rp.meta.as<int32_t>();
  g.first_node = first_node;
  g.status = h->d_rowstatus.as<unsigned long long>() + kind;
  g.status_call = ++h->row_call[kind];
  return g;
}
// Groups the demand rows of one scoring call by key and cuts them into work items (rows_group_kernel, which
// also puts first_node[] to rest).
int plan_rows(nosplan_t* h, int kind, int D, int n_keys, int chunk, const int32_t* dem_off, const int32_t* dem_key,
              const int32_t* dem_qty, const int32_t* key_cost, int32_t* first_node, cudaStream_t st, int* max_work) {
  int rc = reserve_rows(h, kind, D, n_keys, chunk, st, max_work);
  if (rc) return rc;
  rows_group_kernel<<<(D + 255) / 256, 256, 0, st>>>(group_args(h, kind, D, n_keys, chunk, dem_off, dem_key, dem_qty, key_cost, first_node, false));
  CU(cudaGetLastError());
  return NOSPLAN_OK;
}

// Scorers in input order: the rows are checked by a pre-pass instead (rows_check_kernel), which also rests first_node[].
int check_rows(nosplan_t* h, int kind, int D, int n_keys, const int32_t* dem_off, const int32_t* dem_key, const int32_t* dem_qty,
               int32_t* first_node, cudaStream_t st, const uint8_t** row_ok) {
  CU(h->d_rowok[kind].reserve(std::max(D, 1)));
  if (!h->d_rowstatus.p) {
    CU(h->d_rowstatus.reserve(16));
    CU(cudaMemsetAsync(h->d_rowstatus.p, 0, 16, st));
  }
  RowsGroupArgs g{};
  g.D = D;
  g.P = n_keys;
  g.max_keys = NOSPLAN_MAX_DEMAND_KEYS;
  g.max_qty = NOSPLAN_MAX_DEMAND_QTY;
  g.dem_off = dem_off;
  g.dem_key = dem_key;
  g.dem_qty = dem_qty;
  g.first_node = first_node;
  g.status = h->d_rowstatus.as<unsigned long long>() + kind;
  g.status_call = ++h->row_call[kind];
  rows_check_kernel<<<(D + 255) / 256, 256, 0, st>>>(g, h->d_rowok[kind].as<uint8_t>());
  CU(cudaGetLastError());
  *row_ok = h->d_rowok[kind].as<uint8_t>();
  return NOSPLAN_OK;
}

// ---- the ordered MIG scorer: instantiation, shared memory and row chunk for the handle's tables ------------------------
struct Mig16Launch {
  void (*kern)(MigScore16Args) = nullptr;
  size_t smem = 0;
  int chunk = 64, tiles = 0, tn = kScoreTN;
};
// nodes per CTA of the ordered MIG scorer (the CTA always has 512 threads: 512 / tile of them share a node's rows)
int mig16_tile() {  // measurement knob: NOSPLAN_MIG_TILE=32|64|128
  static const int t = [] {
    const char* e = std::getenv("NOSPLAN_MIG_TILE");
    const int v = e ? std::atoi(e) : kMig16Tile;
    return (v == 32 || v == 64 || v == 128) ? v : kMig16Tile;
  }();
  return t;
}
template <int TN, int CW>
void (*mig16_kernel_of(bool out, bool cnt))(MigScore16Args) {
  if (out) return mig_score16_kernel<TN, 512 / TN, CW, true, NOSPLAN_RB, false>;
  if (cnt) return mig_score16_kernel<TN, 512 / TN, CW, false, NOSPLAN_RB, true>;
  return mig_score16_kernel<TN, 512 / TN, CW, false, NOSPLAN_RB, false>;
}
template <int TN>
void (*mig16_kernel_cw(int cw, bool out, bool cnt))(MigScore16Args) {
  switch (cw) {
    case 1: return mig16_kernel_of<TN, 1>(out, cnt);
    case 2: return mig16_kernel_of<TN, 2>(out, cnt);
    case 3: return mig16_kernel_of<TN, 3>(out, cnt);
    default: return mig16_kernel_of<TN, 4>(out, cnt);
  }
}
int mig16_prepare(nosplan_t* h, int D, bool out, bool cnt, Mig16Launch* L, bool both_kinds = false) {
  const int cw = std::max(1, std::min(4, (h->cmax + 3) / 4));
  const int tn = L->tn = mig16_tile();
  L->kern = tn == 32 ? mig16_kernel_cw<32>(cw, out, cnt) : tn == 64 ? mig16_kernel_cw<64>(cw, out, cnt) : mig16_kernel_cw<128>(cw, out, cnt);
  L->smem = mig_score16_smem_bytes(h->h_tables.n_models, h->mig_gs, h->cmax, tn);
  if (L->smem > 200 * 1024) return fail(h, NOSPLAN_ERR_RANGE, "MIG scorer tile needs %zu bytes of shared memory", L->smem);
  L->tiles = (h->mig_n + tn - 1) / tn;
  // (the attribute belongs to the function, not to the handle: another handle may have lowered it)
  CU(cudaFuncSetAttribute(L->kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L->smem));
  int& occ = h->occ_cache[0][out ? 1 : 0][cw];
  if (occ == 0) {  // resident CTAs per SM; reset when a snapshot changes the tile size
    int o = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, L->kern, 512, L->smem));
    occ = std::max(o, 1);
  }
  int gy = 1;
  // at least 64 rows per 128 nodes of a work item: on a small shard the tile load would otherwise outweigh the rows
  pick_chunks(h->sm_count, L->tiles, D, std::min(min_chunk() * 128 / tn, kRowCap), &L->chunk, &gy, occ * score_waves(both_kinds));
  L->chunk = std::min(L->chunk, kRowCap);
  return NOSPLAN_OK;
}
MigScore16Args mig16_args(nosplan_t* h, const MigScoreArgs& a) {
  auto& rp = h->rp[0];
  MigScore16Args A{};
  A.base = a;
  A.perm = h->d_mig_perm.as<int32_t>();
  A.rankg = h->d_mig_rankg.as<uint32_t>();
  A.ufg = h->d_mig_ufg.as<uint16_t>();
  A.rows = rp.rows.as<int32_t>();
  A.work = rp.work.as<int32_t>();
  A.meta = rp.meta.as<int32_t>();
  return A;
}
// The no-reuse regime (nosplan_stream.cuh): so few rows that a node's state is not worth staging for re-use.  The
// rule (measured, profiles/r2_stream_sweep.md): winners only, at most 8 GPU slots per node, and at most
// stream_rows() = 96 demand rows in the call — every kStreamRows rows of a profile cost one pass over the node list at
// HBM speed, which beats the ordered kernel's tile staging until a profile has about two dozen rows (1 M nodes, six
// profiles: 0.41 against 1.50 ms at 24 rows, 1.32 against 1.76 ms at 96, 2.51 against 2.42 ms at 192; at 2 048 nodes
// both forms take the same 20-30 us).
int stream_rows() {  // measurement knob: NOSPLAN_STREAM_ROWS=n (0 = never stream)
  static const int v = [] {
    const char* e = std::getenv("NOSPLAN_STREAM_ROWS");
    const int x = e ? std::atoi(e) : 96;
    return x >= 0 ? x : 96;
  }();
  return v;
}
bool mig_streaming(const nosplan_t* h, const MigScoreArgs& a) {
  return !a.updated && !a.choice && a.D <= stream_rows() && h->mig_gs <= kStreamMaxGpus && !plain_order_forced();
}
template <int CW, int RS>
void (*stream_kernel_gs(int gs))(MigStreamArgs) { return gs == 8 ? mig_stream_kernel<CW, RS, 8> : mig_stream_kernel<CW, RS, 0>; }
template <int CW>
void (*stream_kernel_rs(int rs, int gs))(MigStreamArgs) {
  return rs == 1 ? stream_kernel_gs<CW, 1>(gs) : rs == 2 ? stream_kernel_gs<CW, 2>(gs) : stream_kernel_gs<CW, kStreamRows>(gs);
}
void (*stream_kernel_of(int cw, int rs, int gs))(MigStreamArgs) {
  return cw == 1 ? stream_kernel_rs<1>(rs, gs) : cw == 2 ? stream_kernel_rs<2>(rs, gs) : cw == 3 ? stream_kernel_rs<3>(rs, gs) : stream_kernel_rs<4>(rs, gs);
}
int launch_mig_stream(nosplan_t* h, const MigScoreArgs& a, cudaStream_t st) {
  const int cw = std::max(1, std::min(4, (h->cmax + 3) / 4));
  int max_work = 0;
  const int rs = a.D <= 1 ? 1 : a.D <= 2 ? 2 : kStreamRows;  // rows per work item = rows a thread carries in registers
  int rc = plan_rows(h, 0, a.D, h->P, rs, a.dem_off, a.dem_key, a.dem_qty, h->d_profcost.as<int32_t>(), a.first_node, st, &max_work);
  if (rc) return rc;
  MigStreamArgs A{};
  A.T = a.T;
  A.rankg = h->d_mig_rankg.as<uint32_t>();
  A.ufg = h->d_mig_ufg.as<uint16_t>();
  A.modelg = h->d_mig_modelg.as<uint8_t>();
  A.s_cnt = a.s_cnt;
  A.N = a.N;
  A.npad = a.npad;
  A.gs = a.gs;
  A.dem_off = a.dem_off;
  A.dem_qty = a.dem_qty;
  A.rows = h->rp[0].rows.as<int32_t>();
  A.work = h->rp[0].work.as<int32_t>();
  A.meta = h->rp[0].meta.as<int32_t>();
  A.first_node = a.first_node;
  A.node_base = a.node_base;
  const size_t smem = mig_stream_smem_bytes(h->h_tables.n_models, a.gs, cw);
  if (smem > 200 * 1024) return fail(h, NOSPLAN_ERR_RANGE, "streaming scorer needs %zu bytes of shared memory", smem);
  auto kern = stream_kernel_of(cw, rs, a.gs);
  CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // persistent: as many CTAs as fit at once (or fewer when the node list is short); the warps stride over the work
  int occ = 0;
  CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kStreamWarps * 32, smem));
  // The warps stride over (work item, block of 32 nodes) pairs.  With the grid at its maximum the last round would be
  // partly empty (31 250 pairs over 4 736 warps: 6.6 rounds); size the grid so that every warp gets the same
  // number of rounds, counting the items a call with D rows can have at most (the real count is on the device)
  const long long items = std::min<long long>(max_work, std::max(1, (a.D + rs - 1) / rs));
  const long long pairs = (long long)(a.npad / 32) * items;
  const long long max_warps = (long long)h->sm_count * std::max(occ, 1) * kStreamWarps;
  const long long rounds = std::max<long long>(1, (pairs + max_warps - 1) / max_warps);
  const int grid = (int)std::max<long long>(1, (pairs + rounds * kStreamWarps - 1) / (rounds * kStreamWarps));
  kern<<<grid, kStreamWarps * 32, smem, st>>>(A);
  CU(cudaGetLastError());
  return NOSPLAN_OK;
}

int launch_mig_score16(nosplan_t* h, const MigScoreArgs& a, cudaStream_t st) {
  if (mig_streaming(h, a)) return launch_mig_stream(h, a, st);
  Mig16Launch L;
  int rc = mig16_prepare(h, a.D, a.updated || a.choice, false, &L);
  if (rc) return rc;
  int max_work = 0;
  rc = plan_rows(h, 0, a.D, h->P, L.chunk, a.dem_off, a.dem_key, a.dem_qty, h->d_profcost.as<int32_t>(), a.first_node, st, &max_work);
  if (rc) return rc;
  L.kern<<<dim3(max_work, L.tiles), 512, L.smem, st>>>(mig16_args(h, a));
  CU(cudaGetLastError());
  return NOSPLAN_OK;
}

// the rows-by-key path is taken (and with it the demand rows are validated on the device)
constexpr int kMaxOrderedRows = 60 * 1000 * 1000;  // rows[] positions (key * D + i) stay 32-bit
bool mig_ordered(const nosplan_t* h, int D) {
  return D <= kMaxOrderedRows && h->mig_n > 0 && h->k16ok && (h->cmax + 3) / 4 <= 4 && (h->mig_n + mig16_tile() - 1) / mig16_tile() <= 65535 &&
         !wide_keys_forced() && !plain_order_forced();
}
bool mps_ordered(const nosplan_t* h, int D) {
  return D <= kMaxOrderedRows && h->mps_n > 0 && (h->mps_n + kScoreTN - 1) / kScoreTN <= 65535 && !plain_order_forced();
}

// What the host walk (check_demands) would have said about the row the device flagged.
int describe_bad_row(nosplan_t* h, unsigned long long code, const int32_t* off, const int32_t* key, const int32_t* qty, int n_keys) {
  const int d = (int)(code / 64);
  const int k = off[d + 1] - off[d];
  if (k < 0 || off[d] < 0 || (code % 64) == 0) return fail(h, NOSPLAN_ERR_ARG, "dem_off not ascending at row %d", d);
  if (k > NOSPLAN_MAX_DEMAND_KEYS)
    return fail(h, NOSPLAN_ERR_RANGE, "demand row %d has %d keys (max %d)", d, k, NOSPLAN_MAX_DEMAND_KEYS);
  for (int e = off[d]; e < off[d + 1]; ++e) {
    if (key[e] < 0 || key[e] >= n_keys) return fail(h, NOSPLAN_ERR_RANGE, "demand row %d: key %d out of range", d, key[e]);
    if (e > off[d] && key[e] <= key[e - 1]) return fail(h, NOSPLAN_ERR_ARG, "demand row %d: keys must be strictly ascending", d);
    if (qty[e] < 0 || qty[e] > NOSPLAN_MAX_DEMAND_QTY)
      return fail(h, NOSPLAN_ERR_RANGE, "demand row %d: quantity %d out of range", d, qty[e]);
  }
  return fail(h, NOSPLAN_ERR_ARG, "demand row %d rejected on the device", d);
}
// head of check_demands: what does not need a walk over the rows
int check_demands_head(nosplan_t* h, int D, const int32_t* off, const int32_t* key, const int32_t* qty) {
  if (D < 0) return fail(h, NOSPLAN_ERR_ARG, "negative demand count");
  if (D == 0) return NOSPLAN_OK;
  if (!off) return fail(h, NOSPLAN_ERR_ARG, "dem_off is null");
  if (off[0] != 0) return fail(h, NOSPLAN_ERR_ARG, "dem_off[0] must be 0");
  if (off[D] > 0 && (!key || !qty)) return fail(h, NOSPLAN_ERR_ARG, "null demand arrays");
  return NOSPLAN_OK;
}

int launch_mig_score(nosplan_t* h, int D, const int32_t* d_off, const int32_t* d_key, const int32_t* d_qty,
                     int32_t* d_first, uint8_t* d_updated, int8_t* d_choice, cudaStream_t st, int base = 0) {
  if (D == 0) return NOSPLAN_OK;
  int rc = NOSPLAN_OK;
  const bool ordered = mig_ordered(h, D);
  const uint8_t* row_ok = nullptr;
  if (!ordered && (rc = check_rows(h, 0, D, h->P, d_off, d_key, d_qty, d_first, st, &row_ok))) return rc;  // else rows_group_kernel does it
  if (d_choice) CU(cudaMemsetAsync(d_choice, 0xff, (size_t)D * h->mig_tg, st));
  if (h->mig_n > 0) {
    MigScoreArgs a{};
    a.T = h->d_tables.as<MigTables>();
    a.s_used = h->d_mig_sused.as<uint4>();
    a.s_free = h->d_mig_sfree.as<uint4>();
    a.s_model = h->d_mig_smodel.as<uint8_t>();
    a.s_cnt = h->d_mig_scnt.as<uint8_t>();
    a.gpu_off = h->d_mig_off.as<int32_t>();
    a.N = h->mig_n;
    a.npad = h->mig_npad;
    a.gs = h->mig_gs;
    a.cmax = h->cmax;
    a.k16 = h->k16ok ? 1 : 0;
    a.tg = h->mig_tg;
    a.D = D;
    a.dem_off = d_off;
    a.dem_key = d_key;
    a.dem_qty = d_qty;
    a.first_node = d_first;
    a.node_base = base;
    a.updated = d_updated;
    a.choice = d_choice;
    a.row_list = nullptr;
    a.row_meta = nullptr;
    a.row_ok = row_ok;
    int tiles = (h->mig_n + kScoreTN - 1) / kScoreTN;
    int gy;
    pick_chunks(h->sm_count, tiles, D, 16, &a.d_chunk, &gy);
    size_t smem = mig_score_smem(h, kScoreTN);
    if (smem > 200 * 1024) return fail(h, NOSPLAN_ERR_RANGE, "MIG scorer tile needs %zu bytes of shared memory", smem);
    if (ordered) {
      // common tables: rows grouped by profile, nodes ordered per profile, 16-bit keys (nosplan_score16.cuh);
      // rows with several keys go through the general kernel below, restricted to their list
      if ((rc = launch_mig_score16(h, a, st))) return rc;
      a.row_list = h->rp[0].rows.as<int32_t>();
      a.row_meta = h->rp[0].meta.as<int32_t>();
    }
    // after the ordered / streaming kernel only the rows with several keys are left, usually none: a small grid
    // that strides over the tiles, so that an empty pass costs a few hundred CTAs whatever the node count
    const int gx = ordered ? std::min(tiles, 4 * h->sm_count) : tiles;
    if ((rc = launch_mig_score_kernel(h, a, dim3(gx, gy), smem, st))) return rc;
  }
  CU(cudaGetLastError());
  return NOSPLAN_OK;
}

int launch_mps_score(nosplan_t* h, int D, const int32_t* d_off, const int32_t* d_key, const int32_t* d_qty,
                     int32_t* d_first, uint8_t* d_updated, uint8_t* d_free, uint32_t* d_mask, cudaStream_t st, int base = 0) {
  if (D == 0) return NOSPLAN_OK;
  const bool ordered = mps_ordered(h, D);
  const uint8_t* row_ok = nullptr;
  if (!ordered) {  // else rows_group_kernel does it
    int rc = check_rows(h, 1, D, h->S, d_off, d_key, d_qty, d_first, st, &row_ok);
    if (rc) return rc;
  }
  if (h->mps_n > 0) {
    MpsScoreArgs a{};
    a.sizes = h->d_mps_sizes.as<int32_t>();
    a.S = h->S;
    a.gp = h->d_mps_gpu0.as<MpsGpu>();
    a.gpu_off = h->d_mps_off.as<int32_t>();
    a.valid = h->d_mps_valid.as<uint8_t>();
    a.N = h->mps_n;
    a.tg = h->mps_tg;
    a.D = D;
    a.dem_off = d_off;
    a.dem_key = d_key;
    a.dem_qty = d_qty;
    a.first_node = d_first;
    a.node_base = base;
    a.updated = d_updated;
    a.out_free = d_free;
    a.out_mask = d_mask;
    a.gs = h->mps_gs;
    a.row_ok = row_ok;
    int tiles = (h->mps_n + kScoreTN - 1) / kScoreTN;
    int gy;
    pick_chunks(h->sm_count, tiles, D, (ordered && !d_free) ? 128 : 16, &a.d_chunk, &gy,
                ((ordered && !d_free) ? 4 : 3) * score_waves());  // CTAs per SM that fit; >= 128 rows per item
    const size_t smem = mps_score_smem_bytes(h->mps_gs, h->S, kScoreTN);
    if (smem > 200 * 1024) return fail(h, NOSPLAN_ERR_RANGE, "MPS scorer tile needs %zu bytes of shared memory", smem);
    CU(cudaFuncSetAttribute(mps_score_kernel<kScoreTN, kScoreDS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    a.mode = 0;
    a.perm = a.rows = a.work = a.meta = nullptr;
    if (ordered) {
      // single-size rows grouped by size over nodes ordered per size; the rest over plain tiles
      int max_work = 0;
      int rc = plan_rows(h, 1, D, h->S, std::min(a.d_chunk, kRowCap), d_off, d_key, d_qty, h->d_mps_sizecost.as<int32_t>(), d_first, st, &max_work);
      if (rc) return rc;
      a.perm = h->d_mps_perm.as<int32_t>();
      a.rows = h->rp[1].rows.as<int32_t>();
      a.work = h->rp[1].work.as<int32_t>();
      a.meta = h->rp[1].meta.as<int32_t>();
      a.mode = 1;
      if (!d_free) {  // first_node / updated only: the per-size kernel
        const size_t smem1 = mps_score1_smem_bytes(h->mps_gs, h->S, kScoreTN);
        if (smem1 > 200 * 1024) return fail(h, NOSPLAN_ERR_RANGE, "MPS scorer tile needs %zu bytes of shared memory", smem1);
        if (d_updated) {
          CU(cudaFuncSetAttribute(mps_score1_kernel<kScoreTN, kScoreDS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1));
          mps_score1_kernel<kScoreTN, kScoreDS, true><<<dim3(max_work, tiles), kScoreTN * kScoreDS, smem1, st>>>(a);
        } else {
          CU(cudaFuncSetAttribute(mps_score1_kernel<kScoreTN, kScoreDS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1));
          mps_score1_kernel<kScoreTN, kScoreDS, false><<<dim3(max_work, tiles), kScoreTN * kScoreDS, smem1, st>>>(a);
        }
      } else {
        mps_score_kernel<kScoreTN, kScoreDS><<<dim3(max_work, tiles), kScoreTN * kScoreDS, smem, st>>>(a);
      }
      a.mode = 2;
    }
    mps_score_kernel<kScoreTN, kScoreDS><<<dim3(tiles, gy), kScoreTN * kScoreDS, smem, st>>>(a);
  }
  CU(cudaGetLastError());
  return NOSPLAN_OK;
}

}  // namespace

extern "C" {

const char* nosplan_version(void) { return "nosplan 0.1 (sm_100a)"; }

int nosplan_create(int device, nosplan_t** out) {
  if (!out) return NOSPLAN_ERR_ARG;
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || device < 0 || device >= count) return NOSPLAN_ERR_CUDA;  // no device, no planner
  nosplan_t* h = new nosplan();
  h->device = device;
  if (cudaSetDevice(device) != cudaSuccess || cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithFlags(&h->stream2, cudaStreamNonBlocking) != cudaSuccess) {
    delete h;
    return NOSPLAN_ERR_CUDA;
  }
  cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, device);
  if (cudaHostAlloc(reinterpret_cast<void**>(&h->h_status), 128, cudaHostAllocDefault) != cudaSuccess ||
      cudaEventCreateWithFlags(&h->status_ev, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&h->status_ev2, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming) != cudaSuccess) {
    nosplan_destroy(h);
    return NOSPLAN_ERR_CUDA;
  }
  h->h_limit = reinterpret_cast<int32_t*>(h->h_status + 8);              // bytes 64..71
  h->h_range = reinterpret_cast<int32_t(*)[2]>(h->h_status + 10);       // bytes 80..95
  *out = h;
  return NOSPLAN_OK;
}

void nosplan_destroy(nosplan_t* h) {
  if (!h) return;
  lanes_release(h);
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  if (h->xch.created) {
    for (int r = 0; r < h->xch.world; ++r)
      if (h->xch.open && r != h->xch.rank && h->xch.mapped[r]) cudaIpcCloseMemHandle(h->xch.mapped[r]);
    if (h->xch.local) cudaFree(h->xch.local);
    h->xch.done.release();
    h->xch.err.release();
  }
  DevBuf* all[] = {&h->d_tables, &h->d_mig_off, &h->d_mig_valid, &h->d_mig_model, &h->d_mig_used0, &h->d_mig_free0,
                   &h->d_mig_used, &h->d_mig_free, &h->d_mig_sused, &h->d_mig_sfree, &h->d_mig_smodel, &h->d_mig_scnt,
                   &h->d_mps_sizes, &h->d_mps_off, &h->d_mps_valid, &h->d_mps_gpu0, &h->d_mps_ulo0, &h->d_mps_uhi0,
                   &h->d_mps_gpu, &h->d_mps_ulo, &h->d_mps_uhi, &h->d_mig_gnode, &h->d_mps_gnode, &h->d_mig_gslot, &h->d_mps_gslot, &h->d_status, &h->d_rowstatus,
                   &h->d_mig_cost, &h->d_mig_perm, &h->d_mig_rankg, &h->d_mig_ufg, &h->d_mig_maskg, &h->d_mig_modelg, &h->d_profcost, &h->d_mps_cost, &h->d_mps_perm, &h->d_mps_sizecost,
                   &h->rp[0].rows, &h->rp[0].work, &h->rp[0].meta, &h->rp[0].cnt,
                   &h->rp[1].rows, &h->rp[1].work, &h->rp[1].meta, &h->rp[1].cnt};
  for (DevBuf* b : all) b->release();
  for (auto& b : h->d_in) b.release();
  for (auto& b : h->d_dem) b.release();
  for (auto& b : h->d_out) b.release();
  for (auto& b : h->d_plan) b.release();
  if (h->h_status) cudaFreeHost(h->h_status);
  if (h->status_ev) cudaEventDestroy(h->status_ev);
  if (h->status_ev2) cudaEventDestroy(h->status_ev2);
  if (h->h_pin) cudaFreeHost(h->h_pin);
  for (auto& b : h->d_inp) b.release();
  for (auto& b : h->d_rowok) b.release();
  if (h->ev_fork) cudaEventDestroy(h->ev_fork);
  if (h->ev_join) cudaEventDestroy(h->ev_join);
  for (auto& kind : h->kev)
    for (auto& pair : kind)
      for (cudaEvent_t e : pair)
        if (e) cudaEventDestroy(e);
  h->d_counter.release();
  h->d_passbuf.release();
  if (h->stream) cudaStreamDestroy(h->stream);
  if (h->stream2) cudaStreamDestroy(h->stream2);
  delete h;
}

const char* nosplan_last_error(const nosplan_t* h) { return h ? h->err.c_str() : "null handle"; }

int nosplan_set_mig_tables(nosplan_t* h, int n_profiles, const char* const* profile_names, int n_models,
                           const int32_t* geom_off, const int32_t* geom_counts) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  if (!profile_names || !geom_off) return fail(h, NOSPLAN_ERR_ARG, "null argument");
  if (n_profiles < 1 || n_profiles > NOSPLAN_MAX_MIG_PROFILES)
    return fail(h, NOSPLAN_ERR_RANGE, "n_profiles %d out of range [1,%d]", n_profiles, NOSPLAN_MAX_MIG_PROFILES);
  if (n_models < 1) return fail(h, NOSPLAN_ERR_CONFIG, "no known configs provided");  // known_configs.go:165-167
  if (n_models > NOSPLAN_MAX_MODELS) return fail(h, NOSPLAN_ERR_RANGE, "too many models (%d > %d)", n_models, NOSPLAN_MAX_MODELS);
  if (!geom_counts && geom_off[n_models] > 0) return fail(h, NOSPLAN_ERR_ARG, "null argument");  // (n_models is in range here)
  std::vector<std::string> names(n_profiles);
  for (int p = 0; p < n_profiles; ++p) {
    names[p] = profile_names[p] ? profile_names[p] : "";
    if (p > 0 && !(names[p - 1] < names[p])) return fail(h, NOSPLAN_ERR_ARG, "profile names must be strictly ascending");
  }
  MigTables& T = h->h_tables;
  std::memset(&T, 0, sizeof(T));
  T.n_models = n_models;
  T.n_prof = n_profiles;
  if (geom_off[0] != 0) return fail(h, NOSPLAN_ERR_ARG, "geom_off[0] must be 0");
  int cmax = 0;
  for (int m = 0; m < n_models; ++m) {
    int c = geom_off[m + 1] - geom_off[m];
    if (c < 0) return fail(h, NOSPLAN_ERR_ARG, "geom_off not ascending");
    if (c > NOSPLAN_MAX_GEOMS) return fail(h, NOSPLAN_ERR_RANGE, "model %d has %d geometries (max %d)", m, c, NOSPLAN_MAX_GEOMS);
    cmax = std::max(cmax, c);
    T.geom_off[m] = geom_off[m];
  }
  T.geom_off[n_models] = geom_off[n_models];
  T.n_geoms = geom_off[n_models];
  for (int m = 0; m < n_models; ++m) {
    std::vector<std::string> ids;
    int fewest = -1, fewest_kinds = 0;
    for (int gi = geom_off[m]; gi < geom_off[m + 1]; ++gi) {
      unsigned char bytes[16] = {0};
      int total = 0, kinds = 0;
      std::string id;  // Geometry.Id(): "%s:%d, " per profile, names ascending (partitioning.go:41-56)
      for (int p = 0; p < n_profiles; ++p) {
        int q = geom_counts[(size_t)gi * n_profiles + p];
        if (q == 0) continue;
        if (q < 1) return fail(h, NOSPLAN_ERR_CONFIG, "invalid quantity %d for profile %s", q, names[p].c_str());
        if (!mig_name_valid(names[p])) return fail(h, NOSPLAN_ERR_CONFIG, "invalid profile %s", names[p].c_str());
        if (q > NOSPLAN_MAX_COUNT) return fail(h, NOSPLAN_ERR_RANGE, "quantity %d exceeds %d", q, NOSPLAN_MAX_COUNT);
        bytes[p] = (unsigned char)q;
        total += q;
        kinds++;
        id += names[p] + ":" + std::to_string(q) + ", ";
      }
      std::memcpy(&T.cand[gi], bytes, 16);
      T.total[gi] = (uint16_t)total;
      ids.push_back(id);
      if (fewest < 0 || kinds < fewest_kinds) {  // partitioning.go:67-79: first of the fewest kinds
        fewest = gi - geom_off[m];
        fewest_kinds = kinds;
      }
    }
    T.init_choice[m] = fewest;
    std::vector<std::string> sorted = ids;
    std::sort(sorted.begin(), sorted.end());
    sorted.erase(std::unique(sorted.begin(), sorted.end()), sorted.end());
    for (size_t i = 0; i < ids.size(); ++i)
      T.rank[geom_off[m] + i] = (uint8_t)(std::lower_bound(sorted.begin(), sorted.end(), ids[i]) - sorted.begin());
  }
  int rc = bind(h);
  if (rc) return rc;
  rc = upload(h, h->d_tables, &T, sizeof(T), h->stream);
  if (rc) return rc;
  CU(cudaStreamSynchronize(h->stream));
  h->P = n_profiles;
  h->names = names;
  h->tab_models = n_models;
  h->tab_geom_off.assign(geom_off, geom_off + n_models + 1);
  h->tab_geom_counts.assign(geom_counts, geom_counts + (size_t)geom_off[n_models] * n_profiles);
  ++h->tab_version;
  h->cmax = std::max(cmax, 1);
  int maxq = 0;
  for (size_t i = 0; i < (size_t)geom_off[n_models] * n_profiles; ++i) maxq = std::max(maxq, (int)geom_counts[i]);
  h->k16ok = h->cmax <= 16 && maxq <= 63;
  h->have_tables = true;
  std::memset(h->occ_cache, 0, sizeof(h->occ_cache));
  h->have_mig = false;  // a snapshot refers to model ids of the table it was uploaded against
  return NOSPLAN_OK;
}

}  // extern "C"

namespace {
// What the scorers derive from the packed snapshot (scheduling data and the demand-independent half of the scores):
// re-run whenever the packed state the scorers read changes (a new snapshot; every round of the batch-demand mode).
int mig_derive_tables(nosplan_t* h, int n_nodes, int npad, int gsl, cudaStream_t st) {
  if (!h->k16ok) return NOSPLAN_OK;
  // 16-bit-key scorer: the demand-independent candidate ranks of every GPU, then the node order per
  // profile (scheduling data, not results)
  const int P = h->P;
  const int cwk = (h->cmax + 3) / 4;
  CU(h->d_mig_rankg.reserve(sizeof(uint32_t) * (size_t)npad * gsl * 2 * cwk));
  CU(h->d_mig_ufg.reserve(sizeof(uint16_t) * (size_t)npad * 16 * gsl));
  CU(h->d_mig_maskg.reserve(sizeof(uint32_t) * (size_t)npad * gsl));
  CU(h->d_mig_modelg.reserve((size_t)npad * gsl));
  CU(h->d_mig_cost.reserve((size_t)P * npad));
  CU(h->d_mig_perm.reserve(sizeof(int32_t) * (size_t)P * npad));
  CU(h->d_profcost.reserve(sizeof(int32_t) * 32));
  mig_rank16_kernel<<<(npad * gsl + 127) / 128, 128, 0, st>>>(
      h->d_tables.as<MigTables>(), npad, gsl, cwk, h->d_mig_sused.as<uint4>(), h->d_mig_sfree.as<uint4>(),
      h->d_mig_smodel.as<uint8_t>(), h->d_mig_scnt.as<uint8_t>(), h->d_mig_rankg.as<uint32_t>(), h->d_mig_ufg.as<uint16_t>(),
      h->d_mig_maskg.as<uint32_t>(), h->d_mig_modelg.as<uint8_t>());
  mig_cost_kernel<<<(npad + 127) / 128, 128, 0, st>>>(n_nodes, npad, gsl, P, h->d_mig_scnt.as<uint8_t>(),
                                                      h->d_mig_maskg.as<uint32_t>(), h->d_mig_cost.as<uint8_t>());
  mig_perm_kernel<<<P, 1024, 0, st>>>(npad, h->d_mig_cost.as<uint8_t>(), h->d_mig_perm.as<int32_t>(),
                                      h->d_profcost.as<int32_t>());
  CU(cudaGetLastError());
  return NOSPLAN_OK;
}
int mps_derive_tables(nosplan_t* h, int n_nodes, int S, cudaStream_t st) {
  // node order per size for the scorer (scheduling data, see nosplan_score16.cuh)
  CU(h->d_mps_cost.reserve((size_t)S * n_nodes));
  CU(h->d_mps_perm.reserve(sizeof(int32_t) * (size_t)S * n_nodes));
  CU(h->d_mps_sizecost.reserve(sizeof(int32_t) * 32));
  mps_cost_kernel<<<(n_nodes + 127) / 128, 128, 0, st>>>(n_nodes, S, h->d_mps_sizes.as<int32_t>(), h->d_mps_off.as<int32_t>(),
                                                         h->d_mps_valid.as<uint8_t>(), h->d_mps_gpu0.as<MpsGpu>(),
                                                         h->d_mps_cost.as<uint8_t>());
  mig_perm_kernel<<<S, 1024, 0, st>>>(n_nodes, h->d_mps_cost.as<uint8_t>(), h->d_mps_perm.as<int32_t>(),
                                      h->d_mps_sizecost.as<int32_t>());
  CU(cudaGetLastError());
  return NOSPLAN_OK;
}

// Everything of a snapshot upload except waiting for the device's verdict (the range checks run in the pack
// kernels): host-side checks of the node list, copies, pack, scheduling tables, read-back of the status words.
// The handle's sizes are set at once so that work can be enqueued behind it; *_snapshot_finish waits for the
// status and either confirms the snapshot or drops it.
int mig_snapshot_enqueue(nosplan_t* h, int n_nodes, const int32_t* gpu_off, const uint8_t* valid, const int32_t* model,
                         const int32_t* used, const int32_t* free_, cudaStream_t st) {
  NSP_RANGE("nosplan: MIG snapshot upload + pack");
  if (!h->have_tables) return fail(h, NOSPLAN_ERR_ARG, "nosplan_set_mig_tables must be called first");
  if (n_nodes < 0 || !gpu_off) return fail(h, NOSPLAN_ERR_ARG, "bad node list");
  const int P = h->P;
  if (gpu_off[0] != 0) return fail(h, NOSPLAN_ERR_ARG, "gpu_off[0] must be 0");
  int gs = 0;
  for (int n = 0; n < n_nodes; ++n) {
    int c = gpu_off[n + 1] - gpu_off[n];
    if (c < 0) return fail(h, NOSPLAN_ERR_ARG, "gpu_off not ascending at node %d", n);
    if (c > NOSPLAN_MAX_GPUS_PER_NODE) return fail(h, NOSPLAN_ERR_RANGE, "node %d has %d GPUs (max %d)", n, c, NOSPLAN_MAX_GPUS_PER_NODE);
    gs = std::max(gs, c);
  }
  const int tg = gpu_off[n_nodes];
  if (tg > 0 && (!model || !used || !free_)) return fail(h, NOSPLAN_ERR_ARG, "null state arrays");
  int rc = bind(h);
  if (rc) return rc;
  const int npad = std::max(32, (n_nodes + 31) / 32 * 32);
  const int gsl = std::max(gs, 1);
  h->have_mig = false;  // the device copy is being replaced: a failed call leaves no snapshot behind
  // the arrays go up as they are; range checks, the zeroing of invalid nodes' rows and the per-GPU
  // node/slot tables are the pack kernel's work (the host only walked gpu_off)
  if ((rc = upload(h, h->d_mig_off, gpu_off, sizeof(int32_t) * (n_nodes + 1), st))) return rc;
  CU(h->d_mig_valid.reserve(std::max(n_nodes, 1)));
  if (valid) {
    if ((rc = upload(h, h->d_mig_valid, valid, n_nodes, st))) return rc;
  } else {
    CU(cudaMemsetAsync(h->d_mig_valid.p, 1, std::max(n_nodes, 1), st));
  }
  if ((rc = upload(h, h->d_in[0], model, sizeof(int32_t) * (size_t)tg, st))) return rc;
  if ((rc = upload(h, h->d_in[1], used, sizeof(int32_t) * (size_t)tg * P, st))) return rc;
  if ((rc = upload(h, h->d_in[2], free_, sizeof(int32_t) * (size_t)tg * P, st))) return rc;
  CU(h->d_mig_gnode.reserve(sizeof(int32_t) * std::max(tg, 1)));
  CU(h->d_mig_gslot.reserve(sizeof(int32_t) * std::max(tg, 1)));
  CU(h->d_mig_model.reserve(std::max(tg, 1)));
  CU(h->d_mig_used0.reserve(sizeof(uint4) * std::max(tg, 1)));
  CU(h->d_mig_free0.reserve(sizeof(uint4) * std::max(tg, 1)));
  CU(h->d_mig_used.reserve(sizeof(uint4) * std::max(tg, 1)));
  CU(h->d_mig_free.reserve(sizeof(uint4) * std::max(tg, 1)));
  CU(h->d_mig_sused.reserve(sizeof(uint4) * (size_t)gsl * npad));
  CU(h->d_mig_sfree.reserve(sizeof(uint4) * (size_t)gsl * npad));
  CU(h->d_mig_smodel.reserve((size_t)gsl * npad));
  CU(h->d_mig_scnt.reserve(npad));
  if ((rc = status_reset(h, 0, st))) return rc;
  mig_pack_kernel<<<(npad * gsl + 127) / 128, 128, 0, st>>>(
      n_nodes, P, gsl, npad, NOSPLAN_MAX_COUNT, h->d_tables.as<MigTables>(), h->d_mig_off.as<int32_t>(),
      h->d_mig_valid.as<uint8_t>(), h->d_in[0].as<int32_t>(), h->d_in[1].as<int32_t>(), h->d_in[2].as<int32_t>(),
      h->d_mig_used0.as<uint4>(), h->d_mig_free0.as<uint4>(), h->d_mig_model.as<uint8_t>(), h->d_mig_sused.as<uint4>(),
      h->d_mig_sfree.as<uint4>(), h->d_mig_smodel.as<uint8_t>(), h->d_mig_scnt.as<uint8_t>(), h->d_mig_gnode.as<int32_t>(),
      h->d_mig_gslot.as<int32_t>(), status_words(h, 0));
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(h->d_mig_used.p, h->d_mig_used0.p, sizeof(uint4) * tg, cudaMemcpyDeviceToDevice, st));
  CU(cudaMemcpyAsync(h->d_mig_free.p, h->d_mig_free0.p, sizeof(uint4) * tg, cudaMemcpyDeviceToDevice, st));
  if ((rc = status_enqueue(h, 0, st))) return rc;
  if ((rc = mig_derive_tables(h, n_nodes, npad, gsl, st))) return rc;
  if (h->mig_gs != gsl) std::memset(h->occ_cache[0], 0, sizeof(h->occ_cache[0]));  // the tile follows the GPUs per node
  h->mig_n = n_nodes;
  h->mig_tg = tg;
  h->mig_gs = gsl;
  h->mig_npad = npad;
  h->mig_evals = 0;
  h->mig_state_dirty = false;
  return NOSPLAN_OK;
}
int mig_snapshot_finish(nosplan_t* h, const int32_t* model, const int32_t* used, const int32_t* free_) {
  NSP_RANGE("nosplan: MIG snapshot verdict");
  unsigned long long st[2] = {~0ull, 0ull};
  int rc = status_wait(h, 0, st);
  if (rc) return rc;
  const int P = h->P;
  if (st[0] != ~0ull) {  // first bad entry in GPU order, reported as the host walk used to
    const int g = (int)(st[0] / 64), sub = (int)(st[0] % 64);
    if (sub == 0) return fail(h, NOSPLAN_ERR_RANGE, "GPU %d: model %d unknown (mark the node invalid instead)", g, model[g]);
    const int p = sub - 1;
    return fail(h, NOSPLAN_ERR_RANGE, "GPU %d profile %s: used %d free %d out of range [0,%d]", g, h->names[p].c_str(),
                used[(size_t)g * P + p], free_[(size_t)g * P + p], NOSPLAN_MAX_COUNT);
  }
  h->mig_evals = (int64_t)st[1];
  h->have_mig = true;
  return NOSPLAN_OK;
}

int mps_snapshot_enqueue(nosplan_t* h, int n_sizes, const int32_t* sizes_gb, int n_nodes, const int32_t* gpu_off,
                         const uint8_t* valid, const int32_t* mem_gb, const int32_t* used, const int32_t* free_,
                         const uint32_t* free_mask, cudaStream_t st) {
  NSP_RANGE("nosplan: MPS snapshot upload + pack");
  if (n_sizes < 1 || n_sizes > NOSPLAN_MAX_MPS_SIZES || !sizes_gb)
    return fail(h, NOSPLAN_ERR_RANGE, "n_sizes %d out of range [1,%d]", n_sizes, NOSPLAN_MAX_MPS_SIZES);
  for (int s = 0; s < n_sizes; ++s) {
    if (sizes_gb[s] < 1 || sizes_gb[s] > NOSPLAN_MAX_MEM_GB)  // MinSliceMemoryGB, slicing/constant.go:24
      return fail(h, NOSPLAN_ERR_RANGE, "slice size %d GB out of range [1,%d]", sizes_gb[s], NOSPLAN_MAX_MEM_GB);
    if (s > 0 && sizes_gb[s] <= sizes_gb[s - 1]) return fail(h, NOSPLAN_ERR_ARG, "sizes must be strictly ascending");
  }
  if (n_nodes < 0 || !gpu_off || gpu_off[0] != 0) return fail(h, NOSPLAN_ERR_ARG, "bad node list");
  int gs = 0;
  for (int n = 0; n < n_nodes; ++n) {
    int c = gpu_off[n + 1] - gpu_off[n];
    if (c < 0) return fail(h, NOSPLAN_ERR_ARG, "gpu_off not ascending at node %d", n);
    if (c > NOSPLAN_MAX_GPUS_PER_NODE) return fail(h, NOSPLAN_ERR_RANGE, "node %d has %d GPUs (max %d)", n, c, NOSPLAN_MAX_GPUS_PER_NODE);
    gs = std::max(gs, c);
  }
  const int tg = gpu_off[n_nodes];
  const int S = n_sizes;
  if (tg > 0 && (!mem_gb || !used || !free_)) return fail(h, NOSPLAN_ERR_ARG, "null state arrays");
  int rc = bind(h);
  if (rc) return rc;
  h->have_mps = false;  // a failed call leaves no snapshot behind
  if ((rc = upload(h, h->d_mps_sizes, sizes_gb, sizeof(int32_t) * S, st))) return rc;
  if ((rc = upload(h, h->d_mps_off, gpu_off, sizeof(int32_t) * (n_nodes + 1), st))) return rc;
  CU(h->d_mps_valid.reserve(std::max(n_nodes, 1)));
  if (valid) {
    if ((rc = upload(h, h->d_mps_valid, valid, n_nodes, st))) return rc;
  } else {
    CU(cudaMemsetAsync(h->d_mps_valid.p, 1, std::max(n_nodes, 1), st));
  }
  if ((rc = upload(h, h->d_inp[0], mem_gb, sizeof(int32_t) * (size_t)tg, st))) return rc;
  if ((rc = upload(h, h->d_inp[1], used, sizeof(int32_t) * (size_t)tg * S, st))) return rc;
  if ((rc = upload(h, h->d_inp[2], free_, sizeof(int32_t) * (size_t)tg * S, st))) return rc;
  if (free_mask && (rc = upload(h, h->d_inp[3], free_mask, sizeof(uint32_t) * (size_t)tg, st))) return rc;
  CU(h->d_mps_gnode.reserve(sizeof(int32_t) * std::max(tg, 1)));
  CU(h->d_mps_gslot.reserve(sizeof(int32_t) * std::max(tg, 1)));
  CU(h->d_mps_gpu0.reserve(sizeof(MpsGpu) * std::max(tg, 1)));
  CU(h->d_mps_ulo0.reserve(sizeof(uint4) * std::max(tg, 1)));
  CU(h->d_mps_uhi0.reserve(sizeof(uint4) * std::max(tg, 1)));
  CU(h->d_mps_gpu.reserve(sizeof(MpsGpu) * std::max(tg, 1)));
  CU(h->d_mps_ulo.reserve(sizeof(uint4) * std::max(tg, 1)));
  CU(h->d_mps_uhi.reserve(sizeof(uint4) * std::max(tg, 1)));
  if ((rc = status_reset(h, 1, st))) return rc;
  if (n_nodes > 0) {
    mps_pack_kernel<<<(n_nodes + 127) / 128, 128, 0, st>>>(
        n_nodes, S, NOSPLAN_MAX_COUNT, NOSPLAN_MAX_MEM_GB, h->d_mps_sizes.as<int32_t>(), h->d_mps_off.as<int32_t>(),
        h->d_mps_valid.as<uint8_t>(), h->d_inp[0].as<int32_t>(), h->d_inp[1].as<int32_t>(), h->d_inp[2].as<int32_t>(),
        free_mask ? h->d_inp[3].as<uint32_t>() : nullptr, h->d_mps_gpu0.as<MpsGpu>(), h->d_mps_ulo0.as<uint4>(),
        h->d_mps_uhi0.as<uint4>(), h->d_mps_gnode.as<int32_t>(), h->d_mps_gslot.as<int32_t>(), status_words(h, 1));
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(h->d_mps_gpu.p, h->d_mps_gpu0.p, sizeof(MpsGpu) * tg, cudaMemcpyDeviceToDevice, st));
    CU(cudaMemcpyAsync(h->d_mps_ulo.p, h->d_mps_ulo0.p, sizeof(uint4) * tg, cudaMemcpyDeviceToDevice, st));
    CU(cudaMemcpyAsync(h->d_mps_uhi.p, h->d_mps_uhi0.p, sizeof(uint4) * tg, cudaMemcpyDeviceToDevice, st));
    if ((rc = status_enqueue(h, 1, st))) return rc;
    if ((rc = mps_derive_tables(h, n_nodes, S, st))) return rc;
  }
  if (n_nodes == 0 && (rc = status_enqueue(h, 1, st))) return rc;
  h->S = S;
  h->mps_n = n_nodes;
  h->mps_tg = tg;
  h->mps_gs = std::max(gs, 1);
  h->mps_evals = 0;
  h->mps_state_dirty = false;
  return NOSPLAN_OK;
}
int mps_snapshot_finish(nosplan_t* h, const int32_t* sizes_gb, const int32_t* mem_gb, const int32_t* used, const int32_t* free_) {
  NSP_RANGE("nosplan: MPS snapshot verdict");
  unsigned long long st[2] = {~0ull, 0ull};
  int rc = status_wait(h, 1, st);
  if (rc) return rc;
  const int S = h->S;
  if (st[0] != ~0ull) {
    const int g = (int)(st[0] / 128), sub = (int)(st[0] % 128);
    if (sub == 0) return fail(h, NOSPLAN_ERR_RANGE, "GPU %d: memory %d GB out of range", g, mem_gb[g]);
    if (sub == 127) {
      long long tot = 0;
      for (int s = 0; s < S; ++s) tot += (long long)sizes_gb[s] * ((long long)used[(size_t)g * S + s] + free_[(size_t)g * S + s]);
      // slicing.NewGPU rejects such a GPU (slicing/gpu.go:92-95): its node must be marked invalid
      return fail(h, NOSPLAN_ERR_RANGE, "GPU %d: slices total %lld GB exceed memory %d GB", g, tot, mem_gb[g]);
    }
    const int s = (sub - 1) / 2;
    if ((sub - 1) % 2 == 0)
      return fail(h, NOSPLAN_ERR_RANGE, "GPU %d size %dgb: used %d free %d out of range", g, sizes_gb[s],
                  used[(size_t)g * S + s], free_[(size_t)g * S + s]);
    return fail(h, NOSPLAN_ERR_ARG, "GPU %d: free_mask misses a key with a non-zero value", g);
  }
  h->mps_evals = (int64_t)st[1];
  h->have_mps = true;
  return NOSPLAN_OK;
}
}  // namespace

extern "C" {

int nosplan_set_mig_snapshot(nosplan_t* h, int n_nodes, const int32_t* gpu_off, const uint8_t* valid,
                             const int32_t* model, const int32_t* used, const int32_t* free_) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  int rc = mig_snapshot_enqueue(h, n_nodes, gpu_off, valid, model, used, free_, h->stream);
  if (rc) return rc;
  return mig_snapshot_finish(h, model, used, free_);
}

int nosplan_set_mps_snapshot(nosplan_t* h, int n_sizes, const int32_t* sizes_gb, int n_nodes, const int32_t* gpu_off,
                             const uint8_t* valid, const int32_t* mem_gb, const int32_t* used, const int32_t* free_,
                             const uint32_t* free_mask) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  int rc = mps_snapshot_enqueue(h, n_sizes, sizes_gb, n_nodes, gpu_off, valid, mem_gb, used, free_, free_mask, h->stream);
  if (rc) return rc;
  return mps_snapshot_finish(h, sizes_gb, mem_gb, used, free_);
}

int64_t nosplan_mig_evals_per_demand(const nosplan_t* h) { return (h && h->have_mig) ? h->mig_evals : 0; }
int64_t nosplan_mps_evals_per_demand(const nosplan_t* h) { return (h && h->have_mps) ? h->mps_evals : 0; }

int nosplan_score_mig_dev(nosplan_t* h, int n_demands, const int32_t* dem_off, const int32_t* dem_key,
                          const int32_t* dem_qty, int32_t* first_node, uint8_t* updated, int8_t* choice, void* stream) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  if (!h->have_mig) return fail(h, NOSPLAN_ERR_ARG, "no MIG snapshot");
  if (!first_node) return fail(h, NOSPLAN_ERR_ARG, "first_node is required");
  int rc = bind(h);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  rc = launch_mig_score(h, n_demands, dem_off, dem_key, dem_qty, first_node, updated, choice, st, h->mig_base);
  if (rc) return rc;
  CU(cudaGetLastError());
  return NOSPLAN_OK;
}

int nosplan_score_mps_dev(nosplan_t* h, int n_demands, const int32_t* dem_off, const int32_t* dem_key,
                          const int32_t* dem_qty, int32_t* first_node, uint8_t* updated, uint8_t* out_free,
                          uint32_t* out_free_mask, void* stream) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  if (!h->have_mps) return fail(h, NOSPLAN_ERR_ARG, "no MPS snapshot");
  if (!first_node) return fail(h, NOSPLAN_ERR_ARG, "first_node is required");
  if ((out_free == nullptr) != (out_free_mask == nullptr)) return fail(h, NOSPLAN_ERR_ARG, "out_free and out_free_mask go together");
  int rc = bind(h);
  if (rc) return rc;
  return launch_mps_score(h, n_demands, dem_off, dem_key, dem_qty, first_node, updated, out_free, out_free_mask,
                          (cudaStream_t)stream, h->mps_base);
}

}  // extern "C"

// ---- a whole device-resident scoring pass in one call (nosplan_score_pass_dev) ---------------------------------------
namespace {

MigScoreArgs mig_base_args(nosplan_t* h, int D, const int32_t* off, const int32_t* key, const int32_t* qty, int32_t* first, int base) {
  MigScoreArgs a{};
  a.T = h->d_tables.as<MigTables>();
  a.s_used = h->d_mig_sused.as<uint4>();
  a.s_free = h->d_mig_sfree.as<uint4>();
  a.s_model = h->d_mig_smodel.as<uint8_t>();
  a.s_cnt = h->d_mig_scnt.as<uint8_t>();
  a.gpu_off = h->d_mig_off.as<int32_t>();
  a.N = h->mig_n;
  a.npad = h->mig_npad;
  a.gs = h->mig_gs;
  a.cmax = h->cmax;
  a.k16 = h->k16ok ? 1 : 0;
  a.tg = h->mig_tg;
  a.D = D;
  a.dem_off = off;
  a.dem_key = key;
  a.dem_qty = qty;
  a.first_node = first;
  a.node_base = base;
  return a;
}
MpsScoreArgs mps_base_args(nosplan_t* h, int D, const int32_t* off, const int32_t* key, const int32_t* qty, int32_t* first, int base) {
  MpsScoreArgs a{};
  a.sizes = h->d_mps_sizes.as<int32_t>();
  a.S = h->S;
  a.gp = h->d_mps_gpu0.as<MpsGpu>();
  a.gpu_off = h->d_mps_off.as<int32_t>();
  a.valid = h->d_mps_valid.as<uint8_t>();
  a.N = h->mps_n;
  a.tg = h->mps_tg;
  a.D = D;
  a.dem_off = off;
  a.dem_key = key;
  a.dem_qty = qty;
  a.first_node = first;
  a.node_base = base;
  a.gs = h->mps_gs;
  a.mode = 1;
  a.perm = h->d_mps_perm.as<int32_t>();
  a.rows = h->rp[1].rows.as<int32_t>();
  a.work = h->rp[1].work.as<int32_t>();
  a.meta = h->rp[1].meta.as<int32_t>();
  return a;
}

// CUDA events around a dominant scorer launch (nosplan_kernel_timing): a pool of pairs, summed on request
cudaEvent_t kev_next(nosplan_t* h, int kind, int which) {
  if (!h->kernel_timing) return nullptr;
  const int slot = h->kev_n[kind] % nosplan::kKevPool;
  cudaEvent_t& e = h->kev[kind][slot][which];
  if (!e && cudaEventCreate(&e) != cudaSuccess) return nullptr;
  if (which == 1) ++h->kev_n[kind];
  return e;
}

template <int TN>
void (*both_kernel_tn(int cw, bool cnt))(MigScore16Args, MpsScoreArgs, int, int, int, int) {
  switch (cw) {
    case 1: return cnt ? score_both_kernel<TN, 1, true> : score_both_kernel<TN, 1, false>;
    case 2: return cnt ? score_both_kernel<TN, 2, true> : score_both_kernel<TN, 2, false>;
    case 3: return cnt ? score_both_kernel<TN, 3, true> : score_both_kernel<TN, 3, false>;
    default: return cnt ? score_both_kernel<TN, 4, true> : score_both_kernel<TN, 4, false>;
  }
}
void (*both_kernel_of(int tn, int cw, bool cnt))(MigScore16Args, MpsScoreArgs, int, int, int, int) {
  return tn == 32 ? both_kernel_tn<32>(cw, cnt) : tn == 64 ? both_kernel_tn<64>(cw, cnt) : both_kernel_tn<128>(cw, cnt);
}

// x = the handle that owns the exchange (the caller's), h = the handle that owns the snapshots
int score_pass_enqueue(nosplan_t* h, nosplan_t* x, int Dm, const int32_t* m_off, const int32_t* m_key, const int32_t* m_qty, int Dp,
                       const int32_t* p_off, const int32_t* p_key, const int32_t* p_qty, int32_t* first_node, int flags,
                       cudaStream_t st) {
  NSP_RANGE("nosplan: scoring pass (group, score, exchange)");
  const bool have_m = Dm > 0 && h->have_mig, have_p = Dp > 0 && h->have_mps;
  int32_t* first_m = first_node;
  int32_t* first_p = first_node + Dm;
  const int Dall = Dm + Dp;
  if (Dall == 0) return NOSPLAN_OK;
  auto& xc = x->xch;
  const bool exchange = xc.open;
  if (exchange && xc.broken) return fail(h, NOSPLAN_ERR_CUDA, "the exchange timed out earlier: destroy and re-create it on every rank");
  if (exchange && Dall > xc.max_keys) return fail(h, NOSPLAN_ERR_RANGE, "%d rows exceed the exchange buffer (%d keys)", Dall, xc.max_keys);
  const bool single = (flags & NOSPLAN_PASS_SINGLE_KEY) != 0;
  const bool count = (flags & NOSPLAN_PASS_COUNT) != 0;
  const bool ord_m = !have_m || mig_ordered(h, Dm), ord_p = !have_p || mps_ordered(h, Dp);
  // fused: the scorers fold their winners into the rank's own half of the exchange buffer and signal the step themselves
  const bool fused = exchange && single && ord_m && ord_p && (Dm == 0 || have_m) && (Dp == 0 || have_p);
  // rows of a kind without a snapshot have no node to offer
  if (Dm > 0 && !have_m && !fused) fill_i32_kernel<<<(Dm + 255) / 256, 256, 0, st>>>(first_m, Dm, -1);
  if (Dp > 0 && !have_p && !fused) fill_i32_kernel<<<(Dp + 255) / 256, 256, 0, st>>>(first_p, Dp, -1);
  int rc;
  if (!ord_m || !ord_p || (exchange && !fused)) {
    // general tables / multi-key rows under an exchange: score into a local buffer, then the two exchange kernels
    int32_t* loc = first_node;
    if (exchange) {
      CU(h->d_passbuf.reserve(sizeof(int32_t) * (size_t)Dall));
      loc = h->d_passbuf.as<int32_t>();
      if (Dm > 0 && !have_m) fill_i32_kernel<<<(Dm + 255) / 256, 256, 0, st>>>(loc, Dm, -1);
      if (Dp > 0 && !have_p) fill_i32_kernel<<<(Dp + 255) / 256, 256, 0, st>>>(loc + Dm, Dp, -1);
    }
    if (have_m && (rc = launch_mig_score(h, Dm, m_off, m_key, m_qty, loc, nullptr, nullptr, st, h->mig_base))) return rc;
    if (have_p && (rc = launch_mps_score(h, Dp, p_off, p_key, p_qty, loc + Dm, nullptr, nullptr, nullptr, st, h->mps_base))) return rc;
    if (exchange) {
      const uint32_t step = ++xc.step;
      const int blocks = std::max(1, std::min((Dall + 255) / 256, h->sm_count));
      xchg_publish_kernel<<<blocks, 256, 0, st>>>(xc.peers, xc.world, xc.rank, xc.stride, step, Dall, loc, xc.done.as<unsigned int>(),
                                                  kXchgTimeoutNs, xc.err.as<int>());
      xchg_finish_kernel<<<blocks, 256, 0, st>>>(xc.peers, xc.world, xc.rank, xc.stride, step, Dall, first_node, xc.done.as<unsigned int>(),
                                                 kXchgTimeoutNs, xc.err.as<int>());
    }
    CU(cudaGetLastError());
    return NOSPLAN_OK;
  }
  // ---- both kinds on the ordered kernels: one grouping launch, the two scorers side by side ----
  Mig16Launch L;
  int mw_m = 0, mw_p = 0, chunk_p = 128, tiles_p = 0;
  // both kinds, single-key rows, nobody asked for the kinds one after the other: ONE launch scores both
  const bool both = have_m && have_p && single && !(flags & NOSPLAN_PASS_SERIAL);
  if (have_m) {
    if ((rc = mig16_prepare(h, Dm, false, count, &L, both))) return rc;
    if ((rc = reserve_rows(h, 0, Dm, h->P, L.chunk, st, &mw_m))) return rc;
  }
  size_t smem_p = 0;
  if (have_p) {
    tiles_p = (h->mps_n + kScoreTN - 1) / kScoreTN;
    int gy;
    pick_chunks(h->sm_count, tiles_p, Dp, mps_min_chunk(), &chunk_p, &gy, 4 * score_waves(both));
    chunk_p = std::min(chunk_p, kRowCap);
    smem_p = mps_score1_smem_bytes(h->mps_gs, h->S, kScoreTN);
    if (smem_p > 200 * 1024) return fail(h, NOSPLAN_ERR_RANGE, "MPS scorer tile needs %zu bytes of shared memory", smem_p);
    if ((rc = reserve_rows(h, 1, Dp, h->S, chunk_p, st, &mw_p))) return rc;
  }
  RowsGroupArgs g0{}, g1{};
  int nb0 = 0, nb1 = 0;
  uint32_t step = 0;
  if (fused) {
    step = ++xc.step;
    int32_t* own = xc.peers.keys[xc.rank] + (size_t)(step & 1u) * xc.stride;
    first_m = own;
    first_p = own + Dm;
  }
  if (have_m) {
    g0 = group_args(h, 0, Dm, h->P, L.chunk, m_off, m_key, m_qty, h->d_profcost.as<int32_t>(), first_m, single);
    nb0 = (Dm + 255) / 256;
  }
  if (have_p) {
    g1 = group_args(h, 1, Dp, h->S, chunk_p, p_off, p_key, p_qty, h->d_mps_sizecost.as<int32_t>(), first_p, single);
    nb1 = (Dp + 255) / 256;
  }
  if (fused && step > 2u) {  // the half may be overwritten once every rank has read what it held two steps ago
    for (RowsGroupArgs* g : {&g0, &g1}) {
      g->wait_read = xc.peers.read[xc.rank];
      g->wait_world = xc.world;
      g->wait_need = step - 2u;
      g->wait_err = xc.err.as<int>();
    }
  }
  if (nb0 + nb1 > 0) rows_group2_kernel<<<nb0 + nb1, 256, 0, st>>>(g0, g1, nb0);
  CU(cudaGetLastError());
  PushArgs push{};
  if (fused) {
    push.world = xc.world;
    push.rank = xc.rank;
    push.step = step;
    push.total = (have_m ? (unsigned)mw_m * (unsigned)L.tiles : 0u) + (have_p ? (unsigned)mw_p * (unsigned)tiles_p : 0u);
    push.ticket = xc.done.as<unsigned int>();
    for (int r = 0; r < xc.world; ++r) push.flags[r] = xc.peers.pushed[r];
  }
  cudaStream_t st_p = st;
  if (both) {
    MigScore16Args A = mig16_args(h, mig_base_args(h, Dm, m_off, m_key, m_qty, first_m, h->mig_base));
    MpsScoreArgs a = mps_base_args(h, Dp, p_off, p_key, p_qty, first_p, h->mps_base);
    A.push = a.push = push;
    A.counter = count ? h->d_counter.as<unsigned long long>() : nullptr;
    a.counter = count ? h->d_counter.as<unsigned long long>() + 1 : nullptr;
    const int cw = std::max(1, std::min(4, (h->cmax + 3) / 4));
    auto kern = both_kernel_of(L.tn, cw, count);
    const size_t smem = std::max(L.smem, smem_p);
    CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int n_mig = mw_m * L.tiles, n_mps = mw_p * tiles_p;
    if (cudaEvent_t e = kev_next(h, 0, 0)) CU(cudaEventRecord(e, st));
    static const int both_order = [] { const char* e = std::getenv("NOSPLAN_BOTH_ORDER"); return e ? std::atoi(e) : 0; }();
    kern<<<n_mig + n_mps, 512, smem, st>>>(A, a, mw_m, n_mig, mw_p, both_order == 1 ? n_mps : 0);
    if (cudaEvent_t e = kev_next(h, 0, 1)) CU(cudaEventRecord(e, st));
    CU(cudaGetLastError());
  }
  if (have_m && !both) {
    MigScore16Args A = mig16_args(h, mig_base_args(h, Dm, m_off, m_key, m_qty, first_m, h->mig_base));
    A.push = push;
    A.counter = count ? h->d_counter.as<unsigned long long>() : nullptr;
    if (cudaEvent_t e = kev_next(h, 0, 0)) CU(cudaEventRecord(e, st));
    L.kern<<<dim3(mw_m, L.tiles), 512, L.smem, st>>>(A);
    if (cudaEvent_t e = kev_next(h, 0, 1)) CU(cudaEventRecord(e, st));
    if (!single) {  // rows with several keys: the general kernel restricted to their list (leaves at once when there are none)
      MigScoreArgs a = A.base;
      a.row_list = h->rp[0].rows.as<int32_t>();
      a.row_meta = h->rp[0].meta.as<int32_t>();
      int gy;
      const int tiles128 = std::min((h->mig_n + kScoreTN - 1) / kScoreTN, 4 * h->sm_count);  // the kernel strides over the rest
      pick_chunks(h->sm_count, tiles128, Dm, 16, &a.d_chunk, &gy);
      const size_t smem = mig_score_smem(h, kScoreTN);
      if (smem > 200 * 1024) return fail(h, NOSPLAN_ERR_RANGE, "MIG scorer tile needs %zu bytes of shared memory", smem);
      if ((rc = launch_mig_score_kernel(h, a, dim3(tiles128, gy), smem, st))) return rc;
    }
  }
  if (have_p && !both) {
    MpsScoreArgs a = mps_base_args(h, Dp, p_off, p_key, p_qty, first_p, h->mps_base);
    a.push = push;
    a.counter = count ? h->d_counter.as<unsigned long long>() + 1 : nullptr;
    auto kern = count ? mps_score1_kernel<kScoreTN, kScoreDS, false, true> : mps_score1_kernel<kScoreTN, kScoreDS, false, false>;
    CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_p));
    if (cudaEvent_t e = kev_next(h, 1, 0)) CU(cudaEventRecord(e, st_p));
    kern<<<dim3(mw_p, tiles_p), kScoreTN * kScoreDS, smem_p, st_p>>>(a);
    if (cudaEvent_t e = kev_next(h, 1, 1)) CU(cudaEventRecord(e, st_p));
    if (!single) {
      a.mode = 2;
      int gy;
      pick_chunks(h->sm_count, tiles_p, Dp, 16, &a.d_chunk, &gy, 3 * score_waves());
      const size_t smem = mps_score_smem_bytes(h->mps_gs, h->S, kScoreTN);
      if (smem > 200 * 1024) return fail(h, NOSPLAN_ERR_RANGE, "MPS scorer tile needs %zu bytes of shared memory", smem);
      CU(cudaFuncSetAttribute(mps_score_kernel<kScoreTN, kScoreDS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      mps_score_kernel<kScoreTN, kScoreDS><<<dim3(tiles_p, gy), kScoreTN * kScoreDS, smem, st_p>>>(a);
    }
  }
  CU(cudaGetLastError());
  if (fused) {
    const int blocks = std::max(1, std::min((Dall / 4 + 255) / 256, h->sm_count));
    xchg_finish_kernel<<<blocks, 256, 0, st>>>(xc.peers, xc.world, xc.rank, xc.stride, step, Dall, first_node, xc.done.as<unsigned int>(),
                                               kXchgTimeoutNs, xc.err.as<int>());
    CU(cudaGetLastError());
  }
  return NOSPLAN_OK;
}

}  // namespace

extern "C" {

int nosplan_score_pass_dev(nosplan_t* h, int n_mig_rows, const int32_t* mig_off, const int32_t* mig_key, const int32_t* mig_qty,
                           int n_mps_rows, const int32_t* mps_off, const int32_t* mps_key, const int32_t* mps_qty,
                           int32_t* first_node, int flags, void* stream) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  if (n_mig_rows < 0 || n_mps_rows < 0) return fail(h, NOSPLAN_ERR_ARG, "negative row count");
  if (n_mig_rows + n_mps_rows > 0 && !first_node) return fail(h, NOSPLAN_ERR_ARG, "first_node is required");
  if ((n_mig_rows > 0 && (!mig_off || !mig_key || !mig_qty)) || (n_mps_rows > 0 && (!mps_off || !mps_key || !mps_qty)))
    return fail(h, NOSPLAN_ERR_ARG, "null demand arrays");
  int rc = bind(h);
  if (rc) return rc;
  if ((flags & NOSPLAN_PASS_COUNT) && !h->d_counter.p) {
    CU(h->d_counter.reserve(16));
    CU(cudaMemset(h->d_counter.p, 0, 16));
  }
  return score_pass_enqueue(h, h, n_mig_rows, mig_off, mig_key, mig_qty, n_mps_rows, mps_off, mps_key, mps_qty, first_node, flags,
                            (cudaStream_t)stream);
}

int nosplan_pass_status(nosplan_t* h) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  int rc = bind(h);
  if (rc) return rc;
  CU(cudaDeviceSynchronize());
  if (h->d_rowstatus.p) {
    unsigned long long st[2] = {0, 0};
    CU(cudaMemcpy(st, h->d_rowstatus.p, 16, cudaMemcpyDeviceToHost));
    for (int kind = 0; kind < 2; ++kind)
      if ((st[kind] = row_status_code(h, kind, st[kind])) != ~0ull) {
        const int sub = (int)(st[kind] % 64);
        return fail(h, sub == 0 || sub == kSubNotSingle ? NOSPLAN_ERR_ARG : NOSPLAN_ERR_RANGE, "%s demand row %llu rejected on the device (%s)",
                    kind ? "MPS" : "MIG", st[kind] / 64,
                    sub == 0 ? "dem_off not ascending" : sub == 1 ? "too many keys" : sub == kSubNotSingle ? "NOSPLAN_PASS_SINGLE_KEY set but the row does not have exactly one key"
                    : (sub - 2) % 3 == 0 ? "key out of range" : (sub - 2) % 3 == 1 ? "keys not ascending" : "quantity out of range");
      }
  }
  if (h->xch.created) {
    int e = 0;
    CU(cudaMemcpy(&e, h->xch.err.p, sizeof(int), cudaMemcpyDeviceToHost));
    if (e) {
      h->xch.broken = true;
      return fail(h, NOSPLAN_ERR_CUDA, "exchange: rank %d never signalled a step (waited 2 s); the results of that pass read -2", e - 1);
    }
  }
  return NOSPLAN_OK;
}

int nosplan_kernel_timing(nosplan_t* h, int enable) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  h->kernel_timing = enable != 0;
  h->kev_n[0] = h->kev_n[1] = 0;
  return NOSPLAN_OK;
}

int nosplan_kernel_time_ms(nosplan_t* h, int kind, double* total_ms, int* launches) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  if ((kind != 0 && kind != 1) || !total_ms || !launches) return fail(h, NOSPLAN_ERR_ARG, "bad argument");
  int rc = bind(h);
  if (rc) return rc;
  CU(cudaDeviceSynchronize());
  const int n = std::min(h->kev_n[kind], (int)nosplan::kKevPool);
  double tot = 0;
  for (int i = 0; i < n; ++i) {
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, h->kev[kind][i][0], h->kev[kind][i][1]));
    tot += ms;
  }
  *total_ms = tot;
  *launches = n;
  h->kev_n[kind] = 0;
  return NOSPLAN_OK;
}

int nosplan_executed_evals(nosplan_t* h, int64_t* mig, int64_t* mps) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  int rc = bind(h);
  if (rc) return rc;
  unsigned long long c[2] = {0, 0};
  if (h->d_counter.p) {
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpy(c, h->d_counter.p, 16, cudaMemcpyDeviceToHost));
    CU(cudaMemset(h->d_counter.p, 0, 16));
  }
  if (mig) *mig = (int64_t)c[0];
  if (mps) *mps = (int64_t)c[1];
  return NOSPLAN_OK;
}

int nosplan_score_mig(nosplan_t* h, int n_demands, const int32_t* dem_off, const int32_t* dem_key,
                      const int32_t* dem_qty, int32_t* first_node, int8_t* first_choice, uint8_t* updated, int8_t* choice,
                      int64_t* n_evals) {
  if (!h) return NOSPLAN_ERR_ARG;
  NSP_RANGE("nosplan_score_mig");
  h->err.clear();
  if (!h->have_mig) return fail(h, NOSPLAN_ERR_ARG, "no MIG snapshot");
  // with the rows-by-key scorer the rows are validated on the device while they are grouped
  const bool dev_check = mig_ordered(h, n_demands);
  int rc = dev_check ? check_demands_head(h, n_demands, dem_off, dem_key, dem_qty)
                     : check_demands(h, n_demands, dem_off, dem_key, dem_qty, h->P);
  if (rc) return rc;
  if ((rc = bind(h))) return rc;
  if (n_evals) *n_evals = (int64_t)n_demands * h->mig_evals;
  const int D = n_demands;
  if (D == 0) return NOSPLAN_OK;
  cudaStream_t st = h->stream;
  const int nnz = dem_off[D];
  if ((rc = upload(h, h->d_dem[0], dem_off, sizeof(int32_t) * (D + 1), st))) return rc;
  if ((rc = upload(h, h->d_dem[1], dem_key, sizeof(int32_t) * std::max(nnz, 1), st))) return rc;
  if ((rc = upload(h, h->d_dem[2], dem_qty, sizeof(int32_t) * std::max(nnz, 1), st))) return rc;
  CU(h->d_out[0].reserve(sizeof(int32_t) * D));
  if (updated) CU(h->d_out[1].reserve((size_t)D * std::max(h->mig_n, 1)));
  if (choice) CU(h->d_out[2].reserve((size_t)D * std::max(h->mig_tg, 1)));
  if (first_choice) CU(h->d_out[3].reserve((size_t)D * h->mig_gs));
  rc = launch_mig_score(h, D, h->d_dem[0].as<int32_t>(), h->d_dem[1].as<int32_t>(), h->d_dem[2].as<int32_t>(),
                        h->d_out[0].as<int32_t>(), updated ? h->d_out[1].as<uint8_t>() : nullptr,
                        choice ? h->d_out[2].as<int8_t>() : nullptr, st);
  if (rc) return rc;
  if (first_choice && dev_check) {  // 16-bit-key tables: re-walk the winners with the snapshot-time ranks
    mig_finalize16_kernel<<<(D + 127) / 128, 128, 0, st>>>(
        h->d_tables.as<MigTables>(), h->d_mig_rankg.as<uint32_t>(), h->d_mig_ufg.as<uint16_t>(), h->d_mig_used0.as<uint4>(),
        h->d_mig_free0.as<uint4>(), h->d_mig_model.as<uint8_t>(), h->d_mig_off.as<int32_t>(), D, h->mig_gs, (h->cmax + 3) / 4,
        h->d_dem[0].as<int32_t>(), h->d_dem[1].as<int32_t>(), h->d_dem[2].as<int32_t>(), h->d_out[0].as<int32_t>(),
        h->d_out[3].as<int8_t>(), 0, h->mig_n, h->mig_npad);
  } else if (first_choice) {
    mig_finalize_kernel<<<(D + 127) / 128, 128, 0, st>>>(
        h->d_tables.as<MigTables>(), h->d_mig_used0.as<uint4>(), h->d_mig_free0.as<uint4>(), h->d_mig_model.as<uint8_t>(),
        h->d_mig_off.as<int32_t>(), D, h->mig_gs, h->d_dem[0].as<int32_t>(), h->d_dem[1].as<int32_t>(),
        h->d_dem[2].as<int32_t>(), h->d_out[0].as<int32_t>(), h->d_out[3].as<int8_t>(), 0, h->mig_n);
  }
  CU(cudaGetLastError());
  if (first_node) CU(cudaMemcpyAsync(first_node, h->d_out[0].p, sizeof(int32_t) * D, cudaMemcpyDeviceToHost, st));
  if (first_choice) CU(cudaMemcpyAsync(first_choice, h->d_out[3].p, (size_t)D * h->mig_gs, cudaMemcpyDeviceToHost, st));
  if (updated && h->mig_n) CU(cudaMemcpyAsync(updated, h->d_out[1].p, (size_t)D * h->mig_n, cudaMemcpyDeviceToHost, st));
  if (choice && h->mig_tg) CU(cudaMemcpyAsync(choice, h->d_out[2].p, (size_t)D * h->mig_tg, cudaMemcpyDeviceToHost, st));
  if (dev_check) CU(cudaMemcpyAsync(h->h_status + 4, h->d_rowstatus.p, 8, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  if (dev_check && row_status_code(h, 0, h->h_status[4]) != ~0ull)
    return describe_bad_row(h, row_status_code(h, 0, h->h_status[4]), dem_off, dem_key, dem_qty, h->P);
  return NOSPLAN_OK;
}

int nosplan_score_mps(nosplan_t* h, int n_demands, const int32_t* dem_off, const int32_t* dem_key,
                      const int32_t* dem_qty, int32_t* first_node, uint8_t* updated, uint8_t* out_free,
                      uint32_t* out_free_mask, int64_t* n_evals) {
  if (!h) return NOSPLAN_ERR_ARG;
  NSP_RANGE("nosplan_score_mps");
  h->err.clear();
  if (!h->have_mps) return fail(h, NOSPLAN_ERR_ARG, "no MPS snapshot");
  if ((out_free == nullptr) != (out_free_mask == nullptr)) return fail(h, NOSPLAN_ERR_ARG, "out_free and out_free_mask go together");
  const bool dev_check = mps_ordered(h, n_demands);
  int rc = dev_check ? check_demands_head(h, n_demands, dem_off, dem_key, dem_qty)
                     : check_demands(h, n_demands, dem_off, dem_key, dem_qty, h->S);
  if (rc) return rc;
  if ((rc = bind(h))) return rc;
  if (n_evals) *n_evals = (int64_t)n_demands * h->mps_evals;
  const int D = n_demands;
  if (D == 0) return NOSPLAN_OK;
  cudaStream_t st = h->stream;
  const int nnz = dem_off[D];
  if ((rc = upload(h, h->d_dem[0], dem_off, sizeof(int32_t) * (D + 1), st))) return rc;
  if ((rc = upload(h, h->d_dem[1], dem_key, sizeof(int32_t) * std::max(nnz, 1), st))) return rc;
  if ((rc = upload(h, h->d_dem[2], dem_qty, sizeof(int32_t) * std::max(nnz, 1), st))) return rc;
  CU(h->d_out[0].reserve(sizeof(int32_t) * D));
  if (updated) CU(h->d_out[1].reserve((size_t)D * std::max(h->mps_n, 1)));
  if (out_free) {
    CU(h->d_out[2].reserve((size_t)D * std::max(h->mps_tg, 1) * h->S));
    CU(h->d_out[3].reserve(sizeof(uint32_t) * (size_t)D * std::max(h->mps_tg, 1)));
  }
  rc = launch_mps_score(h, D, h->d_dem[0].as<int32_t>(), h->d_dem[1].as<int32_t>(), h->d_dem[2].as<int32_t>(),
                        h->d_out[0].as<int32_t>(), updated ? h->d_out[1].as<uint8_t>() : nullptr,
                        out_free ? h->d_out[2].as<uint8_t>() : nullptr, out_free ? h->d_out[3].as<uint32_t>() : nullptr, st);
  if (rc) return rc;
  if (first_node) CU(cudaMemcpyAsync(first_node, h->d_out[0].p, sizeof(int32_t) * D, cudaMemcpyDeviceToHost, st));
  if (updated && h->mps_n) CU(cudaMemcpyAsync(updated, h->d_out[1].p, (size_t)D * h->mps_n, cudaMemcpyDeviceToHost, st));
  if (out_free && h->mps_tg) {
    CU(cudaMemcpyAsync(out_free, h->d_out[2].p, (size_t)D * h->mps_tg * h->S, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(out_free_mask, h->d_out[3].p, sizeof(uint32_t) * (size_t)D * h->mps_tg, cudaMemcpyDeviceToHost, st));
  }
  if (dev_check) CU(cudaMemcpyAsync(h->h_status + 5, h->d_rowstatus.as<unsigned long long>() + 1, 8, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  if (dev_check && row_status_code(h, 1, h->h_status[5]) != ~0ull)
    return describe_bad_row(h, row_status_code(h, 1, h->h_status[5]), dem_off, dem_key, dem_qty, h->S);
  return NOSPLAN_OK;
}

}  // extern "C"

namespace {
constexpr int kMaxHybridRounds = 64;
// one kind's share of a nosplan_plan call
struct PlanKind {
  int kd = 0, M = 0, maxq = 1;
  bool hybrid = false;
  cudaStream_t st = nullptr;
  PlanArgs pa;
  ChainArgs ca;
  std::function<int(int, int, int32_t*, int32_t*)> run_seq;
  // the chains over the pods from `first` up to the next pod the gate does not stop; that pod's index comes back in
  // h->h_limit[kd] (pinned) once the stream has drained
  int chains_from(nosplan_t* h, int first) {
    const int32_t range[2] = {first, M};
    h->h_range[kd][0] = range[0];
    h->h_range[kd][1] = range[1];
    CU(cudaMemcpyAsync(ca.limit_out, h->h_range[kd], sizeof(range), cudaMemcpyHostToDevice, st));
    if (ca.sched) {
      if (kd == 0) plan_winmax_kernel<false><<<ca.n_win, 32, 0, st>>>(ca);
      else plan_winmax_kernel<true><<<ca.n_win, 32, 0, st>>>(ca);
    }
    plan_find_miss_kernel<<<1, 1024, 0, st>>>(ca);
    if (maxq <= 2) {
      if (kd == 0) plan_chain_kernel<false, 2><<<ca.n_chains, 32, 0, st>>>(ca);
      else plan_chain_kernel<true, 2><<<ca.n_chains, 32, 0, st>>>(ca);
    } else {
      if (kd == 0) plan_chain_kernel<false, kChainQ><<<ca.n_chains, 32, 0, st>>>(ca);
      else plan_chain_kernel<true, kChainQ><<<ca.n_chains, 32, 0, st>>>(ca);
    }
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(&h->h_limit[kd], ca.limit_out + 1, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    return NOSPLAN_OK;
  }
};
}  // namespace

namespace {
int pin_reserve(nosplan_t* h, size_t bytes);  // defined with nosplan_score_snapshot
}

extern "C" {

int nosplan_plan(nosplan_t* h, int n_pods, const uint8_t* kind, const int32_t* dem_off, const int32_t* dem_key,
                 const int32_t* dem_qty, int flags, uint8_t* gate_hit, int32_t* repart_node, int32_t* sched_node,
                 int32_t* sched_gpu) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  if (n_pods < 0) return fail(h, NOSPLAN_ERR_ARG, "negative pod count");
  if (n_pods > 0 && (!kind || !dem_off)) return fail(h, NOSPLAN_ERR_ARG, "null pod arrays");
  int rc = bind(h);
  if (rc) return rc;
  const bool trace = trace_on();
  const auto t_begin = std::chrono::steady_clock::now();
  NSP_RANGE("nosplan_plan");
  nvtxRangePushA("nosplan_plan: host split of the pod list");
  // split by kind (order kept) and deduplicate request vectors: pods with the same vector share a row
  using Side = nosplan::PlanSide;
  Side* const side = h->plan_side;
  int n_side[2] = {0, 0};  // pods of each kind so far (the arrays are sized for all of them and cut afterwards)
  for (int kd = 0; kd < 2; ++kd) {
    Side& s = side[kd];
    s.pod_idx.resize(n_pods);
    s.pod_sig.resize(n_pods);
    s.off.assign(1, 0);
    s.key.clear();
    s.qty.clear();
    s.sigs.clear();
    s.single.assign((size_t)(kd ? NOSPLAN_MAX_MPS_SIZES : NOSPLAN_MAX_MIG_PROFILES) * (kChainQ + 1), -1);
    s.n_sigs = 0;
    s.chains_ok = true;
  }
  // (tables indexed by the kind, not branches on it: the kinds alternate at random and a branch on them mispredicts)
  const bool have_kind[2] = {h->have_mig, h->have_mps};
  const int keys_of_kind[2] = {h->P, h->S};
  for (int i = 0; i < n_pods; ++i) {
    const int kd = kind[i] ? 1 : 0;
    const int n_keys = keys_of_kind[kd];
    if (!have_kind[kd]) {
      // no nodes of that kind: the controller returns before doing anything (mig_controller.go:78-81)
      if (gate_hit) gate_hit[i] = 0;
      if (repart_node) repart_node[i] = -1;
      if (sched_node) sched_node[i] = -1;
      if (sched_gpu) sched_gpu[i] = -1;
      continue;
    }
    const int e0 = dem_off[i], e1 = dem_off[i + 1];
    if (e1 < e0 || e1 - e0 > NOSPLAN_MAX_DEMAND_KEYS) return fail(h, NOSPLAN_ERR_RANGE, "pod %d: bad demand row", i);
    for (int e = e0; e < e1; ++e) {
      if (dem_key[e] < 0 || dem_key[e] >= n_keys) return fail(h, NOSPLAN_ERR_RANGE, "pod %d: key %d out of range", i, dem_key[e]);
      if (e > e0 && dem_key[e] <= dem_key[e - 1]) return fail(h, NOSPLAN_ERR_ARG, "pod %d: keys must be strictly ascending", i);
      if (dem_qty[e] < 0 || dem_qty[e] > NOSPLAN_MAX_DEMAND_QTY) return fail(h, NOSPLAN_ERR_RANGE, "pod %d: quantity out of range", i);
    }
    Side& s = side[kd];
    int id = -1;
    int* slot = nullptr;
    if (e1 - e0 == 1 && dem_qty[e0] >= 1 && dem_qty[e0] <= kChainQ) {
      slot = &s.single[(size_t)dem_key[e0] * (kChainQ + 1) + dem_qty[e0]];
      id = *slot;
    } else {
      s.chains_ok = false;
      std::vector<int32_t> sig;
      for (int e = e0; e < e1; ++e) {
        sig.push_back(dem_key[e]);
        sig.push_back(dem_qty[e]);
      }
      auto it = s.sigs.find(sig);
      if (it != s.sigs.end()) id = it->second;
      else s.sigs.emplace(sig, s.n_sigs);
    }
    if (id < 0) {
      id = s.n_sigs++;
      if (slot) *slot = id;
      for (int e = e0; e < e1; ++e) {
        s.key.push_back(dem_key[e]);
        s.qty.push_back(dem_qty[e]);
      }
      s.off.push_back((int32_t)s.key.size());
    }
    const int c = n_side[kd]++;
    s.pod_idx[c] = i;
    s.pod_sig[c] = id;
  }
  for (int kd = 0; kd < 2; ++kd) {
    side[kd].pod_idx.resize(n_side[kd]);
    side[kd].pod_sig.resize(n_side[kd]);
  }
  const auto t_split = std::chrono::steady_clock::now();
  nvtxRangePop();
  NvtxRange enqueue_range("nosplan_plan: chains / sequencer");
  // reset the working state to the snapshot
  if (h->have_mig && h->mig_tg) {
    CU(cudaMemcpyAsync(h->d_mig_used.p, h->d_mig_used0.p, sizeof(uint4) * h->mig_tg, cudaMemcpyDeviceToDevice, h->stream));
    CU(cudaMemcpyAsync(h->d_mig_free.p, h->d_mig_free0.p, sizeof(uint4) * h->mig_tg, cudaMemcpyDeviceToDevice, h->stream));
  }
  if (h->have_mps && h->mps_tg) {
    CU(cudaMemcpyAsync(h->d_mps_gpu.p, h->d_mps_gpu0.p, sizeof(MpsGpu) * h->mps_tg, cudaMemcpyDeviceToDevice, h->stream2));
    CU(cudaMemcpyAsync(h->d_mps_ulo.p, h->d_mps_ulo0.p, sizeof(uint4) * h->mps_tg, cudaMemcpyDeviceToDevice, h->stream2));
    CU(cudaMemcpyAsync(h->d_mps_uhi.p, h->d_mps_uhi0.p, sizeof(uint4) * h->mps_tg, cudaMemcpyDeviceToDevice, h->stream2));
  }
  const int32_t* out_ptr[2] = {nullptr, nullptr};
  PlanKind pk[2];
  for (int kd = 0; kd < 2; ++kd) {
    Side& s = side[kd];
    const int M = (int)s.pod_idx.size();
    if (M == 0) continue;
    cudaStream_t st = kd ? h->stream2 : h->stream;
    DevBuf* b = h->d_plan + kd * 16;
    const int D = s.n_sigs;
    PlanArgs a;
    std::memset(&a, 0, sizeof(a));
    a.M = M;
    a.D = D;
    a.flags = flags;
    if ((rc = upload(h, b[0], s.pod_sig.data(), sizeof(int32_t) * M, st))) return rc;
    if ((rc = upload(h, b[1], s.off.data(), sizeof(int32_t) * (D + 1), st))) return rc;
    if (s.key.empty()) {  // only empty request vectors: keep the device arrays non-null
      s.key.push_back(0);
      s.qty.push_back(0);
    }
    if ((rc = upload(h, b[2], s.key.data(), sizeof(int32_t) * s.key.size(), st))) return rc;
    if ((rc = upload(h, b[3], s.qty.data(), sizeof(int32_t) * s.qty.size(), st))) return rc;
    a.pod_sig = b[0].as<int32_t>();
    a.dem_off = b[1].as<int32_t>();
    a.dem_key = b[2].as<int32_t>();
    a.dem_qty = b[3].as<int32_t>();
    if (kd == 0) {
      a.N = h->mig_n;
      a.tg = h->mig_tg;
      a.n_keys = h->P;
      a.gpu_off = h->d_mig_off.as<int32_t>();
      a.valid = h->d_mig_valid.as<uint8_t>();
      a.T = h->d_tables.as<MigTables>();
      a.m_used = h->d_mig_used.as<uint4>();
      a.m_free = h->d_mig_free.as<uint4>();
      a.m_model = h->d_mig_model.as<uint8_t>();
      h->mig_state_dirty = true;
    } else {
      a.N = h->mps_n;
      a.tg = h->mps_tg;
      a.n_keys = h->S;
      a.gpu_off = h->d_mps_off.as<int32_t>();
      a.valid = h->d_mps_valid.as<uint8_t>();
      a.sizes = h->d_mps_sizes.as<int32_t>();
      a.p_gpu = h->d_mps_gpu.as<MpsGpu>();
      a.p_used_lo = h->d_mps_ulo.as<uint4>();
      a.p_used_hi = h->d_mps_uhi.as<uint4>();
      h->mps_state_dirty = true;
    }
    a.words = (a.tg + 31) / 32 + 1;
    CU(b[4].reserve(sizeof(uint32_t) * (size_t)a.words * (a.n_keys + 1)));
    CU(b[5].reserve(sizeof(int32_t) * 2 * (size_t)D));
    CU(b[6].reserve((size_t)M * (1 + 3 * sizeof(int32_t)) + 64));
    a.bitmaps = b[4].as<uint32_t>();
    a.bounds = b[5].as<int32_t>();
    // outputs packed in one buffer: rep | node | gpu (int32 each) | gate (u8)
    a.repart_node = b[6].as<int32_t>();
    a.sched_node = a.repart_node + M;
    a.sched_gpu = a.sched_node + M;
    a.gate_hit = reinterpret_cast<uint8_t*>(a.sched_gpu + M);
    const bool want_chains = s.chains_ok && (flags & NOSPLAN_PLAN_GATE) && a.tg > 0;
    PlanKind& K = pk[kd];
    K.M = M;
    K.st = st;
    K.pa = a;
    K.kd = kd;
    // the general path: derive the masks/bitmaps from the working state, then one sequencer warp over pods
    // [first, first + count) of the kind
    K.run_seq = [h, kd, b, &K](int first, int count, int32_t* present_out, int32_t* dead) -> int {
      const PlanArgs& a = K.pa;
      cudaStream_t st = K.st;
      SeqArgs q;
      std::memset(&q, 0, sizeof(q));
      q.M = count; q.pod_sig = a.pod_sig + first; q.D = a.D; q.dem_off = a.dem_off; q.dem_key = a.dem_key; q.dem_qty = a.dem_qty;
      q.flags = a.flags; q.N = a.N; q.tg = a.tg; q.n_keys = a.n_keys; q.gpu_off = a.gpu_off; q.valid = a.valid;
      q.gpu_node = kd ? h->d_mps_gnode.as<int32_t>() : h->d_mig_gnode.as<int32_t>();
      q.T = a.T; q.m_used = a.m_used; q.m_free = a.m_free; q.m_model = a.m_model;
      q.sizes = a.sizes; q.p_gpu = a.p_gpu; q.p_used_lo = a.p_used_lo; q.p_used_hi = a.p_used_hi;
      q.nw = (a.N + 31) / 32 + 1;
      q.gw = (a.tg + 31) / 32 + 1;
      const size_t n_gm = (size_t)std::max(a.tg, 1), n_nb = (size_t)a.n_keys * q.nw, n_fb = (size_t)(a.n_keys + 1) * q.gw;
      const size_t words = 2 * n_gm + 2 * n_nb + n_fb + kMaxSizes;
      CU(b[13].reserve(words * 4));
      q.gm = b[13].as<uint32_t>();
      q.f1 = q.gm + n_gm;
      q.nbt = q.f1 + n_gm;
      q.nu1t = q.nbt + n_nb;
      q.fbt = q.nu1t + n_nb;
      q.present = reinterpret_cast<int32_t*>(q.fbt + n_fb);
      q.gate_hit = a.gate_hit + first; q.repart_node = a.repart_node + first; q.sched_node = a.sched_node + first;
      q.sched_gpu = a.sched_gpu + first;
      q.present_out = present_out;
      q.dead = dead;
      CU(cudaMemsetAsync(q.nbt, 0, (2 * n_nb + n_fb + kMaxSizes) * 4, st));
      const size_t smem = (2 * n_nb + n_fb) * 4;
      q.use_smem = smem <= 160 * 1024 ? 1 : 0;
      if (a.tg > 0) {
        if (kd == 0) seq_prep_gpu_kernel<false><<<(a.tg + 127) / 128, 128, 0, st>>>(q);
        else seq_prep_gpu_kernel<true><<<(a.tg + 127) / 128, 128, 0, st>>>(q);
      }
      if (a.N > 0) seq_prep_node_kernel<<<(a.N + 127) / 128, 128, 0, st>>>(q);
      if (kd == 0) {
        CU(cudaFuncSetAttribute(plan_seq_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        plan_seq_kernel<false><<<1, kSeqThreads, q.use_smem ? smem : 0, st>>>(q);
      } else {
        CU(cudaFuncSetAttribute(plan_seq_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        plan_seq_kernel<true><<<1, kSeqThreads, q.use_smem ? smem : 0, st>>>(q);
      }
      CU(cudaGetLastError());
      return NOSPLAN_OK;
    };
    if (want_chains) {
      // Reference behaviour (gate on), every pod one (profile, quantity): the chains own every pod the gate stops;
      // a pod whose profile is absent ends the chains' range, goes through tryRepartition on the sequencer (one pod),
      // and the chains go on behind it (the host loop below, one round per such pod).
      // pods grouped by requested key, planner order kept inside a group
      std::vector<int32_t> ckey, coff{0}, cpod(M), cq(M);
      std::vector<int> cnt(a.n_keys, 0), chain_of(a.n_keys, -1);
      // every row is one (key, quantity) here: a small per-row table instead of two indirections per pod
      std::vector<int32_t> row_key(D), row_qty(D), row_cnt(D, 0);
      for (int r = 0; r < D; ++r) {
        row_key[r] = s.key[s.off[r]];
        row_qty[r] = s.qty[s.off[r]];
      }
      const int32_t* sig = s.pod_sig.data();
      for (int j = 0; j < M; ++j) row_cnt[sig[j]]++;
      for (int r = 0; r < D; ++r) cnt[row_key[r]] += row_cnt[r];
      for (int k = 0; k < a.n_keys; ++k)
        if (cnt[k]) {
          chain_of[k] = (int)ckey.size();
          ckey.push_back(k);
          coff.push_back(coff.back() + cnt[k]);
        }
      std::vector<int32_t> fill(coff.begin(), coff.end() - 1), row_chain(D);
      for (int r = 0; r < D; ++r) row_chain[r] = chain_of[row_key[r]];
      for (int j = 0; j < M; ++j) {
        const int r = sig[j];
        const int at = fill[row_chain[r]]++;
        cpod[at] = j;
        cq[at] = row_qty[r];
      }
      ChainArgs& ca = K.ca;
      std::memset(&ca, 0, sizeof(ca));
      ca.n_chains = (int)ckey.size();
      if ((rc = upload(h, b[8], ckey.data(), sizeof(int32_t) * ckey.size(), st))) return rc;
      if ((rc = upload(h, b[9], coff.data(), sizeof(int32_t) * coff.size(), st))) return rc;
      if ((rc = upload(h, b[10], cpod.data(), sizeof(int32_t) * M, st))) return rc;
      if ((rc = upload(h, b[11], cq.data(), sizeof(int32_t) * M, st))) return rc;
      // control block: present[kMaxSizes] | dead[kMaxSizes] | range[2]
      CU(b[12].reserve(sizeof(int32_t) * (2 * kMaxSizes + 2)));
      CU(cudaMemsetAsync(b[12].p, 0, sizeof(int32_t) * (2 * kMaxSizes + 2), st));
      ca.chain_key = b[8].as<int32_t>();
      ca.chain_off = b[9].as<int32_t>();
      ca.chain_pod = b[10].as<int32_t>();
      ca.chain_q = b[11].as<int32_t>();
      ca.present = b[12].as<int32_t>();
      ca.dead = ca.present + kMaxSizes;
      ca.limit_out = ca.present + 2 * kMaxSizes;
      ca.range = ca.limit_out;
      ca.pod_sig = a.pod_sig;
      ca.dem_off = a.dem_off;
      ca.dem_key = a.dem_key;
      ca.M = M;
      ca.N = a.N;
      ca.tg = a.tg;
      ca.n_keys = a.n_keys;
      ca.gpu_off = a.gpu_off;
      ca.gpu_node = kd ? h->d_mps_gnode.as<int32_t>() : h->d_mig_gnode.as<int32_t>();
      ca.gpu_slot = kd ? h->d_mps_gslot.as<int32_t>() : h->d_mig_gslot.as<int32_t>();
      ca.m_used = a.m_used;
      ca.m_free = a.m_free;
      ca.sizes = a.sizes;
      ca.p_gpu = a.p_gpu;
      ca.p_used_lo = a.p_used_lo;
      ca.p_used_hi = a.p_used_hi;
      ca.sched = (flags & NOSPLAN_PLAN_SCHED) ? 1 : 0;
      ca.gate_hit = a.gate_hit;
      ca.repart_node = a.repart_node;
      ca.sched_node = a.sched_node;
      ca.sched_gpu = a.sched_gpu;
      ca.n_win = (a.tg + 31) / 32;
      CU(b[14].reserve((size_t)a.n_keys * ca.n_win));
      ca.winmax = b[14].as<uint8_t>();
      K.maxq = 1;
      for (int q : s.qty) K.maxq = std::max(K.maxq, q);
      if (kd == 0) plan_presence_kernel<false><<<(a.tg + 255) / 256, 256, 0, st>>>(ca);
      else plan_presence_kernel<true><<<(a.tg + 255) / 256, 256, 0, st>>>(ca);
      K.hybrid = true;
      if ((rc = K.chains_from(h, 0))) return rc;
    } else {
      if ((rc = K.run_seq(0, M, nullptr, nullptr))) return rc;
    }
    CU(cudaGetLastError());
    out_ptr[kd] = a.repart_node;
  }
  // Hybrid kinds: one round per pod that the gate does not stop.  Both kinds' first rounds are in flight by now; the
  // later rounds (rare: a profile has to be absent from the whole cluster) run kind after kind.
  for (int kd = 0; kd < 2; ++kd) {
    PlanKind& K = pk[kd];
    if (!K.hybrid) continue;
    int rounds = 0;
    for (;;) {
      CU(cudaStreamSynchronize(K.st));
      const int pos = h->h_limit[kd];
      if (pos >= K.M) break;
      if (++rounds > kMaxHybridRounds) {  // many absent profiles: the sequencer takes the rest in one go
        if ((rc = K.run_seq(pos, K.M - pos, nullptr, nullptr))) return rc;
        break;
      }
      if ((rc = K.run_seq(pos, 1, K.ca.present, const_cast<int32_t*>(K.ca.dead)))) return rc;
      if ((rc = K.chains_from(h, pos + 1))) return rc;
    }
  }
  // results back only after both kinds are in flight: a device-to-host copy into pageable memory blocks
  // the host until its stream drains, which would serialise the MIG and MPS chains
  // (the four result arrays of a kind lie back to back on the device: repart_node, sched_node, sched_gpu as int32 [M],
  // then gate_hit as bytes [M]; one copy per kind into pinned memory)
  size_t pin_off[2] = {0, 0}, pin_bytes = 0;
  for (int kd = 0; kd < 2; ++kd) {
    pin_off[kd] = pin_bytes;
    pin_bytes += ((size_t)side[kd].pod_idx.size() * 13 + 15) / 16 * 16;
  }
  if ((rc = pin_reserve(h, std::max<size_t>(pin_bytes, 16)))) return rc;
  char* const pin = static_cast<char*>(h->h_pin);
  for (int kd = 0; kd < 2; ++kd) {
    const int M = (int)side[kd].pod_idx.size();
    if (M == 0) continue;
    cudaStream_t st = kd ? h->stream2 : h->stream;
    CU(cudaMemcpyAsync(pin + pin_off[kd], out_ptr[kd], (size_t)M * 13, cudaMemcpyDeviceToHost, st));
  }
  const auto t_enq = std::chrono::steady_clock::now();
  CU(cudaStreamSynchronize(h->stream));
  CU(cudaStreamSynchronize(h->stream2));
  const auto t_sync = std::chrono::steady_clock::now();
  if (trace) {  // the device-to-host copies above block until their stream drains: "enqueue" includes the kernels
    auto us = [](auto a, auto b) { return std::chrono::duration<double, std::micro>(b - a).count(); };
    std::fprintf(stderr, "nosplan_plan: host split %.0f us, enqueue + kernels %.0f us, wait %.0f us\n", us(t_begin, t_split),
                 us(t_split, t_enq), us(t_enq, t_sync));
  }
  for (int kd = 0; kd < 2; ++kd) {
    Side& s = side[kd];
    const size_t M = s.pod_idx.size();
    const int32_t* o_rep = reinterpret_cast<const int32_t*>(pin + pin_off[kd]);
    const int32_t* o_node = o_rep + M;
    const int32_t* o_gpu = o_rep + 2 * M;
    const uint8_t* o_gate = reinterpret_cast<const uint8_t*>(o_rep + 3 * M);
    for (size_t j = 0; j < M; ++j) {
      const int i = s.pod_idx[j];
      if (gate_hit) gate_hit[i] = o_gate[j];
      if (repart_node) repart_node[i] = o_rep[j];
      if (sched_node) sched_node[i] = o_node[j];
      if (sched_gpu) sched_gpu[i] = o_gpu[j];
    }
  }
  return NOSPLAN_OK;
}

int nosplan_get_mig_state(nosplan_t* h, int32_t* used, int32_t* free_) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  if (!h->have_mig) return fail(h, NOSPLAN_ERR_ARG, "no MIG snapshot");
  if (!used || !free_) return fail(h, NOSPLAN_ERR_ARG, "null output");
  int rc = bind(h);
  if (rc) return rc;
  const int tg = h->mig_tg, P = h->P;
  if (tg == 0) return NOSPLAN_OK;
  CU(h->d_in[4].reserve(sizeof(int32_t) * (size_t)tg * P));
  CU(h->d_in[5].reserve(sizeof(int32_t) * (size_t)tg * P));
  mig_unpack_kernel<<<(tg + 127) / 128, 128, 0, h->stream>>>(tg, P, h->d_mig_used.as<uint4>(), h->d_mig_free.as<uint4>(),
                                                            h->d_in[4].as<int32_t>(), h->d_in[5].as<int32_t>());
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(used, h->d_in[4].p, sizeof(int32_t) * (size_t)tg * P, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaMemcpyAsync(free_, h->d_in[5].p, sizeof(int32_t) * (size_t)tg * P, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return NOSPLAN_OK;
}

int nosplan_get_mps_state(nosplan_t* h, int32_t* used, int32_t* free_, uint32_t* free_mask) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  if (!h->have_mps) return fail(h, NOSPLAN_ERR_ARG, "no MPS snapshot");
  if (!used || !free_) return fail(h, NOSPLAN_ERR_ARG, "null output");
  int rc = bind(h);
  if (rc) return rc;
  const int tg = h->mps_tg, S = h->S;
  if (tg == 0) return NOSPLAN_OK;
  CU(h->d_in[4].reserve(sizeof(int32_t) * (size_t)tg * S));
  CU(h->d_in[5].reserve(sizeof(int32_t) * (size_t)tg * S));
  CU(h->d_in[3].reserve(sizeof(uint32_t) * (size_t)tg));
  mps_unpack_kernel<<<(tg + 127) / 128, 128, 0, h->stream>>>(tg, S, h->d_mps_gpu.as<MpsGpu>(), h->d_mps_ulo.as<uint4>(),
                                                            h->d_mps_uhi.as<uint4>(), h->d_in[4].as<int32_t>(),
                                                            h->d_in[5].as<int32_t>(), h->d_in[3].as<uint32_t>());
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(used, h->d_in[4].p, sizeof(int32_t) * (size_t)tg * S, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaMemcpyAsync(free_, h->d_in[5].p, sizeof(int32_t) * (size_t)tg * S, cudaMemcpyDeviceToHost, h->stream));
  if (free_mask) CU(cudaMemcpyAsync(free_mask, h->d_in[3].p, sizeof(uint32_t) * (size_t)tg, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return NOSPLAN_OK;
}

int nosplan_init_geometry_mig(nosplan_t* h, uint8_t* err, int32_t* out_free) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  if (!h->have_mig) return fail(h, NOSPLAN_ERR_ARG, "no MIG snapshot");
  if (!err || !out_free) return fail(h, NOSPLAN_ERR_ARG, "null output");
  int rc = bind(h);
  if (rc) return rc;
  const int tg = h->mig_tg, P = h->P;
  if (tg == 0) return NOSPLAN_OK;
  CU(h->d_out[4].reserve(tg));
  CU(h->d_out[5].reserve(sizeof(int32_t) * (size_t)tg * P));
  mig_init_geometry_kernel<<<(tg + 127) / 128, 128, 0, h->stream>>>(
      h->d_tables.as<MigTables>(), tg, P, h->d_mig_used0.as<uint4>(), h->d_mig_free0.as<uint4>(),
      h->d_mig_model.as<uint8_t>(), h->d_out[4].as<uint8_t>(), h->d_out[5].as<int32_t>());
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(err, h->d_out[4].p, tg, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaMemcpyAsync(out_free, h->d_out[5].p, sizeof(int32_t) * (size_t)tg * P, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return NOSPLAN_OK;
}

}  // extern "C"

// =================================================================================================================
// Allocation-free MIG scorer.
//
// Some callers of this library are not allowed to allocate device memory at the moment they need a
// score: a planner sidecar that shares its CUDA context with a latency-critical process and has
// been handed a fixed set of buffers, or a control loop that must not hit cudaMalloc's implicit
// device synchronisation.  For them this entry point does the same work as nosplan_score_mig_dev
// but takes every input and output as a caller-owned device pointer and keeps its one piece of
// scratch — the byte-packed, slot-major copy of the snapshot that the scorer kernel reads — in a
// fixed array that lives inside the library image.  It calls no cudaMalloc / cudaFree and no other
// allocating routine.  It has no counterpart in the reference.
//
// Limits (rejected with NOSPLAN_ERR_RANGE beyond them): at most NOSPLAN_STATIC_MAX_NODES nodes, every
// node with the same number of GPUs, at most NOSPLAN_STATIC_MAX_GPUS_PER_NODE of them.  The array
// costs about 1.1 MB of device memory on every GPU that loads the library, until the process exits.
//
// The array is shared by all calls on a device and keeps the previous call's contents.  Each call
// rewrites every element the scorer will read before it reads it (nothing in it needs to start at
// zero; the caller's first_node output is reset by the call itself), so stale data cannot leak into a
// result — but two calls whose kernels overlap on the GPU (different streams, or different host
// threads) race on it.  Serialise them.
//
// The first call on a device looks up the array's device address (cudaGetSymbolAddress); make that
// first call before starting any stream capture.
namespace {

constexpr int kStaticMaxNodes = NOSPLAN_STATIC_MAX_NODES;
constexpr int kStaticMaxGpus = NOSPLAN_STATIC_MAX_GPUS_PER_NODE;

struct StaticScratch {
  uint4 s_used[kStaticMaxGpus][kStaticMaxNodes];
  uint4 s_free[kStaticMaxGpus][kStaticMaxNodes];
  uint8_t s_model[kStaticMaxGpus][kStaticMaxNodes];
  uint8_t s_cnt[kStaticMaxNodes];
  int32_t gpu_off[kStaticMaxNodes + 1];
};

// In host code this name is only a stand-in for the symbol; the device address is fetched once per
// device with cudaGetSymbolAddress (static_scratch_for) and that pointer is what kernels receive.
static __device__ __align__(256) unsigned char g_static_scratch[sizeof(StaticScratch)];

constexpr int kMaxDevices = 64;
struct StaticSetup {
  std::once_flag once;  // first calls on a device may come from different handles on different threads
  cudaError_t rc = cudaSuccess;
  StaticScratch* ptr = nullptr;
};
StaticSetup g_static_setup[kMaxDevices];

// packs int32 rows (device) into the static scratch; same byte layout as mig_pack_kernel's
// slot-major outputs, uniform GPUs per node
__global__ void static_pack_kernel(StaticScratch* sc, int N, int P, int G, int npad, const uint8_t* __restrict__ valid,
                                   const int32_t* __restrict__ model, const int32_t* __restrict__ used,
                                   const int32_t* __restrict__ fr) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n <= N) sc->gpu_off[n] = n * G;
  if (n >= npad) return;
  const bool real = n < N;
  const bool ok = real && (valid == nullptr || valid[n]);
  sc->s_cnt[n] = ok ? (uint8_t)G : 0;
  for (int s = 0; s < G; ++s) {
    uint32_t uw[4] = {0, 0, 0, 0}, fw[4] = {0, 0, 0, 0};
    uint8_t m = 0xff;
    if (ok) {
      const size_t g = (size_t)n * G + s;
      for (int p = 0; p < P; ++p) {
        uw[p >> 2] |= (uint32_t)(used[g * P + p] & 0xff) << ((p & 3) * 8);
        fw[p >> 2] |= (uint32_t)(fr[g * P + p] & 0xff) << ((p & 3) * 8);
      }
      m = (uint8_t)model[g];
    }
    sc->s_used[s][n] = make_uint4(uw[0], uw[1], uw[2], uw[3]);
    sc->s_free[s][n] = make_uint4(fw[0], fw[1], fw[2], fw[3]);
    sc->s_model[s][n] = m;
  }
}

int static_scratch_for(nosplan_t* h, StaticScratch** out) {
  if (h->device < 0 || h->device >= kMaxDevices) return fail(h, NOSPLAN_ERR_RANGE, "device ordinal %d not supported", h->device);
  StaticSetup& st = g_static_setup[h->device];
  // once per device, with that device current (bind() ran): the symbol's address on THIS device
  std::call_once(st.once, [&st] {
    void* p = nullptr;
    st.rc = cudaGetSymbolAddress(&p, g_static_scratch);
    st.ptr = reinterpret_cast<StaticScratch*>(p);
  });
  if (st.rc != cudaSuccess || !st.ptr)
    return fail(h, NOSPLAN_ERR_CUDA, "cudaGetSymbolAddress(static scratch) failed: %s", cudaGetErrorString(st.rc));
  *out = st.ptr;
  return NOSPLAN_OK;
}

}  // namespace

extern "C" int nosplan_score_mig_static(nosplan_t* h, int n_nodes, int gpus_per_node, const uint8_t* valid,
                                        const int32_t* model, const int32_t* used, const int32_t* free_, int n_demands,
                                        const int32_t* dem_off, const int32_t* dem_key, const int32_t* dem_qty,
                                        int32_t* first_node, void* stream) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  if (!h->have_tables) return fail(h, NOSPLAN_ERR_ARG, "nosplan_set_mig_tables must be called first");
  if (n_nodes < 0 || gpus_per_node < 0 || n_demands < 0) return fail(h, NOSPLAN_ERR_ARG, "negative size");
  if (n_nodes > kStaticMaxNodes)
    return fail(h, NOSPLAN_ERR_RANGE, "nosplan_score_mig_static: %d nodes exceed the fixed scratch (max %d nodes); use nosplan_score_mig_dev",
                n_nodes, kStaticMaxNodes);
  if (gpus_per_node > kStaticMaxGpus)
    return fail(h, NOSPLAN_ERR_RANGE, "nosplan_score_mig_static: %d GPUs per node exceed the fixed scratch (max %d); use nosplan_score_mig_dev",
                gpus_per_node, kStaticMaxGpus);
  if (n_demands > 0 && (!dem_off || !dem_key || !dem_qty || !first_node)) return fail(h, NOSPLAN_ERR_ARG, "null demand/output pointer");
  if (n_nodes > 0 && gpus_per_node > 0 && (!model || !used || !free_)) return fail(h, NOSPLAN_ERR_ARG, "null snapshot pointer");
  int rc = bind(h);
  if (rc) return rc;
  StaticScratch* sc = nullptr;
  if ((rc = static_scratch_for(h, &sc))) return rc;
  if (n_demands == 0) return NOSPLAN_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int D = n_demands;
  const int gs = std::max(gpus_per_node, 1);
  const int npad = std::max(32, (n_nodes + 31) / 32 * 32);
  fill_i32_kernel<<<(D + 255) / 256, 256, 0, st>>>(first_node, D, -1);
  if (n_nodes > 0) {
    const int threads = std::max(npad, n_nodes + 1);
    static_pack_kernel<<<(threads + 127) / 128, 128, 0, st>>>(sc, n_nodes, h->P, gpus_per_node, npad, valid, model, used, free_);
    MigScoreArgs a{};
    a.T = h->d_tables.as<MigTables>();
    a.s_used = &sc->s_used[0][0];
    a.s_free = &sc->s_free[0][0];
    a.s_model = &sc->s_model[0][0];
    a.s_cnt = sc->s_cnt;
    a.gpu_off = sc->gpu_off;
    a.N = n_nodes;
    a.npad = kStaticMaxNodes;  // row stride of the fixed arrays
    a.gs = gs;
    a.cmax = h->cmax;
    a.k16 = h->k16ok ? 1 : 0;
    a.tg = n_nodes * gpus_per_node;
    a.D = D;
    a.dem_off = dem_off;
    a.dem_key = dem_key;
    a.dem_qty = dem_qty;
    a.first_node = first_node;
    a.node_base = 0;
    a.updated = nullptr;
    a.choice = nullptr;
    a.row_list = nullptr;
    a.row_meta = nullptr;
    const int tiles = (n_nodes + kScoreTN - 1) / kScoreTN;
    int gy;
    pick_chunks(h->sm_count, tiles, D, 16, &a.d_chunk, &gy);
    const size_t smem = mig_score_smem_bytes(h->h_tables.n_models, gs, h->cmax, kScoreTN);
    if (smem > 200 * 1024) return fail(h, NOSPLAN_ERR_RANGE, "MIG scorer tile needs %zu bytes of shared memory", smem);
    if ((rc = launch_mig_score_kernel(h, a, dim3(tiles, gy), smem, st))) return rc;  // sets a smem limit, allocates nothing
  }
  CU(cudaGetLastError());
  return NOSPLAN_OK;
}

// =================================================================================================================
// Agent-side diff planner, count level: plan.NewMigConfigPlan
// (internal/controllers/migagent/plan/plan.go:31-92) over every (GPU, profile) of a batch of nodes.
namespace {
// 16 lanes per GPU (one per profile column): coalesced rows, "any create on this GPU" by half-warp ballot
__global__ void mig_diff_plan_kernel(int G, int P, const int32_t* __restrict__ used, const int32_t* __restrict__ fr,
                                     const int32_t* __restrict__ spec, const uint32_t* __restrict__ spec_mask,
                                     int32_t* __restrict__ create, int32_t* __restrict__ del_free, int32_t* __restrict__ del_used) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int g = idx >> 4, p = idx & 15;
  const bool live = g < G && p < P;
  int c = 0, df = 0, du = 0, f = 0;
  if (live) {
    const size_t i = (size_t)g * P + p;
    const int u = used[i];
    f = fr[i];
    if (!((spec_mask[g] >> p) & 1u)) {  // not named by a spec annotation: delete everything (plan.go:37-41)
      df = f;
      du = u;
    } else {
      const int diff = spec[i] - (u + f);  // plan.go:60
      if (diff > 0) c = diff;
      else if (diff < 0) {  // free devices first, then used ones (plan.go:112-137)
        df = min(-diff, f);
        du = -diff - df;
      }
    }
  }
  const uint32_t creating = __ballot_sync(0xffffffffu, c > 0);
  const uint32_t half = (threadIdx.x & 16) ? 0xffff0000u : 0x0000ffffu;
  if (live) {
    if (creating & half) {  // plan.go:73-88: re-create the free devices that are not already being deleted
      const int again = f - df;
      df += again;
      c += again;
    }
    const size_t i = (size_t)g * P + p;
    create[i] = c;
    del_free[i] = df;
    del_used[i] = du;
  }
}
}  // namespace

extern "C" int nosplan_diff_plan_mig(nosplan_t* h, int n_gpus, int n_profiles, const int32_t* used, const int32_t* free_,
                                     const int32_t* spec, const uint32_t* spec_mask, int32_t* create, int32_t* del_free,
                                     int32_t* del_used) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  if (n_gpus < 0 || n_profiles < 1 || n_profiles > NOSPLAN_MAX_MIG_PROFILES)
    return fail(h, NOSPLAN_ERR_RANGE, "n_profiles %d out of range [1,%d]", n_profiles, NOSPLAN_MAX_MIG_PROFILES);
  if (n_gpus == 0) return NOSPLAN_OK;
  if (!used || !free_ || !spec || !spec_mask || !create || !del_free || !del_used) return fail(h, NOSPLAN_ERR_ARG, "null argument");
  int rc = bind(h);
  if (rc) return rc;
  const size_t n = (size_t)n_gpus * n_profiles, bytes = n * sizeof(int32_t);
  cudaStream_t st = h->stream;
  if ((rc = upload(h, h->d_in[0], used, bytes, st))) return rc;
  if ((rc = upload(h, h->d_in[1], free_, bytes, st))) return rc;
  if ((rc = upload(h, h->d_in[2], spec, bytes, st))) return rc;
  if ((rc = upload(h, h->d_in[3], spec_mask, sizeof(uint32_t) * n_gpus, st))) return rc;
  CU(h->d_out[0].reserve(bytes));
  CU(h->d_out[1].reserve(bytes));
  CU(h->d_out[2].reserve(bytes));
  const long long threads = (long long)n_gpus * 16;
  mig_diff_plan_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(
      n_gpus, n_profiles, h->d_in[0].as<int32_t>(), h->d_in[1].as<int32_t>(), h->d_in[2].as<int32_t>(),
      h->d_in[3].as<uint32_t>(), h->d_out[0].as<int32_t>(), h->d_out[1].as<int32_t>(), h->d_out[2].as<int32_t>());
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(create, h->d_out[0].p, bytes, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(del_free, h->d_out[1].p, bytes, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(del_used, h->d_out[2].p, bytes, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return NOSPLAN_OK;
}

// ---- agent-side diff planner, device level ------------------------------------------------------------------------
// plan.NewMigConfigPlan (internal/controllers/migagent/plan/plan.go:31-92) down to which DEVICES go: one thread per
// GPU walks the GPU's devices in the order the reference walks them (state.Flatten().SortByDeviceId(), :45 — the
// caller sorts, ids are strings and never reach the device) and marks every device with the operation that deletes it.
namespace {
__global__ void mig_diff_devices_kernel(int G, int P, const int32_t* __restrict__ dev_off, const int32_t* __restrict__ dev_profile,
                                        const uint8_t* __restrict__ dev_status, const int32_t* __restrict__ spec,
                                        const uint32_t* __restrict__ spec_mask, uint8_t* __restrict__ dev_op,
                                        int32_t* __restrict__ create_new, int32_t* __restrict__ create_again,
                                        uint8_t* __restrict__ matches) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= G) return;
  const int d0 = dev_off[g], d1 = dev_off[g + 1];
  const uint32_t mask = spec_mask[g];
  int cnt[kMaxProf], fre[kMaxProf], del_free[kMaxProf], del_used[kMaxProf];
#pragma unroll
  for (int p = 0; p < kMaxProf; ++p) cnt[p] = fre[p] = del_free[p] = del_used[p] = 0;
  for (int d = d0; d < d1; ++d) {
    const int p = dev_profile[d];
    for (int q = 0; q < kMaxProf; ++q)  // (register arrays: no dynamic indexing)
      if (q == p) {
        cnt[q]++;
        if (dev_status[d] == 1) fre[q]++;
      }
  }
  bool any_create = false, same = true;
#pragma unroll
  for (int p = 0; p < kMaxProf; ++p) {
    int c = 0;
    if (p < P) {
      const bool named = (mask >> p) & 1u;
      const int want = named ? spec[(size_t)g * P + p] : 0;
      // MigState.Matches (mig_state.go:40-58): both maps hold the same keys with the same counts; a spec
      // annotation creates its key even with quantity 0, the state only has keys of devices that exist
      same &= named ? (want == cnt[p] && cnt[p] > 0) : (cnt[p] == 0);
      if (named) {
        const int diff = want - cnt[p];  // plan.go:60
        if (diff > 0) c = diff;
        else if (diff < 0) {  // extractCandidatesForDeletion (:112-137): free devices first, then the others
          del_free[p] = min(-diff, fre[p]);
          del_used[p] = -diff - del_free[p];
        }
      }
      create_new[(size_t)g * P + p] = c;
    }
    any_create |= c > 0;
  }
  int again[kMaxProf];
#pragma unroll
  for (int p = 0; p < kMaxProf; ++p) again[p] = 0;
  for (int d = d0; d < d1; ++d) {
    const int p = dev_profile[d];
    const bool is_free = dev_status[d] == 1;
    uint8_t op = 0;
    if (!((mask >> p) & 1u)) {
      op = 1;  // its profile is not named by any spec annotation of the GPU (:37-41, getResourcesNotIncludedInSpec :168-182)
    } else {
      for (int q = 0; q < kMaxProf; ++q)
        if (q == p) {
          if (is_free && del_free[q] > 0) {
            del_free[q]--;
            op = 2;
          } else if (!is_free && del_used[q] > 0) {
            del_used[q]--;
            op = 2;
          } else if (is_free && any_create) {  // re-created so that more creation orders are possible (:73-88)
            again[q]++;
            op = 3;
          }
        }
    }
    dev_op[d] = op;
  }
#pragma unroll
  for (int p = 0; p < kMaxProf; ++p)
    if (p < P) create_again[(size_t)g * P + p] = again[p];
  matches[g] = same ? 1 : 0;
}
}  // namespace

extern "C" int nosplan_diff_plan_mig_devices(nosplan_t* h, int n_gpus, int n_profiles, const int32_t* dev_off, const int32_t* dev_profile,
                                             const uint8_t* dev_status, const int32_t* spec, const uint32_t* spec_mask, uint8_t* dev_op,
                                             int32_t* create_new, int32_t* create_again, uint8_t* matches) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  if (n_gpus < 0 || n_profiles < 1 || n_profiles > NOSPLAN_MAX_MIG_PROFILES)
    return fail(h, NOSPLAN_ERR_RANGE, "n_profiles %d out of range [1,%d]", n_profiles, NOSPLAN_MAX_MIG_PROFILES);
  if (n_gpus == 0) return NOSPLAN_OK;
  if (!dev_off || !spec || !spec_mask || !create_new || !create_again || !matches) return fail(h, NOSPLAN_ERR_ARG, "null argument");
  if (dev_off[0] != 0) return fail(h, NOSPLAN_ERR_ARG, "dev_off[0] must be 0");
  for (int g = 0; g < n_gpus; ++g)
    if (dev_off[g + 1] < dev_off[g]) return fail(h, NOSPLAN_ERR_ARG, "dev_off not ascending at GPU %d", g);
  const int nd = dev_off[n_gpus];
  if (nd > 0 && (!dev_profile || !dev_status || !dev_op)) return fail(h, NOSPLAN_ERR_ARG, "null device arrays");
  for (int d = 0; d < nd; ++d) {
    if (dev_profile[d] < 0 || dev_profile[d] >= n_profiles) return fail(h, NOSPLAN_ERR_RANGE, "device %d: profile %d out of range", d, dev_profile[d]);
    if (dev_status[d] > 2) return fail(h, NOSPLAN_ERR_RANGE, "device %d: status %d (0 used, 1 free, 2 unknown)", d, dev_status[d]);
  }
  int rc = bind(h);
  if (rc) return rc;
  cudaStream_t st = h->stream;
  const size_t gp = (size_t)n_gpus * n_profiles * sizeof(int32_t);
  if ((rc = upload(h, h->d_in[0], dev_off, sizeof(int32_t) * (n_gpus + 1), st))) return rc;
  if ((rc = upload(h, h->d_in[1], dev_profile, sizeof(int32_t) * std::max(nd, 1), st))) return rc;
  if ((rc = upload(h, h->d_in[2], spec, gp, st))) return rc;
  if ((rc = upload(h, h->d_in[3], spec_mask, sizeof(uint32_t) * n_gpus, st))) return rc;
  if ((rc = upload(h, h->d_in[4], dev_status, std::max(nd, 1), st))) return rc;
  CU(h->d_out[0].reserve(gp));
  CU(h->d_out[1].reserve(gp));
  CU(h->d_out[2].reserve(std::max(nd, 1)));
  CU(h->d_out[4].reserve(n_gpus));
  mig_diff_devices_kernel<<<(n_gpus + 127) / 128, 128, 0, st>>>(
      n_gpus, n_profiles, h->d_in[0].as<int32_t>(), h->d_in[1].as<int32_t>(), h->d_in[4].as<uint8_t>(), h->d_in[2].as<int32_t>(),
      h->d_in[3].as<uint32_t>(), h->d_out[2].as<uint8_t>(), h->d_out[0].as<int32_t>(), h->d_out[1].as<int32_t>(), h->d_out[4].as<uint8_t>());
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(create_new, h->d_out[0].p, gp, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(create_again, h->d_out[1].p, gp, cudaMemcpyDeviceToHost, st));
  if (nd > 0) CU(cudaMemcpyAsync(dev_op, h->d_out[2].p, nd, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(matches, h->d_out[4].p, n_gpus, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return NOSPLAN_OK;
}

// ---- batch-demand planning mode (NEW BEHAVIOUR: the reference only documents it; definition in oracle/batch_plan.py) ----
namespace {
// slot-major copies of the MIG working state for the scorers (what mig_pack_kernel makes from the host rows)
__global__ void mig_repack_kernel(int N, int gs, int npad, const int32_t* __restrict__ gpu_off, const uint8_t* __restrict__ s_cnt,
                                  const uint4* __restrict__ a_used, const uint4* __restrict__ a_free, uint4* __restrict__ s_used,
                                  uint4* __restrict__ s_free) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int n = i / gs, s = i - n * gs;
  if (n >= N || s >= s_cnt[n]) return;  // slots without a GPU and invalid nodes keep their zeros
  const int g = gpu_off[n] + s;
  s_used[(size_t)s * npad + n] = a_used[g];
  s_free[(size_t)s * npad + n] = a_free[g];
}

struct BatchCtx {
  nosplan_t* h;
  int kd, M, N, tg, K;
  cudaStream_t st;
  const int32_t* key;
  const int32_t* qty;
  DevBuf* b;  // h->d_plan + kd * 16
  ChainArgs ca;
};

// the scorers read the snapshot copies (*0) and the tables derived from them: make them describe the working state
int batch_sync_scorer_state(BatchCtx& c) {
  nosplan_t* h = c.h;
  cudaStream_t st = c.st;
  if (c.kd == 0) {
    CU(cudaMemcpyAsync(h->d_mig_used0.p, h->d_mig_used.p, sizeof(uint4) * c.tg, cudaMemcpyDeviceToDevice, st));
    CU(cudaMemcpyAsync(h->d_mig_free0.p, h->d_mig_free.p, sizeof(uint4) * c.tg, cudaMemcpyDeviceToDevice, st));
    mig_repack_kernel<<<(c.N * h->mig_gs + 127) / 128, 128, 0, st>>>(c.N, h->mig_gs, h->mig_npad, h->d_mig_off.as<int32_t>(),
                                                                     h->d_mig_scnt.as<uint8_t>(), h->d_mig_used0.as<uint4>(),
                                                                     h->d_mig_free0.as<uint4>(), h->d_mig_sused.as<uint4>(),
                                                                     h->d_mig_sfree.as<uint4>());
    CU(cudaGetLastError());
    return mig_derive_tables(h, c.N, h->mig_npad, h->mig_gs, st);
  }
  CU(cudaMemcpyAsync(h->d_mps_gpu0.p, h->d_mps_gpu.p, sizeof(MpsGpu) * c.tg, cudaMemcpyDeviceToDevice, st));
  CU(cudaMemcpyAsync(h->d_mps_ulo0.p, h->d_mps_ulo.p, sizeof(uint4) * c.tg, cudaMemcpyDeviceToDevice, st));
  CU(cudaMemcpyAsync(h->d_mps_uhi0.p, h->d_mps_uhi.p, sizeof(uint4) * c.tg, cudaMemcpyDeviceToDevice, st));
  return mps_derive_tables(h, c.N, h->S, st);
}

// Node.AddPod first-fit (the chain kernels, gate ignored) for the pods `todo` (ascending indices); results into node/gpu
int batch_first_fit(BatchCtx& c, const std::vector<int32_t>& todo, std::vector<int32_t>& node, std::vector<int32_t>& gpu) {
  nosplan_t* h = c.h;
  cudaStream_t st = c.st;
  const int m = (int)todo.size();
  if (m == 0 || c.tg == 0) return NOSPLAN_OK;
  std::vector<int32_t> ckey, coff{0}, cpod(m), cq(m), cnt(c.K, 0), at(c.K, 0);
  for (int j = 0; j < m; ++j) cnt[c.key[todo[j]]]++;
  for (int k = 0; k < c.K; ++k)
    if (cnt[k]) {
      at[k] = coff.back();
      ckey.push_back(k);
      coff.push_back(coff.back() + cnt[k]);
    }
  int maxq = 1;
  for (int j = 0; j < m; ++j) {
    const int i = todo[j];
    const int pos = at[c.key[i]]++;
    cpod[pos] = j;  // kind-local index = position in todo (ascending: planner order kept inside a chain)
    cq[pos] = c.qty[i];
    maxq = std::max(maxq, (int)c.qty[i]);
  }
  DevBuf* b = c.b;
  int rc;
  if ((rc = upload(h, b[8], ckey.data(), sizeof(int32_t) * ckey.size(), st))) return rc;
  if ((rc = upload(h, b[9], coff.data(), sizeof(int32_t) * coff.size(), st))) return rc;
  if ((rc = upload(h, b[10], cpod.data(), sizeof(int32_t) * m, st))) return rc;
  if ((rc = upload(h, b[11], cq.data(), sizeof(int32_t) * m, st))) return rc;
  CU(b[12].reserve(sizeof(int32_t) * (2 * kMaxSizes + 2)));
  CU(b[6].reserve((size_t)m * (1 + 3 * sizeof(int32_t)) + 64));
  CU(b[14].reserve((size_t)c.K * ((c.tg + 31) / 32)));
  ChainArgs& ca = c.ca;
  std::memset(&ca, 0, sizeof(ca));
  ca.n_chains = (int)ckey.size();
  ca.chain_key = b[8].as<int32_t>();
  ca.chain_off = b[9].as<int32_t>();
  ca.chain_pod = b[10].as<int32_t>();
  ca.chain_q = b[11].as<int32_t>();
  ca.present = b[12].as<int32_t>();
  ca.dead = ca.present + kMaxSizes;
  ca.limit_out = ca.present + 2 * kMaxSizes;
  ca.range = ca.limit_out;
  ca.ignore_gate = 1;
  ca.M = m;
  ca.N = c.N;
  ca.tg = c.tg;
  ca.n_keys = c.K;
  ca.gpu_off = c.kd ? h->d_mps_off.as<int32_t>() : h->d_mig_off.as<int32_t>();
  ca.gpu_node = c.kd ? h->d_mps_gnode.as<int32_t>() : h->d_mig_gnode.as<int32_t>();
  ca.gpu_slot = c.kd ? h->d_mps_gslot.as<int32_t>() : h->d_mig_gslot.as<int32_t>();
  ca.m_used = h->d_mig_used.as<uint4>();
  ca.m_free = h->d_mig_free.as<uint4>();
  ca.sizes = h->d_mps_sizes.as<int32_t>();
  ca.p_gpu = h->d_mps_gpu.as<MpsGpu>();
  ca.p_used_lo = h->d_mps_ulo.as<uint4>();
  ca.p_used_hi = h->d_mps_uhi.as<uint4>();
  ca.sched = 1;
  ca.repart_node = b[6].as<int32_t>();
  ca.sched_node = ca.repart_node + m;
  ca.sched_gpu = ca.sched_node + m;
  ca.gate_hit = reinterpret_cast<uint8_t*>(ca.sched_gpu + m);
  ca.n_win = (c.tg + 31) / 32;
  ca.winmax = b[14].as<uint8_t>();
  const int32_t range[2] = {0, m};
  h->h_range[c.kd][0] = range[0];
  h->h_range[c.kd][1] = range[1];
  CU(cudaMemcpyAsync(ca.limit_out, h->h_range[c.kd], sizeof(range), cudaMemcpyHostToDevice, st));
  CU(cudaMemsetAsync(ca.sched_node, 0xff, sizeof(int32_t) * 2 * (size_t)m, st));  // -1: a pod no chain step reaches stays pending
  if (c.kd == 0) plan_winmax_kernel<false><<<ca.n_win, 32, 0, st>>>(ca);
  else plan_winmax_kernel<true><<<ca.n_win, 32, 0, st>>>(ca);
  if (maxq <= 2) {
    if (c.kd == 0) plan_chain_kernel<false, 2><<<ca.n_chains, 32, 0, st>>>(ca);
    else plan_chain_kernel<true, 2><<<ca.n_chains, 32, 0, st>>>(ca);
  } else {
    if (c.kd == 0) plan_chain_kernel<false, kChainQ><<<ca.n_chains, 32, 0, st>>>(ca);
    else plan_chain_kernel<true, kChainQ><<<ca.n_chains, 32, 0, st>>>(ca);
  }
  CU(cudaGetLastError());
  std::vector<int32_t> rn(m), rg(m);
  CU(cudaMemcpyAsync(rn.data(), ca.sched_node, sizeof(int32_t) * m, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(rg.data(), ca.sched_gpu, sizeof(int32_t) * m, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  for (int j = 0; j < m; ++j) {
    node[todo[j]] = rn[j];
    gpu[todo[j]] = rg[j];
  }
  return NOSPLAN_OK;
}
}  // namespace

extern "C" int nosplan_plan_batch(nosplan_t* h, int kind, int n_pods, const int32_t* pod_key, const int32_t* pod_qty, int max_rounds,
                                  int32_t* sched_node, int32_t* sched_gpu, uint8_t* phase, uint8_t* repartitioned,
                                  int32_t* order_out, int32_t* rounds_out) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  NSP_RANGE("nosplan_plan_batch");
  if (kind != 0 && kind != 1) return fail(h, NOSPLAN_ERR_ARG, "kind must be 0 (MIG) or 1 (MPS)");
  if (kind ? !h->have_mps : !h->have_mig) return fail(h, NOSPLAN_ERR_ARG, "no snapshot of that kind");
  if (n_pods < 0 || (n_pods > 0 && (!pod_key || !pod_qty || !sched_node || !sched_gpu || !phase))) return fail(h, NOSPLAN_ERR_ARG, "null pod arrays");
  if (max_rounds < 1) max_rounds = 8;
  BatchCtx c{};
  c.h = h;
  c.kd = kind;
  c.M = n_pods;
  c.N = kind ? h->mps_n : h->mig_n;
  c.tg = kind ? h->mps_tg : h->mig_tg;
  c.K = kind ? h->S : h->P;
  c.st = h->stream;
  c.key = pod_key;
  c.qty = pod_qty;
  c.b = h->d_plan + kind * 16;
  if (!repartitioned && c.N > 0) return fail(h, NOSPLAN_ERR_ARG, "repartitioned is required");
  for (int i = 0; i < n_pods; ++i) {
    if (pod_key[i] < 0 || pod_key[i] >= c.K) return fail(h, NOSPLAN_ERR_RANGE, "pod %d: key %d out of range", i, pod_key[i]);
    if (pod_qty[i] < 1 || pod_qty[i] > kChainQ) return fail(h, NOSPLAN_ERR_RANGE, "pod %d: quantity %d outside [1,%d]", i, pod_qty[i], kChainQ);
  }
  int rc = bind(h);
  if (rc) return rc;
  cudaStream_t st = c.st;
  const int M = n_pods, N = c.N, tg = c.tg, K = c.K;
  if (order_out) *order_out = 0;
  if (rounds_out) *rounds_out = 0;
  for (int n = 0; n < N; ++n) repartitioned[n] = 0;
  for (int i = 0; i < M; ++i) {
    sched_node[i] = sched_gpu[i] = -1;
    phase[i] = 0;
  }
  if (M == 0 || N == 0 || tg == 0) return NOSPLAN_OK;
  // the snapshot is kept aside: the scorers' view (*0 copies, derived tables) follows the working state during the call
  const size_t sb = kind ? sizeof(MpsGpu) * tg : sizeof(uint4) * tg;
  DevBuf* b = c.b;
  // b[0..2]: the snapshot; b[3..5]: state after phase A; b[15], b[7], b[13]: best final state
  auto copy_state = [&](DevBuf& d0, DevBuf& d1, DevBuf& d2, bool from_working, bool orig) -> int {
    // direction: working -> buffers (from_working) or buffers -> working; orig selects the *0 arrays instead of the working ones
    void* s0 = kind ? (orig ? h->d_mps_gpu0.p : h->d_mps_gpu.p) : (orig ? h->d_mig_used0.p : h->d_mig_used.p);
    void* s1 = kind ? (orig ? h->d_mps_ulo0.p : h->d_mps_ulo.p) : (orig ? h->d_mig_free0.p : h->d_mig_free.p);
    void* s2 = kind ? (orig ? h->d_mps_uhi0.p : h->d_mps_uhi.p) : nullptr;
    const size_t z1 = sizeof(uint4) * tg;
    CU(d0.reserve(sb));
    CU(d1.reserve(z1));
    if (s2) CU(d2.reserve(z1));
    if (from_working) {
      CU(cudaMemcpyAsync(d0.p, s0, sb, cudaMemcpyDeviceToDevice, st));
      CU(cudaMemcpyAsync(d1.p, s1, z1, cudaMemcpyDeviceToDevice, st));
      if (s2) CU(cudaMemcpyAsync(d2.p, s2, z1, cudaMemcpyDeviceToDevice, st));
    } else {
      CU(cudaMemcpyAsync(s0, d0.p, sb, cudaMemcpyDeviceToDevice, st));
      CU(cudaMemcpyAsync(s1, d1.p, z1, cudaMemcpyDeviceToDevice, st));
      if (s2) CU(cudaMemcpyAsync(s2, d2.p, z1, cudaMemcpyDeviceToDevice, st));
    }
    return NOSPLAN_OK;
  };
  if ((rc = copy_state(b[0], b[1], b[2], true, true))) return rc;          // snapshot aside
  if ((rc = copy_state(b[0], b[1], b[2], false, false))) return rc;        // working := snapshot
  (kind ? h->mps_state_dirty : h->mig_state_dirty) = true;
  // restores the handle's snapshot view whatever happens below
  auto restore = [&]() -> int {
    int r2 = copy_state(b[0], b[1], b[2], false, true);
    if (r2) return r2;
    if (!kind) {
      mig_repack_kernel<<<(N * h->mig_gs + 127) / 128, 128, 0, st>>>(N, h->mig_gs, h->mig_npad, h->d_mig_off.as<int32_t>(),
                                                                     h->d_mig_scnt.as<uint8_t>(), h->d_mig_used0.as<uint4>(),
                                                                     h->d_mig_free0.as<uint4>(), h->d_mig_sused.as<uint4>(),
                                                                     h->d_mig_sfree.as<uint4>());
      if ((r2 = mig_derive_tables(h, N, h->mig_npad, h->mig_gs, st))) return r2;
    } else if ((r2 = mps_derive_tables(h, N, h->S, st))) {
      return r2;
    }
    CU(cudaStreamSynchronize(st));
    return NOSPLAN_OK;
  };
  // ---- phase A: what fits on the devices that are free now
  std::vector<int32_t> all(M), node_a(M, -1), gpu_a(M, -1);
  for (int i = 0; i < M; ++i) all[i] = i;
  if ((rc = batch_first_fit(c, all, node_a, gpu_a))) { restore(); return rc; }
  if ((rc = copy_state(b[3], b[4], b[5], true, false))) { restore(); return rc; }
  std::vector<int32_t> unplaced0;
  for (int i = 0; i < M; ++i)
    if (node_a[i] < 0) unplaced0.push_back(i);
  // ---- phase B, both profile orders from the same state; the better one is kept
  int best_sched = -1, best_order = 0, best_rounds = 0;
  std::vector<int32_t> best_node, best_gpu;
  std::vector<uint8_t> best_rep;
  CU(b[12].reserve(sizeof(int32_t) * (2 * kMaxSizes + 2)));
  for (int order = 0; order < 2; ++order) {
    if ((rc = copy_state(b[3], b[4], b[5], false, false))) { restore(); return rc; }
    std::vector<int32_t> unplaced = unplaced0, node_b(M, -1), gpu_b(M, -1);
    std::vector<uint8_t> rep(N, 0);
    CU(h->d_out[4].reserve(2 * (size_t)N));
    uint8_t* d_taken = h->d_out[4].as<uint8_t>();
    uint8_t* d_rep = d_taken + N;
    CU(cudaMemsetAsync(d_rep, 0, N, st));
    int rounds = 0;
    for (int round = 0; round < max_rounds && !unplaced.empty(); ++round) {
      std::vector<long long> lacking(K, 0);
      for (int i : unplaced) lacking[pod_key[i]] += pod_qty[i];
      std::vector<int32_t> off{0}, rkey, rqty, okey, oidx, olack;
      for (int k = 0; k < K; ++k)
        if (lacking[k] > 0) {
          rkey.push_back(k);
          rqty.push_back((int32_t)std::min<long long>(lacking[k], NOSPLAN_MAX_DEMAND_QTY));
          off.push_back((int32_t)rkey.size());
        }
      const int D = (int)rkey.size();
      for (int r = 0; r < D; ++r) {
        const int rr = order ? D - 1 - r : r;
        okey.push_back(rkey[rr]);
        oidx.push_back(rr);
        olack.push_back((int32_t)std::min<long long>(lacking[rkey[rr]], 0x7fffffff));
      }
      if ((rc = batch_sync_scorer_state(c))) { restore(); return rc; }
      // the batched scorer: every row against every node, full tables out
      if ((rc = upload(h, h->d_dem[0], off.data(), sizeof(int32_t) * (D + 1), st))) { restore(); return rc; }
      if ((rc = upload(h, h->d_dem[1], rkey.data(), sizeof(int32_t) * D, st))) { restore(); return rc; }
      if ((rc = upload(h, h->d_dem[2], rqty.data(), sizeof(int32_t) * D, st))) { restore(); return rc; }
      CU(h->d_out[0].reserve(sizeof(int32_t) * D));
      CU(h->d_out[1].reserve((size_t)D * N));
      BatchCommitArgs ba{};
      if (!kind) {
        CU(h->d_out[2].reserve((size_t)D * tg));
        rc = launch_mig_score(h, D, h->d_dem[0].as<int32_t>(), h->d_dem[1].as<int32_t>(), h->d_dem[2].as<int32_t>(), h->d_out[0].as<int32_t>(),
                              h->d_out[1].as<uint8_t>(), h->d_out[2].as<int8_t>(), st);
        ba.choice = h->d_out[2].as<int8_t>();
      } else {
        CU(h->d_out[2].reserve((size_t)D * tg * K));
        CU(h->d_out[3].reserve(sizeof(uint32_t) * (size_t)D * tg));
        rc = launch_mps_score(h, D, h->d_dem[0].as<int32_t>(), h->d_dem[1].as<int32_t>(), h->d_dem[2].as<int32_t>(), h->d_out[0].as<int32_t>(),
                              h->d_out[1].as<uint8_t>(), h->d_out[2].as<uint8_t>(), h->d_out[3].as<uint32_t>(), st);
        ba.out_free = h->d_out[2].as<uint8_t>();
        ba.out_mask = h->d_out[3].as<uint32_t>();
      }
      if (rc) { restore(); return rc; }
      // commit, profile by profile in this candidate's order
      if ((rc = upload(h, b[8], okey.data(), sizeof(int32_t) * D, st))) { restore(); return rc; }
      if ((rc = upload(h, b[9], oidx.data(), sizeof(int32_t) * D, st))) { restore(); return rc; }
      if ((rc = upload(h, b[10], olack.data(), sizeof(int32_t) * D, st))) { restore(); return rc; }
      CU(cudaMemsetAsync(d_taken, 0, N, st));
      ba.D = D;
      ba.N = N;
      ba.tg = tg;
      ba.n_keys = K;
      ba.row_key = b[8].as<int32_t>();
      ba.row_index = b[9].as<int32_t>();
      ba.row_lacking = b[10].as<int32_t>();
      ba.updated = h->d_out[1].as<uint8_t>();
      ba.gpu_off = kind ? h->d_mps_off.as<int32_t>() : h->d_mig_off.as<int32_t>();
      ba.T = h->d_tables.as<MigTables>();
      ba.m_model = h->d_mig_model.as<uint8_t>();
      ba.m_used = h->d_mig_used.as<uint4>();
      ba.m_free = h->d_mig_free.as<uint4>();
      ba.p_gpu = h->d_mps_gpu.as<MpsGpu>();
      ba.taken = d_taken;
      ba.repartitioned = d_rep;
      if (kind) batch_commit_kernel<true><<<1, 32, 0, st>>>(ba);
      else batch_commit_kernel<false><<<1, 32, 0, st>>>(ba);
      CU(cudaGetLastError());
      ++rounds;
      // the round's gains are locked in: the unplaced pods are first-fitted on the new state
      const size_t before = unplaced.size();
      if ((rc = batch_first_fit(c, unplaced, node_b, gpu_b))) { restore(); return rc; }
      std::vector<int32_t> still;
      for (int i : unplaced)
        if (node_b[i] < 0) still.push_back(i);
      unplaced.swap(still);
      if (unplaced.size() == before) break;
    }
    int sched = 0;
    for (int i = 0; i < M; ++i) sched += node_b[i] >= 0;
    if (sched > best_sched) {
      best_sched = sched;
      best_order = order;
      best_rounds = rounds;
      best_node = node_b;
      best_gpu = gpu_b;
      best_rep.resize(N);
      CU(cudaMemcpyAsync(best_rep.data(), d_rep, N, cudaMemcpyDeviceToHost, st));
      if ((rc = copy_state(b[15], b[7], b[13], true, false))) { restore(); return rc; }
      CU(cudaStreamSynchronize(st));
    }
  }
  if ((rc = copy_state(b[15], b[7], b[13], false, false))) { restore(); return rc; }  // working := the kept candidate's final state
  if ((rc = restore())) return rc;
  for (int i = 0; i < M; ++i) {
    if (node_a[i] >= 0) {
      sched_node[i] = node_a[i];
      sched_gpu[i] = gpu_a[i];
      phase[i] = 1;
    } else if (best_node[i] >= 0) {
      sched_node[i] = best_node[i];
      sched_gpu[i] = best_gpu[i];
      phase[i] = 2;
    }
  }
  for (int n = 0; n < N; ++n) repartitioned[n] = best_rep[n];
  if (order_out) *order_out = best_order;
  if (rounds_out) *rounds_out = best_rounds;
  return NOSPLAN_OK;
}

// ---- sharded scoring: node base + the first-node exchange over peer memory (nosplan_xchg.cuh) ---------------
namespace {
int xchg_release(nosplan_t* h) {
  auto& x = h->xch;
  if (!x.created) return NOSPLAN_OK;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  for (int r = 0; r < x.world; ++r)
    if (x.open && r != x.rank && x.mapped[r]) cudaIpcCloseMemHandle(x.mapped[r]);
  if (x.local) cudaFree(x.local);
  x.done.release();
  x.err.release();
  x = {};
  return NOSPLAN_OK;
}
}  // namespace

extern "C" int nosplan_set_node_base(nosplan_t* h, int mig_first_node, int mps_first_node) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  if (mig_first_node < 0 || mps_first_node < 0) return fail(h, NOSPLAN_ERR_ARG, "negative node base");
  h->mig_base = mig_first_node;
  h->mps_base = mps_first_node;
  return NOSPLAN_OK;
}

extern "C" int nosplan_exchange_create(nosplan_t* h, int rank, int world, int max_keys, unsigned char* handle_out) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  if (world < 1 || world > kMaxWorld || rank < 0 || rank >= world)
    return fail(h, NOSPLAN_ERR_RANGE, "rank %d / world %d out of range (max %d ranks)", rank, world, kMaxWorld);
  if (max_keys < 1) return fail(h, NOSPLAN_ERR_ARG, "max_keys must be positive");
  if (!handle_out) return fail(h, NOSPLAN_ERR_ARG, "handle_out is null");
  int rc = bind(h);
  if (rc) return rc;
  xchg_release(h);
  auto& x = h->xch;
  x.rank = rank;
  x.world = world;
  x.max_keys = max_keys;
  x.stride = (max_keys + 3) / 4 * 4;
  x.flags_off = ((size_t)2 * x.stride * sizeof(int32_t) + 255) / 256 * 256;
  const size_t bytes = x.flags_off + 256;  // pushed[16] + read[16] (+ padding)
  CU(cudaMalloc(&x.local, bytes));
  x.created = true;
  CU(x.done.reserve(2 * sizeof(unsigned int)));
  CU(x.err.reserve(sizeof(int)));
  CU(cudaMemsetAsync(x.local, 0xff, (size_t)2 * x.stride * sizeof(int32_t), h->stream));  // -1: no node
  CU(cudaMemsetAsync(static_cast<char*>(x.local) + x.flags_off, 0, 256, h->stream));
  CU(cudaMemsetAsync(x.done.p, 0, 2 * sizeof(unsigned int), h->stream));
  CU(cudaMemsetAsync(x.err.p, 0, sizeof(int), h->stream));
  CU(cudaStreamSynchronize(h->stream));  // at rest before any peer can learn the handle
  static_assert(sizeof(cudaIpcMemHandle_t) == NOSPLAN_IPC_HANDLE_BYTES, "IPC handle size");
  cudaIpcMemHandle_t ipc;
  std::memset(&ipc, 0, sizeof(ipc));
  if (world > 1) CU(cudaIpcGetMemHandle(&ipc, x.local));
  std::memcpy(handle_out, &ipc, sizeof(ipc));
  return NOSPLAN_OK;
}

extern "C" int nosplan_exchange_open(nosplan_t* h, const unsigned char* handles) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  auto& x = h->xch;
  if (!x.created) return fail(h, NOSPLAN_ERR_ARG, "nosplan_exchange_create must be called first");
  if (x.open) return fail(h, NOSPLAN_ERR_ARG, "exchange already open");
  if (!handles && x.world > 1) return fail(h, NOSPLAN_ERR_ARG, "handles is null");
  int rc = bind(h);
  if (rc) return rc;
  for (int r = 0; r < x.world; ++r) {
    if (r == x.rank) {
      x.mapped[r] = x.local;
    } else {
      cudaIpcMemHandle_t ipc;
      std::memcpy(&ipc, handles + (size_t)r * NOSPLAN_IPC_HANDLE_BYTES, sizeof(ipc));
      CU(cudaIpcOpenMemHandle(&x.mapped[r], ipc, cudaIpcMemLazyEnablePeerAccess));
    }
    x.peers.keys[r] = static_cast<int32_t*>(x.mapped[r]);
    x.peers.pushed[r] = reinterpret_cast<uint32_t*>(static_cast<char*>(x.mapped[r]) + x.flags_off);
    x.peers.read[r] = x.peers.pushed[r] + kMaxWorld;
  }
  x.open = true;
  x.step = 0;
  return NOSPLAN_OK;
}

extern "C" int nosplan_exchange_min_dev(nosplan_t* h, int n_keys, const int32_t* keys_in, int32_t* keys_out, void* stream) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  auto& x = h->xch;
  if (!x.open) return fail(h, NOSPLAN_ERR_ARG, "exchange is not open");
  if (x.broken) return fail(h, NOSPLAN_ERR_CUDA, "the exchange timed out earlier: destroy and re-create it on every rank");
  if (n_keys < 0 || n_keys > x.max_keys) return fail(h, NOSPLAN_ERR_RANGE, "%d keys exceed the exchange buffer (%d)", n_keys, x.max_keys);
  if (n_keys > 0 && (!keys_in || !keys_out)) return fail(h, NOSPLAN_ERR_ARG, "null key arrays");
  int rc = bind(h);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  NSP_RANGE("nosplan: first-node exchange");
  const uint32_t step = ++x.step;
  const int blocks = std::max(1, std::min((n_keys + 255) / 256, h->sm_count));
  xchg_publish_kernel<<<blocks, 256, 0, st>>>(x.peers, x.world, x.rank, x.stride, step, n_keys, keys_in, x.done.as<unsigned int>(),
                                              kXchgTimeoutNs, x.err.as<int>());
  xchg_finish_kernel<<<blocks, 256, 0, st>>>(x.peers, x.world, x.rank, x.stride, step, n_keys, keys_out, x.done.as<unsigned int>(),
                                             kXchgTimeoutNs, x.err.as<int>());
  CU(cudaGetLastError());
  return NOSPLAN_OK;
}

extern "C" int nosplan_exchange_status(nosplan_t* h) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  auto& x = h->xch;
  if (!x.created) return fail(h, NOSPLAN_ERR_ARG, "no exchange");
  int rc = bind(h);
  if (rc) return rc;
  int e = 0;
  CU(cudaMemcpy(&e, x.err.p, sizeof(int), cudaMemcpyDeviceToHost));
  if (e) {
    x.broken = true;
    return fail(h, NOSPLAN_ERR_CUDA, "exchange: rank %d never signalled a step (waited 2 s); the results of that pass read -2", e - 1);
  }
  return NOSPLAN_OK;
}

extern "C" int nosplan_exchange_destroy(nosplan_t* h) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  return xchg_release(h);
}

// ---- one call for a whole scoring pass: nosplan_score_snapshot ---------------------------------------------------
// Snapshot upload, packing, scheduling tables, row grouping, both scorers, the exchange (when one is open on the
// caller's handle), the winners' per-GPU choices and the copy back are all enqueued from the calling thread on a
// private child handle — MIG work on its first stream, MPS work on its second, joined where one needs the other —
// and the call waits exactly once, at the end.  The device's verdicts on the inputs (range checks of both snapshots
// and of the demand rows run inside the kernels) are read after that one wait.  Nothing is installed on the caller's handle.
struct Lanes {
  nosplan_t* pass = nullptr;
  uint64_t tab_version = 0;
  DevBuf d_demp[3];  // MPS demand rows (the MIG rows use the child's d_dem)
};

void lanes_release(nosplan_t* h) {
  Lanes* L = h->lanes;
  if (!L) return;
  for (auto& b : L->d_demp) b.release();
  if (L->pass) nosplan_destroy(L->pass);
  delete L;
  h->lanes = nullptr;
}

namespace {
int lanes_get(nosplan_t* h, Lanes** out) {
  if (!h->lanes) {
    Lanes* L = new Lanes();
    if (nosplan_create(h->device, &L->pass) != NOSPLAN_OK) {
      delete L;
      return fail(h, NOSPLAN_ERR_CUDA, "could not create the scoring lane");
    }
    h->lanes = L;
  }
  *out = h->lanes;
  return NOSPLAN_OK;
}
bool rows_single_key(const nosplan_demands_t* r) {
  const int32_t* off = r->off;
  int bad = 0;
  for (int d = 0; d < r->n_rows; ++d) bad |= (off[d + 1] - off[d]) ^ 1;
  return bad == 0;
}
int pin_reserve(nosplan_t* h, size_t bytes) {
  if (bytes <= h->h_pin_cap) return NOSPLAN_OK;
  if (h->h_pin) cudaFreeHost(h->h_pin);
  h->h_pin = nullptr;
  h->h_pin_cap = 0;
  CU(cudaHostAlloc(&h->h_pin, bytes, cudaHostAllocDefault));
  h->h_pin_cap = bytes;
  return NOSPLAN_OK;
}
}  // namespace

extern "C" int nosplan_score_snapshot(nosplan_t* h, const nosplan_mig_snapshot_t* mig, const nosplan_demands_t* mig_rows,
                                      int32_t* mig_first_node, int8_t* mig_first_choice, const nosplan_mps_snapshot_t* mps,
                                      const nosplan_demands_t* mps_rows, int32_t* mps_first_node, int64_t* n_evals) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  if (n_evals) *n_evals = 0;
  NSP_RANGE("nosplan_score_snapshot");
  if (mig && !h->have_tables) return fail(h, NOSPLAN_ERR_ARG, "nosplan_set_mig_tables must be called first");
  if ((mig && (!mig_rows || (mig_rows->n_rows > 0 && !mig_first_node))) || (mps && (!mps_rows || (mps_rows->n_rows > 0 && !mps_first_node))))
    return fail(h, NOSPLAN_ERR_ARG, "a snapshot needs its demand rows and an output array");
  Lanes* L = nullptr;
  int rc = lanes_get(h, &L);
  if (rc) return rc;
  nosplan_t* c = L->pass;
  auto child_fail = [&](int code, const char* who) { return fail(h, code, "%s%s", who, nosplan_last_error(c)); };
  if (mig && L->tab_version != h->tab_version) {  // the lane scores against the caller's tables
    std::vector<const char*> names;
    for (auto& n : h->names) names.push_back(n.c_str());
    rc = nosplan_set_mig_tables(c, h->P, names.data(), h->tab_models, h->tab_geom_off.data(), h->tab_geom_counts.data());
    if (rc) return child_fail(rc, "MIG: ");
    L->tab_version = h->tab_version;
  }
  c->err.clear();
  c->mig_base = h->mig_base;
  c->mps_base = h->mps_base;
  const int Dm = mig ? mig_rows->n_rows : 0, Dp = mps ? mps_rows->n_rows : 0, Dall = Dm + Dp;
  if (Dm < 0 || Dp < 0) return fail(h, NOSPLAN_ERR_ARG, "negative demand count");
  if (mig && (rc = check_demands_head(c, Dm, mig_rows->off, mig_rows->key, mig_rows->qty))) return child_fail(rc, "MIG: ");
  if (mps && (rc = check_demands_head(c, Dp, mps_rows->off, mps_rows->key, mps_rows->qty))) return child_fail(rc, "MPS: ");
  cudaStream_t sm = c->stream, sp = c->stream2;
  c->have_mig = c->have_mps = false;
  // NOSPLAN_TRACE=1: where a pass spends its time (host clock + CUDA events on the MIG stream)
  const bool trace = trace_on();
  static cudaEvent_t tev[5] = {};
  const auto t_begin = std::chrono::steady_clock::now();
  auto mark = [&](int i) {
    if (!trace) return;
    if (!tev[i]) cudaEventCreate(&tev[i]);
    cudaEventRecord(tev[i], sm);
  };
  mark(0);
  if (mig) {
    if ((rc = mig_snapshot_enqueue(c, mig->n_nodes, mig->gpu_off, mig->valid, mig->model, mig->used, mig->free_, sm))) return child_fail(rc, "MIG: ");
    if (Dm > 0) {
      const int nnz = mig_rows->off[Dm];
      if ((rc = upload(c, c->d_dem[0], mig_rows->off, sizeof(int32_t) * (Dm + 1), sm))) return child_fail(rc, "MIG: ");
      if ((rc = upload(c, c->d_dem[1], mig_rows->key, sizeof(int32_t) * std::max(nnz, 1), sm))) return child_fail(rc, "MIG: ");
      if ((rc = upload(c, c->d_dem[2], mig_rows->qty, sizeof(int32_t) * std::max(nnz, 1), sm))) return child_fail(rc, "MIG: ");
    }
  }
  if (mps) {
    if ((rc = mps_snapshot_enqueue(c, mps->n_sizes, mps->sizes_gb, mps->n_nodes, mps->gpu_off, mps->valid, mps->mem_gb, mps->used,
                                   mps->free_, mps->free_mask, sp)))
      return child_fail(rc, "MPS: ");
    if (Dp > 0) {
      const int nnz = mps_rows->off[Dp];
      if ((rc = upload(c, L->d_demp[0], mps_rows->off, sizeof(int32_t) * (Dp + 1), sp))) return child_fail(rc, "MPS: ");
      if ((rc = upload(c, L->d_demp[1], mps_rows->key, sizeof(int32_t) * std::max(nnz, 1), sp))) return child_fail(rc, "MPS: ");
      if ((rc = upload(c, L->d_demp[2], mps_rows->qty, sizeof(int32_t) * std::max(nnz, 1), sp))) return child_fail(rc, "MPS: ");
    }
    // the pass starts on the MIG stream: it needs the MPS snapshot and rows from here on
    if (cudaEventRecord(c->ev_join, sp) != cudaSuccess || cudaStreamWaitEvent(sm, c->ev_join, 0) != cudaSuccess)
      return fail(h, NOSPLAN_ERR_CUDA, "stream join failed");
  }
  mark(1);
  const int gs = mig ? c->mig_gs : 1;
  const size_t out_bytes = sizeof(int32_t) * (size_t)std::max(Dall, 1) + (mig_first_choice ? (size_t)Dm * gs : 0);
  if ((rc = pin_reserve(c, out_bytes))) return child_fail(rc, "");
  if (Dall > 0) {
    if (cudaSuccess != c->d_out[0].reserve(sizeof(int32_t) * (size_t)Dall)) return fail(h, NOSPLAN_ERR_CUDA, "out of device memory");
    // the snapshots are confirmed after the one wait at the end; until then the kernels run on what was uploaded
    // (every value is masked or clamped while it is packed, so unchecked inputs cannot take a kernel out of bounds)
    c->have_mig = mig != nullptr;
    c->have_mps = mps != nullptr;
    int flags = 0;
    if ((!mig || rows_single_key(mig_rows)) && (!mps || rows_single_key(mps_rows))) flags |= NOSPLAN_PASS_SINGLE_KEY;
    rc = score_pass_enqueue(c, h, Dm, c->d_dem[0].as<int32_t>(), c->d_dem[1].as<int32_t>(), c->d_dem[2].as<int32_t>(), Dp,
                            L->d_demp[0].as<int32_t>(), L->d_demp[1].as<int32_t>(), L->d_demp[2].as<int32_t>(), c->d_out[0].as<int32_t>(),
                            flags, sm);
    c->have_mig = c->have_mps = false;
    if (rc) return child_fail(rc, "");
    mark(2);
    if (mig_first_choice && Dm > 0) {
      if (cudaSuccess != c->d_out[3].reserve((size_t)Dm * gs)) return fail(h, NOSPLAN_ERR_CUDA, "out of device memory");
      if (mig_ordered(c, Dm)) {
        mig_finalize16_kernel<<<(Dm + 127) / 128, 128, 0, sm>>>(
            c->d_tables.as<MigTables>(), c->d_mig_rankg.as<uint32_t>(), c->d_mig_ufg.as<uint16_t>(), c->d_mig_used0.as<uint4>(),
            c->d_mig_free0.as<uint4>(), c->d_mig_model.as<uint8_t>(), c->d_mig_off.as<int32_t>(), Dm, gs, (c->cmax + 3) / 4,
            c->d_dem[0].as<int32_t>(), c->d_dem[1].as<int32_t>(), c->d_dem[2].as<int32_t>(), c->d_out[0].as<int32_t>(),
            c->d_out[3].as<int8_t>(), c->mig_base, c->mig_n, c->mig_npad);
      } else {
        mig_finalize_kernel<<<(Dm + 127) / 128, 128, 0, sm>>>(
            c->d_tables.as<MigTables>(), c->d_mig_used0.as<uint4>(), c->d_mig_free0.as<uint4>(), c->d_mig_model.as<uint8_t>(),
            c->d_mig_off.as<int32_t>(), Dm, gs, c->d_dem[0].as<int32_t>(), c->d_dem[1].as<int32_t>(), c->d_dem[2].as<int32_t>(),
            c->d_out[0].as<int32_t>(), c->d_out[3].as<int8_t>(), c->mig_base, c->mig_n);
      }
    }
    char* pin = static_cast<char*>(c->h_pin);
    if (cudaMemcpyAsync(pin, c->d_out[0].p, sizeof(int32_t) * (size_t)Dall, cudaMemcpyDeviceToHost, sm) != cudaSuccess)
      return fail(h, NOSPLAN_ERR_CUDA, "copy back failed");
    if (mig_first_choice && Dm > 0 &&
        cudaMemcpyAsync(pin + sizeof(int32_t) * (size_t)Dall, c->d_out[3].p, (size_t)Dm * gs, cudaMemcpyDeviceToHost, sm) != cudaSuccess)
      return fail(h, NOSPLAN_ERR_CUDA, "copy back failed");
    if (c->d_rowstatus.p && cudaMemcpyAsync(c->h_status + 4, c->d_rowstatus.p, 16, cudaMemcpyDeviceToHost, sm) != cudaSuccess)
      return fail(h, NOSPLAN_ERR_CUDA, "copy back failed");
  }
  c->h_status[6] = 0;
  if (h->xch.open && cudaMemcpyAsync(c->h_status + 6, h->xch.err.p, sizeof(int), cudaMemcpyDeviceToHost, sm) != cudaSuccess)
    return fail(h, NOSPLAN_ERR_CUDA, "copy back failed");
  mark(3);
  const auto t_enq = std::chrono::steady_clock::now();
  if (cudaStreamSynchronize(sm) != cudaSuccess || cudaStreamSynchronize(sp) != cudaSuccess)
    return fail(h, NOSPLAN_ERR_CUDA, "scoring pass failed: %s", cudaGetErrorString(cudaGetLastError()));
  const auto t_sync = std::chrono::steady_clock::now();
  // the device's verdicts
  if (mig && (rc = mig_snapshot_finish(c, mig->model, mig->used, mig->free_))) return child_fail(rc, "MIG: ");
  if (mps && (rc = mps_snapshot_finish(c, mps->sizes_gb, mps->mem_gb, mps->used, mps->free_))) return child_fail(rc, "MPS: ");
  c->have_mig = c->have_mps = false;  // the lane keeps no snapshot between calls
  if (Dall > 0 && c->d_rowstatus.p) {
    const unsigned long long cm = row_status_code(c, 0, c->h_status[4]), cp = row_status_code(c, 1, c->h_status[5]);
    if (Dm > 0 && cm != ~0ull) return child_fail(describe_bad_row(c, cm, mig_rows->off, mig_rows->key, mig_rows->qty, c->P), "MIG: ");
    if (Dp > 0 && cp != ~0ull) return child_fail(describe_bad_row(c, cp, mps_rows->off, mps_rows->key, mps_rows->qty, c->S), "MPS: ");
  }
  if ((int)c->h_status[6] != 0) {
    h->xch.broken = true;
    return fail(h, NOSPLAN_ERR_CUDA, "exchange: rank %d never signalled a step (waited 2 s)", (int)c->h_status[6] - 1);
  }
  const char* pin = static_cast<const char*>(c->h_pin);
  if (Dm > 0) std::memcpy(mig_first_node, pin, sizeof(int32_t) * (size_t)Dm);
  if (Dp > 0) std::memcpy(mps_first_node, pin + sizeof(int32_t) * (size_t)Dm, sizeof(int32_t) * (size_t)Dp);
  if (mig_first_choice && Dm > 0) std::memcpy(mig_first_choice, pin + sizeof(int32_t) * (size_t)Dall, (size_t)Dm * gs);
  if (n_evals) *n_evals = (int64_t)Dm * c->mig_evals + (int64_t)Dp * c->mps_evals;
  if (trace && Dall > 0) {
    auto us = [](auto a, auto b) { return std::chrono::duration<double, std::micro>(b - a).count(); };
    float g01 = 0, g12 = 0, g23 = 0;
    cudaEventElapsedTime(&g01, tev[0], tev[1]);
    cudaEventElapsedTime(&g12, tev[1], tev[2]);
    cudaEventElapsedTime(&g23, tev[2], tev[3]);
    std::fprintf(stderr, "nosplan_score_snapshot: host enqueue %.0f us, wait %.0f us, copy out %.0f us | device: uploads + packing %.0f us, "
                 "grouping + scoring + exchange %.0f us, choices + copy back %.0f us\n", us(t_begin, t_enq), us(t_enq, t_sync),
                 us(t_sync, std::chrono::steady_clock::now()), g01 * 1e3, g12 * 1e3, g23 * 1e3);
  }
  return NOSPLAN_OK;
}

// ---- presence counts (the gate's ingredient; per shard when the node list is sharded) ------------------------
namespace {
template <bool MPS>
__global__ void presence_kernel(int tg, int n_keys, const int32_t* __restrict__ gpu_node, const uint4* __restrict__ m_used,
                                const uint4* __restrict__ m_free, const MpsGpu* __restrict__ p_gpu,
                                const uint4* __restrict__ p_ulo, const uint4* __restrict__ p_uhi,
                                unsigned long long* __restrict__ count) {
  __shared__ unsigned int sh[kMaxSizes];
  if (threadIdx.x < kMaxSizes) sh[threadIdx.x] = 0;
  __syncthreads();
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g < tg && gpu_node[g] >= 0) {  // GPUs of invalid nodes are skipped, as the controller skips those nodes
    for (int key = 0; key < n_keys; ++key) {
      unsigned int v;
      if constexpr (MPS) v = mps_byte(p_gpu[g].f0, p_gpu[g].f1, key) + mps_byte(p_ulo[g], p_uhi[g], key);
      else v = byte_of(m_free[g], key) + byte_of(m_used[g], key);
      if (v) atomicAdd(&sh[key], v);
    }
  }
  __syncthreads();
  if (threadIdx.x < n_keys && sh[threadIdx.x]) atomicAdd(&count[threadIdx.x], (unsigned long long)sh[threadIdx.x]);
}
}  // namespace

extern "C" int nosplan_presence(nosplan_t* h, int kind, int64_t* count) {
  if (!h) return NOSPLAN_ERR_ARG;
  h->err.clear();
  if (!count) return fail(h, NOSPLAN_ERR_ARG, "count is null");
  if (kind != 0 && kind != 1) return fail(h, NOSPLAN_ERR_ARG, "kind must be 0 (MIG) or 1 (MPS)");
  if (kind ? !h->have_mps : !h->have_mig) return fail(h, NOSPLAN_ERR_ARG, "no snapshot of that kind");
  int rc = bind(h);
  if (rc) return rc;
  const int n_keys = kind ? h->S : h->P, tg = kind ? h->mps_tg : h->mig_tg;
  CU(h->d_out[5].reserve(sizeof(unsigned long long) * kMaxSizes));
  CU(cudaMemsetAsync(h->d_out[5].p, 0, sizeof(unsigned long long) * kMaxSizes, h->stream));
  if (tg > 0) {
    if (kind) presence_kernel<true><<<(tg + 255) / 256, 256, 0, h->stream>>>(
        tg, n_keys, h->d_mps_gnode.as<int32_t>(), nullptr, nullptr, h->d_mps_gpu.as<MpsGpu>(), h->d_mps_ulo.as<uint4>(),
        h->d_mps_uhi.as<uint4>(), h->d_out[5].as<unsigned long long>());
    else presence_kernel<false><<<(tg + 255) / 256, 256, 0, h->stream>>>(
        tg, n_keys, h->d_mig_gnode.as<int32_t>(), h->d_mig_used.as<uint4>(), h->d_mig_free.as<uint4>(), nullptr, nullptr, nullptr,
        h->d_out[5].as<unsigned long long>());
    CU(cudaGetLastError());
  }
  CU(cudaMemcpyAsync(count, h->d_out[5].p, sizeof(int64_t) * n_keys, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return NOSPLAN_OK;
}
