// nos_oracle.cpp — CPU restatement of the walkai-nos GPU-partitioning planner hot path.
//
// THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py may load it.  The product (libnosplan.so) never
// links, loads or calls anything in this directory.
//
// Parity status: the per-GPU and per-node functions are PINNED by the reference's own table tests
// (transcribed to tests/golden/*.json, checked by tests/test_oracle_kat.py).  The cluster-level
// composition orc_plan() restates the reference controller (mig_controller.go:56-198) and the
// SURVEY §A.5 composition; the reference has 0 tests for that package, so that level is
// "parity unpinned" (it can only be pinned against the functions it is composed of).
// The reference is Go and no Go toolchain exists in this image, so the reference itself cannot be
// compiled here; every function below cites the reference file:line (relative to the reference
// root) that it follows.
//
// Go maps are restated as PMap: a dense value array plus a "key present" bitmask, so zero-valued
// keys, deletion and len() behave like the Go maps.  Iteration over a PMap is ascending key id; ids
// are assigned in the canonical order of SURVEY §D (MIG: ascending profile-name string; MPS:
// ascending slice size), which is how Go's unordered map iteration is canonicalised.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <string>
#include <thread>
#include <vector>

namespace {

constexpr int PMAX = 32;

struct PMap {
  int32_t v[PMAX];
  uint32_t mask;
  PMap() : mask(0) { std::memset(v, 0, sizeof(v)); }
  bool has(int p) const { return (mask >> p) & 1u; }
  int get(int p) const { return has(p) ? v[p] : 0; }  // Go: missing key reads as 0
  void set(int p, int x) { v[p] = x; mask |= (1u << p); }
  void del(int p) { v[p] = 0; mask &= ~(1u << p); }
  int len() const { return __builtin_popcount(mask); }
  bool operator==(const PMap& o) const {
    if (mask != o.mask) return false;
    for (int p = 0; p < PMAX; ++p)
      if (has(p) && v[p] != o.v[p]) return false;
    return true;
  }
};

static PMap pmap_from(const int32_t* vals, uint32_t mask, int n) {
  PMap m;
  for (int p = 0; p < n; ++p)
    if ((mask >> p) & 1u) m.set(p, vals[p]);
  return m;
}
static void pmap_to(const PMap& m, int32_t* vals, uint32_t* mask, int n) {
  for (int p = 0; p < n; ++p) vals[p] = m.has(p) ? m.v[p] : 0;
  *mask = m.mask;
}

struct Ctx {
  // MIG vocabulary: profile names, strictly ascending bytewise (canonical order)
  std::vector<std::string> mig_names;
  // model -> ordered list of allowed geometries (pkg/gpu/mig/known_configs.go:24-142)
  std::vector<std::vector<PMap>> models;
  std::vector<std::vector<std::string>> model_ids;  // Geometry.Id() of each, precomputed
  // MPS vocabulary: slice sizes in GB, strictly ascending (profile "<k>gb", slicing/profile.go:48-59)
  std::vector<int> mps_sizes;
  std::string err;
};

// ---------------------------------------------------------------------------------------------
// pkg/gpu/partitioning.go:37-56  Geometry.Id()/String(): profiles sorted by name, "%s:%d, " each.
static std::string geometry_id(const Ctx& c, const PMap& g) {
  std::string s;
  for (int p = 0; p < (int)c.mig_names.size(); ++p)
    if (g.has(p)) {
      s += c.mig_names[p];
      s += ":";
      s += std::to_string(g.v[p]);
      s += ", ";
    }
  return s;
}

// pkg/gpu/partitioning.go:67-79  GetFewestSlicesGeometry: fewest distinct kinds, first wins on ties.
static int fewest_slices(const std::vector<PMap>& geoms) {
  int min = -1;
  for (int i = 0; i < (int)geoms.size(); ++i) {
    if (min < 0) { min = i; continue; }
    if (geoms[i].len() < geoms[min].len()) min = i;
  }
  return min;
}

// ---------------------------------------------------------------------------------------------
// MIG GPU  (pkg/gpu/mig/gpu.go)
struct MigGPU {
  int model;
  PMap used, free;
};

// gpu.go:84-95 GetGeometry
static PMap mig_get_geometry(const MigGPU& g) {
  PMap r;
  for (int p = 0; p < PMAX; ++p)
    if (g.used.has(p)) r.set(p, r.get(p) + g.used.v[p]);
  for (int p = 0; p < PMAX; ++p)
    if (g.free.has(p)) r.set(p, r.get(p) + g.free.v[p]);
  return r;
}

// gpu.go:271-278 AllowsGeometry (cmp.Equal on maps: same key set, same values)
static bool mig_allows(const Ctx& c, const MigGPU& g, const PMap& geom) {
  for (const PMap& a : c.models[g.model])
    if (a == geom) return true;
  return false;
}

// gpu.go:99-112 CanApplyGeometry
static bool mig_can_apply(const Ctx& c, const MigGPU& g, const PMap& geom) {
  if (!mig_allows(c, g, geom)) return false;
  for (int p = 0; p < PMAX; ++p)
    if (g.used.has(p) && geom.get(p) < g.used.v[p]) return false;
  return true;
}

// gpu.go:134-154 ApplyGeometry; returns false where the reference returns an error
static bool mig_apply(const Ctx& c, MigGPU& g, const PMap& geom) {
  if (!mig_can_apply(c, g, geom)) return false;
  for (int p = 0; p < PMAX; ++p)
    if (geom.has(p)) g.free.set(p, geom.v[p] - g.used.get(p));
  for (int p = 0; p < PMAX; ++p)
    if (g.free.has(p) && !geom.has(p)) g.free.del(p);
  return true;
}

// gpu.go:120-129 InitGeometry
static bool mig_init_geometry(const Ctx& c, MigGPU& g) {
  const auto& allowed = c.models[g.model];
  int i = fewest_slices(allowed);
  PMap largest = i >= 0 ? allowed[i] : PMap();  // nil geometry when the list is empty
  if (!mig_can_apply(c, g, largest)) return false;
  return mig_apply(c, g, largest);
}

// gpu.go:198-223 countProvidedProfiles
static int mig_count_provided(const MigGPU& g, const PMap& cand, const PMap& req, const PMap& cur) {
  int provided = 0;
  for (int p = 0; p < PMAX; ++p) {
    if (!req.has(p)) continue;
    int rq = req.v[p];
    if (g.free.get(p) >= rq) continue;
    int needed = rq - g.free.get(p);
    if (needed <= 0) continue;
    int add = cand.get(p) - cur.get(p);
    if (add <= 0) continue;
    provided += std::min(add, needed);
  }
  return provided;
}

// gpu.go:245-260 geometryDistance
static int mig_distance(const PMap& cur, const PMap& cand) {
  int sum = 0;
  for (int p = 0; p < PMAX; ++p)
    if (cur.has(p)) sum += std::abs(cand.get(p) - cur.v[p]);
  for (int p = 0; p < PMAX; ++p)
    if (cand.has(p) && !cur.has(p)) sum += std::abs(cand.v[p]);
  return sum;
}

// gpu.go:262-268 totalSliceCount
static int mig_total(const PMap& g) {
  int s = 0;
  for (int p = 0; p < PMAX; ++p)
    if (g.has(p)) s += g.v[p];
  return s;
}

struct Score {
  int provided, total, distance;
  const std::string* id;
};

// gpu.go:232-243 isBetterGeometryScore
static bool is_better(const Score& a, const Score& b) {
  if (a.provided != b.provided) return a.provided > b.provided;
  if (a.total != b.total) return a.total > b.total;
  if (a.distance != b.distance) return a.distance < b.distance;
  return *a.id < *b.id;  // Go string '<' is bytewise, as std::string's
}

// gpu.go:160-196 UpdateGeometryFor. *choice = index of the applied geometry in the model's list, -1 if none.
static bool mig_gpu_update(const Ctx& c, MigGPU& g, const PMap& req, int* choice) {
  PMap cur = mig_get_geometry(g);
  int best = -1;
  Score bs{0, 0, 0, nullptr};
  const auto& allowed = c.models[g.model];
  for (int ci = 0; ci < (int)allowed.size(); ++ci) {
    const PMap& cand = allowed[ci];
    if (!mig_can_apply(c, g, cand)) continue;
    int provided = mig_count_provided(g, cand, req, cur);
    if (provided <= 0) continue;
    Score sc{provided, mig_total(cand), mig_distance(cur, cand), &c.model_ids[g.model][ci]};
    if (best < 0 || is_better(sc, bs)) {
      best = ci;
      bs = sc;
    }
  }
  if (choice) *choice = best;
  if (best < 0) return false;
  mig_apply(c, g, allowed[best]);
  return true;
}

// gpu.go:289-303 AddPod: incremental mutation, error keeps earlier deductions. Ascending profile order (§D).
static bool mig_gpu_add_pod(MigGPU& g, const PMap& req) {
  for (int p = 0; p < PMAX; ++p) {
    if (!req.has(p)) continue;
    int q = req.v[p];
    if (g.free.get(p) < q) return false;
    g.free.set(p, g.free.get(p) - q);
    g.used.set(p, g.used.get(p) + q);
  }
  return true;
}

// ---------------------------------------------------------------------------------------------
// MIG Node (pkg/gpu/mig/node.go)

// node.go:145-177 UpdateGeometryFor; choices[i] = applied geometry index per GPU (-1 none)
static bool mig_node_update(const Ctx& c, MigGPU* gpus, int n, const PMap& slices, int* choices) {
  for (int i = 0; i < n; ++i)
    if (choices) choices[i] = -1;
  if (n == 0) return false;
  if (slices.len() == 0) return false;
  PMap required = slices;
  bool any = false;
  for (int i = 0; i < n; ++i) {
    int ch = -1;
    bool updated = mig_gpu_update(c, gpus[i], required, &ch);
    if (choices) choices[i] = ch;
    any = any || updated;
    for (int p = 0; p < PMAX; ++p) {
      if (!gpus[i].free.has(p)) continue;
      required.set(p, required.get(p) - gpus[i].free.v[p]);
      if (required.v[p] <= 0) required.del(p);
    }
  }
  return any;
}

// node.go:201-210 AddPod: first GPU (slice order) whose GPU.AddPod succeeds; returns slot or -1.
static int mig_node_add_pod(MigGPU* gpus, int n, const PMap& req) {
  for (int i = 0; i < n; ++i)
    if (mig_gpu_add_pod(gpus[i], req)) return i;
  return -1;
}

// node.go:122-139 HasFreeCapacity
static bool mig_node_has_free_capacity(const Ctx& c, const MigGPU* gpus, int n) {
  if (n == 0) return false;
  for (int i = 0; i < n; ++i) {
    if (gpus[i].free.len() > 0) return true;  // gpu.go:305-307 HasFreeMigDevices
    if (!mig_allows(c, gpus[i], mig_get_geometry(gpus[i]))) return true;
  }
  return false;
}

// ---------------------------------------------------------------------------------------------
// MPS ("slicing") GPU (pkg/gpu/slicing/gpu.go). Keys are ids into Ctx::mps_sizes.
struct MpsGPU {
  int mem;
  PMap used, free;
};
constexpr int kMinSliceMemoryGB = 1;  // slicing/constant.go:24

// gpu.go:67-97 Validate
static bool mps_validate(const Ctx& c, const MpsGPU& g) {
  int tot = 0;
  for (int p = 0; p < PMAX; ++p)
    if (g.used.has(p)) {
      if (c.mps_sizes[p] < kMinSliceMemoryGB) return false;
      tot += c.mps_sizes[p] * g.used.v[p];
    }
  for (int p = 0; p < PMAX; ++p)
    if (g.free.has(p)) {
      if (c.mps_sizes[p] < kMinSliceMemoryGB) return false;
      tot += c.mps_sizes[p] * g.free.v[p];
    }
  return tot <= g.mem;
}

// gpu.go:256-265 getTotSlicesMemory
static int mps_tot(const Ctx& c, const MpsGPU& g) {
  int t = 0;
  for (int p = 0; p < PMAX; ++p) {
    if (g.used.has(p)) t += c.mps_sizes[p] * g.used.v[p];
    if (g.free.has(p)) t += c.mps_sizes[p] * g.free.v[p];
  }
  return t;
}
// gpu.go:250-254 canCreateMoreSlices
static bool mps_can_create_more(const Ctx& c, const MpsGPU& g) { return g.mem - mps_tot(c, g) >= kMinSliceMemoryGB; }
// gpu.go:238-247 createSlices (all-or-nothing); key = NewProfile(sizeGb) i.e. the id of that size
static bool mps_create(const Ctx& c, MpsGPU& g, int id, int num) {
  int spare = g.mem - mps_tot(c, g);
  if (spare < c.mps_sizes[id] * num) return false;
  g.free.set(id, g.free.get(id) + num);
  return true;
}

// gpu.go:162-220 UpdateGeometryFor (with getMissingSlices :222-232)
static bool mps_gpu_update(const Ctx& c, MpsGPU& g, const PMap& req) {
  PMap missing;
  for (int p = 0; p < PMAX; ++p) {
    if (!req.has(p)) continue;
    int diff = req.v[p] - g.free.get(p);
    if (diff > 0) missing.set(p, diff);
  }
  if (missing.len() == 0) return false;
  bool updated = false;
  PMap orig = g.free;  // util.CopyMap, taken once
  // ascending size = ascending id (stable; canonical names only, so no equal sizes)
  for (int s = 0; s < PMAX; ++s) {
    if (!missing.has(s)) continue;
    if (mps_can_create_more(c, g)) {
      int q = missing.v[s];
      for (int i = 0; i < q; ++i) {
        if (!mps_create(c, g, s, 1)) break;
        missing.v[s]--;
        updated = true;
      }
    }
    int q = missing.v[s];
    for (int k = 0; k < PMAX; ++k)
      if (orig.has(k)) g.free.del(k);
    for (int i = 0; i < q; ++i) {
      if (!mps_can_create_more(c, g)) break;
      if (!mps_create(c, g, s, 1)) break;
      missing.v[s]--;
      updated = true;
    }
    for (int k = 0; k < PMAX; ++k)  // restore, ascending size (§D), result ignored
      if (orig.has(k)) (void)mps_create(c, g, k, orig.v[k]);
  }
  return updated;
}

// gpu.go:142-156 AddPod
static bool mps_gpu_add_pod(MpsGPU& g, const PMap& req) {
  for (int p = 0; p < PMAX; ++p) {
    if (!req.has(p)) continue;
    int q = req.v[p];
    if (g.free.get(p) < q) return false;
    g.free.set(p, g.free.get(p) - q);
    g.used.set(p, g.used.get(p) + q);
  }
  return true;
}

// slicing/node.go:120-152 UpdateGeometryFor
static bool mps_node_update(const Ctx& c, MpsGPU* gpus, int n, const PMap& slices, uint8_t* gpu_updated) {
  for (int i = 0; i < n; ++i)
    if (gpu_updated) gpu_updated[i] = 0;
  if (n == 0) return false;
  if (slices.len() == 0) return false;
  PMap required = slices;
  bool any = false;
  for (int i = 0; i < n; ++i) {
    bool updated = mps_gpu_update(c, gpus[i], required);
    if (gpu_updated) gpu_updated[i] = updated;
    any = any || updated;
    for (int p = 0; p < PMAX; ++p) {
      if (!gpus[i].free.has(p)) continue;
      required.set(p, required.get(p) - gpus[i].free.v[p]);
      if (required.v[p] <= 0) required.del(p);
    }
  }
  return any;
}

// slicing/node.go:196-205 AddPod
static int mps_node_add_pod(MpsGPU* gpus, int n, const PMap& req) {
  for (int i = 0; i < n; ++i)
    if (mps_gpu_add_pod(gpus[i], req)) return i;
  return -1;
}

// slicing/gpu.go:129-134 + node.go:208-215 HasFreeCapacity
static bool mps_node_has_free_capacity(const Ctx& c, const MpsGPU* gpus, int n) {
  for (int i = 0; i < n; ++i) {
    if (gpus[i].free.len() > 0) return true;
    if (mps_can_create_more(c, gpus[i])) return true;
  }
  return false;
}

// ---------------------------------------------------------------------------------------------
// strconv.Atoi: optional sign, decimal digits only, no spaces, error on overflow/empty.
static bool go_atoi(const std::string& s, long long* out) {
  size_t i = 0;
  bool neg = false;
  if (s.empty()) return false;
  if (s[0] == '+' || s[0] == '-') {
    neg = s[0] == '-';
    i = 1;
  }
  if (i >= s.size()) return false;
  long long v = 0;
  for (; i < s.size(); ++i) {
    if (s[i] < '0' || s[i] > '9') return false;
    if (v > (INT64_MAX - (s[i] - '0')) / 10) return false;
    v = v * 10 + (s[i] - '0');
  }
  *out = neg ? -v : v;
  return true;
}
static std::vector<std::string> split_dash(const std::string& s) {
  std::vector<std::string> parts;
  size_t b = 0;
  for (;;) {
    size_t e = s.find('-', b);
    if (e == std::string::npos) {
      parts.push_back(s.substr(b));
      break;
    }
    parts.push_back(s.substr(b, e - b));
    b = e + 1;
  }
  return parts;
}
static bool has_prefix(const std::string& s, const char* p) { return s.rfind(p, 0) == 0; }
static std::string lower(std::string s) {
  for (auto& ch : s)
    if (ch >= 'A' && ch <= 'Z') ch = ch - 'A' + 'a';
  return s;
}

}  // namespace

// =============================================================================================
// C API (ctypes).  Maps travel as (int32 values[n], uint32 presence-mask).
extern "C" {

typedef struct orc orc_t;
struct orc {
  Ctx c;
};

orc_t* orc_new(void) { return new orc(); }
void orc_free(orc_t* o) { delete o; }
const char* orc_error(orc_t* o) { return o->c.err.c_str(); }

int orc_set_mig_vocab(orc_t* o, int n, const char* const* names) {
  if (n < 0 || n > PMAX) { o->c.err = "too many MIG profiles"; return -1; }
  o->c.mig_names.clear();
  for (int i = 0; i < n; ++i) {
    std::string s = names[i];
    if (i > 0 && !(o->c.mig_names.back() < s)) { o->c.err = "MIG vocabulary must be strictly ascending"; return -1; }
    o->c.mig_names.push_back(s);
  }
  return 0;
}

// pkg/gpu/mig/known_configs.go:144-185 SetKnownGeometries + ValidateConfigs (quantity >= 1, so a
// key is present iff its count is non-zero; negative counts are rejected like the reference does).
int orc_set_mig_models(orc_t* o, int n_models, const int32_t* geom_off, const int32_t* counts) {
  int P = (int)o->c.mig_names.size();
  if (n_models == 0) { o->c.err = "no known configs provided"; return -1; }
  std::vector<std::vector<PMap>> models(n_models);
  for (int m = 0; m < n_models; ++m)
    for (int gi = geom_off[m]; gi < geom_off[m + 1]; ++gi) {
      PMap g;
      for (int p = 0; p < P; ++p) {
        int q = counts[(size_t)gi * P + p];
        if (q < 0) { o->c.err = "invalid quantity"; return -1; }
        if (q > 0) g.set(p, q);
      }
      models[m].push_back(g);
    }
  o->c.models = models;
  o->c.model_ids.assign(n_models, {});
  for (int m = 0; m < n_models; ++m)
    for (const PMap& g : models[m]) o->c.model_ids[m].push_back(geometry_id(o->c, g));
  return 0;
}

int orc_set_mps_sizes(orc_t* o, int n, const int32_t* sizes) {
  if (n < 0 || n > PMAX) { o->c.err = "too many MPS sizes"; return -1; }
  o->c.mps_sizes.assign(sizes, sizes + n);
  for (int i = 1; i < n; ++i)
    if (!(sizes[i - 1] < sizes[i])) { o->c.err = "MPS sizes must be strictly ascending"; return -1; }
  return 0;
}

int orc_geometry_id(orc_t* o, const int32_t* counts, uint32_t mask, char* buf, int buflen) {
  std::string s = geometry_id(o->c, pmap_from(counts, mask, (int)o->c.mig_names.size()));
  std::snprintf(buf, buflen, "%s", s.c_str());
  return (int)s.size();
}

int orc_model_geometry_id(orc_t* o, int model, int ci, char* buf, int buflen) {
  const std::string& s = o->c.model_ids[model][ci];
  std::snprintf(buf, buflen, "%s", s.c_str());
  return (int)s.size();
}

int orc_fewest_slices(int n, const uint32_t* masks) {
  std::vector<PMap> g(n);
  for (int i = 0; i < n; ++i) g[i].mask = masks[i];
  return fewest_slices(g);
}

// ---- MIG, single GPU -------------------------------------------------------------------------
static MigGPU load_mig(orc_t* o, int model, const int32_t* used, uint32_t um, const int32_t* fr, uint32_t fm) {
  int P = (int)o->c.mig_names.size();
  MigGPU g;
  g.model = model;
  g.used = pmap_from(used, um, P);
  g.free = pmap_from(fr, fm, P);
  return g;
}

int orc_mig_gpu_update(orc_t* o, int model, const int32_t* used, uint32_t used_mask, int32_t* fr, uint32_t* free_mask,
                       const int32_t* req, uint32_t req_mask, int32_t* choice) {
  int P = (int)o->c.mig_names.size();
  MigGPU g = load_mig(o, model, used, used_mask, fr, *free_mask);
  int ch = -1;
  bool u = mig_gpu_update(o->c, g, pmap_from(req, req_mask, P), &ch);
  pmap_to(g.free, fr, free_mask, P);
  if (choice) *choice = ch;
  return u;
}

// returns 0 on success, 1 where the reference returns an error
int orc_mig_gpu_apply_geometry(orc_t* o, int model, const int32_t* used, uint32_t used_mask, int32_t* fr,
                               uint32_t* free_mask, const int32_t* geom, uint32_t geom_mask) {
  int P = (int)o->c.mig_names.size();
  MigGPU g = load_mig(o, model, used, used_mask, fr, *free_mask);
  bool ok = mig_apply(o->c, g, pmap_from(geom, geom_mask, P));
  pmap_to(g.free, fr, free_mask, P);
  return ok ? 0 : 1;
}

int orc_mig_gpu_init_geometry(orc_t* o, int model, const int32_t* used, uint32_t used_mask, int32_t* fr,
                              uint32_t* free_mask) {
  int P = (int)o->c.mig_names.size();
  MigGPU g = load_mig(o, model, used, used_mask, fr, *free_mask);
  bool ok = mig_init_geometry(o->c, g);
  pmap_to(g.free, fr, free_mask, P);
  return ok ? 0 : 1;
}

int orc_mig_gpu_add_pod(orc_t* o, int32_t* used, uint32_t* used_mask, int32_t* fr, uint32_t* free_mask,
                        const int32_t* req, uint32_t req_mask) {
  int P = (int)o->c.mig_names.size();
  MigGPU g = load_mig(o, 0, used, *used_mask, fr, *free_mask);
  bool ok = mig_gpu_add_pod(g, pmap_from(req, req_mask, P));
  pmap_to(g.used, used, used_mask, P);
  pmap_to(g.free, fr, free_mask, P);
  return ok ? 0 : 1;
}

// ---- MIG, node (G GPUs, row-major [G][P]) ------------------------------------------------------
static std::vector<MigGPU> load_mig_node(orc_t* o, int G, const int32_t* models, const int32_t* used,
                                         const uint32_t* um, const int32_t* fr, const uint32_t* fm) {
  int P = (int)o->c.mig_names.size();
  std::vector<MigGPU> v(G);
  for (int i = 0; i < G; ++i) v[i] = load_mig(o, models[i], used + (size_t)i * P, um[i], fr + (size_t)i * P, fm[i]);
  return v;
}
static void store_mig_node(orc_t* o, const std::vector<MigGPU>& v, int32_t* used, uint32_t* um, int32_t* fr,
                           uint32_t* fm) {
  int P = (int)o->c.mig_names.size();
  for (size_t i = 0; i < v.size(); ++i) {
    if (used) pmap_to(v[i].used, used + i * P, um + i, P);
    pmap_to(v[i].free, fr + i * P, fm + i, P);
  }
}

int orc_mig_node_update(orc_t* o, int G, const int32_t* models, const int32_t* used, const uint32_t* used_mask,
                        int32_t* fr, uint32_t* free_mask, const int32_t* req, uint32_t req_mask, int32_t* choices) {
  int P = (int)o->c.mig_names.size();
  auto v = load_mig_node(o, G, models, used, used_mask, fr, free_mask);
  std::vector<int> ch(G > 0 ? G : 1, -1);
  bool u = mig_node_update(o->c, v.data(), G, pmap_from(req, req_mask, P), ch.data());
  store_mig_node(o, v, nullptr, nullptr, fr, free_mask);
  if (choices)
    for (int i = 0; i < G; ++i) choices[i] = ch[i];
  return u;
}

// returns the GPU slot that took the pod, -1 if none (reference: error)
int orc_mig_node_add_pod(orc_t* o, int G, int32_t* used, uint32_t* used_mask, int32_t* fr, uint32_t* free_mask,
                         const int32_t* req, uint32_t req_mask) {
  int P = (int)o->c.mig_names.size();
  std::vector<int32_t> models(G > 0 ? G : 1, 0);
  auto v = load_mig_node(o, G, models.data(), used, used_mask, fr, free_mask);
  int slot = mig_node_add_pod(v.data(), G, pmap_from(req, req_mask, P));
  store_mig_node(o, v, used, used_mask, fr, free_mask);
  return slot;
}

int orc_mig_node_has_free_capacity(orc_t* o, int G, const int32_t* models, const int32_t* used,
                                   const uint32_t* used_mask, const int32_t* fr, const uint32_t* free_mask) {
  auto v = load_mig_node(o, G, models, used, used_mask, fr, free_mask);
  return mig_node_has_free_capacity(o->c, v.data(), G);
}

// ---- MPS -----------------------------------------------------------------------------------------
static MpsGPU load_mps(orc_t* o, int mem, const int32_t* used, uint32_t um, const int32_t* fr, uint32_t fm) {
  int S = (int)o->c.mps_sizes.size();
  MpsGPU g;
  g.mem = mem;
  g.used = pmap_from(used, um, S);
  g.free = pmap_from(fr, fm, S);
  return g;
}

// slicing/gpu.go:46-58 NewGPU validation: 0 ok, 1 error
int orc_mps_gpu_validate(orc_t* o, int mem, const int32_t* used, uint32_t um, const int32_t* fr, uint32_t fm) {
  return mps_validate(o->c, load_mps(o, mem, used, um, fr, fm)) ? 0 : 1;
}

int orc_mps_gpu_update(orc_t* o, int mem, const int32_t* used, uint32_t used_mask, int32_t* fr, uint32_t* free_mask,
                       const int32_t* req, uint32_t req_mask) {
  int S = (int)o->c.mps_sizes.size();
  MpsGPU g = load_mps(o, mem, used, used_mask, fr, *free_mask);
  bool u = mps_gpu_update(o->c, g, pmap_from(req, req_mask, S));
  pmap_to(g.free, fr, free_mask, S);
  return u;
}

int orc_mps_node_update(orc_t* o, int G, const int32_t* mem, const int32_t* used, const uint32_t* used_mask,
                        int32_t* fr, uint32_t* free_mask, const int32_t* req, uint32_t req_mask, uint8_t* gpu_updated) {
  int S = (int)o->c.mps_sizes.size();
  std::vector<MpsGPU> v(G);
  for (int i = 0; i < G; ++i) v[i] = load_mps(o, mem[i], used + (size_t)i * S, used_mask[i], fr + (size_t)i * S, free_mask[i]);
  bool u = mps_node_update(o->c, v.data(), G, pmap_from(req, req_mask, S), gpu_updated);
  for (int i = 0; i < G; ++i) pmap_to(v[i].free, fr + (size_t)i * S, free_mask + i, S);
  return u;
}

int orc_mps_node_add_pod(orc_t* o, int G, const int32_t* mem, int32_t* used, uint32_t* used_mask, int32_t* fr,
                         uint32_t* free_mask, const int32_t* req, uint32_t req_mask) {
  int S = (int)o->c.mps_sizes.size();
  std::vector<MpsGPU> v(G);
  for (int i = 0; i < G; ++i) v[i] = load_mps(o, mem[i], used + (size_t)i * S, used_mask[i], fr + (size_t)i * S, free_mask[i]);
  int slot = mps_node_add_pod(v.data(), G, pmap_from(req, req_mask, S));
  for (int i = 0; i < G; ++i) {
    pmap_to(v[i].used, used + (size_t)i * S, used_mask + i, S);
    pmap_to(v[i].free, fr + (size_t)i * S, free_mask + i, S);
  }
  return slot;
}

int orc_mps_node_has_free_capacity(orc_t* o, int G, const int32_t* mem, const int32_t* used, const uint32_t* used_mask,
                                   const int32_t* fr, const uint32_t* free_mask) {
  int S = (int)o->c.mps_sizes.size();
  std::vector<MpsGPU> v(G);
  for (int i = 0; i < G; ++i) v[i] = load_mps(o, mem[i], used + (size_t)i * S, used_mask[i], fr + (size_t)i * S, free_mask[i]);
  return mps_node_has_free_capacity(o->c, v.data(), G);
}

// slicing/profile.go:52-59 GetMemorySizeGB: trim "nvidia.com/gpu-" prefix and "gb" suffix, Atoi, else 0
int orc_mps_profile_size_gb(const char* name) {
  std::string s = name;
  const char* pre = "nvidia.com/gpu-";
  if (has_prefix(s, pre)) s = s.substr(std::strlen(pre));
  if (s.size() >= 2 && s.compare(s.size() - 2, 2, "gb") == 0) s = s.substr(0, s.size() - 2);
  long long v;
  if (go_atoi(s, &v) && v >= INT32_MIN && v <= INT32_MAX) return (int)v;
  return 0;
}

// ---- wire format ---------------------------------------------------------------------------------
// pkg/gpu/annotation.go:56-85 ParseStatusAnnotation. status: 0 used, 1 free, 2 unknown. 0 ok / 1 error.
int orc_parse_status_annotation(const char* key_, const char* value_, int32_t* index, char* profile, int profile_len,
                                int32_t* status, int32_t* quantity) {
  std::string key = key_, value = value_;
  if (!has_prefix(key, "nos.nebuly.com/status-gpu")) return 1;
  auto parts = split_dash(key);
  if (parts.size() != 5) return 1;
  long long q, idx;
  if (!go_atoi(value, &q)) return 1;
  if (!go_atoi(parts[2], &idx)) return 1;
  std::string st = lower(parts[4]);  // pkg/resource/device.go:28-39 ParseStatus
  int s;
  if (st == "free") s = 1;
  else if (st == "used") s = 0;
  else if (st == "unknown") s = 2;
  else return 1;
  *index = (int32_t)idx;
  *quantity = (int32_t)q;
  *status = s;
  std::snprintf(profile, profile_len, "%s", parts[3].c_str());
  return 0;
}

// pkg/gpu/annotation.go:29-54 ParseSpecAnnotation
int orc_parse_spec_annotation(const char* key_, const char* value_, int32_t* index, char* profile, int profile_len,
                              int32_t* quantity) {
  std::string key = key_, value = value_;
  if (!has_prefix(key, "nos.nebuly.com/spec-gpu")) return 1;
  auto parts = split_dash(key);
  if (parts.size() != 4) return 1;
  long long q, idx;
  if (!go_atoi(value, &q)) return 1;
  if (!go_atoi(parts[2], &idx)) return 1;
  *index = (int32_t)idx;
  *quantity = (int32_t)q;
  std::snprintf(profile, profile_len, "%s", parts[3].c_str());
  return 0;
}

// pkg/gpu/util.go:59-73 GetMemoryGB: Atoi the label, ceil(float64(MB)/1000) — the path's only FP op.
int orc_memory_gb(const char* label, int32_t* out) {
  long long mb;
  if (!go_atoi(label, &mb)) return 1;
  *out = (int32_t)std::ceil((double)mb / 1000.0);
  return 0;
}

// =============================================================================================
// Cluster level.  Flattened snapshot ("SoA" arrays), same meaning as include/nosplan.h.
//   MIG: n nodes; gpu_off[n+1] (CSR over GPUs in canonical order); valid[n] (0 = NewNode failed:
//        node skipped everywhere); model[totalG]; used/free [totalG][P]; masks [totalG].
//   MPS: same with mem_gb[totalG] and S sizes.
//   demands: rows of [P] (or [S]) values + presence mask.
struct orc_mig_snapshot {
  int32_t n_nodes;
  const int32_t* gpu_off;
  const uint8_t* valid;
  const int32_t* model;
  int32_t* used;
  uint32_t* used_mask;
  int32_t* free_;
  uint32_t* free_mask;
};
struct orc_mps_snapshot {
  int32_t n_nodes;
  const int32_t* gpu_off;
  const uint8_t* valid;
  const int32_t* mem_gb;
  int32_t* used;
  uint32_t* used_mask;
  int32_t* free_;
  uint32_t* free_mask;
};

}  // extern "C"

namespace {
void run_threads_impl(int n_threads, int n_items, const std::function<void(int)>& fn) {
  if (n_threads <= 1) {
    for (int i = 0; i < n_items; ++i) fn(i);
    return;
  }
  std::atomic<int> next(0);
  std::vector<std::thread> th;
  for (int t = 0; t < n_threads; ++t)
    th.emplace_back([&] {
      for (;;) {
        int i = next.fetch_add(1);
        if (i >= n_items) break;
        fn(i);
      }
    });
  for (auto& t : th) t.join();
}
}  // namespace

extern "C" {

// Batched speculative scorer: for every demand row d and every valid node n, run
// mig.Node.UpdateGeometryFor (node.go:145-177) on a FRESH COPY of the node (as the controller does,
// mig_controller.go:173-178). Outputs (any may be NULL):
//   updated[d][n]           1 if the node would update
//   choice[d][totalG]       applied geometry index per GPU (-1 none)
//   first_node[d]           min n with updated, -1 if none   (tryRepartition's winner, :146-166)
// Candidate evaluations performed = sum over (d, valid node, gpu) of |allowed(model)|.
int orc_score_mig(orc_t* o, const orc_mig_snapshot* s, int n_demands, const int32_t* req, const uint32_t* req_mask,
                  uint8_t* updated, int8_t* choice, int32_t* first_node, int n_threads) {
  int P = (int)o->c.mig_names.size();
  int N = s->n_nodes;
  int totalG = s->gpu_off[N];
  // demands are chunked so threads stay busy even when D is small
  const int chunk = std::max(1, std::min(8, n_demands / std::max(1, n_threads * 4)));
  int n_chunks = (n_demands + chunk - 1) / chunk;
  run_threads_impl(n_threads, n_chunks, [&](int ci) {
    std::vector<MigGPU> node;
    std::vector<int> ch;
    for (int d = ci * chunk; d < std::min(n_demands, (ci + 1) * chunk); ++d) {
      PMap rq = pmap_from(req + (size_t)d * P, req_mask[d], P);
      int first = -1;
      for (int n = 0; n < N; ++n) {
        int g0 = s->gpu_off[n], G = s->gpu_off[n + 1] - g0;
        bool u = false;
        if (s->valid[n]) {
          node.resize(G);
          ch.assign(G > 0 ? G : 1, -1);
          for (int i = 0; i < G; ++i)
            node[i] = load_mig(o, s->model[g0 + i], s->used + (size_t)(g0 + i) * P, s->used_mask[g0 + i],
                               s->free_ + (size_t)(g0 + i) * P, s->free_mask[g0 + i]);
          u = mig_node_update(o->c, node.data(), G, rq, ch.data());
          if (choice)
            for (int i = 0; i < G; ++i) choice[(size_t)d * totalG + g0 + i] = (int8_t)ch[i];
        } else if (choice) {
          for (int i = 0; i < G; ++i) choice[(size_t)d * totalG + g0 + i] = -1;
        }
        if (updated) updated[(size_t)d * N + n] = u;
        if (u && first < 0) first = n;
      }
      if (first_node) first_node[d] = first;
    }
  });
  return 0;
}

// MPS twin: slicing.Node.UpdateGeometryFor (slicing/node.go:120-152) on fresh copies.
//   out_free[d][totalG][S], out_free_mask[d][totalG]: the free layout after the (trial) update
int orc_score_mps(orc_t* o, const orc_mps_snapshot* s, int n_demands, const int32_t* req, const uint32_t* req_mask,
                  uint8_t* updated, int32_t* out_free, uint32_t* out_free_mask, int32_t* first_node, int n_threads) {
  int S = (int)o->c.mps_sizes.size();
  int N = s->n_nodes;
  int totalG = s->gpu_off[N];
  const int chunk = std::max(1, std::min(8, n_demands / std::max(1, n_threads * 4)));
  int n_chunks = (n_demands + chunk - 1) / chunk;
  run_threads_impl(n_threads, n_chunks, [&](int ci) {
    std::vector<MpsGPU> node;
    for (int d = ci * chunk; d < std::min(n_demands, (ci + 1) * chunk); ++d) {
      PMap rq = pmap_from(req + (size_t)d * S, req_mask[d], S);
      int first = -1;
      for (int n = 0; n < N; ++n) {
        int g0 = s->gpu_off[n], G = s->gpu_off[n + 1] - g0;
        node.resize(G);
        for (int i = 0; i < G; ++i)
          node[i] = load_mps(o, s->mem_gb[g0 + i], s->used + (size_t)(g0 + i) * S, s->used_mask[g0 + i],
                             s->free_ + (size_t)(g0 + i) * S, s->free_mask[g0 + i]);
        bool u = false;
        if (s->valid[n]) u = mps_node_update(o->c, node.data(), G, rq, nullptr);
        if (out_free)
          for (int i = 0; i < G; ++i)
            pmap_to(node[i].free, out_free + ((size_t)d * totalG + g0 + i) * S, out_free_mask + (size_t)d * totalG + g0 + i, S);
        if (updated) updated[(size_t)d * N + n] = u;
        if (u && first < 0) first = n;
      }
      if (first_node) first_node[d] = first;
    }
  });
  return 0;
}

// Plan composition, replaying Controller.Reconcile (mig_controller.go:56-98) pod by
// pod on the evolving in-memory state.  Pods arrive already filtered by shouldConsiderPod (:100-111)
// and sorted (priority desc, creation asc, ns/name asc) by the host.
//   pod_kind[i]  0 = MIG, 1 = MPS; pod_req = [M][W] values with W = max(P, S) columns; pod_req_mask[M]
//   flags bit0: apply the profileAlreadyPresent gate (:121-144)   (reference behaviour: on)
//         bit1: simulate scheduling with Node.AddPod first-fit (§A.5 3c)
// Outputs per pod: gate_hit, repart_node (-1 none), sched_node / sched_gpu (-1 none).
// The snapshots are mutated in place into the final state.
enum { ORC_FLAG_GATE = 1, ORC_FLAG_SCHED = 2 };

int orc_plan(orc_t* o, orc_mig_snapshot* mig, orc_mps_snapshot* mps, int n_pods, const uint8_t* pod_kind,
             const int32_t* pod_req, const uint32_t* pod_req_mask, int req_stride, int flags, uint8_t* gate_hit,
             int32_t* repart_node, int32_t* sched_node, int32_t* sched_gpu) {
  int P = (int)o->c.mig_names.size();
  int S = (int)o->c.mps_sizes.size();
  std::vector<MigGPU> mg;
  std::vector<MpsGPU> sg;
  if (mig) {
    int tg = mig->gpu_off[mig->n_nodes];
    mg.resize(tg);
    for (int i = 0; i < tg; ++i)
      mg[i] = load_mig(o, mig->model[i], mig->used + (size_t)i * P, mig->used_mask[i], mig->free_ + (size_t)i * P,
                       mig->free_mask[i]);
  }
  if (mps) {
    int tg = mps->gpu_off[mps->n_nodes];
    sg.resize(tg);
    for (int i = 0; i < tg; ++i)
      sg[i] = load_mps(o, mps->mem_gb[i], mps->used + (size_t)i * S, mps->used_mask[i], mps->free_ + (size_t)i * S,
                       mps->free_mask[i]);
  }
  std::vector<MigGPU> mtmp;
  std::vector<MpsGPU> stmp;
  for (int i = 0; i < n_pods; ++i) {
    gate_hit[i] = 0;
    repart_node[i] = sched_node[i] = sched_gpu[i] = -1;
    bool is_mps = pod_kind[i] != 0;
    int W = is_mps ? S : P;
    PMap rq = pmap_from(pod_req + (size_t)i * req_stride, pod_req_mask[i], W);
    if (rq.len() == 0) continue;  // mig_controller.go:68-71
    int N = is_mps ? (mps ? mps->n_nodes : 0) : (mig ? mig->n_nodes : 0);
    if (N == 0) continue;  // :78-81
    const int32_t* off = is_mps ? mps->gpu_off : mig->gpu_off;
    const uint8_t* valid = is_mps ? mps->valid : mig->valid;
    // profileAlreadyPresent (:121-144): every requested profile has (used+free) > 0 on some valid node
    bool present = false;
    if (flags & ORC_FLAG_GATE) {
      present = true;
      for (int p = 0; p < W && present; ++p) {
        if (!rq.has(p)) continue;
        bool found = false;
        for (int n = 0; n < N && !found; ++n) {
          if (!valid[n]) continue;
          int q = 0;  // Node.Geometry()[profile], node.go:113-120
          for (int g = off[n]; g < off[n + 1]; ++g)
            q += is_mps ? sg[g].used.get(p) + sg[g].free.get(p) : mg[g].used.get(p) + mg[g].free.get(p);
          if (q > 0) found = true;
        }
        if (!found) present = false;
      }
    }
    gate_hit[i] = present;
    if (!present) {
      // tryRepartition (:146-166): nodes in list order, each on a fresh copy; first updated wins
      for (int n = 0; n < N; ++n) {
        if (!valid[n]) continue;
        int g0 = off[n], G = off[n + 1] - g0;
        bool u;
        if (is_mps) {
          stmp.assign(sg.begin() + g0, sg.begin() + g0 + G);
          u = mps_node_update(o->c, stmp.data(), G, rq, nullptr);
          if (u) std::copy(stmp.begin(), stmp.end(), sg.begin() + g0);
        } else {
          mtmp.assign(mg.begin() + g0, mg.begin() + g0 + G);
          u = mig_node_update(o->c, mtmp.data(), G, rq, nullptr);
          if (u) std::copy(mtmp.begin(), mtmp.end(), mg.begin() + g0);
        }
        if (u) {
          repart_node[i] = n;
          break;
        }
      }
    }
    if (flags & ORC_FLAG_SCHED) {
      // §A.5 3c: first node whose Node.AddPod succeeds is committed; a failing node is a discarded
      // trial copy.  Node.AddPod succeeds iff some GPU fits every key, so the (side-effect free)
      // fit check below decides success and the literal AddPod then runs on the real node only.
      for (int n = 0; n < N; ++n) {
        if (!valid[n]) continue;
        int g0 = off[n], G = off[n + 1] - g0;
        bool fits = false;
        for (int g = g0; g < g0 + G && !fits; ++g) {
          bool ok = true;
          for (int p = 0; p < W && ok; ++p)
            if (rq.has(p)) ok = (is_mps ? sg[g].free.get(p) : mg[g].free.get(p)) >= rq.v[p];
          fits = ok;
        }
        int slot = -1;
        if (fits) slot = is_mps ? mps_node_add_pod(sg.data() + g0, G, rq) : mig_node_add_pod(mg.data() + g0, G, rq);
        if (slot >= 0) {
          sched_node[i] = n;
          sched_gpu[i] = slot;
          break;
        }
      }
    }
  }
  if (mig)
    for (size_t i = 0; i < mg.size(); ++i) {
      pmap_to(mg[i].used, mig->used + i * P, mig->used_mask + i, P);
      pmap_to(mg[i].free, mig->free_ + i * P, mig->free_mask + i, P);
    }
  if (mps)
    for (size_t i = 0; i < sg.size(); ++i) {
      pmap_to(sg[i].used, mps->used + i * S, mps->used_mask + i, S);
      pmap_to(sg[i].free, mps->free_ + i * S, mps->free_mask + i, S);
    }
  return 0;
}

// internal/partitioning/mig/state.go:25-45 BuildNodePartitioning for every GPU of a snapshot:
// out[totalG][P] = used+free where > 0, else 0 (a zero means "resource key absent").
int orc_build_partitioning(int total_gpus, int P, const int32_t* used, const uint32_t* used_mask, const int32_t* fr,
                           const uint32_t* free_mask, int32_t* out) {
  for (int g = 0; g < total_gpus; ++g)
    for (int p = 0; p < P; ++p) {
      int q = 0;
      if ((used_mask[g] >> p) & 1u) q += used[(size_t)g * P + p];
      if ((free_mask[g] >> p) & 1u) q += fr[(size_t)g * P + p];
      out[(size_t)g * P + p] = q > 0 ? q : 0;
    }
  return 0;
}

}  // extern "C"

// ---- agent-side diff planner, count level ----------------------------------------------------
// plan.NewMigConfigPlan (internal/controllers/migagent/plan/plan.go:31-92) reduced to counts per
// (GPU, profile); the device-level restatement pinned by the reference's table tests is
// oracle/diff_plan.py, and tests/test_diff_plan.py checks the two against each other.
//   used/free_ [G][P]   devices of the profile on the GPU ("unknown" status counts as used: not free)
//   spec [G][P], spec_mask[G] bit p = a spec annotation names (GPU, profile)
//   create[G][P]     devices to create (new ones + re-created free ones)
//   del_free[G][P]   free devices to delete,  del_used[G][P]  non-free devices to delete
extern "C" int orc_mig_diff_counts(int G, int P, const int32_t* used, const int32_t* free_, const int32_t* spec,
                                   const uint32_t* spec_mask, int32_t* create, int32_t* del_free, int32_t* del_used) {
  for (int g = 0; g < G; ++g) {
    bool any_create = false;
    for (int p = 0; p < P; ++p) {
      const size_t i = (size_t)g * P + p;
      create[i] = del_free[i] = del_used[i] = 0;
      if (!((spec_mask[g] >> p) & 1u)) {  // not included in spec: everything goes (plan.go:37-41, :168-182)
        del_free[i] = free_[i];
        del_used[i] = used[i];
        continue;
      }
      const int diff = spec[i] - (used[i] + free_[i]);  // plan.go:60
      if (diff > 0) {
        create[i] = diff;
        any_create = true;
      } else if (diff < 0) {  // free devices first, then used (plan.go:112-137)
        del_free[i] = std::min(-diff, free_[i]);
        del_used[i] = -diff - del_free[i];
      }
    }
    if (any_create)  // plan.go:73-88: re-create the free devices that are not already being deleted
      for (int p = 0; p < P; ++p) {
        const size_t i = (size_t)g * P + p;
        const int again = free_[i] - del_free[i];
        del_free[i] += again;
        create[i] += again;
      }
  }
  return 0;
}
