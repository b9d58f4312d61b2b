// noshost.cpp — host side of the planner drop-in (include/noshost.h).  Plain C++17, no CUDA.
// Mirrors the Go code on either side of the hot path; citations are relative to the reference root.
#include "noshost.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <set>
#include <string>
#include <vector>

namespace {

const char kSpecPrefix[] = "nos.nebuly.com/spec-gpu";      // v1alpha1/annotations.go:22
const char kStatusPrefix[] = "nos.nebuly.com/status-gpu";  // v1alpha1/annotations.go:23
const char kMigResourcePrefix[] = "nvidia.com/mig-";       // pkg/constant/constants.go:46
const char kGpuResourcePrefix[] = "nvidia.com/gpu-";       // slicing/profile.go:30

constexpr long long kMaxLabelGpuCount = 4096;  // nvidia.com/gpu.count above this marks the node invalid (see snapshot_build)

bool has_prefix(const std::string& s, const char* p) { return s.rfind(p, 0) == 0; }

// The C ABI carries indexes and quantities as int32.  strconv.Atoi accepts up to 64 bits; a number that does not fit
// 32 bits is treated here as a parse failure (the annotation is ignored) instead of being truncated.
bool fits_i32(long long v) { return v >= INT32_MIN && v <= INT32_MAX; }

// strconv.Atoi: optional sign, decimal digits, nothing else; error on overflow
bool go_atoi(const std::string& s, long long* out) {
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

std::vector<std::string> split(const std::string& s, char sep) {
  std::vector<std::string> parts;
  size_t b = 0;
  for (;;) {
    size_t e = s.find(sep, b);
    if (e == std::string::npos) {
      parts.push_back(s.substr(b));
      return parts;
    }
    parts.push_back(s.substr(b, e - b));
    b = e + 1;
  }
}

std::string lower(std::string s) {
  for (auto& c : s)
    if (c >= 'A' && c <= 'Z') c = (char)(c - 'A' + 'a');
  return s;
}

size_t digits(const std::string& s, size_t i) {
  size_t j = i;
  while (j < s.size() && s[j] >= '0' && s[j] <= '9') ++j;
  return j;
}

// unanchored match of  nvidia\.com\/mig-\d+g\.\d+gb  (pkg/constant/constants.go:39)
bool matches_mig_resource(const std::string& s) {
  for (size_t p = s.find(kMigResourcePrefix); p != std::string::npos; p = s.find(kMigResourcePrefix, p + 1)) {
    size_t i = p + sizeof(kMigResourcePrefix) - 1;
    size_t j = digits(s, i);
    if (j == i || j + 1 >= s.size() || s[j] != 'g' || s[j + 1] != '.') continue;
    size_t k = digits(s, j + 2);
    if (k == j + 2) continue;
    if (k + 1 < s.size() && s[k] == 'g' && s[k + 1] == 'b') return true;
  }
  return false;
}
// unanchored match of  nvidia\.com/gpu-\d+gb  (pkg/gpu/slicing/profile.go:31)
bool matches_mps_resource(const std::string& s) {
  for (size_t p = s.find(kGpuResourcePrefix); p != std::string::npos; p = s.find(kGpuResourcePrefix, p + 1)) {
    size_t i = p + sizeof(kGpuResourcePrefix) - 1;
    size_t j = digits(s, i);
    if (j == i) continue;
    if (j + 1 < s.size() && s[j] == 'g' && s[j + 1] == 'b') return true;
  }
  return false;
}

int mps_size_of(const std::string& name) {  // slicing/profile.go:52-59
  std::string t = name;
  if (has_prefix(t, kGpuResourcePrefix)) t = t.substr(sizeof(kGpuResourcePrefix) - 1);
  if (t.size() >= 2 && t.compare(t.size() - 2, 2, "gb") == 0) t = t.substr(0, t.size() - 2);
  long long v;
  if (go_atoi(t, &v) && v >= INT32_MIN && v <= INT32_MAX) return (int)v;
  return 0;
}

// a parsed name either fits the caller's buffer with its terminator or the call fails: never a truncated name
bool fits_out(const std::string& s, const char* out, int cap) { return out && cap > 0 && (int)s.size() < cap; }

int copy_out(const std::string& s, char* out, int cap) {
  if (out && cap > 0) std::snprintf(out, cap, "%s", s.c_str());
  return (int)s.size();
}

struct StatusAnn {
  int index;
  std::string profile;
  int status;
  int qty;
};

bool parse_status(const std::string& key, const std::string& value, StatusAnn* a) {
  if (!has_prefix(key, kStatusPrefix)) return false;
  auto parts = split(key, '-');
  if (parts.size() != 5) return false;
  long long q, idx;
  if (!go_atoi(value, &q) || !fits_i32(q)) return false;
  if (!go_atoi(parts[2], &idx) || !fits_i32(idx)) return false;
  std::string st = lower(parts[4]);  // pkg/resource/device.go:28-39
  int s;
  if (st == "free") s = NOSHOST_STATUS_FREE;
  else if (st == "used") s = NOSHOST_STATUS_USED;
  else if (st == "unknown") s = NOSHOST_STATUS_UNKNOWN;
  else return false;
  a->index = (int)idx;
  a->profile = parts[3];
  a->status = s;
  a->qty = (int)q;
  return true;
}

bool parse_spec(const std::string& key, const std::string& value, int* index, std::string* profile, int* qty) {
  if (!has_prefix(key, kSpecPrefix)) return false;
  auto parts = split(key, '-');
  if (parts.size() != 4) return false;
  long long q, idx;
  if (!go_atoi(value, &q) || !fits_i32(q)) return false;
  if (!go_atoi(parts[2], &idx) || !fits_i32(idx)) return false;
  *index = (int)idx;
  *profile = parts[3];
  *qty = (int)q;
  return true;
}

}  // namespace

struct noshost_snapshot {
  int kind = 0;
  std::vector<std::string> key_names;  // vocabulary in canonical order
  std::vector<int> key_sizes;
  std::map<std::string, int> name_id;  // MIG
  std::map<int, int> size_id;          // MPS
  std::vector<int32_t> gpu_off{0}, model, mem, gpu_index, used, free_;
  std::vector<uint8_t> valid;
  std::vector<uint32_t> used_mask, free_mask;
};

struct noshost_pods {
  std::vector<int32_t> source, off{0}, key, qty;
  std::vector<uint8_t> kind;
  std::string err;
};

extern "C" {

int noshost_parse_status_annotation(const char* key, const char* value, int32_t* index, char* profile, int profile_cap,
                                    int32_t* status, int32_t* quantity) {
  StatusAnn a;
  if (!key || !value || !parse_status(key, value, &a)) return 1;
  if (!fits_out(a.profile, profile, profile_cap)) return 2;
  *index = a.index;
  *status = a.status;
  *quantity = a.qty;
  copy_out(a.profile, profile, profile_cap);
  return 0;
}

int noshost_parse_spec_annotation(const char* key, const char* value, int32_t* index, char* profile, int profile_cap,
                                  int32_t* quantity) {
  int idx, q;
  std::string p;
  if (!key || !value || !parse_spec(key, value, &idx, &p, &q)) return 1;
  if (!fits_out(p, profile, profile_cap)) return 2;
  *index = idx;
  *quantity = q;
  copy_out(p, profile, profile_cap);
  return 0;
}

int noshost_format_status_annotation(int32_t index, const char* profile, int32_t status, char* out, int cap) {
  const char* st = status == NOSHOST_STATUS_FREE ? "free" : status == NOSHOST_STATUS_USED ? "used" : "unknown";
  return copy_out(std::string(kStatusPrefix) + "-" + std::to_string(index) + "-" + (profile ? profile : "") + "-" + st, out, cap);
}

int noshost_format_spec_annotation(int32_t index, const char* profile, char* out, int cap) {
  return copy_out(std::string(kSpecPrefix) + "-" + std::to_string(index) + "-" + (profile ? profile : ""), out, cap);
}

int noshost_memory_gb(const char* label, int32_t* out) {
  long long mb;
  if (!label || !go_atoi(label, &mb)) return 1;
  const double gb = std::ceil((double)mb / 1000.0);  // the path's only floating-point operation
  if (gb < (double)INT32_MIN || gb > (double)INT32_MAX) return 1;  // does not fit the ABI's int32
  *out = (int32_t)gb;
  return 0;
}

int noshost_mig_profile_of_resource(const char* resource, char* profile, int cap) {
  std::string r = resource ? resource : "";
  if (!matches_mig_resource(r)) return 1;
  const std::string name = has_prefix(r, kMigResourcePrefix) ? r.substr(sizeof(kMigResourcePrefix) - 1) : r;
  if (!fits_out(name, profile, cap)) return 2;
  copy_out(name, profile, cap);
  return 0;
}

int noshost_mps_profile_of_resource(const char* resource, char* profile, int cap) {
  std::string r = resource ? resource : "";
  if (!matches_mps_resource(r)) return 1;
  const std::string name = has_prefix(r, kGpuResourcePrefix) ? r.substr(sizeof(kGpuResourcePrefix) - 1) : r;
  if (!fits_out(name, profile, cap)) return 2;
  copy_out(name, profile, cap);
  return 0;
}

int noshost_mps_profile_size_gb(const char* profile) { return mps_size_of(profile ? profile : ""); }

noshost_snapshot* noshost_snapshot_build(int kind, int n_nodes, const char* const* product, const char* const* count,
                                         const char* const* memory, const int32_t* ann_off, const char* const* ann_key,
                                         const char* const* ann_val, int n_models, const char* const* model_names,
                                         int n_extra_names, const char* const* extra_names, int n_extra_sizes,
                                         const int32_t* extra_sizes) try {
  // arguments first: nothing below indexes through an array that was not looked at here
  if ((kind != NOSHOST_KIND_MIG && kind != NOSHOST_KIND_MPS) || n_nodes < 0 || n_models < 0 || n_extra_names < 0 || n_extra_sizes < 0)
    return nullptr;
  if (n_nodes > 0 && (!product || !count || !ann_off || (kind == NOSHOST_KIND_MPS && !memory))) return nullptr;
  if ((n_models > 0 && !model_names) || (n_extra_names > 0 && !extra_names) || (n_extra_sizes > 0 && !extra_sizes)) return nullptr;
  if (n_nodes > 0) {
    if (ann_off[0] < 0) return nullptr;
    for (int n = 0; n < n_nodes; ++n)
      if (ann_off[n + 1] < ann_off[n]) return nullptr;
    if (ann_off[n_nodes] > 0 && (!ann_key || !ann_val)) return nullptr;
  }
  struct Gpu {
    int index;
    std::map<std::string, int> used, free_;
  };
  struct Node {
    bool valid = false;
    int model = -1, mem = 0;
    std::vector<Gpu> gpus;
  };
  std::map<std::string, int> model_id;
  for (int m = 0; m < n_models; ++m)
    if (model_names[m]) model_id[model_names[m]] = m;
  std::vector<Node> nodes(n_nodes);
  std::set<std::string> names;
  std::set<int> sizes;
  for (int i = 0; i < n_extra_names; ++i)
    if (extra_names[i]) names.insert(extra_names[i]);
  for (int i = 0; i < n_extra_sizes; ++i) sizes.insert(extra_sizes[i]);
  for (int n = 0; n < n_nodes; ++n) {
    Node& nd = nodes[n];
    // gpu.GetModel / GetCount (pkg/gpu/util.go:30-56), GetMemoryGB for slicing (:59-73)
    if (!product[n] || !count[n]) continue;
    long long cnt;
    if (!go_atoi(count[n], &cnt)) continue;
    // the loop below materialises every "missing" GPU up to the label's count: a label no machine can carry is refused
    // (the node stays invalid) rather than allowed to allocate without bound
    if (cnt > kMaxLabelGpuCount) continue;
    int32_t mem = 0;
    if (kind == NOSHOST_KIND_MPS && (!memory[n] || noshost_memory_gb(memory[n], &mem))) continue;
    // ParseNodeAnnotations (annotation.go:87-101): spec first, then status; the rest ignored
    std::map<int, Gpu> by_index;  // GroupByGpuIndex; canonical order = ascending index
    for (int a = ann_off[n]; a < ann_off[n + 1]; ++a) {
      if (!ann_key[a] || !ann_val[a]) continue;  // not an annotation
      int idx, q;
      std::string p;
      if (parse_spec(ann_key[a], ann_val[a], &idx, &p, &q)) continue;
      StatusAnn sa;
      if (!parse_status(ann_key[a], ann_val[a], &sa)) continue;
      Gpu& g = by_index[sa.index];
      g.index = sa.index;
      if (sa.status == NOSHOST_STATUS_USED) g.used[sa.profile] = sa.qty;
      if (sa.status == NOSHOST_STATUS_FREE) g.free_[sa.profile] = sa.qty;
    }
    bool ok = true;
    const bool known = model_id.count(product[n]) > 0;
    std::vector<Gpu> gpus;
    for (auto& kv : by_index) {
      if (kind == NOSHOST_KIND_MIG) {
        if (!known) ok = false;  // NewGPU: unknown model (mig/gpu.go:46-49)
      } else {
        long long tot = 0;  // slicing.NewGPU -> Validate (slicing/gpu.go:67-97)
        for (auto* mp : {&kv.second.used, &kv.second.free_})
          for (auto& e : *mp) {
            int sz = mps_size_of(e.first);
            if (sz < 1) ok = false;
            tot += (long long)sz * e.second;
          }
        if (tot > mem) ok = false;
      }
      gpus.push_back(kv.second);
    }
    for (long long i = (long long)gpus.size(); i < cnt; ++i) {  // "missing" GPUs, index = len..count-1
      if (kind == NOSHOST_KIND_MIG && !known) ok = false;
      Gpu g;
      g.index = (int)i;
      gpus.push_back(g);
    }
    if (!ok) continue;
    nd.valid = true;
    nd.model = known ? model_id[product[n]] : -1;
    nd.mem = mem;
    nd.gpus = gpus;
    for (auto& g : gpus)
      for (auto* mp : {&g.used, &g.free_})
        for (auto& e : *mp) {
          if (kind == NOSHOST_KIND_MIG) names.insert(e.first);
          else sizes.insert(mps_size_of(e.first));
        }
  }
  auto* s = new noshost_snapshot();
  s->kind = kind;
  if (kind == NOSHOST_KIND_MIG) {
    for (auto& nm : names) {  // std::set<string> iterates bytewise ascending: the canonical order
      s->name_id[nm] = (int)s->key_names.size();
      s->key_names.push_back(nm);
      s->key_sizes.push_back(0);
    }
  } else {
    for (int sz : sizes) {
      s->size_id[sz] = (int)s->key_names.size();
      s->key_names.push_back(std::to_string(sz) + "gb");  // NewProfile (slicing/profile.go:48-50)
      s->key_sizes.push_back(sz);
    }
  }
  const int K = (int)s->key_names.size();
  for (int n = 0; n < n_nodes; ++n) {
    const Node& nd = nodes[n];
    s->valid.push_back(nd.valid ? 1 : 0);
    if (nd.valid)
      for (const auto& g : nd.gpus) {
        s->model.push_back(nd.model);
        s->mem.push_back(nd.mem);
        s->gpu_index.push_back(g.index);
        std::vector<int32_t> u(K, 0), f(K, 0);
        uint32_t um = 0, fm = 0;
        for (auto& e : g.used) {
          int k = kind == NOSHOST_KIND_MIG ? s->name_id[e.first] : s->size_id[mps_size_of(e.first)];
          u[k] = e.second;
          if (k < 32) um |= 1u << k;
        }
        for (auto& e : g.free_) {
          int k = kind == NOSHOST_KIND_MIG ? s->name_id[e.first] : s->size_id[mps_size_of(e.first)];
          f[k] = e.second;
          if (k < 32) fm |= 1u << k;
        }
        s->used.insert(s->used.end(), u.begin(), u.end());
        s->free_.insert(s->free_.end(), f.begin(), f.end());
        s->used_mask.push_back(um);
        s->free_mask.push_back(fm);
      }
    s->gpu_off.push_back((int32_t)s->model.size());
  }
  return s;
} catch (...) {  // out of memory while building: no C++ exception crosses the C boundary
  return nullptr;
}

void noshost_snapshot_free(noshost_snapshot* s) { delete s; }
int noshost_snapshot_n_nodes(const noshost_snapshot* s) { return s ? (int)s->valid.size() : 0; }
int noshost_snapshot_total_gpus(const noshost_snapshot* s) { return s ? (int)s->model.size() : 0; }
int noshost_snapshot_n_keys(const noshost_snapshot* s) { return s ? (int)s->key_names.size() : 0; }
const char* noshost_snapshot_key_name(const noshost_snapshot* s, int k) {
  return s && k >= 0 && k < (int)s->key_names.size() ? s->key_names[k].c_str() : "";
}
int noshost_snapshot_key_size(const noshost_snapshot* s, int k) { return s && k >= 0 && k < (int)s->key_sizes.size() ? s->key_sizes[k] : 0; }
const int32_t* noshost_snapshot_gpu_off(const noshost_snapshot* s) { return s ? s->gpu_off.data() : nullptr; }
const uint8_t* noshost_snapshot_valid(const noshost_snapshot* s) { return s ? s->valid.data() : nullptr; }
const int32_t* noshost_snapshot_model(const noshost_snapshot* s) { return s ? s->model.data() : nullptr; }
const int32_t* noshost_snapshot_mem_gb(const noshost_snapshot* s) { return s ? s->mem.data() : nullptr; }
const int32_t* noshost_snapshot_gpu_index(const noshost_snapshot* s) { return s ? s->gpu_index.data() : nullptr; }
const int32_t* noshost_snapshot_used(const noshost_snapshot* s) { return s ? s->used.data() : nullptr; }
const int32_t* noshost_snapshot_free_(const noshost_snapshot* s) { return s ? s->free_.data() : nullptr; }
const uint32_t* noshost_snapshot_used_mask(const noshost_snapshot* s) { return s ? s->used_mask.data() : nullptr; }
const uint32_t* noshost_snapshot_free_mask(const noshost_snapshot* s) { return s ? s->free_mask.data() : nullptr; }

noshost_pods* noshost_pods_build(int n_pods, const uint8_t* pending, const uint8_t* scheduled, const uint8_t* unschedulable,
                                 const int32_t* priority, const int64_t* created, const char* const* name,
                                 const int32_t* cont_off, const uint8_t* cont_init, const int32_t* res_off,
                                 const char* const* res_name, const int64_t* res_qty, const noshost_snapshot* mig,
                                 const noshost_snapshot* mps) {
  auto* out = new noshost_pods();
  struct Row {
    int pod;
    int kind;
    std::vector<std::pair<int, int>> kv;  // (key id, qty), ascending key
  };
  // the CSR arrays come from the caller: walk them once before anything is indexed through them
  auto bad_input = [&](const std::string& why) {
    out->err = why;
    out->key.push_back(0);
    out->qty.push_back(0);
    return out;
  };
  if (n_pods < 0) return bad_input("negative pod count");
  if (n_pods > 0 && (!pending || !scheduled || !unschedulable || !priority || !created || !name || !cont_off))
    return bad_input("null pod arrays");
  if (n_pods > 0) {
    if (cont_off[0] < 0) return bad_input("cont_off[0] is negative");
    for (int i = 0; i < n_pods; ++i)
      if (cont_off[i + 1] < cont_off[i]) return bad_input("cont_off not ascending at pod " + std::to_string(i));
    const int n_cont = cont_off[n_pods];
    if (n_cont > 0 && (!cont_init || !res_off)) return bad_input("null container arrays");
    if (n_cont > 0) {
      if (res_off[0] < 0) return bad_input("res_off[0] is negative");
      for (int c = 0; c < n_cont; ++c)
        if (res_off[c + 1] < res_off[c]) return bad_input("res_off not ascending at container " + std::to_string(c));
      const int n_res = res_off[n_cont];
      if (n_res > 0 && (!res_name || !res_qty)) return bad_input("null resource arrays");
      for (int r = 0; r < n_res; ++r)
        if (!res_name[r]) return bad_input("resource " + std::to_string(r) + " has no name");
    }
  }
  auto name_of = [&](int i) { return name[i] ? name[i] : ""; };
  // quantities are int64 on the wire (resource.Quantity.Value() saturates there too); sums saturate instead of wrapping
  auto sat_add = [](long long a, long long b) {
    long long r;
    if (__builtin_add_overflow(a, b, &r)) return b > 0 ? (long long)INT64_MAX : (long long)INT64_MIN;
    return r;
  };
  std::vector<int> order;
  for (int i = 0; i < n_pods; ++i)
    if (pending[i] && !scheduled[i] && unschedulable[i]) order.push_back(i);  // shouldConsiderPod
  // IsMoreImportant (priority desc) — then creation time asc, then namespace/name asc
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) {
    if (priority[a] != priority[b]) return priority[a] > priority[b];
    if (created[a] != created[b]) return created[a] < created[b];
    return std::strcmp(name_of(a), name_of(b)) < 0;
  });
  for (int i : order) {
    // ComputePodRequest: sum over containers, element-wise max with the init containers
    std::map<std::string, long long> sum, initmax;
    for (int c = cont_off[i]; c < cont_off[i + 1]; ++c)
      for (int r = res_off[c]; r < res_off[c + 1]; ++r) {
        if (cont_init[c]) {
          auto it = initmax.find(res_name[r]);
          if (it == initmax.end()) initmax[res_name[r]] = res_qty[r];  // quota.Max: a key on one side only keeps its value
          else it->second = std::max<long long>(it->second, res_qty[r]);
        } else {
          sum[res_name[r]] = sat_add(sum[res_name[r]], res_qty[r]);
        }
      }
    std::map<std::string, long long> total = sum;
    for (auto& e : initmax) {
      auto it = total.find(e.first);
      if (it == total.end()) total[e.first] = e.second;
      else it->second = std::max(it->second, e.second);
    }
    // GetRequestedProfiles, MIG then MPS
    std::map<int, long long> mreq, preq;
    bool bad = false;
    for (auto& e : total) {
      if (matches_mig_resource(e.first)) {
        std::string p = has_prefix(e.first, kMigResourcePrefix) ? e.first.substr(sizeof(kMigResourcePrefix) - 1) : e.first;
        if (!mig || !mig->name_id.count(p)) {
          out->err = "MIG profile " + p + " is not in the snapshot vocabulary (pass it as an extra name)";
          bad = true;
          break;
        }
        mreq[mig->name_id.at(p)] = sat_add(mreq[mig->name_id.at(p)], e.second);
      } else if (matches_mps_resource(e.first)) {
        std::string p = has_prefix(e.first, kGpuResourcePrefix) ? e.first.substr(sizeof(kGpuResourcePrefix) - 1) : e.first;
        int sz = mps_size_of(p);
        if (!mps || !mps->size_id.count(sz)) {
          out->err = "MPS size " + std::to_string(sz) + "gb is not in the snapshot vocabulary (pass it as an extra size)";
          bad = true;
          break;
        }
        preq[mps->size_id.at(sz)] = sat_add(preq[mps->size_id.at(sz)], e.second);
      }
    }
    for (auto* rq : {&mreq, &preq})
      for (auto& e : *rq)
        if (!bad && !fits_i32(e.second)) {  // int(quantity.Value()) is 64 bits in Go; the planner ABI carries int32
          out->err = std::string("pod ") + name_of(i) + ": a requested quantity does not fit 32 bits";
          bad = true;
        }
    if (bad) break;
    for (int kd = 0; kd < 2; ++kd) {
      auto& rq = kd ? preq : mreq;
      if (rq.empty()) continue;  // len(requestedProfiles) == 0 -> the controller returns (mig_controller.go:68-71)
      out->source.push_back(i);
      out->kind.push_back((uint8_t)kd);
      for (auto& e : rq) {
        out->key.push_back(e.first);
        out->qty.push_back((int32_t)e.second);
      }
      out->off.push_back((int32_t)out->key.size());
    }
  }
  if (!out->err.empty()) {  // a refusal returns no rows at all, not the ones built before the offending pod
    out->source.clear();
    out->kind.clear();
    out->key.clear();
    out->qty.clear();
    out->off.assign(1, 0);
  }
  if (out->key.empty()) {
    out->key.push_back(0);
    out->qty.push_back(0);
  }
  return out;
}

void noshost_pods_free(noshost_pods* p) { delete p; }
// (a null handle answers like an empty result)
int noshost_pods_count(const noshost_pods* p) { return p ? (int)p->source.size() : 0; }
const int32_t* noshost_pods_source(const noshost_pods* p) { return p ? p->source.data() : nullptr; }
const uint8_t* noshost_pods_kind(const noshost_pods* p) { return p ? p->kind.data() : nullptr; }
const int32_t* noshost_pods_off(const noshost_pods* p) { return p ? p->off.data() : nullptr; }
const int32_t* noshost_pods_key(const noshost_pods* p) { return p ? p->key.data() : nullptr; }
const int32_t* noshost_pods_qty(const noshost_pods* p) { return p ? p->qty.data() : nullptr; }
const char* noshost_pods_error(const noshost_pods* p) { return p ? p->err.c_str() : "null handle"; }

int noshost_spec_annotations(const noshost_snapshot* s, int node, const int32_t* used, const int32_t* free_, char* out,
                             int cap) {
  if (!s || !used || !free_ || node < 0 || node >= (int)s->valid.size()) return -1;
  const int K = (int)s->key_names.size();
  std::string txt;
  for (int g = s->gpu_off[node]; g < s->gpu_off[node + 1]; ++g)
    for (int k = 0; k < K; ++k) {
      const long long q = (long long)used[(size_t)g * K + k] + free_[(size_t)g * K + k];  // GetGeometry
      if (q <= 0) continue;                                             // state.go:33-35
      txt += std::string(kSpecPrefix) + "-" + std::to_string(s->gpu_index[g]) + "-" + s->key_names[k] + "=" +
             std::to_string(q) + "\n";
    }
  if (out && cap > 0) std::snprintf(out, cap, "%s", txt.c_str());
  return (int)txt.size();
}

int noshost_node_partitioning_equal(int n_keys, int n_a, const int32_t* idx_a, const int32_t* res_a, int n_b,
                                    const int32_t* idx_b, const int32_t* res_b) {
  if (n_a != n_b) return 0;  // partitioning.go:34-36
  if (n_a <= 0) return 1;    // two empty lists
  if (n_keys < 0 || !idx_a || !idx_b || (n_keys > 0 && (!res_a || !res_b))) return 0;
  auto rows = [&](int n, const int32_t* idx, const int32_t* res) {
    std::vector<std::vector<int32_t>> v;
    for (int i = 0; i < n; ++i) {
      std::vector<int32_t> r{idx[i]};
      r.insert(r.end(), res + (size_t)i * n_keys, res + (size_t)(i + 1) * n_keys);
      v.push_back(r);
    }
    std::sort(v.begin(), v.end());
    return v;
  };
  return rows(n_a, idx_a, res_a) == rows(n_b, idx_b, res_b) ? 1 : 0;  // util.UnorderedEqual
}

}  // extern "C"
