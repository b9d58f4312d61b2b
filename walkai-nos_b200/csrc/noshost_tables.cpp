// noshost_tables.cpp — the MIG geometry-table loader of the drop-in's host side (libnoshost.so, no CUDA).
//
// The reference reads config/gpupartitioner/manager/known_mig_geometries.yaml with sigs.k8s.io/yaml into
// mig.AllowedMigGeometriesList (cmd/gpupartitioner/gpupartitioner.go:122-132), whose elements unmarshal through
// AllowedMigGeometries.UnmarshalJSON (pkg/gpu/mig/allowed_geometries.go:30-59: both fields required, `models` a list
// of strings, `allowedGeometries` a list of {profile: quantity} maps), and flattens it with GroupByModel (:74-82:
// a model named by several entries keeps the LAST entry's list).  This file restates that for the YAML the
// reference ships and tests with (a block sequence of mappings; `models` as a flow or block sequence of scalars;
// `allowedGeometries` as a block sequence of block mappings) and emits the flat arrays nosplan_set_mig_tables takes.
// Validation of names and quantities is not the loader's job in the reference either (ValidateConfigs,
// known_configs.go:164-185, restated in nosplan_set_mig_tables).
#include <algorithm>
#include <cctype>
#include <cstdint>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "noshost.h"

namespace {

struct Line {
  int indent = 0;
  std::string text;  // without indentation, comment and trailing blanks
  int no = 0;
};

struct Entry {
  std::vector<std::string> models;
  std::vector<std::vector<std::pair<std::string, long long>>> geoms;
};

std::string rstrip(std::string s) {
  while (!s.empty() && (s.back() == ' ' || s.back() == '\t' || s.back() == '\r')) s.pop_back();
  return s;
}
std::string strip(const std::string& s) {
  size_t a = 0;
  while (a < s.size() && (s[a] == ' ' || s[a] == '\t')) ++a;
  return rstrip(s.substr(a));
}

// comment: '#' at the start of the text or after a blank, outside quotes
std::string drop_comment(const std::string& s) {
  char q = 0;
  for (size_t i = 0; i < s.size(); ++i) {
    const char c = s[i];
    if (q) {
      if (c == q) q = 0;
    } else if (c == '"' || c == '\'') {
      q = c;
    } else if (c == '#' && (i == 0 || s[i - 1] == ' ' || s[i - 1] == '\t')) {
      return s.substr(0, i);
    }
  }
  return s;
}

std::vector<Line> split_lines(const char* text) {
  std::vector<Line> out;
  std::string cur;
  int no = 0;
  auto flush = [&] {
    ++no;
    std::string t = rstrip(drop_comment(cur));
    cur.clear();
    size_t a = 0;
    while (a < t.size() && t[a] == ' ') ++a;
    if (a == t.size()) return;
    if (t.compare(a, 3, "---") == 0 && a == 0 && t.size() == 3) return;  // document marker
    Line l;
    l.indent = (int)a;
    l.text = t.substr(a);
    l.no = no;
    out.push_back(l);
  };
  for (const char* p = text; *p; ++p) {
    if (*p == '\n') flush();
    else cur.push_back(*p);
  }
  flush();
  return out;
}

// a YAML scalar as the string JSON would carry; quoted says whether it was quoted (then it is a string, never a number)
bool scalar(const std::string& raw, std::string* out, bool* quoted) {
  const std::string s = strip(raw);
  *quoted = false;
  if (s.size() >= 2 && (s.front() == '"' || s.front() == '\'') && s.back() == s.front()) {
    *out = s.substr(1, s.size() - 2);
    *quoted = true;
    return true;
  }
  if (!s.empty() && (s.front() == '"' || s.front() == '\'')) return false;  // unterminated
  *out = s;
  return true;
}

bool parse_int(const std::string& s, long long* v) {
  if (s.empty()) return false;
  size_t i = (s[0] == '-' || s[0] == '+') ? 1 : 0;
  if (i == s.size()) return false;
  long long x = 0;
  for (; i < s.size(); ++i) {
    if (!isdigit((unsigned char)s[i])) return false;
    x = x * 10 + (s[i] - '0');
    if (x > (1ll << 40)) return false;
  }
  *v = s[0] == '-' ? -x : x;
  return true;
}

// "key: value" -> key, value ("" when nothing follows the colon); false when the text is no mapping entry
bool split_entry(const std::string& t, std::string* key, std::string* value) {
  char q = 0;
  for (size_t i = 0; i < t.size(); ++i) {
    const char c = t[i];
    if (q) {
      if (c == q) q = 0;
      continue;
    }
    if (c == '"' || c == '\'') q = c;
    else if (c == ':' && (i + 1 == t.size() || t[i + 1] == ' ')) {
      bool quoted;
      if (!scalar(t.substr(0, i), key, &quoted)) return false;
      *value = strip(t.substr(i + 1));
      return !key->empty();
    }
  }
  return false;
}

// "[ a, "b", 'c' ]" -> strings; false when it is no flow sequence of scalars
bool flow_strings(const std::string& v, std::vector<std::string>* out) {
  if (v.size() < 2 || v.front() != '[' || v.back() != ']') return false;
  const std::string body = v.substr(1, v.size() - 2);
  std::string cur;
  char q = 0;
  std::vector<std::string> raw;
  for (char c : body) {
    if (q) {
      cur.push_back(c);
      if (c == q) q = 0;
    } else if (c == '"' || c == '\'') {
      q = c;
      cur.push_back(c);
    } else if (c == ',') {
      raw.push_back(cur);
      cur.clear();
    } else if (c == '[' || c == '{') {
      return false;
    } else {
      cur.push_back(c);
    }
  }
  if (q) return false;
  if (!strip(cur).empty() || !raw.empty()) raw.push_back(cur);
  for (auto& r : raw) {
    std::string s;
    bool quoted;
    if (!scalar(r, &s, &quoted) || (s.empty() && !quoted)) return false;
    out->push_back(s);
  }
  return true;
}

struct Parser {
  std::vector<Line> L;
  size_t i = 0;
  std::string err;
  bool fail(const Line* l, const std::string& m) {
    err = l ? ("line " + std::to_string(l->no) + ": " + m) : m;
    return false;
  }

  // block sequence of scalars deeper than `indent`
  bool block_strings(int indent, std::vector<std::string>* out) {
    while (i < L.size() && L[i].indent > indent) {
      const Line& l = L[i];
      if (l.text.size() < 1 || l.text[0] != '-') return fail(&l, "expected a list item");
      std::string s;
      bool quoted;
      if (!scalar(l.text.substr(1), &s, &quoted) || (s.empty() && !quoted)) return fail(&l, "expected a string");
      if (!quoted && s.find(": ") != std::string::npos) return fail(&l, "cannot unmarshal object into a model name");
      out->push_back(s);
      ++i;
    }
    return true;
  }

  // block sequence of {profile: quantity} mappings deeper than `indent` (or at it: "key:\n- item" is valid YAML)
  bool block_geometries(int indent, std::vector<std::vector<std::pair<std::string, long long>>>* out) {
    while (i < L.size() && (L[i].indent > indent || (L[i].indent == indent && L[i].text[0] == '-' && false))) {
      const Line& l = L[i];
      if (l.text[0] != '-') return fail(&l, "expected a list item");
      const int item_indent = l.indent;
      std::vector<std::pair<std::string, long long>> g;
      std::string rest = strip(l.text.substr(1));
      int body_indent = item_indent + 1 + (int)(l.text.size() - 1 - strip(l.text.substr(1)).size());  // column of the first key
      ++i;
      auto add = [&](const Line& at, const std::string& t) -> bool {
        std::string k, v;
        if (!split_entry(t, &k, &v)) return fail(&at, "cannot unmarshal into map[ProfileName]int");
        std::string sv;
        bool quoted;
        long long q;
        if (!scalar(v, &sv, &quoted) || quoted || !parse_int(sv, &q)) return fail(&at, "quantity of " + k + " is not an integer");
        for (auto& kv : g)
          if (kv.first == k) {  // a repeated key: the last value stays (encoding/json map semantics)
            kv.second = q;
            return true;
          }
        g.emplace_back(k, q);
        return true;
      };
      if (rest == "{}") {
        // empty geometry
      } else if (!rest.empty()) {
        if (!add(l, rest)) return false;
      }
      while (i < L.size() && L[i].indent > item_indent && L[i].text[0] != '-') {
        if (L[i].indent != body_indent && !rest.empty()) return fail(&L[i], "bad indentation inside a geometry");
        if (rest.empty()) {
          body_indent = L[i].indent;
          rest = "x";
        }
        if (!add(L[i], L[i].text)) return false;
        ++i;
      }
      out->push_back(g);
    }
    return true;
  }

  bool entry(int indent, const Line& first, Entry* e) {
    // first.text starts with '-': its mapping's keys sit at the column of the text after "- "
    std::string rest = strip(first.text.substr(1));
    const int key_indent = indent + 1 + (int)(first.text.size() - 1 - rest.size());
    bool have_models = false, have_geoms = false;
    auto handle = [&](const Line& at, const std::string& t) -> bool {
      std::string k, v;
      if (!split_entry(t, &k, &v)) return fail(&at, "expected `key: value`");
      if (k == "models") {
        e->models.clear();
        if (v.empty()) {
          if (!block_strings(key_indent, &e->models)) return false;
        } else if (v == "[]" ) {
        } else if (!flow_strings(v, &e->models)) {
          return fail(&at, "field 'models' must be a list of strings");  // json: cannot unmarshal string into []gpu.Model
        }
        have_models = true;
      } else if (k == "allowedGeometries") {
        e->geoms.clear();
        if (v.empty()) {
          if (!block_geometries(key_indent, &e->geoms)) return false;
        } else if (v != "[]") {
          return fail(&at, "field 'allowedGeometries' must be a list of {profile: quantity} maps");
        }
        have_geoms = true;
      } else if (v.empty()) {  // an unknown key with a nested value: skip it (map[string]json.RawMessage keeps and ignores it)
        while (i < L.size() && L[i].indent > key_indent) ++i;
      }
      return true;
    };
    ++i;
    if (!rest.empty() && !handle(first, rest)) return false;
    while (i < L.size() && L[i].indent == key_indent && L[i].text[0] != '-') {
      const Line& l = L[i];
      ++i;
      if (!handle(l, l.text)) return false;
    }
    if (i < L.size() && L[i].indent > indent && !(L[i].indent == key_indent)) return fail(&L[i], "bad indentation");
    if (!have_models) return fail(&first, "missing field 'models'");                     // allowed_geometries.go:39-41
    if (!have_geoms) return fail(&first, "missing field 'allowedGeometries'");          // allowed_geometries.go:48-50
    return true;
  }

  bool document(std::vector<Entry>* out) {
    if (L.empty()) return true;  // empty document: an empty list (allowed_geometries_test.go:35-40)
    if (L.size() == 1 && L[0].text == "[]") return true;
    const int indent = L[0].indent;
    while (i < L.size()) {
      const Line& l = L[i];
      if (l.indent != indent || l.text[0] != '-') return fail(&l, "expected a list of {models, allowedGeometries} entries");
      Entry e;
      if (!entry(indent, l, &e)) return false;
      out->push_back(e);
    }
    return true;
  }
};

}  // namespace

struct noshost_tables {
  std::vector<Entry> entries;
  // GroupByModel, flattened: models in order of first appearance, the vocabulary ascending bytewise
  std::vector<std::string> models, profiles;
  std::vector<int32_t> geom_off, geom_counts;
};

extern "C" {

noshost_tables* noshost_tables_load(const char* text, char* err, int err_cap) {
  if (err && err_cap > 0) err[0] = 0;
  Parser ps;
  ps.L = split_lines(text ? text : "");
  auto* t = new noshost_tables();
  if (!ps.document(&t->entries)) {
    if (err && err_cap > 0) {
      std::strncpy(err, ps.err.c_str(), (size_t)err_cap - 1);
      err[err_cap - 1] = 0;
    }
    delete t;
    return nullptr;
  }
  // GroupByModel (allowed_geometries.go:74-82): res[model] = entry's list, later entries overwrite earlier ones
  std::map<std::string, int> entry_of;
  for (int e = 0; e < (int)t->entries.size(); ++e)
    for (const auto& m : t->entries[e].models) {
      if (!entry_of.count(m)) t->models.push_back(m);
      entry_of[m] = e;
    }
  std::vector<std::string> names;
  for (const auto& m : t->models)
    for (const auto& g : t->entries[entry_of[m]].geoms)
      for (const auto& kv : g) names.push_back(kv.first);
  std::sort(names.begin(), names.end());
  names.erase(std::unique(names.begin(), names.end()), names.end());
  t->profiles = names;
  const int P = (int)names.size();
  t->geom_off.push_back(0);
  for (const auto& m : t->models) {
    for (const auto& g : t->entries[entry_of[m]].geoms) {
      std::vector<int32_t> row(std::max(P, 1), 0);
      for (const auto& kv : g)
        row[std::lower_bound(names.begin(), names.end(), kv.first) - names.begin()] = (int32_t)std::max<long long>(std::min<long long>(kv.second, INT32_MAX), INT32_MIN);
      t->geom_counts.insert(t->geom_counts.end(), row.begin(), row.begin() + P);
    }
    t->geom_off.push_back(P ? (int32_t)(t->geom_counts.size() / P) : t->geom_off.back() + (int32_t)t->entries[entry_of[m]].geoms.size());
  }
  return t;
}

void noshost_tables_free(noshost_tables* t) { delete t; }

// Accessors: a null handle or an index outside the table answers 0 / "" / NULL, never reads outside a vector.
namespace {
const Entry* entry_of(const noshost_tables* t, int e) { return t && e >= 0 && e < (int)t->entries.size() ? &t->entries[e] : nullptr; }
const std::pair<std::string, long long>* cell_of(const noshost_tables* t, int e, int g, int k) {
  const Entry* en = entry_of(t, e);
  if (!en || g < 0 || g >= (int)en->geoms.size() || k < 0 || k >= (int)en->geoms[g].size()) return nullptr;
  return &en->geoms[g][k];
}
}  // namespace
int noshost_tables_n_entries(const noshost_tables* t) { return t ? (int)t->entries.size() : 0; }
int noshost_tables_entry_n_models(const noshost_tables* t, int e) {
  const Entry* en = entry_of(t, e);
  return en ? (int)en->models.size() : 0;
}
const char* noshost_tables_entry_model(const noshost_tables* t, int e, int i) {
  const Entry* en = entry_of(t, e);
  return en && i >= 0 && i < (int)en->models.size() ? en->models[i].c_str() : "";
}
int noshost_tables_entry_n_geometries(const noshost_tables* t, int e) {
  const Entry* en = entry_of(t, e);
  return en ? (int)en->geoms.size() : 0;
}
int noshost_tables_entry_geometry_size(const noshost_tables* t, int e, int g) {
  const Entry* en = entry_of(t, e);
  return en && g >= 0 && g < (int)en->geoms.size() ? (int)en->geoms[g].size() : 0;
}
const char* noshost_tables_entry_geometry_profile(const noshost_tables* t, int e, int g, int k) {
  const auto* c = cell_of(t, e, g, k);
  return c ? c->first.c_str() : "";
}
int64_t noshost_tables_entry_geometry_quantity(const noshost_tables* t, int e, int g, int k) {
  const auto* c = cell_of(t, e, g, k);
  return c ? c->second : 0;
}
int noshost_tables_n_models(const noshost_tables* t) { return t ? (int)t->models.size() : 0; }
const char* noshost_tables_model(const noshost_tables* t, int m) { return t && m >= 0 && m < (int)t->models.size() ? t->models[m].c_str() : ""; }
int noshost_tables_n_profiles(const noshost_tables* t) { return t ? (int)t->profiles.size() : 0; }
const char* noshost_tables_profile(const noshost_tables* t, int p) {
  return t && p >= 0 && p < (int)t->profiles.size() ? t->profiles[p].c_str() : "";
}
const int32_t* noshost_tables_geom_off(const noshost_tables* t) { return t ? t->geom_off.data() : nullptr; }
const int32_t* noshost_tables_geom_counts(const noshost_tables* t) { return t ? t->geom_counts.data() : nullptr; }

}  // extern "C"
