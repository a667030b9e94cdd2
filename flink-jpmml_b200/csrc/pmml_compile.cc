// pmml_compile.cc -- the PMML compiler "in the model reader": PMML bytes -> flat, device-ready arrays.
//
// Replaces the model-load step of the reference,
//   PmmlModel.fromReader  flink-jpmml-scala/src/main/scala/io/radicalbit/flink/pmml/scala/api/PmmlModel.scala:53-61
//   (JAXBUtil.unmarshalPMML + ModelEvaluatorFactory.newModelEvaluator from org.jpmml:pmml-evaluator:1.3.9),
// and pre-computes everything the per-event pipeline (PmmlModel.scala:109-174, Pipeline.scala:61-98) decides from the
// document alone: active-field order, per-field input preparation, and how the first target field turns into a Double.
//
// This file is host-only C++; it shares no code with oracle/.
#include <algorithm>
#include <atomic>
#include <charconv>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <deque>
#include <functional>
#include <limits>
#include <map>
#include <optional>
#include <stdexcept>
#include <tuple>
#include <unordered_map>

#include "model.h"
#include "xml.h"

namespace fjp {
namespace {

using sv = std::string_view;

[[noreturn]] void unsupported(const std::string& what) { throw CompileError("unsupported PMML feature: " + what); }
[[noreturn]] void malformed(const std::string& what) { throw CompileError("malformed PMML: " + what); }

sv trim(sv s) {
  while (!s.empty() && (unsigned char)s.front() <= ' ') s.remove_prefix(1);
  while (!s.empty() && (unsigned char)s.back() <= ' ') s.remove_suffix(1);
  return s;
}

// Text -> double with java.lang.Double.parseDouble's grammar (what String.toDouble uses, Pipeline.scala:95).
bool parse_double(sv raw, double* out) {
  sv s = trim(raw);
  if (s.empty()) return false;
  bool neg = false;
  if (s[0] == '+' || s[0] == '-') { neg = s[0] == '-'; s.remove_prefix(1); }
  if (s.empty()) return false;
  if (s == "NaN") { *out = std::numeric_limits<double>::quiet_NaN(); return true; }
  if (s == "Infinity") { *out = neg ? -INFINITY : INFINITY; return true; }
  bool hex = s.size() > 2 && s[0] == '0' && (s[1] == 'x' || s[1] == 'X');
  if (!hex) { char l = s.back(); if (l == 'd' || l == 'D' || l == 'f' || l == 'F') s.remove_suffix(1); }
  if (s.empty()) return false;
  double v;
  if (hex) {
    std::string tmp(s); char* e = nullptr; v = strtod(tmp.c_str(), &e);
    if (!e || *e) return false;
  } else {
    if (!((s[0] >= '0' && s[0] <= '9') || s[0] == '.')) return false;  // no "inf"/"nan" spellings
    auto r = std::from_chars(s.data(), s.data() + s.size(), v, std::chars_format::general);
    if (r.ptr != s.data() + s.size()) return false;
    if (r.ec == std::errc::result_out_of_range) { std::string tmp(s); v = strtod(tmp.c_str(), nullptr); }
  }
  *out = neg ? -v : v;
  return true;
}
bool parse_float(sv raw, double* out) {  // Float.parseFloat: one rounding from decimal text to float
  sv s = trim(raw);
  double d;
  if (!parse_double(s, &d)) return false;
  bool neg = false; sv t = s;
  if (!t.empty() && (t[0] == '+' || t[0] == '-')) { neg = t[0] == '-'; t.remove_prefix(1); }
  if (!t.empty()) { char l = t.back(); if (l == 'd' || l == 'D' || l == 'f' || l == 'F') t.remove_suffix(1); }
  float f;
  auto r = std::from_chars(t.data(), t.data() + t.size(), f, std::chars_format::general);
  if (r.ptr == t.data() + t.size() && r.ec == std::errc()) { *out = neg ? -(double)f : (double)f; return true; }
  *out = (double)(float)d;
  return true;
}
double need_double(sv s, const char* what) {
  double v; if (!parse_double(s, &v)) malformed(std::string("bad number for ") + what + ": '" + std::string(s) + "'");
  return v;
}
uint8_t dtype_of(sv s) {
  if (s == "string") return DT_STRING; if (s == "integer") return DT_INTEGER; if (s == "float") return DT_FLOAT;
  if (s == "double") return DT_DOUBLE; if (s == "boolean") return DT_BOOLEAN; return DT_OTHER;
}
// parse `text` the way a value of a field with dataType dt is parsed; false when it does not parse
bool parse_in_dtype(uint8_t dt, sv text, double* out) {
  switch (dt) {
    case DT_DOUBLE: return parse_double(text, out);
    case DT_FLOAT: return parse_float(text, out);
    case DT_INTEGER: { double v; if (!parse_double(text, &v) || v != std::floor(v) || std::fabs(v) > 2147483647.0) return false; *out = v; return true; }
    case DT_BOOLEAN: { sv t = trim(text); if (t == "true" || t == "TRUE" || t == "True") { *out = 1; return true; }
                       if (t == "false" || t == "FALSE" || t == "False") { *out = 0; return true; }
                       double v; if (parse_double(t, &v) && (v == 0 || v == 1)) { *out = v; return true; } return false; }
    default: return false;
  }
}

std::vector<std::string> split_array(const XNode& arr) {
  std::vector<std::string> out; sv t = arr.text; size_t i = 0;
  while (i < t.size()) {
    while (i < t.size() && (unsigned char)t[i] <= ' ') i++;
    if (i >= t.size()) break;
    std::string tok;
    if (t[i] == '"') { i++; while (i < t.size() && t[i] != '"') { if (t[i] == '\\' && i + 1 < t.size()) i++; tok += t[i++]; } i++; }
    else while (i < t.size() && (unsigned char)t[i] > ' ') tok += t[i++];
    out.push_back(std::move(tok));
  }
  return out;
}

struct Field { std::string name; uint8_t dtype; bool continuous; const XNode* node; };
bool sg_add_layer(const double* W, const double* bias, int N, int K, SgLayer& ly, std::vector<int8_t>& SW, std::vector<int32_t>& TAU);

struct Builder {
  CompiledModel cm;
  std::vector<Field> dict; std::unordered_map<sv, int> dict_idx;   // keys are views of names_ entries (a deque never moves them)
  std::unordered_map<sv, int> active_slot;   // field name -> vector slot
  std::deque<std::string> names_;
  sv keep(sv name) { names_.emplace_back(name); return names_.back(); }
  std::vector<uint8_t> slot_dtype;
  std::string target_name; bool has_target = false; uint8_t target_dtype = DT_DOUBLE;
  // LocalTransformations of models nested in Segments: every DerivedField has a slot of its own in the augmented row (its
  // expression is a pure function of fields known before any model runs, so the pre-pass evaluates it with all the others); its
  // NAME is bound to that slot only while the nested model is compiled (two segments may use one name for different things)
  struct ScopedName { sv name; int dict_index, slot; };
  std::unordered_map<const XNode*, std::vector<ScopedName>> nested_scopes;
  struct Scope {
    Builder& b; std::vector<std::tuple<sv, int, int>> saved;   // name, previous dict index (-1: none), previous slot (-2: none)
    Scope(Builder& bb, const XNode* model) : b(bb) {
      auto it = b.nested_scopes.find(model); if (it == b.nested_scopes.end()) return;
      for (auto& sn : it->second) bind(sn);
    }
    void bind(const ScopedName& sn) {
      auto di = b.dict_idx.find(sn.name); auto si = b.active_slot.find(sn.name);
      saved.emplace_back(sn.name, di == b.dict_idx.end() ? -1 : di->second, si == b.active_slot.end() ? -2 : si->second);
      b.dict_idx[sn.name] = sn.dict_index; b.active_slot[sn.name] = sn.slot;
    }
    ~Scope() {
      for (auto it = saved.rbegin(); it != saved.rend(); ++it) {
        const sv n = std::get<0>(*it);
        if (std::get<1>(*it) < 0) b.dict_idx.erase(n); else b.dict_idx[n] = std::get<1>(*it);
        if (std::get<2>(*it) == -2) b.active_slot.erase(n); else b.active_slot[n] = std::get<2>(*it);
      }
    }
  };

  template <class T> void put(const T* const* field_addr, const std::vector<T>& v) {
    size_t off = (cm.blob.size() + 255) & ~size_t(255);
    cm.blob.resize(off + std::max<size_t>(v.size() * sizeof(T), 16));
    if (!v.empty()) memcpy(cm.blob.data() + off, v.data(), v.size() * sizeof(T));
    cm.fixups.push_back({off, (size_t)((const char*)field_addr - (const char*)&cm.dev)});
  }
  int slot_of(sv name) const {  // -1: not an active field of the top-level schema -> always missing
    auto it = active_slot.find(name);
    if (it != active_slot.end()) return it->second;
    if (!dict_idx.count(name)) malformed("reference to unknown field '" + std::string(name) + "'");
    return -1;
  }
  const Field& dict_field(sv name) const {
    auto it = dict_idx.find(name);
    if (it == dict_idx.end()) malformed("unknown field '" + std::string(name) + "'");
    return dict[it->second];
  }

  // class label / cluster id -> what Pipeline.extractTargetValue (Pipeline.scala:94-98) makes of it
  void set_labels(const std::vector<std::string>& labels, bool clustering) {
    cm.labels = labels;
    std::vector<double> val(labels.size()); std::vector<uint8_t> st(labels.size());
    uint8_t tdt = clustering ? (uint8_t)DT_STRING : target_dtype;  // entity ids are always strings
    for (size_t i = 0; i < labels.size(); i++) {
      double v = std::numeric_limits<double>::quiet_NaN(); uint8_t s = ST_EXTRACT;
      if ((tdt == DT_STRING || tdt == DT_DOUBLE) && parse_double(labels[i], &v)) s = ST_OK;  // String.toDouble / Double
      val[i] = s == ST_OK ? v : std::numeric_limits<double>::quiet_NaN(); st[i] = s;
    }
    cm.dev.n_labels = (int)labels.size();
    put(&cm.dev.label_value, val); put(&cm.dev.label_status, st);
  }
};

// ---------------------------------------------------------------------------------------------- schema + prep
void read_dictionary(const XNode& root, Builder& b) {
  const XNode* dd = root.child("DataDictionary");
  if (!dd) malformed("no DataDictionary");
  dd->each("DataField", [&](const XNode& f) {
    Field fl{std::string(f.attr("name")), dtype_of(f.attr("dataType")), f.attr("optype") == "continuous", &f};
    b.dict_idx[b.keep(fl.name)] = (int)b.dict.size(); b.dict.push_back(fl);
  });
}

void check_nested_schema(const XNode& model, bool transformations_ok = false) {
  if (const XNode* lt = model.child("LocalTransformations")) if (lt->first && !transformations_ok) unsupported("LocalTransformations in this nested model");
  if (const XNode* ms = model.child("MiningSchema"))
    ms->each("MiningField", [&](const XNode& f) {
      if (f.has("missingValueReplacement") || f.attr("outliers", "asIs") != "asIs" ||
          (f.has("invalidValueTreatment") && f.attr("invalidValueTreatment") != "returnInvalid" && f.attr("invalidValueTreatment") != "asIs"))
        unsupported("value treatments on a nested MiningSchema");
    });
}

void read_schema(const XNode& model, Builder& b) {
  const XNode* ms = model.child("MiningSchema");
  if (!ms) malformed("no MiningSchema");
  std::vector<PrepField> prep; std::vector<double> pool; bool simple = true, identity = true;
  ms->each("MiningField", [&](const XNode& f) {
    sv usage = f.attr("usageType", "active");
    std::string name(f.attr("name"));
    if (usage == "target" || usage == "predicted") {
      if (!b.has_target) { b.has_target = true; b.target_name = name; b.target_dtype = b.dict_field(name).dtype; }
      return;
    }
    if (usage != "active") return;
    const Field& df = b.dict_field(name);
    PrepField p{}; p.dtype = df.dtype; p.continuous = df.continuous; p.pool_off = (uint32_t)pool.size();
    sv ivt = f.attr("invalidValueTreatment", "returnInvalid");
    p.ivt = ivt == "asIs" ? IVT_AS_IS : ivt == "asMissing" ? IVT_AS_MISSING : IVT_RETURN_INVALID;
    sv ot = f.attr("outliers", "asIs");
    p.outlier = ot == "asMissingValues" ? OT_AS_MISSING : ot == "asExtremeValues" ? OT_AS_EXTREME : OT_AS_IS;
    if (f.has("lowValue")) { p.has_low = 1; p.low = need_double(f.attr("lowValue"), "lowValue"); }
    if (f.has("highValue")) { p.has_high = 1; p.high = need_double(f.attr("highValue"), "highValue"); }
    if (f.has("missingValueReplacement")) {
      double v;
      if (parse_in_dtype(df.dtype, f.attr("missingValueReplacement"), &v)) { p.has_repl = 1; p.repl = v; }
      else if (df.dtype == DT_STRING) unsupported("missingValueReplacement on a string field");
      else p.has_repl = 2;  // replacement text does not parse in the field's dataType -> preparation fails
    }
    bool numeric = df.dtype == DT_DOUBLE || df.dtype == DT_FLOAT || df.dtype == DT_INTEGER || df.dtype == DT_BOOLEAN;
    if (numeric) {
      if (df.continuous) df.node->each("Interval", [&](const XNode& iv) {
        sv c = iv.attr("closure"); double flags = 0, lo = 0, hi = 0;
        if (iv.has("leftMargin")) { flags += 1; lo = need_double(iv.attr("leftMargin"), "leftMargin"); }
        if (iv.has("rightMargin")) { flags += 2; hi = need_double(iv.attr("rightMargin"), "rightMargin"); }
        if (c.substr(0, 4) == "open") flags += 4;
        if (c.size() >= 4 && c.substr(c.size() - 4) == "Open") flags += 8;
        pool.push_back(lo); pool.push_back(hi); pool.push_back(flags); p.n_intervals++;
      });
      // a listed value that does not parse in the field's dataType can never equal an input: dropped, but a
      // non-empty valid list still defines the valid space
      std::vector<double> valid, invalid, missing; bool any_valid = false;
      df.node->each("Value", [&](const XNode& v) {
        sv prop = v.attr("property", "valid"); double x; bool ok = parse_in_dtype(df.dtype, v.attr("value"), &x);
        if (prop == "valid") { any_valid = true; if (ok) valid.push_back(x); }
        else if (ok) (prop == "invalid" ? invalid : missing).push_back(x);
      });
      p.has_valid_space = any_valid ? 1 : 0;
      p.n_valid = (uint16_t)valid.size(); p.n_invalid = (uint16_t)invalid.size(); p.n_missing = (uint16_t)missing.size();
      pool.insert(pool.end(), valid.begin(), valid.end());
      pool.insert(pool.end(), invalid.begin(), invalid.end());
      pool.insert(pool.end(), missing.begin(), missing.end());
    }
    bool plain = (df.dtype == DT_DOUBLE || df.dtype == DT_FLOAT) && p.n_intervals == 0 && !p.has_valid_space && p.n_invalid == 0 &&
                 p.n_missing == 0 && p.has_repl == 0 && p.outlier == OT_AS_IS;
    simple = simple && plain; identity = identity && plain && df.dtype == DT_DOUBLE;
    b.active_slot[b.keep(name)] = (int)prep.size(); b.slot_dtype.push_back(df.dtype);
    b.cm.active_names.push_back(name); prep.push_back(p);
  });
  b.cm.dev.F = (int)prep.size(); b.cm.dev.prep_simple = simple ? 1 : 0; b.cm.dev.prep_identity = identity ? 1 : 0;
  b.cm.dev.F_in = b.cm.dev.F; b.cm.dev.prep_in_simple = simple ? 1 : 0;
  b.put(&b.cm.dev.prep, prep); b.put(&b.cm.dev.prep_pool, pool);
  b.put(&b.cm.dev.prep_in, prep); b.put(&b.cm.dev.prep_in_pool, pool);
  // <Targets>
  TargetXform tx{1.0, 0.0, 0.0, 0.0, 0, 0, 0};
  if (const XNode* ts = model.child("Targets")) {
    bool done = false;
    ts->each("Target", [&](const XNode& t) {
      if (done) return;
      if (t.has("field") && b.has_target && t.attr("field") != b.target_name) return;
      if (t.has("rescaleFactor")) tx.factor = need_double(t.attr("rescaleFactor"), "rescaleFactor");
      if (t.has("rescaleConstant")) tx.constant = need_double(t.attr("rescaleConstant"), "rescaleConstant");
      if (t.has("min")) { tx.has_min = 1; tx.mn = need_double(t.attr("min"), "min"); }
      if (t.has("max")) { tx.has_max = 1; tx.mx = need_double(t.attr("max"), "max"); }
      sv c = t.attr("castInteger"); tx.cast = c == "round" ? 1 : c == "ceiling" ? 2 : c == "floor" ? 3 : 0;
      done = true;
    });
  }
  b.cm.dev.tx = tx;
}

// ---------------------------------------------------------------------------------------------- derived fields
struct DerivedCompiler {
  Builder& b; std::vector<DxOp> ops; std::vector<double> pool;
  // DefineFunction: calls are expanded in place.  Inside a body only the ParameterFields are visible; a FieldRef to one
  // re-emits the argument's expression in the caller's scope (expressions are pure), then applies its own mapMissingTo.
  struct Scope { std::map<std::string, const XNode*> params; const Scope* caller; int depth; };
  std::map<std::string, const XNode*> fns;
  explicit DerivedCompiler(Builder& bb) : b(bb) {}
  static DxOp mk(int op) { DxOp o{}; o.op = op; return o; }
  int numeric_slot(sv field) {
    uint8_t dt = b.dict_field(field).dtype;
    if (dt == DT_STRING || dt == DT_OTHER) unsupported("transformation of a non-numeric field");
    int s = b.slot_of(field);
    if (s < 0) unsupported("transformation of a field that is not an active field");
    return s;
  }
  void opt(const XNode& e, DxOp& o) {
    if (e.has("mapMissingTo")) { o.flags |= 1; o.x = need_double(e.attr("mapMissingTo"), "mapMissingTo"); }
    if (e.has("defaultValue")) { o.flags |= 2; o.y = need_double(e.attr("defaultValue"), "defaultValue"); }
  }
  void emit(const XNode& e, int depth, const Scope* sc = nullptr) {
    if (depth > DX_STACK - 2) unsupported("transformation nested deeper than 14 levels");
    if (sc) {
      if (e.name == "FieldRef") {
        auto it = sc->params.find(std::string(e.attr("field")));
        if (it == sc->params.end()) unsupported("DefineFunction body refers to a field that is not a ParameterField");
        emit(*it->second, depth, sc->caller);
        if (e.has("mapMissingTo")) { DxOp o = mk(DX_APPLY); o.a = DF_COALESCE; o.b = 1; opt(e, o); ops.push_back(o); }
        return;
      }
      if (e.name != "Constant" && e.name != "Apply") unsupported("DefineFunction body element " + std::string(e.name));
    }
    if (e.name == "Constant") { DxOp o = mk(DX_CONST); o.x = need_double(e.text, "Constant"); ops.push_back(o); return; }
    if (e.name == "FieldRef") { DxOp o = mk(DX_FIELD); o.a = numeric_slot(e.attr("field")); opt(e, o); ops.push_back(o); return; }
    if (e.name == "NormDiscrete") {
      DxOp o = mk(DX_NORMDISC); o.a = numeric_slot(e.attr("field")); opt(e, o);
      if (!parse_in_dtype(b.dict_field(e.attr("field")).dtype, e.attr("value"), &o.z)) unsupported("NormDiscrete value does not parse in the field's dataType");
      ops.push_back(o); return;
    }
    if (e.name == "NormContinuous") {
      DxOp o = mk(DX_NORMCONT); o.a = numeric_slot(e.attr("field")); opt(e, o); o.b = (int32_t)pool.size();
      sv ol = e.attr("outliers", "asIs"); o.outliers = ol == "asMissingValues" ? 1 : ol == "asExtremeValues" ? 2 : 0;
      e.each("LinearNorm", [&](const XNode& ln) { pool.push_back(need_double(ln.attr("orig"), "orig")); pool.push_back(need_double(ln.attr("norm"), "norm")); o.c++; });
      if (o.c < 2) malformed("NormContinuous needs two LinearNorm");
      ops.push_back(o); return;
    }
    if (e.name == "Discretize") {
      DxOp o = mk(DX_DISCRETIZE); o.a = numeric_slot(e.attr("field")); opt(e, o); o.b = (int32_t)pool.size();
      e.each("DiscretizeBin", [&](const XNode& bin) {
        const XNode* iv = bin.child("Interval"); if (!iv) malformed("DiscretizeBin without Interval");
        sv c = iv->attr("closure"); double flags = 0, lo = 0, hi = 0;
        if (iv->has("leftMargin")) { flags += 1; lo = need_double(iv->attr("leftMargin"), "leftMargin"); }
        if (iv->has("rightMargin")) { flags += 2; hi = need_double(iv->attr("rightMargin"), "rightMargin"); }
        if (c.substr(0, 4) == "open") flags += 4;
        if (c.size() >= 4 && c.substr(c.size() - 4) == "Open") flags += 8;
        pool.push_back(lo); pool.push_back(hi); pool.push_back(flags); pool.push_back(need_double(bin.attr("binValue"), "binValue")); o.c++;
      });
      ops.push_back(o); return;
    }
    if (e.name == "MapValues") {
      DxOp o = mk(DX_MAPVALUES); opt(e, o); o.b = (int32_t)pool.size();
      if (e.count("FieldColumnPair") != 1) unsupported("MapValues with several FieldColumnPairs");
      const XNode& fc = *e.child("FieldColumnPair"); o.a = numeric_slot(fc.attr("field"));
      sv in_col = fc.attr("column"), out_col = e.attr("outputColumn");
      const XNode* tbl = e.child("InlineTable"); if (!tbl) unsupported("MapValues without InlineTable");
      uint8_t dt = b.dict_field(fc.attr("field")).dtype;
      tbl->each("row", [&](const XNode& row) {
        const XNode* i = row.child(in_col); const XNode* ot = row.child(out_col);
        if (!i || !ot) malformed("MapValues row without both columns");
        double iv; if (!parse_in_dtype(dt == DT_FLOAT ? DT_DOUBLE : dt, i->text, &iv)) unsupported("MapValues input does not parse as a number");
        pool.push_back(iv); pool.push_back(need_double(ot->text, "MapValues output")); o.c++;
      });
      ops.push_back(o); return;
    }
    if (e.name == "Apply") {
      static const char* fn[] = {"+", "-", "*", "/", "min", "max", "sum", "avg", "product", "pow", "exp", "ln", "log10", "sqrt", "abs", "floor", "ceil",
                                 "round", "threshold", "equal", "notEqual", "lessThan", "lessOrEqual", "greaterThan", "greaterOrEqual", "and", "or", "not",
                                 "isIn", "isNotIn", "if", "isMissing", "isNotMissing"};
      sv f = e.attr("function"); int id = -1;
      auto uf = fns.find(std::string(f));
      if (uf != fns.end()) {   // user function: [arguments (only when the Apply's mapMissingTo has to look at them)] body -> DF_USERCALL
        const XNode& df = *uf->second; const int level = sc ? sc->depth + 1 : 1;
        if (level > 8) unsupported("DefineFunction calls nested deeper than 8");
        uint8_t fdt = dtype_of(df.attr("dataType", "double"));
        if (fdt != DT_DOUBLE && fdt != DT_FLOAT) unsupported("DefineFunction dataType other than double / float");
        std::vector<const XNode*> args; for (const XNode* k = e.first; k; k = k->next) if (k->name != "Extension") args.push_back(k);
        Scope inner{{}, sc, level}; size_t i = 0; const XNode* body = nullptr;
        for (const XNode* k = df.first; k; k = k->next) {
          if (k->name == "Extension") continue;
          if (k->name == "ParameterField") {
            if (k->has("dataType") && k->attr("dataType") != "double") unsupported("ParameterField dataType other than double");
            if (i >= args.size()) unsupported("too few arguments for " + std::string(f));
            inner.params[std::string(k->attr("name"))] = args[i++];
          } else body = k;
        }
        if (i != args.size() || !body) unsupported("wrong number of arguments for " + std::string(f));
        int pushed = 0;
        if (e.has("mapMissingTo")) for (const XNode* a : args) { emit(*a, depth + pushed, sc); pushed++; }
        emit(*body, depth + pushed, &inner); pushed++;
        if (depth + pushed > DX_STACK) unsupported("Apply " + std::string(f) + " with " + std::to_string(args.size()) + " arguments");
        DxOp o = mk(DX_APPLY); o.a = DF_USERCALL; o.b = pushed; o.outliers = fdt == DT_FLOAT ? 1 : 0; opt(e, o); ops.push_back(o); return;
      }
      for (int i = 0; i < 33; i++) if (f == fn[i]) id = i;
      if (id < 0) unsupported("Apply function " + std::string(f));
      int nargs = 0;
      for (const XNode* k = e.first; k; k = k->next) if (k->name != "Extension") { emit(*k, depth + nargs, sc); nargs++; }
      static const int arity[] = {2, 2, 2, 2, -1, -1, -1, -1, -1, 2, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 2, 2, 2, -1, -1, 1, -2, -2, -3, 1, 1};
      int ar = arity[id];
      if ((ar > 0 && nargs != ar) || (ar == -1 && nargs < 1) || (ar == -2 && nargs < 2) || (ar == -3 && (nargs < 2 || nargs > 3)) || depth + nargs > DX_STACK)
        unsupported("Apply " + std::string(f) + " with " + std::to_string(nargs) + " arguments");
      DxOp o = mk(DX_APPLY); o.a = id; o.b = nargs; opt(e, o); ops.push_back(o); return;
    }
    unsupported("transformation element " + std::string(e.name));
  }
};

// TransformationDictionary + the top-level model's LocalTransformations.  Derived field d becomes slot F_in + d of the
// augmented row; everything compiled after this sees it like any other numeric field.
void read_derived(const XNode& root, const XNode& model, Builder& b) {
  std::vector<const XNode*> fields;
  if (const XNode* td = root.child("TransformationDictionary")) {
    td->each("DerivedField", [&](const XNode& f) { fields.push_back(&f); });
  }
  if (const XNode* lt = model.child("LocalTransformations")) lt->each("DerivedField", [&](const XNode& f) { fields.push_back(&f); });
  DerivedDesc& dd = b.cm.dev.derived; dd.n_derived = (int)fields.size();
  DerivedCompiler dc(b); std::vector<int32_t> prog_off{0}; std::vector<uint8_t> dtypes;
  if (const XNode* td = root.child("TransformationDictionary")) td->each("DefineFunction", [&](const XNode& f) { dc.fns[std::string(f.attr("name"))] = &f; });
  for (const XNode* f : fields) {
    uint8_t dt = dtype_of(f->attr("dataType", "double"));
    if (dt != DT_DOUBLE && dt != DT_FLOAT && dt != DT_INTEGER) unsupported("non-numeric DerivedField");
    const XNode* ex = nullptr;
    for (const XNode* k = f->first; k && !ex; k = k->next) if (k->name != "Extension" && k->name != "Value") ex = k;
    if (!ex) malformed("DerivedField without expression");
    dc.emit(*ex, 0);
    prog_off.push_back((int32_t)dc.ops.size()); dtypes.push_back(dt);
    std::string name(f->attr("name"));
    Field fl{name, dt, f->attr("optype", "continuous") == "continuous", nullptr};
    b.dict_idx[b.keep(name)] = (int)b.dict.size(); b.dict.push_back(fl);
    b.active_slot[b.keep(name)] = b.cm.dev.F; b.slot_dtype.push_back(dt); b.cm.dev.F++;
  }
  {  // nested models (TreeModels of a Segmentation, the MiningModels inside a modelChain): their LocalTransformations, in document order
    std::function<void(const XNode&)> scan = [&](const XNode& parent) {
      const XNode* sg = parent.child("Segmentation"); if (!sg) return;
      sg->each("Segment", [&](const XNode& seg) {
        for (const XNode* nm = seg.first; nm; nm = nm->next) {
          if (nm->name != "TreeModel" && nm->name != "MiningModel") continue;
          Builder::Scope scope(b, nullptr);   // names of this model's fields become visible one by one, to the fields after them
          if (const XNode* lt = nm->child("LocalTransformations")) lt->each("DerivedField", [&](const XNode& f) {
            uint8_t dt = dtype_of(f.attr("dataType", "double"));
            if (dt != DT_DOUBLE && dt != DT_FLOAT && dt != DT_INTEGER) unsupported("non-numeric DerivedField");
            const XNode* ex = nullptr;
            for (const XNode* k = f.first; k && !ex; k = k->next) if (k->name != "Extension" && k->name != "Value") ex = k;
            if (!ex) malformed("DerivedField without expression");
            dc.emit(*ex, 0);
            prog_off.push_back((int32_t)dc.ops.size()); dtypes.push_back(dt);
            const sv user = b.keep(f.attr("name"));
            const sv unique = b.keep(std::string(user) + "\x01" + std::to_string(b.dict.size()));
            Field fl{std::string(unique), dt, f.attr("optype", "continuous") == "continuous", nullptr};
            const int di = (int)b.dict.size(), slot = b.cm.dev.F;
            b.dict_idx[unique] = di; b.dict.push_back(fl); b.slot_dtype.push_back(dt); b.cm.dev.F++;
            Builder::ScopedName sn{user, di, slot};
            b.nested_scopes[nm].push_back(sn); scope.bind(sn);
          });
          if (nm->name == "MiningModel") scan(*nm);   // (with this model's names still bound)
        }
      });
    };
    scan(model);
    dd.n_derived = (int)dtypes.size();
  }
  b.put(&dd.prog_off, prog_off); b.put(&dd.ops, dc.ops); b.put(&dd.pool, dc.pool); b.put(&dd.dtype, dtypes);
  if (dd.n_derived) {   // the family kernel reads already prepared values: identity preparation over the augmented row
    std::vector<PrepField> ident(b.cm.dev.F); std::vector<double> nopool;
    for (auto& p : ident) { memset(&p, 0, sizeof(p)); p.dtype = DT_DOUBLE; p.continuous = 1; }
    b.put(&b.cm.dev.prep, ident); b.put(&b.cm.dev.prep_pool, nopool);
    b.cm.dev.prep_simple = 1; b.cm.dev.prep_identity = 1;
  }
}

// ---------------------------------------------------------------------------------------------- clustering
void compile_clustering(const XNode& m, Builder& b) {
  b.cm.family = "ClusteringModel"; b.cm.dev.kind = MK_CLUSTERING;
  if (m.attr("modelClass") != "centerBased") unsupported("distribution-based ClusteringModel");
  const XNode* cmz = m.child("ComparisonMeasure"); if (!cmz) malformed("no ComparisonMeasure");
  if (cmz->attr("kind") != "distance") unsupported("similarity ComparisonMeasure");
  sv dflt_cmp = cmz->attr("compareFunction", "absDiff");
  ClusterDesc& d = b.cm.dev.cluster; d.p = 2; d.measure = -1;
  for (const XNode* k = cmz->first; k; k = k->next) {
    if (k->name == "Extension") continue;
    if (k->name == "euclidean") d.measure = CM_EUCLIDEAN; else if (k->name == "squaredEuclidean") d.measure = CM_SQEUCLIDEAN;
    else if (k->name == "chebychev") d.measure = CM_CHEBYCHEV; else if (k->name == "cityBlock") d.measure = CM_CITYBLOCK;
    else if (k->name == "minkowski") { d.measure = CM_MINKOWSKI; d.p = need_double(k->attr("p-parameter"), "p-parameter"); }
    else unsupported("comparison measure " + std::string(k->name));
  }
  if (d.measure < 0) malformed("ComparisonMeasure without a measure");
  std::vector<int32_t> slot, cmp; std::vector<double> w, q, scale;
  m.each("ClusteringField", [&](const XNode& cf) {
    if (cf.attr("isCenterField", "true") != "true") return;
    slot.push_back(b.slot_of(cf.attr("field")));
    w.push_back(need_double(cf.attr("fieldWeight", "1"), "fieldWeight"));
    sv c = cf.attr("compareFunction", dflt_cmp);
    cmp.push_back(c == "absDiff" ? CF_ABSDIFF : c == "gaussSim" ? CF_GAUSSSIM : c == "delta" ? CF_DELTA : c == "equal" ? CF_EQUAL : -1);
    if (cmp.back() < 0) unsupported("compareFunction " + std::string(c));
    scale.push_back(cf.has("similarityScale") ? need_double(cf.attr("similarityScale"), "similarityScale") : 0.0);
  });
  size_t nf = slot.size(); if (!nf) malformed("no ClusteringField");
  if (const XNode* mw = m.child("MissingValueWeights")) { if (const XNode* a = mw->child("Array")) for (auto& t : split_array(*a)) q.push_back(need_double(t, "MissingValueWeights")); }
  if (q.empty()) q.assign(nf, 1.0);
  if (q.size() != nf) malformed("MissingValueWeights size mismatch");
  std::vector<double> centers; std::vector<std::string> ids; int pos = 0;
  m.each("Cluster", [&](const XNode& c) {
    pos++; ids.push_back(c.has("id") ? std::string(c.attr("id")) : std::to_string(pos));  // EntityUtil.getId: 1-based position
    const XNode* a = c.child("Array"); if (!a) malformed("Cluster without Array");
    auto toks = split_array(*a); if (toks.size() != nf) malformed("Cluster Array size mismatch");
    for (auto& t : toks) centers.push_back(need_double(t, "Cluster Array"));
  });
  if (ids.empty()) malformed("no Cluster");
  d.k = (int)ids.size(); d.nf = (int)nf;
  bool fast = d.measure == CM_SQEUCLIDEAN || d.measure == CM_EUCLIDEAN;
  for (size_t i = 0; i < nf; i++) fast = fast && cmp[i] == CF_ABSDIFF && w[i] == 1.0 && q[i] == 1.0 && slot[i] >= 0;
  d.fast = fast ? 1 : 0;
  d.slots_identity = (int)nf == b.cm.dev.F ? 1 : 0;
  for (size_t i = 0; i < nf; i++) if (slot[i] != (int)i) d.slots_identity = 0;
  b.put(&d.slot, slot); b.put(&d.weight, w); b.put(&d.q, q); b.put(&d.compare, cmp); b.put(&d.sim_scale, scale); b.put(&d.centers, centers);
  {  // FMA form for the dense tile kernel (model.h): -2 c transposed, |c|^2
    std::vector<double> fw, fi, fq; d.cmax = 0; bool finite = true;
    if (fast && d.slots_identity && d.k <= 64) {
      size_t tcs = 8, bestpad = ~size_t(0); for (size_t cand : {size_t(8), size_t(10), size_t(12)}) { size_t pad = ((size_t)d.k + cand - 1) / cand * cand - (size_t)d.k; if (pad < bestpad) { bestpad = pad; tcs = cand; } }
      const size_t KP = ((size_t)d.k + tcs - 1) / tcs * tcs; fw.assign(KP * nf, 0.0); fi.assign(KP, 0.0); fq.assign(KP * (nf + 1), 0.0);   // fq: one more, all-zero row
      for (int t = 0; t < d.k; t++) {
        double s2 = 0;
        for (size_t f = 0; f < nf; f++) { const double c = centers[(size_t)t * nf + f]; finite = finite && std::isfinite(c); fw[f * KP + t] = -2.0 * c; fq[f * KP + t] = c * c; s2 += c * c; d.cmax = std::max(d.cmax, std::fabs(c)); }
        fi[t] = s2;
      }
      if (!finite || !(d.cmax < 1e150)) { fw.clear(); fi.clear(); fq.clear(); }
    }
    if (fw.empty()) { d.fast_wt = nullptr; d.fast_icpt = nullptr; d.fast_csq = nullptr; }
    else { b.put(&d.fast_wt, fw); b.put(&d.fast_icpt, fi); b.put(&d.fast_csq, fq); }
  }
  b.set_labels(ids, true);
  b.cm.dev.target_status = b.has_target ? ST_OK : ST_EXTRACT;  // no named target field -> nothing to extract
}

// ---------------------------------------------------------------------------------------------- trees
struct Thr { int slot; double value; bool strict; };  // test: x <= value (strict: x < value)
struct TNode { int thr; int32_t on_true, on_false, on_missing; };  // thr index; thr<0: "always true when present"
struct TLeaf { double value; int32_t label; int32_t node_id; std::vector<std::pair<int32_t, double>> dist; };

struct ForestBuilder {
  Builder& b; bool classify;
  std::vector<Thr> thr; std::vector<TNode> nodes; std::vector<int> node_slot; std::vector<TLeaf> leaves; std::vector<int32_t> roots;
  std::vector<double> weights; std::vector<std::string> labels; std::unordered_map<std::string, int> label_idx;
  std::vector<int32_t> chain_channels;   // modelChain: channel of every tree (empty otherwise)
  // Segment predicates: a segment whose predicate is not TRUE leaves the aggregation (LEAF_SKIP).  For sum / weightedSum
  // with at least one always-active segment, "leaves" is the same as "adds +0.0" bit for bit (the running sum starts at
  // +0.0 and x + (+-0.0) == x for every x it can hold), so the skip becomes one shared zero leaf with node id -1 and the
  // ensemble keeps its on-chip forms.
  bool skip_as_zero = false; int32_t zero_skip_leaf = 0;
  int32_t skip_successor() {
    if (!skip_as_zero) return LEAF_SKIP;
    if (zero_skip_leaf == 0) { leaves.push_back(TLeaf{0.0, -1, -1, {}}); zero_skip_leaf = ~(int32_t)(leaves.size() - 1); }
    return zero_skip_leaf;
  }
  explicit ForestBuilder(Builder& bb, bool c) : b(bb), classify(c) {}

  int label_of(sv s) { std::string k(s); auto it = label_idx.find(k); if (it != label_idx.end()) return it->second; int i = (int)labels.size(); labels.push_back(k); label_idx[k] = i; return i; }

  struct TreeCtx {
    int mvs;  // 0 none, 1 defaultChild, 2 lastPrediction, 3 nullPrediction
    bool last_on_no_true; int counter = 0; TargetXform sub_tx; bool has_sub_tx = false;
    uint32_t epoch = 0;   // per-Node memos live in XNode::scratch: [0] node number, [1] entered, [2] leaf, [3] chain from this child on
  };
  enum : int { M_NUMBER = 0, M_ENTERED = 1, M_LEAF = 2, M_CHAIN = 3 };
  static uint32_t next_epoch() { static std::atomic<uint32_t> e{0}; return ++e; }   // documents are not shared between threads; the counter is
  static bool memo_get(const XNode& n, const TreeCtx& cx, int slot, int32_t* v) {
    if (n.scratch_epoch != cx.epoch || !(n.scratch_set & (1u << slot))) return false;
    *v = n.scratch[slot]; return true;
  }
  static void memo_put(const XNode& n, const TreeCtx& cx, int slot, int32_t v) {
    if (n.scratch_epoch != cx.epoch) { n.scratch_epoch = cx.epoch; n.scratch_set = 0; }
    n.scratch[slot] = v; n.scratch_set |= 1u << slot;
  }
  static const XNode* next_node(const XNode* k) { for (k = k ? k->next : nullptr; k && k->name != "Node"; k = k->next) {} return k; }
  static const XNode* first_node(const XNode& p) { const XNode* k = p.first; while (k && k->name != "Node") k = k->next; return k; }

  static const XNode* pred_of(const XNode& n) {
    for (const XNode* k = n.first; k; k = k->next)
      if (k->name == "True" || k->name == "False" || k->name == "SimplePredicate" || k->name == "SimpleSetPredicate" || k->name == "CompoundPredicate") return k;
    malformed("Node/Segment without predicate");
  }
  void number_nodes(const XNode& n, TreeCtx& cx) {
    cx.counter++; double v; sv id = n.attr("id");
    memo_put(n, cx, M_NUMBER, (!id.empty() && parse_double(id, &v) && v == std::floor(v) && std::fabs(v) < 2e9) ? (int)v : cx.counter);
    for (const XNode* k = first_node(n); k; k = next_node(k)) number_nodes(*k, cx);
  }
  int32_t leaf_for(const XNode& n, TreeCtx& cx) {  // successor for "the result is node n's score"
    if (!n.has("score")) return LEAF_NULL;
    int32_t memo; if (memo_get(n, cx, M_LEAF, &memo)) return memo;
    int32_t number = 0; memo_get(n, cx, M_NUMBER, &number);
    TLeaf lf{0.0, -1, number, {}};
    if (classify) {
      lf.label = label_of(n.attr("score"));
      // ScoreDistribution: `probability` if given, else recordCount / sum (JPMML; SURVEY App. A TreeModel)
      std::vector<std::pair<int32_t, double>> raw; bool is_prob = false; double tot = 0;
      n.each("ScoreDistribution", [&](const XNode& sd) {
        double v;
        if (sd.has("probability")) { v = need_double(sd.attr("probability"), "probability"); is_prob = true; }
        else v = need_double(sd.attr("recordCount"), "recordCount");
        raw.emplace_back(label_of(sd.attr("value")), v); tot += v;
      });
      if (raw.empty()) lf.dist.emplace_back(lf.label, 1.0);
      else for (auto& r : raw) lf.dist.emplace_back(r.first, is_prob ? r.second : r.second / tot);
    }
    else {
      if (!parse_double(n.attr("score"), &lf.value)) unsupported("non-numeric regression score '" + std::string(n.attr("score")) + "'");
      if (cx.has_sub_tx) {  // a segment model's own <Targets> act on its result before aggregation
        const TargetXform& t = cx.sub_tx; double v = lf.value;
        if (t.has_min && v < t.mn) v = t.mn; if (t.has_max && v > t.mx) v = t.mx; v = v * t.factor + t.constant;
        if (t.cast == 1) v = std::floor(v + 0.5); else if (t.cast == 2) v = std::ceil(v); else if (t.cast == 3) v = std::floor(v);
        lf.value = v;
      }
    }
    leaves.push_back(lf); int32_t s = ~(int32_t)(leaves.size() - 1); memo_put(n, cx, M_LEAF, s); return s;
  }
  // One decision node.  presence_only: "passes whenever the value is there" (thr = -1; the false branch is never taken).
  // Nodes are hash-consed, and a test all of whose outcomes lead to the same place is not emitted at all: compound
  // predicates then collapse to graphs in which a missing value follows one of the two branches wherever the logic
  // allows it, which is what the on-chip tree forms need.
  struct NodeKey { uint64_t vb; int32_t slot, kind, t, f, m; bool operator==(const NodeKey& o) const { return vb == o.vb && slot == o.slot && kind == o.kind && t == o.t && f == o.f && m == o.m; } };
  struct NodeKeyHash { size_t operator()(const NodeKey& k) const {
    uint64_t h = k.vb * 0x9E3779B97F4A7C15ull; for (uint32_t v : {(uint32_t)k.slot, (uint32_t)k.kind, (uint32_t)k.t, (uint32_t)k.f, (uint32_t)k.m}) h = (h ^ v) * 0x100000001B3ull + (h >> 29);
    return (size_t)h; } };
  std::unordered_map<NodeKey, int32_t, NodeKeyHash> node_memo; int cons = 0;   // consing is on inside compound / Segment / root predicates only
  int32_t add_test(int slot, double value, bool strict, int32_t t, int32_t f, int32_t m, bool presence_only = false) {
    if (slot < 0) return m;  // field is never supplied -> always UNKNOWN
    if (presence_only ? t == m : (t == f && f == m)) return t;
    uint64_t vb; memcpy(&vb, &value, 8);
    NodeKey key{presence_only ? uint64_t(0) : vb, slot, presence_only ? 2 : (int)strict, t, f, m};
    if (cons) { auto it = node_memo.find(key); if (it != node_memo.end()) return it->second; }
    thr.push_back({slot, value, strict}); nodes.push_back({presence_only ? -1 : (int)thr.size() - 1, t, f, m}); node_slot.push_back(slot);
    if (cons) node_memo[key] = (int32_t)nodes.size() - 1;
    return (int32_t)nodes.size() - 1;
  }
  // x == v on slot:  (x <= v) and not (x < v)
  int32_t emit_eq(int slot, double v, int32_t sT, int32_t sF, int32_t sU) {
    int32_t n2 = add_test(slot, v, true, sF, sT, sU);
    return add_test(slot, v, false, n2, sF, sU);
  }
  // Decision sub-graph for one predicate with three-valued logic: routes to sT / sF / sU (TRUE / FALSE / UNKNOWN).
  // CompoundPredicate becomes a small automaton over its operands (state = operand index + what was seen so far).
  int32_t emit_pred(const XNode& p, int32_t sT, int32_t sF, int32_t sU) {
    if (p.name == "True") return sT;
    if (p.name == "False") return sF;
    if (p.name == "SimplePredicate") {
      sv op = p.attr("operator"); sv fname = p.attr("field"); int slot = b.slot_of(fname);
      uint8_t dt = b.dict_field(fname).dtype;
      // thr = -1: "passes whenever present"; the missing successor is the failing branch
      if (op == "isMissing") { if (slot < 0) return sT; return add_test(slot, 0, false, sF, sT, sT, true); }
      if (op == "isNotMissing") { if (slot < 0) return sF; return add_test(slot, 0, false, sT, sF, sF, true); }
      // A string-typed field can never hold a value here: vectors carry doubles, and a Double offered to a string field
      // fails input preparation for the whole event before any model code runs (status 2 / EmptyScore, as the reference
      // does for its k-means string fixture, T/api/PmmlModelSpec.scala:131-136).  What reaches the walk is only the
      // absent slot, so the test is always UNKNOWN -- the model loads and keeps the job alive, like in the reference.
      if (dt == DT_STRING || dt == DT_OTHER) return sU;
      double v; if (!parse_in_dtype(dt, p.attr("value"), &v)) unsupported("SimplePredicate value '" + std::string(p.attr("value")) + "' does not parse in the field's dataType");
      if (op == "lessOrEqual") return add_test(slot, v, false, sT, sF, sU);
      if (op == "lessThan") return add_test(slot, v, true, sT, sF, sU);
      if (op == "greaterThan") return add_test(slot, v, false, sF, sT, sU);
      if (op == "greaterOrEqual") return add_test(slot, v, true, sF, sT, sU);
      if (op == "equal") return emit_eq(slot, v, sT, sF, sU);
      if (op == "notEqual") return emit_eq(slot, v, sF, sT, sU);
      malformed("bad operator " + std::string(op));
    }
    if (p.name == "SimpleSetPredicate") {
      sv fname = p.attr("field"); int slot = b.slot_of(fname); uint8_t dt = b.dict_field(fname).dtype;
      sv bop = p.attr("booleanOperator"); bool in = bop == "isIn"; if (!in && bop != "isNotIn") malformed("bad booleanOperator " + std::string(bop));
      if (slot < 0 || dt == DT_STRING || dt == DT_OTHER) return sU;   // string field: see SimplePredicate above
      const XNode* arr = p.child("Array"); std::vector<double> vals;
      if (arr) for (auto& t : split_array(*arr)) { double v; if (!parse_in_dtype(dt, t, &v)) unsupported("set member does not parse in the field's dataType"); vals.push_back(v); }
      int32_t hit = in ? sT : sF, miss = in ? sF : sT, next = miss;
      if (vals.empty()) return add_test(slot, 0, false, miss, miss, sU, true);
      for (size_t k = vals.size(); k-- > 0;) next = emit_eq(slot, vals[k], hit, next, sU);   // membership: a chain of equalities
      return next;
    }
    if (p.name == "CompoundPredicate") {
      std::vector<const XNode*> ops;
      for (const XNode* k = p.first; k; k = k->next) if (k->name != "Extension") ops.push_back(k);
      if (ops.size() < 2) malformed("CompoundPredicate needs two operands");
      sv bop = p.attr("booleanOperator"); const int n = (int)ops.size();
      cons++; struct Leave { int& c; ~Leave() { c--; } } leave{cons};
      std::map<std::pair<int, int>, int32_t> memo;   // (operand index, state) -> successor
      std::function<int32_t(int, int)> go = [&](int i, int state) -> int32_t {
        auto key = std::make_pair(i, state); auto it = memo.find(key); if (it != memo.end()) return it->second;
        int32_t r;
        if (bop == "surrogate") r = i == n ? sU : emit_pred(*ops[i], sT, sF, go(i + 1, 0));
        else if (bop == "and") r = i == n ? (state ? sU : sT) : emit_pred(*ops[i], go(i + 1, state), sF, go(i + 1, 1));      // state: an UNKNOWN was seen
        else if (bop == "or") r = i == n ? (state ? sU : sF) : emit_pred(*ops[i], sT, go(i + 1, state), go(i + 1, 1));
        else if (bop == "xor") r = i == n ? (state & 2 ? sU : (state & 1 ? sT : sF))                                          // bit0 parity, bit1 unknown
                                        : emit_pred(*ops[i], go(i + 1, state ^ 1), go(i + 1, state), go(i + 1, state | 2));
        else malformed("bad booleanOperator " + std::string(bop));
        memo[key] = r; return r;
      };
      return go(0, 0);
    }
    unsupported(std::string(p.name) + " in a TreeModel");
  }
  int32_t enter(const XNode& n, int32_t last, TreeCtx& cx) {
    int32_t r; if (memo_get(n, cx, M_ENTERED, &r)) return r;
    const XNode* kid = first_node(n);
    if (!kid) r = leaf_for(n, cx);
    else { int32_t l2 = n.has("score") ? leaf_for(n, cx) : last; r = chain(n, kid, l2, cx); }
    memo_put(n, cx, M_ENTERED, r); return r;
  }
  // successor for "the children of parent from `kid` on are still to be tried" (kid == nullptr: none left)
  int32_t chain(const XNode& parent, const XNode* kid, int32_t last, TreeCtx& cx) {
    if (!kid) return cx.last_on_no_true ? last : LEAF_NULL;
    int32_t r; if (memo_get(*kid, cx, M_CHAIN, &r)) return r;
    const XNode& p = *pred_of(*kid); const XNode* rest = next_node(kid);
    if (p.name == "True") r = enter(*kid, last, cx);
    else if (p.name == "False") r = chain(parent, rest, last, cx);
    else {
      int32_t s_true = enter(*kid, last, cx), s_false = chain(parent, rest, last, cx), s_miss;
      switch (cx.mvs) {
        case 0: s_miss = s_false; break;
        case 1: { sv dc = parent.attr("defaultChild"); const XNode* d = nullptr;
                  if (!dc.empty()) for (const XNode* k = first_node(parent); k; k = next_node(k)) if (k->attr("id") == dc) d = k;
                  if (!d) unsupported("missingValueStrategy=defaultChild without a resolvable defaultChild");
                  s_miss = enter(*d, last, cx); break; }
        case 2: s_miss = last; break;
        default: s_miss = LEAF_NULL; break;
      }
      r = emit_pred(p, s_true, s_false, s_miss);
    }
    memo_put(*kid, cx, M_CHAIN, r); return r;
  }

  void add_tree(const XNode& tm, double weight, bool nested, const XNode* seg_pred = nullptr) {
    if (nested) check_nested_schema(tm, true);
    std::optional<Builder::Scope> scope; scope.emplace(b, &tm);   // the tree's own LocalTransformations are visible inside the tree only
    TreeCtx cx; cx.epoch = next_epoch(); sv mvs = tm.attr("missingValueStrategy", "none");
    cx.mvs = mvs == "none" ? 0 : mvs == "defaultChild" ? 1 : mvs == "lastPrediction" ? 2 : mvs == "nullPrediction" ? 3 : -1;
    if (cx.mvs < 0) unsupported("missingValueStrategy " + std::string(mvs));
    cx.last_on_no_true = tm.attr("noTrueChildStrategy", "returnNullPrediction") == "returnLastPrediction";
    if (classify != (tm.attr("functionName") == "classification")) unsupported("TreeModel functionName differs from the ensemble's");
    if (nested) if (const XNode* ts = tm.child("Targets")) if (const XNode* t = ts->child("Target")) {
      if (classify) unsupported("Targets on a classification tree");
      TargetXform x{1, 0, 0, 0, 0, 0, 0};
      if (t->has("rescaleFactor")) x.factor = need_double(t->attr("rescaleFactor"), "rescaleFactor");
      if (t->has("rescaleConstant")) x.constant = need_double(t->attr("rescaleConstant"), "rescaleConstant");
      if (t->has("min")) { x.has_min = 1; x.mn = need_double(t->attr("min"), "min"); }
      if (t->has("max")) { x.has_max = 1; x.mx = need_double(t->attr("max"), "max"); }
      sv c = t->attr("castInteger"); x.cast = c == "round" ? 1 : c == "ceiling" ? 2 : c == "floor" ? 3 : 0;
      cx.sub_tx = x; cx.has_sub_tx = true;
    }
    const XNode* root = tm.child("Node"); if (!root) malformed("TreeModel without Node");
    number_nodes(*root, cx);
    const XNode& rp = *pred_of(*root);
    int32_t r;
    if (rp.name == "False") r = LEAF_NULL;
    else { int32_t body = enter(*root, LEAF_NULL, cx); cons++; r = emit_pred(rp, body, LEAF_NULL, LEAF_NULL); cons--; }   // a root test that is not TRUE: null prediction
    scope.reset();   // the Segment's predicate belongs to the enclosing model's scope
    if (seg_pred && seg_pred->name != "True") { int32_t skip = skip_successor(); cons++; r = emit_pred(*seg_pred, r, skip, skip); cons--; }
    roots.push_back(r); weights.push_back(weight);
  }

  // ---- emit
  static uint32_t ordkey(float f) {
    if (f != f) return 0xFFC00000u;
    uint32_t u; memcpy(&u, &f, 4);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
  }
  // total order of java.lang.Double.compareTo
  static bool total_less(double a, double b) {
    if (a < b) return true; if (a > b) return false;
    if (a == b) return a == 0.0 && std::signbit(a) && !std::signbit(b);
    return !std::isnan(a) && std::isnan(b);
  }

  // Compact form for the throughput path (model.h CompactDesc): rank keys per feature + bucket tables for the key pre-pass,
  // 4-byte nodes.  `wide` = false: every threshold is a float and one float image per value decides every test (ENC_FLOAT's
  // argument); `wide` = true: thresholds are compared as doubles, the tables hold 64-bit images (ENC_RANK's order).
  static uint64_t ordkey64(double d) { uint64_t u; memcpy(&u, &d, 8); return (u & 0x8000000000000000ull) ? ~u : (u | 0x8000000000000000ull); }
  template <class FeatOf> void emit_compact(int depth, const std::vector<uint8_t>& used, FeatOf feat_of, bool wide, bool sparse) {
    ForestDesc& fd = b.cm.dev.forest; CompactDesc& cd = fd.compact; const int F = b.cm.dev.F;
    if (F > 256 || (!sparse && depth < 1)) return;
    // per feature: sorted distinct threshold images and the float each one falls on (bucket arithmetic is float)
    auto image = [&](const Thr& t) -> uint64_t { return wide ? ordkey64(t.value) : (uint64_t)ordkey((float)t.value); };
    std::vector<std::vector<std::pair<uint64_t, float>>> img(F);
    for (size_t i = 0; i < thr.size(); i++) if (used[i] && thr[i].slot >= 0) {
      if (std::isnan(thr[i].value)) return;
      img[thr[i].slot].push_back({image(thr[i]), (float)thr[i].value});
    }
    for (auto& v : img) { std::sort(v.begin(), v.end()); v.erase(std::unique(v.begin(), v.end()), v.end()); if (v.size() > 16382) return; }
    // rank key of every threshold
    std::vector<uint32_t> K16(thr.size(), 0);
    for (size_t i = 0; i < thr.size(); i++) {
      if (!used[i] || thr[i].slot < 0) continue;
      const auto& v = img[thr[i].slot];
      uint32_t j = (uint32_t)(std::lower_bound(v.begin(), v.end(), std::make_pair(image(thr[i]), -INFINITY)) - v.begin());
      K16[i] = thr[i].strict ? 2u * j + 1u : 2u * j + 2u;
    }
    // tables, one area per block of CK_FB features
    const int nblocks = (F + CK_FB - 1) / CK_FB; const size_t esz = wide ? 8 : 4;
    std::vector<CkFeat> feats(F); std::vector<uint8_t> tables; std::vector<uint32_t> block_off(nblocks + 1, 0); size_t block_smem = 0;
    for (int fb = 0; fb < nblocks; fb++) {
      const size_t start = tables.size(); block_off[fb] = (uint32_t)start;
      for (int f = fb * CK_FB; f < std::min(F, (fb + 1) * CK_FB); f++) {
        const auto& v = img[f]; CkFeat& cf = feats[f]; const uint32_t n = (uint32_t)v.size();
        uint32_t nb = 16; while (nb < 2 * n && nb < 4096) nb *= 2;
        float lo = n ? v.front().second : 0.0f, hi = n ? v.back().second : 0.0f;
        if (!std::isfinite(lo)) lo = -FLT_MAX; if (!std::isfinite(hi)) hi = FLT_MAX;
        float scale = (n && hi > lo) ? (float)nb / (hi - lo) : 0.0f;
        if (!std::isfinite(scale)) scale = 0.0f;
        cf.n_thr = n; cf.nb = nb; cf.lo = lo; cf.scale = scale;
        tables.resize((tables.size() + 7) & ~size_t(7));
        cf.thr_off = (uint32_t)(tables.size() - start);
        tables.resize(tables.size() + esz * (size_t)(n + 4), 0xFF);   // four all-ones sentinels end the array: the scan needs no bounds
        for (uint32_t j = 0; j < n; j++) {
          if (wide) memcpy(tables.data() + start + cf.thr_off + 8 * (size_t)j, &v[j].first, 8);
          else { uint32_t k = (uint32_t)v[j].first; memcpy(tables.data() + start + cf.thr_off + 4 * (size_t)j, &k, 4); }
        }
        cf.tab_off = (uint32_t)(tables.size() - start);
        std::vector<uint16_t> tab(nb + 1, 0);   // tab[b] = thresholds whose bucket is below b
        for (uint32_t j = 0; j < n; j++) tab[ck_bucket(v[j].second, lo, scale, nb) + 1]++;
        for (uint32_t q = 1; q <= nb; q++) tab[q] = (uint16_t)(tab[q] + tab[q - 1]);
        tables.resize(tables.size() + 2 * (size_t)nb); memcpy(tables.data() + start + cf.tab_off, tab.data(), 2 * (size_t)nb);
        cf.ent_off = 0; cf.nb_e = 0; cf.scale_e = 0.0f; cf.pad = 0;
      }
      // direct tables of the block's features (float thresholds only): ~4 buckets per threshold, at most 1024; dropped for the block
      // when they would push its table area past the 104 KB two resident CTAs can have
      if (!wide) {
        std::vector<uint8_t> ent; std::vector<std::tuple<int, uint32_t, uint32_t, float>> placed;   // feature, offset in ent, nb_e, scale_e
        for (int f = fb * CK_FB; f < std::min(F, (fb + 1) * CK_FB); f++) {
          const auto& v = img[f]; const uint32_t n = (uint32_t)v.size(); const CkFeat& cf = feats[f];
          uint32_t nbe = 16; while (nbe < 4 * n && nbe < 1024) nbe *= 2;
          float hi = n ? v.back().second : 0.0f; if (!std::isfinite(hi)) hi = FLT_MAX;
          float se = (n && hi > cf.lo) ? (float)nbe / (hi - cf.lo) : 0.0f; if (!std::isfinite(se)) se = 0.0f;
          std::vector<uint32_t> cnt(nbe, 0), firstj(nbe, 0);
          for (uint32_t j = n; j-- > 0;) { uint32_t bq = ck_bucket(v[j].second, cf.lo, se, nbe); cnt[bq]++; firstj[bq] = j; }
          const uint32_t off = (uint32_t)ent.size(); ent.resize(ent.size() + 8 * (size_t)nbe);
          uint32_t below = 0;
          for (uint32_t bq = 0; bq < nbe; bq++) {
            const uint32_t tf = cnt[bq] ? (uint32_t)v[firstj[bq]].first : 0xFFFFFFFFu, meta = below | (std::min<uint32_t>(cnt[bq], 0xFFFF) << 16);
            memcpy(ent.data() + off + 8 * (size_t)bq, &tf, 4); memcpy(ent.data() + off + 8 * (size_t)bq + 4, &meta, 4);
            below += cnt[bq];
          }
          placed.emplace_back(f, off, nbe, se);
        }
        tables.resize((tables.size() + 7) & ~size_t(7));
        if (tables.size() - start + ent.size() + 16 <= 104 * 1024) {
          const uint32_t base = (uint32_t)(tables.size() - start);
          tables.insert(tables.end(), ent.begin(), ent.end());
          for (auto& pl : placed) { CkFeat& cf = feats[std::get<0>(pl)]; cf.ent_off = base + std::get<1>(pl); cf.nb_e = std::get<2>(pl); cf.scale_e = std::get<3>(pl); }
        }
      }
      tables.resize((tables.size() + 15) & ~size_t(15));
      block_smem = std::max(block_smem, tables.size() - start);
    }
    block_off[nblocks] = (uint32_t)tables.size();
    if (block_smem > 104 * 1024) return;   // two CTAs of the key pre-pass per SM
    // node word shared by both tree forms
    bool ok = true;
    auto node_word = [&](const TNode& n) -> uint32_t {
      // right iff (a ? key : key & 0x7FFF) > K; a missing value has key 0x8000.  No threshold: never right (a = 0, ~K := 0)
      // or right only when missing (a = 1, K := 0x7FFF)
      uint32_t a = (n.on_missing == n.on_false && n.on_true != n.on_false) ? 1u : 0u, feat = (uint32_t)feat_of(n), y;
      if (n.thr >= 0) { uint32_t kk = K16[n.thr]; if (kk > 0x7FFEu) ok = false; y = ~kk; } else y = a ? ~0x7FFFu : 0u;
      if (feat > 255) ok = false;
      return (feat << 1) | (a << 15) | ((y & 0xFFFFu) << 16);
    };
    // Ensembles that average the leaves' class distributions carry them inline (at most 15 entries over at most 16 labels)
    // (a weighted majority vote is the same sum with one-entry distributions: the leaf's label with probability 1)
    const bool wvote = classify && fd.method == FM_WVOTE;
    const bool dist_leaves = classify && (fd.method == FM_AVG_PROB || fd.method == FM_WAVG_PROB || wvote);
    size_t LB = 8;
    if (dist_leaves) {
      size_t most = 1;
      if (wvote) { for (auto& lf : leaves) if (lf.label > 15) return; }
      else for (auto& lf : leaves) { most = std::max(most, lf.dist.size()); for (auto& d : lf.dist) if (d.first > 15) return; }
      if (most > 15) return;
      LB = 8 + 8 * most;
    }
    auto put_leaf = [&](uint8_t* dst, const TLeaf& lf) {
      // header: label nibbles in document order (up to 15), the entry count in the top nibble
      if (wvote) { uint64_t hdr = (uint64_t)(uint32_t)lf.label | (1ull << 60); double one = 1.0; memcpy(dst, &hdr, 8); memcpy(dst + 8, &one, 8); }
      else if (dist_leaves) {
        uint64_t hdr = (uint64_t)lf.dist.size() << 60;
        for (size_t i = 0; i < lf.dist.size(); i++) { hdr |= (uint64_t)(uint32_t)lf.dist[i].first << (4 * i); memcpy(dst + 8 + 8 * i, &lf.dist[i].second, 8); }
        memcpy(dst, &hdr, 8);
      } else if (classify) { uint64_t l = (uint64_t)(uint32_t)lf.label; memcpy(dst, &l, 8); } else memcpy(dst, &lf.value, 8);
    };
    if (sparse) {
      // explicit children: every node sends a missing value down one of its two branches, no null prediction reachable
      std::vector<uint8_t> out; std::vector<uint32_t> group_off{0}; size_t group_max = 0;
      std::vector<uint8_t> grp; int in_grp = 0; std::vector<uint32_t> dir;
      auto close_group = [&]() {
        if (!in_grp) return;
        std::vector<uint8_t> g(CK_SG_HDR + 8, 0); uint32_t nt = (uint32_t)in_grp; memcpy(g.data(), &nt, 4);
        for (int k = 0; k < in_grp; k++) { uint32_t o = dir[k] + (uint32_t)g.size(); memcpy(g.data() + 8 + 4 * k, &o, 4); }
        g.insert(g.end(), grp.begin(), grp.end()); g.resize((g.size() + 15) & ~size_t(15));
        group_max = std::max(group_max, g.size()); out.insert(out.end(), g.begin(), g.end()); group_off.push_back((uint32_t)out.size());
        grp.clear(); dir.clear(); in_grp = 0;
      };
      std::vector<int32_t> local(nodes.size(), -1), leaf_local(leaves.size(), -1);
      for (size_t t = 0; t < roots.size() && ok; t++) {
        std::vector<int32_t> order, lorder; std::vector<int32_t> stack{roots[t]};
        while (!stack.empty() && ok) {   // reachable nodes / leaves of this tree, numbered in visiting order
          int32_t s = stack.back(); stack.pop_back();
          if (s == LEAF_NULL || s == LEAF_SKIP) { ok = false; break; }
          if (s < 0) { if (leaf_local[~s] < 0) { leaf_local[~s] = (int32_t)lorder.size(); lorder.push_back(~s); } continue; }
          if (local[s] >= 0) continue;
          const TNode& n = nodes[s];
          if (n.on_missing != n.on_true && n.on_missing != n.on_false) { ok = false; break; }
          local[s] = (int32_t)order.size(); order.push_back(s); stack.push_back(n.on_false); stack.push_back(n.on_true);
        }
        if (order.size() > 32767 || lorder.size() > 32767) ok = false;
        if (!ok) break;
        auto code = [&](int32_t s) -> uint32_t { return s < 0 ? 0x8000u | (uint32_t)leaf_local[~s] : (uint32_t)local[s]; };
        const size_t n_slots = std::max<size_t>(order.size(), 1);   // a bare leaf still gets one (all-zero, never visited for real) node:
        std::vector<uint8_t> tb(16 + 8 * n_slots + LB * lorder.size(), 0);   // the walk re-reads node 0 for chains that are already done
        uint32_t nn = (uint32_t)n_slots, rc = code(roots[t]); memcpy(tb.data(), &nn, 4); memcpy(tb.data() + 4, &rc, 4); memcpy(tb.data() + 8, &weights[t], 8);
        if (!chain_channels.empty()) { uint64_t ch = (uint64_t)(uint32_t)chain_channels[t]; memcpy(tb.data() + 8, &ch, 8); }   // modelChain: channel instead of weight
        for (size_t i = 0; i < order.size(); i++) {
          const TNode& n = nodes[order[i]];
          uint64_t w = (uint64_t)node_word(n) | ((uint64_t)(code(n.on_true) | (code(n.on_false) << 16)) << 32);
          memcpy(tb.data() + 16 + 8 * i, &w, 8);
        }
        for (size_t i = 0; i < lorder.size(); i++) {
          put_leaf(tb.data() + 16 + 8 * n_slots + LB * i, leaves[lorder[i]]);
        }
        for (int32_t s : order) local[s] = -1;
        for (int32_t l : lorder) leaf_local[l] = -1;
        if (tb.size() > 100 * 1024) { ok = false; break; }                      // one tree must fit a ring slot (the launch plan checks the rest)
        if (in_grp == CK_SG_MAX || grp.size() + tb.size() > 40 * 1024) close_group();   // a ring slot holds up to 40 KB of trees
        dir.push_back((uint32_t)grp.size()); grp.insert(grp.end(), tb.begin(), tb.end()); in_grp++;
      }
      close_group();
      if (!ok) return;
      std::vector<uint32_t> words((out.size() + 3) / 4, 0); memcpy(words.data(), out.data(), out.size());
      cd.enabled = 1; cd.wide = wide ? 1 : 0; cd.sparse = 1; cd.n_groups = (int)group_off.size() - 1; cd.group_max_bytes = (int)group_max;
      cd.tree_bytes = 0; cd.leaf_bytes = (int)LB; cd.n_blocks = nblocks; cd.block_smem = (int)block_smem;
      b.put(&cd.blobs, words); b.put(&cd.group_off, group_off); b.put(&cd.feat, feats); b.put(&cd.tables, tables); b.put(&cd.block_off, block_off);
      return;
    }
    // trees
    const size_t first_leaf = size_t(1) << depth, TB = (4 * first_leaf + LB * first_leaf + 8 + 15) & ~size_t(15);
    std::vector<uint32_t> blobs(roots.size() * TB / 4, 0);
    for (size_t t = 0; t < roots.size() && ok; t++) {
      uint8_t* bl = reinterpret_cast<uint8_t*>(blobs.data()) + t * TB;
      uint32_t* nd = reinterpret_cast<uint32_t*>(bl);
      memcpy(bl + (4 + LB) * first_leaf, &weights[t], 8);
      if (!chain_channels.empty()) nd[0] = (uint32_t)chain_channels[t];   // modelChain: the unused heap slot 0 carries the channel
      std::function<void(int32_t, size_t)> place = [&](int32_t s, size_t h) {
        if (s < 0) {
          const TLeaf& lf = leaves[~s];
          if (h < first_leaf) { nd[h] = 0; place(s, 2 * h); place(s, 2 * h + 1); return; }   // pass-through: never "right"
          uint8_t* dst = bl + 4 * first_leaf + LB * (h - first_leaf);
          put_leaf(dst, lf);
          return;
        }
        const TNode& n = nodes[s];
        nd[h] = node_word(n);
        place(n.on_true, 2 * h); place(n.on_false, 2 * h + 1);
      };
      place(roots[t], 1);
    }
    if (!ok) return;
    cd.enabled = 1; cd.wide = wide ? 1 : 0; cd.tree_bytes = (int)TB; cd.leaf_bytes = (int)LB; cd.n_blocks = nblocks; cd.block_smem = (int)block_smem;
    std::vector<uint32_t> no_groups{0};
    b.put(&cd.blobs, blobs); b.put(&cd.group_off, no_groups); b.put(&cd.feat, feats); b.put(&cd.tables, tables); b.put(&cd.block_off, block_off);
  }

  void emit(int method) {
    ForestDesc& fd = b.cm.dev.forest; int F = b.cm.dev.F;
    fd.n_trees = (int)roots.size(); fd.method = method; fd.classify = classify ? 1 : 0;
    b.cm.n_nodes = nodes.size();
    // Encoding.  ENC_FLOAT: the key of a feature is the order-preserving integer image of ONE float per value.
    // For a float-typed field that float is the value itself.  For a double/integer field it is the value rounded
    // up (tests are all `x <= t`) or down (tests are all `x < t`) to float, which decides every test exactly iff
    // each threshold is itself a finite float:  x <= t  <=>  RU(x) <= t,   x < t  <=>  RD(x) < t.
    // Anything else (non-float thresholds, both kinds of test on one double feature) falls back to ENC_RANK.
    bool float_ok = true; std::vector<uint8_t> used(thr.size(), 0);
    for (auto& n : nodes) if (n.thr >= 0) used[n.thr] = 1;
    std::vector<uint8_t> has_strict(F, 0), has_loose(F, 0), feat_mode(F + 1, 0);
    for (size_t i = 0; i < thr.size(); i++) {
      if (!used[i]) continue; const Thr& t = thr[i];
      if (!std::isfinite(t.value) || std::fabs(t.value) > 3.0e38) { float_ok = false; break; }
      if (b.slot_dtype[t.slot] != DT_FLOAT && (double)(float)t.value != t.value) { float_ok = false; break; }
      (t.strict ? has_strict : has_loose)[t.slot] = 1;
    }
    for (int f = 0; f < F && float_ok; f++) {
      if (b.slot_dtype[f] == DT_FLOAT) feat_mode[f] = 0;
      else if (has_strict[f] && has_loose[f]) float_ok = false;
      else feat_mode[f] = has_strict[f] ? 2 : 1;
    }
    fd.enc = float_ok ? ENC_FLOAT : ENC_RANK;
    b.put(&fd.feat_mode, feat_mode);
    {  // the key pre-pass converts without a per-lane choice when every feature wants the same rounding
      int um = F > 0 ? feat_mode[0] : -1;
      for (int f = 1; f < F; f++) if (feat_mode[f] != um) um = -1;
      fd.compact.uniform_mode = float_ok ? um : -1;
    }
    // every test of the ensemble sees a feature only through that one float: float rows rounded the same way score identically
    // (fields whose preparation could tell the double from its float image -- intervals, value lists, integers -- keep double rows)
    if (float_ok && b.cm.dev.prep_simple && b.cm.dev.derived.n_derived == 0) {
      bool plain = true; for (int f = 0; f < F; f++) plain = plain && (b.slot_dtype[f] == DT_FLOAT || b.slot_dtype[f] == DT_DOUBLE);
      if (plain) b.cm.float_row_modes.assign(feat_mode.begin(), feat_mode.begin() + F);
    }
    std::vector<uint32_t> K(thr.size());
    std::vector<int32_t> rank_off(F + 1, 0); std::vector<double> rank_thr;
    if (float_ok) {
      for (size_t i = 0; i < thr.size(); i++) { uint32_t k = ordkey((float)thr[i].value); K[i] = thr[i].strict ? k - 1 : k; }
    } else {
      std::vector<std::vector<double>> per(F);
      for (size_t i = 0; i < thr.size(); i++) if (used[i]) per[thr[i].slot].push_back(thr[i].value);
      for (int f = 0; f < F; f++) {
        auto& v = per[f]; std::sort(v.begin(), v.end(), total_less);
        v.erase(std::unique(v.begin(), v.end(), [](double a, double c) { return !total_less(a, c) && !total_less(c, a); }), v.end());
        rank_off[f] = (int32_t)rank_thr.size(); rank_thr.insert(rank_thr.end(), v.begin(), v.end());
      }
      rank_off[F] = (int32_t)rank_thr.size();
      for (size_t i = 0; i < thr.size(); i++) {
        if (!used[i]) continue;
        const double* base = rank_thr.data() + rank_off[thr[i].slot]; int n = rank_off[thr[i].slot + 1] - rank_off[thr[i].slot];
        int j = (int)(std::lower_bound(base, base + n, thr[i].value, total_less) - base);
        K[i] = thr[i].strict ? 2u * j + 1u : 2u * j + 2u;
      }
    }
    b.put(&fd.rank_off, rank_off); b.put(&fd.rank_thr, rank_thr);
    auto feat_of = [&](const TNode& n) { return node_slot[&n - nodes.data()]; };

    // Complete-tree form: every tree is laid out as a complete binary tree of ONE depth D (the deepest path of the
    // ensemble, at most 10).  Shallower branches are padded: a leaf above depth D becomes a subtree of pass-through nodes
    // whose leaves all repeat it (pruned XGBoost / LightGBM trees, trees of different depths).  Needs: every test sends a
    // missing value down one of its two branches, and no null prediction is reachable.  Sub-graphs shared by several
    // parents (compound predicates, equality tests) are simply duplicated while the expansion stays small.
    const bool general_only = method == FM_SELECT_FIRST || method == FM_MEDIAN || method == FM_MAX;
    int depth = -1; bool complete = !roots.empty() && !general_only;
    std::vector<int> dep_memo(nodes.size(), -1);
    std::function<int(int32_t)> dep = [&](int32_t s) -> int {  // longest path below s; -2: cannot be laid out
      if (s < 0) return (s == LEAF_NULL || s == LEAF_SKIP) ? -2 : 0;
      if (dep_memo[s] != -1) return dep_memo[s];
      const TNode& n = nodes[s]; int r;
      if (n.on_missing != n.on_true && n.on_missing != n.on_false) r = -2;
      else { int a = dep(n.on_true), c = dep(n.on_false); r = (a < 0 || c < 0) ? -2 : std::max(a, c) + 1; }
      return dep_memo[s] = r;
    };
    for (int32_t r : roots) { int dd = dep(r); if (dd < 0 || dd > 10) { complete = false; break; } depth = std::max(depth, dd); }
    if (complete && depth < 1) depth = 1;   // a forest of bare leaves still gets one (pass-through) level
    if (complete) {  // padding must not blow the model up: at most 8x the real node count (always fine up to depth 6)
      size_t padded = roots.size() << depth;
      if (depth > 6 && padded > 8 * (nodes.size() + roots.size())) complete = false;
    }
    std::vector<int32_t> leaf_ids;
    if (complete) {
      fd.format = FOREST_COMPLETE; fd.depth = depth;
      size_t W = size_t(1) << (depth + 1); std::vector<uint64_t> blobs(roots.size() * W, 0);
      leaf_ids.assign(roots.size() << depth, 0);
      for (size_t t = 0; t < roots.size(); t++) {
        uint64_t* bl = blobs.data() + t * W; memcpy(&bl[0], &weights[t], 8);
        if (!chain_channels.empty()) bl[0] = (uint64_t)(uint32_t)chain_channels[t];   // modelChain: word 0 carries the channel instead
        std::function<void(int32_t, size_t)> place = [&](int32_t s, size_t h) {
          if (s < 0) {
            const TLeaf& lf = leaves[~s];
            if (h < (size_t(1) << depth)) {   // a leaf above the common depth: pass-through node (every key goes left), repeated below
              bl[h] = (uint64_t)0u | ((uint64_t)0u << 32);   // feat 0, a = 0, ~ka = 0 => ka = 0xFFFFFFFF: never "right"
              place(s, 2 * h); place(s, 2 * h + 1); return;
            }
            if (classify) bl[h] = (uint64_t)(uint32_t)lf.label; else memcpy(&bl[h], &lf.value, 8);
            leaf_ids[(t << depth) + (h - (size_t(1) << depth))] = lf.node_id; return;
          }
          const TNode& n = nodes[s]; uint32_t a = (n.on_missing == n.on_false && n.on_true != n.on_false) ? 1u : 0u;
          uint32_t kk = n.thr >= 0 ? K[n.thr] : 0xFFFFFFFFu; uint32_t lo = (uint32_t)feat_of(n) | (a << 31), hi = ~(kk - a);
          bl[h] = (uint64_t)lo | ((uint64_t)hi << 32);
          place(n.on_true, 2 * h); place(n.on_false, 2 * h + 1);
        };
        place(roots[t], 1);
      }
      b.put(&fd.blobs, blobs);
      if (method != FM_SINGLE) emit_compact(depth, used, feat_of, !float_ok, false);
    } else {
      fd.format = FOREST_GENERAL; fd.depth = 0;
      std::vector<uint64_t> blobs; b.put(&fd.blobs, blobs);
      if (method != FM_SINGLE && !general_only && !roots.empty()) emit_compact(0, used, feat_of, !float_ok, true);
    }
    b.put(&fd.leaf_node_id, leaf_ids);
    {  // the general form is always emitted: the engine falls back to it when a tile of keys does not fit on chip
      std::vector<GNode> g(nodes.size());
      for (size_t i = 0; i < nodes.size(); i++) {
        const TNode& n = nodes[i];
        g[i] = GNode{(uint32_t)feat_of(n), n.thr >= 0 ? K[n.thr] : 0xFFFFFFFFu, n.on_true, n.on_false, n.on_missing};
      }
      std::vector<double> gv(leaves.size()); std::vector<int32_t> gl(leaves.size()), gid(leaves.size());
      std::vector<int32_t> doff{0}, dlab; std::vector<double> dprob;
      for (size_t i = 0; i < leaves.size(); i++) {
        gv[i] = leaves[i].value; gl[i] = leaves[i].label; gid[i] = leaves[i].node_id;
        for (auto& d : leaves[i].dist) { dlab.push_back(d.first); dprob.push_back(d.second); }
        doff.push_back((int32_t)dlab.size());
      }
      b.put(&fd.gdist_off, doff); b.put(&fd.gdist_label, dlab); b.put(&fd.gdist_prob, dprob);
      b.put(&fd.gnodes, g); b.put(&fd.troot, roots); b.put(&fd.gleaf_value, gv); b.put(&fd.gleaf_label, gl); b.put(&fd.tweight, weights);
      b.put(&fd.gleaf_node_id, gid);
    }
  }
};

void finish_target(Builder& b, bool classify) {
  // regression results are delivered in the target field's dataType; only java.lang.Double is matched
  // (Pipeline.scala:94-98); classification labels go through set_labels()
  if (!b.has_target) b.cm.dev.target_status = ST_EXTRACT;
  else if (!classify && b.target_dtype != DT_DOUBLE) b.cm.dev.target_status = ST_EXTRACT;
  else b.cm.dev.target_status = ST_OK;
}

void compile_tree(const XNode& m, Builder& b) {
  b.cm.family = "TreeModel"; b.cm.dev.kind = MK_FOREST;
  bool classify = m.attr("functionName") == "classification";
  ForestBuilder fb(b, classify); fb.add_tree(m, 1.0, false); fb.emit(FM_SINGLE);
  b.set_labels(fb.labels, false); finish_target(b, classify);
}

bool compile_regression_tables(const XNode& m, Builder& b, const std::unordered_map<std::string, int>* channels);
TargetXform read_target_xform(const XNode& model) {
  TargetXform x{1, 0, 0, 0, 0, 0, 0};
  if (const XNode* ts = model.child("Targets")) if (const XNode* t = ts->child("Target")) {
    if (t->has("rescaleFactor")) x.factor = need_double(t->attr("rescaleFactor"), "rescaleFactor");
    if (t->has("rescaleConstant")) x.constant = need_double(t->attr("rescaleConstant"), "rescaleConstant");
    if (t->has("min")) { x.has_min = 1; x.mn = need_double(t->attr("min"), "min"); }
    if (t->has("max")) { x.has_max = 1; x.mx = need_double(t->attr("max"), "max"); }
    sv c = t->attr("castInteger"); x.cast = c == "round" ? 1 : c == "ceiling" ? 2 : c == "floor" ? 3 : 0;
  }
  return x;
}

// multipleModelMethod="modelChain" in the shape boosted-tree exporters write: K regression segments (a TreeModel or a
// MiningModel sum / average of TreeModels), each publishing its result as a predictedValue OutputField, followed by a
// RegressionModel over those outputs (logit / softmax / exp ... link, or a classification label).  Each of the K
// segments becomes a "channel" of one forest; the final tables run in the kernel epilogue.
void compile_model_chain(const XNode& m, const XNode& sg, Builder& b) {
  std::vector<const XNode*> segs;
  sg.each("Segment", [&](const XNode& s) {
    const XNode& p = *ForestBuilder::pred_of(s);
    if (p.name != "True") unsupported("modelChain Segment predicate other than True");
    segs.push_back(&s);
  });
  if (segs.size() < 2) unsupported("modelChain with fewer than two segments");
  ForestBuilder fb(b, false);
  std::unordered_map<std::string, int> channels; std::vector<ChannelSpec> specs; std::vector<int32_t> tree_channel;
  for (size_t k = 0; k + 1 < segs.size(); k++) {
    const XNode& s = *segs[k]; ChannelSpec cs{}; cs.tx = TargetXform{1, 0, 0, 0, 0, 0, 0}; cs.average = 0;
    const XNode* inner = s.child("MiningModel"); const XNode* tm = s.child("TreeModel"); const XNode* model = inner ? inner : tm;
    if (!model) unsupported("modelChain segment that is neither a TreeModel nor a MiningModel of TreeModels");
    if (model->attr("functionName") != "regression") unsupported("modelChain: non-regression inner segment");
    const XNode* out = model->child("Output"); std::string name;
    if (out) out->each("OutputField", [&](const XNode& f) { if (name.empty() && f.attr("feature", "predictedValue") == "predictedValue") name = std::string(f.attr("name")); });
    if (name.empty()) unsupported("modelChain segment without a predictedValue OutputField");
    size_t before = fb.roots.size();
    if (tm && !inner) { fb.add_tree(*tm, 1.0, true); }   // the tree's own Targets are folded into its leaves by add_tree
    else {
      check_nested_schema(*inner, true);
      Builder::Scope inner_scope(b, inner);   // the inner model's LocalTransformations, visible to its trees
      const XNode* isg = inner->child("Segmentation"); if (!isg) malformed("MiningModel without Segmentation");
      sv imm = isg->attr("multipleModelMethod");
      if (imm == "average") cs.average = 1; else if (imm != "sum") unsupported("modelChain: inner multipleModelMethod " + std::string(imm));
      cs.tx = read_target_xform(*inner);
      isg->each("Segment", [&](const XNode& is) {
        const XNode& ip = *ForestBuilder::pred_of(is);
        if (ip.name == "False") return;
        if (ip.name != "True") unsupported("Segment predicate other than True/False");
        const XNode* itm = is.child("TreeModel"); if (!itm) unsupported("Segment model other than TreeModel");
        fb.add_tree(*itm, 1.0, true);
      });
    }
    cs.n_trees = (int)(fb.roots.size() - before);
    if (cs.n_trees == 0) unsupported("modelChain segment without trees");
    for (int i = 0; i < cs.n_trees; i++) tree_channel.push_back((int32_t)specs.size());
    channels[name] = (int)specs.size(); specs.push_back(cs);
  }
  if ((int)specs.size() > MAX_CHANNELS) unsupported("modelChain with more than 16 tree segments");
  const XNode* last = segs.back()->child("RegressionModel");
  if (!last) unsupported("modelChain whose last segment is not a RegressionModel");
  check_nested_schema(*last);
  fb.chain_channels = tree_channel;
  fb.emit(FM_CHAIN);
  ForestDesc& fd = b.cm.dev.forest; fd.n_channels = (int)specs.size();
  b.put(&fd.channels, specs); b.put(&fd.tree_channel, tree_channel);
  compile_regression_tables(*last, b, &channels);   // sets labels / target status from the final model
  if (b.cm.dev.reg.n_tables > 64) unsupported("modelChain with more than 64 RegressionTables");
}

void compile_mining(const XNode& m, Builder& b) {
  b.cm.family = "MiningModel"; b.cm.dev.kind = MK_FOREST;
  const XNode* sg = m.child("Segmentation"); if (!sg) malformed("MiningModel without Segmentation");
  bool classify = m.attr("functionName") == "classification"; sv mm = sg->attr("multipleModelMethod");
  if (mm == "modelChain") { compile_model_chain(m, *sg, b); return; }
  int method = -1;
  if (!classify) method = mm == "sum" ? FM_SUM : mm == "average" ? FM_AVERAGE : mm == "weightedSum" ? FM_WSUM : mm == "weightedAverage" ? FM_WAVERAGE
                        : mm == "median" ? FM_MEDIAN : mm == "max" ? FM_MAX : -1;
  else method = mm == "majorityVote" ? FM_VOTE : mm == "weightedMajorityVote" ? FM_WVOTE : mm == "average" ? FM_AVG_PROB : mm == "weightedAverage" ? FM_WAVG_PROB : -1;
  if (mm == "selectFirst") method = FM_SELECT_FIRST;
  if (method < 0) unsupported("multipleModelMethod " + std::string(mm) + " for " + (classify ? "classification" : "regression"));
  ForestBuilder fb(b, classify);
  bool always_active = false;
  sg->each("Segment", [&](const XNode& s) { if (ForestBuilder::pred_of(s)->name == "True") always_active = true; });
  fb.skip_as_zero = (method == FM_SUM || method == FM_WSUM) && always_active;
  sg->each("Segment", [&](const XNode& s) {
    const XNode& p = *ForestBuilder::pred_of(s);
    if (p.name == "False") return;   // never takes part
    const XNode* tm = s.child("TreeModel");
    if (!tm) unsupported("Segment model other than TreeModel");
    fb.add_tree(*tm, need_double(s.attr("weight", "1"), "weight"), true, &p);
  });
  if (fb.roots.empty()) unsupported("Segmentation without active segments");
  if (method == FM_MEDIAN && (int)fb.roots.size() > MAX_MEDIAN_SEGMENTS) unsupported("median over more than 64 segments");
  fb.emit(method); b.set_labels(fb.labels, false); finish_target(b, classify);
}

// ---------------------------------------------------------------------------------------------- regression
int norm_of(sv s) {
  if (s.empty() || s == "none") return NORM_NONE; if (s == "simplemax") return NORM_SIMPLEMAX; if (s == "softmax") return NORM_SOFTMAX;
  if (s == "logit") return NORM_LOGIT; if (s == "probit") return NORM_PROBIT; if (s == "cloglog") return NORM_CLOGLOG;
  if (s == "exp") return NORM_EXP; if (s == "loglog") return NORM_LOGLOG; if (s == "cauchit") return NORM_CAUCHIT;
  unsupported("normalizationMethod " + std::string(s));
}

// `channels` != nullptr: the last segment of a modelChain -- predictors name earlier segments' output fields and
// RegTerm.slot is a channel index instead of a vector slot
bool compile_regression_tables(const XNode& m, Builder& b, const std::unordered_map<std::string, int>* channels) {
  RegressionDesc& d = b.cm.dev.reg; bool classify = m.attr("functionName") == "classification";
  d.classify = classify; d.norm = norm_of(m.attr("normalizationMethod", "none"));
  std::vector<double> icpt; std::vector<int32_t> toff{0}, coff{0}; std::vector<RegTerm> terms; std::vector<RegCat> cats; std::vector<std::string> labels;
  int F = b.cm.dev.F; bool dense = true;
  m.each("RegressionTable", [&](const XNode& t) {
    icpt.push_back(need_double(t.attr("intercept"), "intercept")); labels.push_back(std::string(t.attr("targetCategory")));
    size_t t0 = terms.size();
    t.each("NumericPredictor", [&](const XNode& n) {
      double e = need_double(n.attr("exponent", "1"), "exponent");
      if (e != std::floor(e)) unsupported("non-integer exponent");
      sv fname = n.attr("name");
      if (channels) {
        auto it = channels->find(std::string(fname));
        if (it == channels->end()) unsupported("modelChain: the final RegressionModel may only use earlier segments' predictedValue outputs");
        terms.push_back({it->second, (int32_t)e, need_double(n.attr("coefficient"), "coefficient")});
        return;
      }
      uint8_t dt = b.dict_field(fname).dtype;
      if (dt == DT_STRING || dt == DT_OTHER) unsupported("NumericPredictor on a non-numeric field");
      terms.push_back({b.slot_of(fname), (int32_t)e, need_double(n.attr("coefficient"), "coefficient")});
    });
    t.each("CategoricalPredictor", [&](const XNode& n) {
      if (channels) unsupported("modelChain: CategoricalPredictor in the final RegressionModel");
      sv fname = n.attr("name"); uint8_t dt = b.dict_field(fname).dtype; double v;
      if (dt == DT_STRING || dt == DT_OTHER) unsupported("CategoricalPredictor on a string field");
      if (!parse_in_dtype(dt, n.attr("value"), &v)) unsupported("CategoricalPredictor value does not parse in the field's dataType");
      cats.push_back({b.slot_of(fname), 0, v, need_double(n.attr("coefficient"), "coefficient")});
    });
    if (t.child("PredictorTerm")) unsupported("PredictorTerm");
    if ((int)(terms.size() - t0) != F) dense = false;
    else for (int i = 0; i < F; i++) if (terms[t0 + i].slot != i || terms[t0 + i].exponent != 1) { dense = false; break; }
    toff.push_back((int32_t)terms.size()); coff.push_back((int32_t)cats.size());
  });
  if (icpt.empty()) malformed("no RegressionTable");
  if (!classify && icpt.size() != 1) unsupported("regression with several RegressionTables");
  d.n_tables = (int)icpt.size(); dense = dense && cats.empty() && !channels; d.dense = dense ? 1 : 0;
  std::vector<double> w;
  // same choice as the launcher: 8, 10 or 12 tables per thread, least padding
  size_t tcs = 8, bestpad = ~size_t(0); for (size_t cand : {size_t(8), size_t(10), size_t(12)}) { size_t pad = (icpt.size() + cand - 1) / cand * cand - icpt.size(); if (pad < bestpad) { bestpad = pad; tcs = cand; } }
  d.dense_amax = 0; d.dense_bmax = 0;
  if (dense) {
    size_t KP = (icpt.size() + tcs - 1) / tcs * tcs; w.assign(KP * (size_t)F, 0.0);
    for (size_t t = 0; t < icpt.size(); t++) {
      double rowsum = 0;
      for (int i = 0; i < F; i++) { w[(size_t)i * KP + t] = terms[toff[t] + i].coef; rowsum += std::fabs(terms[toff[t] + i].coef); }
      d.dense_amax = std::max(d.dense_amax, rowsum * (1.0 + 1e-12)); d.dense_bmax = std::max(d.dense_bmax, std::fabs(icpt[t]));
    }
    if (!std::isfinite(d.dense_amax) || !std::isfinite(d.dense_bmax)) d.dense_amax = d.dense_bmax = std::numeric_limits<double>::infinity();   // the FMA path then decides nothing
  }
  b.put(&d.intercept, icpt); b.put(&d.term_off, toff); b.put(&d.terms, terms); b.put(&d.cat_off, coff); b.put(&d.cats, cats); b.put(&d.dense_wt, w);
  {  // weight digit planes of the fused tensor-core layer (model.h): softmax classification, <= 24 tables, wide rows
    std::vector<int8_t> SW; std::vector<int32_t> TAU; d.sg = SgDesc{};
    if (dense && classify && d.norm == NORM_SOFTMAX && icpt.size() <= 24 && F >= 128 && F <= 8192) {
      std::vector<double> Wm(icpt.size() * (size_t)F);
      for (size_t t = 0; t < icpt.size(); t++) for (int i = 0; i < F; i++) Wm[t * (size_t)F + i] = terms[toff[t] + i].coef;
      SgLayer& sl = d.sg.layers[0];
      if (sg_add_layer(Wm.data(), icpt.data(), (int)icpt.size(), F, sl, SW, TAU)) { sl.act = ACT_IDENTITY; sl.last = 1; sl.norm = NORM_SOFTMAX; sl.b_off = 0; d.sg.enabled = 1; d.sg.n_layers = 1; }
      else { SW.clear(); TAU.clear(); }
    }
    b.put(&d.sg.weights, SW); b.put(&d.sg.tau_exp, TAU);
  }
  if (classify) b.set_labels(labels, false); else b.set_labels({}, false);
  finish_target(b, classify);
  return classify;
}
void compile_regression(const XNode& m, Builder& b) {
  b.cm.family = "RegressionModel"; b.cm.dev.kind = MK_REGRESSION;
  compile_regression_tables(m, b, nullptr);
}

// ---------------------------------------------------------------------------------------------- general regression
void compile_general_regression(const XNode& m, Builder& b) {
  b.cm.family = "GeneralRegressionModel"; b.cm.dev.kind = MK_GENERAL_REGRESSION;
  GeneralRegDesc& d = b.cm.dev.greg; bool classify = m.attr("functionName") == "classification"; d.classify = classify;
  sv mt = m.attr("modelType");
  if (mt == "regression") d.model_type = GR_REGRESSION;
  else if (mt == "generalLinear" || mt == "generalizedLinear") d.model_type = GR_GENERAL_LINEAR;
  else if (mt == "multinomialLogistic") d.model_type = GR_MULTINOMIAL;
  else unsupported("GeneralRegressionModel modelType " + std::string(mt));
  if (d.model_type == GR_MULTINOMIAL && !classify) unsupported("multinomialLogistic regression");
  static const char* links[] = {"identity", "log", "logit", "probit", "cloglog", "loglog", "logc", "cauchit", "power", "negbin"};
  sv lk = m.attr("linkFunction", "identity"); d.link = -1;
  for (int i = 0; i < 10; i++) if (lk == links[i]) d.link = i;
  if (d.link < 0) unsupported("linkFunction " + std::string(lk));
  d.link_param = m.has("linkParameter") ? need_double(m.attr("linkParameter"), "linkParameter") : 0.0;
  d.offset = m.has("offsetValue") ? need_double(m.attr("offsetValue"), "offsetValue") : 0.0;
  if (m.has("offsetVariable") || m.has("trialsVariable") || m.has("trialsValue")) unsupported("offsetVariable / trials");
  std::unordered_map<std::string, int> pidx; std::unordered_map<std::string, bool> is_factor;
  if (const XNode* pl = m.child("ParameterList")) pl->each("Parameter", [&](const XNode& p) { int i = (int)pidx.size(); pidx[std::string(p.attr("name"))] = i; });
  if (pidx.empty()) malformed("empty ParameterList");
  if ((int)pidx.size() > GR_MAX_PARAMS) unsupported("more than 64 parameters in a GeneralRegressionModel");
  if (const XNode* fl = m.child("FactorList")) fl->each("Predictor", [&](const XNode& p) { is_factor[std::string(p.attr("name"))] = true; });
  if (const XNode* cl = m.child("CovariateList")) cl->each("Predictor", [&](const XNode& p) { is_factor[std::string(p.attr("name"))] = false; });
  std::vector<std::vector<GCell>> per(pidx.size());
  if (const XNode* pp = m.child("PPMatrix")) pp->each("PPCell", [&](const XNode& c) {
    auto it = pidx.find(std::string(c.attr("parameterName"))); if (it == pidx.end()) malformed("PPCell for an unknown parameter");
    std::string pn(c.attr("predictorName")); bool fac = is_factor.count(pn) ? is_factor[pn] : false;
    uint8_t dt = b.dict_field(pn).dtype; if (dt == DT_STRING || dt == DT_OTHER) unsupported("GeneralRegression predictor on a non-numeric field");
    GCell cell{b.slot_of(pn), fac ? 1 : 0, 1.0};
    if (fac) { if (!parse_in_dtype(dt, c.attr("value"), &cell.value)) unsupported("factor value does not parse in the field's dataType"); }
    else cell.value = need_double(c.attr("value", "1"), "PPCell value");
    per[it->second].push_back(cell);
  });
  std::vector<int32_t> cell_off{0}; std::vector<GCell> cells;
  for (auto& v : per) { cells.insert(cells.end(), v.begin(), v.end()); cell_off.push_back((int32_t)cells.size()); }
  // target categories: the target DataField's valid values, in document order
  std::vector<std::string> cats;
  if (classify) {
    if (!b.has_target) unsupported("classification GeneralRegressionModel without a target field");
    b.dict_field(b.target_name).node->each("Value", [&](const XNode& v) { if (v.attr("property", "valid") == "valid") cats.push_back(std::string(v.attr("value"))); });
    if (cats.empty()) unsupported("target field declares no categories");
    if ((int)cats.size() > GR_MAX_CATS) unsupported("more than 32 target categories");
  }
  std::vector<std::string> row_names; std::vector<std::vector<std::pair<int, double>>> rows;
  if (const XNode* pm = m.child("ParamMatrix")) pm->each("PCell", [&](const XNode& c) {
    std::string cat(c.attr("targetCategory")); size_t r = 0;
    for (; r < row_names.size(); r++) if (row_names[r] == cat) break;
    if (r == row_names.size()) { row_names.push_back(cat); rows.emplace_back(); }
    auto it = pidx.find(std::string(c.attr("parameterName"))); if (it == pidx.end()) malformed("PCell for an unknown parameter");
    rows[r].emplace_back(it->second, need_double(c.attr("beta"), "beta"));
  });
  if (rows.empty()) malformed("empty ParamMatrix");
  std::vector<int32_t> beta_off{0}, beta_param, row_cat; std::vector<double> beta;
  for (size_t r = 0; r < rows.size(); r++) {
    for (auto& pb : rows[r]) { beta_param.push_back(pb.first); beta.push_back(pb.second); }
    beta_off.push_back((int32_t)beta.size());
    int ci = -1; for (size_t k = 0; k < cats.size(); k++) if (cats[k] == row_names[r]) ci = (int)k;
    row_cat.push_back(ci);
  }
  d.n_params = (int)pidx.size(); d.n_rows = (int)rows.size(); d.n_cats = (int)cats.size();
  b.put(&d.cell_off, cell_off); b.put(&d.cells, cells); b.put(&d.beta_off, beta_off); b.put(&d.beta_param, beta_param); b.put(&d.beta, beta); b.put(&d.row_cat, row_cat);
  b.set_labels(cats, false); finish_target(b, classify);
}

// ---------------------------------------------------------------------------------------------- tensor-core form of dense layers
// model.h "tensor-core form": fixed-point images of a weight matrix, cut into SG_S balanced base-128 digits per weight.
void sg_digits(long long q, int8_t* d /* d[0] = most significant */) {
  long long bias = 0; for (int k = 0; k < SG_S - 1; k++) bias |= 0x80LL << (8 * k);
  const long long t = q + bias;                          // bytes 0 .. S-2 of t are d_k + 128, the rest is the top digit
  for (int k = 0; k < SG_S - 1; k++) d[SG_S - 1 - k] = (int8_t)(int)(((t >> (8 * k)) & 255) - 128);
  d[0] = (int8_t)(t >> (8 * (SG_S - 1)));
}
// W: N x K row-major.  Appends the layer's slices / tau exponents and fills its SgLayer; false when a weight is not finite.
bool sg_add_layer(const double* W, const double* bias, int N, int K, SgLayer& ly, std::vector<int8_t>& SW, std::vector<int32_t>& TAU) {
  ly.K = K; ly.N = N; ly.Kp = (K + SG_KC - 1) / SG_KC * SG_KC; ly.Np = (N + SG_NC - 1) / SG_NC * SG_NC;
  ly.w_off = (int64_t)SW.size(); ly.tau_off = (int64_t)TAU.size();
  const int n_nc = ly.Np / SG_NC, n_kc = ly.Kp / SG_KC;
  SW.resize(SW.size() + (size_t)n_nc * n_kc * SG_S * (SG_NC * SG_KC), 0); TAU.resize(TAU.size() + ly.Np, 0);
  int8_t* base = SW.data() + ly.w_off; double amax = 0, bmax = 0, wmax = 0;
  for (int j = 0; j < N; j++) {
    double rowsum = 0;
    for (int i = 0; i < K; i++) { double w = W[(size_t)j * K + i]; if (!std::isfinite(w)) return false; wmax = std::max(wmax, std::fabs(w)); rowsum += std::fabs(w); }
    if (!std::isfinite(bias[j])) return false;
    amax = std::max(amax, rowsum * (1.0 + 1e-12)); bmax = std::max(bmax, std::fabs(bias[j]));
  }
  int e = 0; if (wmax > 0) std::frexp(wmax, &e);          // wmax = f * 2^e, f in [0.5, 1): tau = 2^e > wmax
  if (e < -400 || e > 400) return false;
  for (int j = 0; j < ly.Np; j++) TAU[ly.tau_off + j] = e;
  for (int j = 0; j < N; j++) {
    const int nc = j / SG_NC, jn = j % SG_NC;
    for (int i = 0; i < K; i++) {
      long long q = std::llrint(std::ldexp(W[(size_t)j * K + i], SG_QBITS - e));   // exact scaling, one rounding: |w / tau - q 2^-Q| <= 2^-(Q+1)
      int8_t d[SG_S]; sg_digits(q, d);
      const int kc = i / SG_KC, kb = i % SG_KC;
      int8_t* stage = base + ((size_t)nc * n_kc + kc) * SG_S * (SG_NC * SG_KC);
      for (int sl = 0; sl < SG_S; sl++) stage[(size_t)sl * (SG_NC * SG_KC) + (kb / 16) * (SG_NC * 16) + jn * 16 + (kb % 16)] = d[sl];
    }
  }
  ly.amax = amax; ly.bmax = bmax; ly.tau_max_exp = e;
  return true;
}

// ---------------------------------------------------------------------------------------------- neural network
int act_of(sv s) {
  static const char* names[] = {"threshold", "logistic", "tanh", "identity", "exponential", "reciprocal", "square", "Gauss", "sine", "cosine", "Elliott", "arctan", "rectifier"};
  for (int i = 0; i < 13; i++) if (s == names[i]) return i;
  if (s == "radialBasis") return ACT_RADIAL;
  unsupported("activationFunction " + std::string(s));
}

void compile_neural(const XNode& m, Builder& b) {
  b.cm.family = "NeuralNetwork"; b.cm.dev.kind = MK_NEURAL;
  NeuralDesc& d = b.cm.dev.nn; bool classify = m.attr("functionName") == "classification"; d.classify = classify;
  sv dflt_act = m.attr("activationFunction", "logistic"), dflt_norm = m.attr("normalizationMethod", "none");
  double dflt_thr = m.has("threshold") ? need_double(m.attr("threshold"), "threshold") : 0.0;
  std::vector<NeuralInput> inputs; std::vector<double> pts; std::vector<std::string> prev_ids;
  auto read_expr = [&](const XNode& df, NeuralInput& in, bool is_output) {
    for (const XNode* k = df.first; k; k = k->next) {
      if (k->name == "FieldRef") { in.kind = NI_FIELDREF; in.slot = is_output ? -1 : b.slot_of(k->attr("field"));
        if (k->has("mapMissingTo")) { in.has_map_missing = 1; in.map_missing = need_double(k->attr("mapMissingTo"), "mapMissingTo"); } return; }
      if (k->name == "Constant") { in.kind = NI_CONSTANT; in.value = need_double(k->text, "Constant"); return; }
      if (k->name == "NormDiscrete") {
        in.kind = NI_NORMDISC;
        if (!is_output) { sv fname = k->attr("field"); in.slot = b.slot_of(fname); uint8_t dt = b.dict_field(fname).dtype;
          if (!parse_in_dtype(dt, k->attr("value"), &in.value)) unsupported("NormDiscrete input on a non-numeric field"); }
        return; }
      if (k->name == "NormContinuous") {
        in.kind = NI_NORMCONT; in.slot = is_output ? -1 : b.slot_of(k->attr("field")); in.pts_off = (int32_t)pts.size() / 2;
        sv o = k->attr("outliers", "asIs"); in.outliers = o == "asMissingValues" ? 1 : o == "asExtremeValues" ? 2 : 0;
        if (k->has("mapMissingTo")) { in.has_map_missing = 1; in.map_missing = need_double(k->attr("mapMissingTo"), "mapMissingTo"); }
        k->each("LinearNorm", [&](const XNode& ln) { pts.push_back(need_double(ln.attr("orig"), "orig")); pts.push_back(need_double(ln.attr("norm"), "norm")); in.n_pts++; });
        if (in.n_pts < 2) malformed("NormContinuous needs two LinearNorm"); return; }
    }
    unsupported("DerivedField expression in NeuralNetwork");
  };
  const XNode* ni = m.child("NeuralInputs"); if (!ni) malformed("no NeuralInputs");
  ni->each("NeuralInput", [&](const XNode& in) {
    const XNode* df = in.child("DerivedField"); if (!df) malformed("NeuralInput without DerivedField");
    NeuralInput x{}; read_expr(*df, x, false); inputs.push_back(x); prev_ids.push_back(std::string(in.attr("id")));
  });
  d.n_inputs = (int)inputs.size(); d.max_width = d.n_inputs;
  std::vector<double> W, B; int nl = 0;
  m.each("NeuralLayer", [&](const XNode& L) {
    if (nl >= MAX_LAYERS) unsupported("more than 8 NeuralLayers");
    NeuralLayer& ly = d.layers[nl]; ly.n_in = (int)prev_ids.size(); ly.act = act_of(L.attr("activationFunction", dflt_act));
    sv nm = L.attr("normalizationMethod", dflt_norm); ly.norm = nm == "softmax" ? NORM_SOFTMAX : nm == "simplemax" ? NORM_SIMPLEMAX : NORM_NONE;
    if (nm != "none" && ly.norm == NORM_NONE) unsupported("layer normalizationMethod " + std::string(nm));
    ly.threshold = L.has("threshold") ? need_double(L.attr("threshold"), "threshold") : dflt_thr;
    ly.w_off = (int64_t)W.size(); ly.b_off = (int64_t)B.size();
    std::vector<std::string> ids; std::vector<double> rbf_den, rbf_alt;
    L.each("Neuron", [&](const XNode& n) {
      ids.push_back(std::string(n.attr("id"))); B.push_back(need_double(n.attr("bias", "0"), "bias"));
      if (ly.act == ACT_RADIAL) {   // width / altitude: the Neuron's, else the layer's, else the network's
        const XNode* src[3] = {&n, &L, &m}; double width = 0, alt = 1; bool has_w = false;
        for (const XNode* x : src) if (!has_w && x->has("width")) { width = need_double(x->attr("width"), "width"); has_w = true; }
        for (const XNode* x : src) if (x->has("altitude")) { alt = need_double(x->attr("altitude"), "altitude"); break; }
        if (!has_w) unsupported("radialBasis neuron without width");
        rbf_den.push_back(2.0 * width * width); rbf_alt.push_back((double)prev_ids.size() * std::log(alt));
      }
      size_t i = 0;
      n.each("Con", [&](const XNode& c) {
        // the accumulation order is the Con order; only "every neuron of the previous layer, in order" is flattened
        if (i >= prev_ids.size() || c.attr("from") != prev_ids[i]) unsupported("NeuralNetwork that is not layer-wise fully connected in neuron order");
        W.push_back(need_double(c.attr("weight"), "weight")); i++;
      });
      if (i != prev_ids.size()) unsupported("NeuralNetwork that is not layer-wise fully connected in neuron order");
    });
    ly.n_out = (int)ids.size(); if (!ly.n_out) malformed("empty NeuralLayer");
    B.insert(B.end(), rbf_den.begin(), rbf_den.end()); B.insert(B.end(), rbf_alt.begin(), rbf_alt.end());
    d.max_width = std::max(d.max_width, ly.n_out); prev_ids = ids; nl++;
  });
  if (!nl) malformed("no NeuralLayer");
  d.n_layers = nl;
  const XNode* no = m.child("NeuralOutputs"); if (!no) malformed("no NeuralOutputs");
  std::vector<int32_t> out_neuron; std::vector<std::string> labels; d.out_denorm_n = 0; d.out_denorm_off = 0;
  no->each("NeuralOutput", [&](const XNode& o) {
    const XNode* df = o.child("DerivedField"); if (!df) malformed("NeuralOutput without DerivedField");
    int idx = -1; for (size_t i = 0; i < prev_ids.size(); i++) if (prev_ids[i] == o.attr("outputNeuron")) idx = (int)i;
    if (idx < 0) unsupported("NeuralOutput not attached to the last layer");
    NeuralInput x{}; read_expr(*df, x, true);
    if (classify) { if (x.kind != NI_NORMDISC) unsupported("classification NeuralOutput without NormDiscrete"); labels.push_back(std::string(df->child("NormDiscrete")->attr("value"))); }
    else if (out_neuron.empty() && x.kind == NI_NORMCONT) { d.out_denorm_off = x.pts_off; d.out_denorm_n = x.n_pts; }
    else if (!classify && x.kind != NI_FIELDREF && x.kind != NI_NORMCONT) unsupported("regression NeuralOutput expression");
    out_neuron.push_back(idx);
  });
  if (out_neuron.empty()) malformed("no NeuralOutput");
  d.n_outputs = (int)out_neuron.size();
  std::vector<double> WT;
  for (int l = 0; l < nl; l++) {
    NeuralLayer& ly = d.layers[l]; ly.n_out_pad = (ly.n_out + 7) / 8 * 8; ly.wt_off = (int64_t)WT.size();
    WT.resize(WT.size() + (size_t)ly.n_in * ly.n_out_pad, 0.0);
    for (int j = 0; j < ly.n_out; j++) for (int i = 0; i < ly.n_in; i++) WT[(size_t)ly.wt_off + (size_t)i * ly.n_out_pad + j] = W[(size_t)ly.w_off + (size_t)j * ly.n_in + i];
  }
  // tensor-core form: classification, layers of at most 256 inputs, activations with Lipschitz constant <= 1 (an input error is
  // not amplified), identity + softmax / no normalisation on the last layer, at most 16 output neurons
  std::vector<int8_t> SW; std::vector<int32_t> TAU; SgDesc& sg = d.sg; sg = SgDesc{};
  {
    bool ok = classify && d.n_inputs <= 256 && nl <= MAX_LAYERS;
    for (int l = 0; l < nl && ok; l++) {
      const NeuralLayer& ly = d.layers[l]; const bool last = l == nl - 1;
      ok = ly.n_in <= 256 && ly.n_out <= 256;   // also what neural_pair_kernel (the re-scoring kernel) takes
      if (!last) ok = ok && ly.norm == NORM_NONE && (ly.act == ACT_RECTIFIER || ly.act == ACT_IDENTITY || ly.act == ACT_TANH || ly.act == ACT_LOGISTIC);
      else ok = ok && ly.act == ACT_IDENTITY && (ly.norm == NORM_SOFTMAX || ly.norm == NORM_NONE) && ly.n_out <= 16;
    }
    for (int l = 0; l < nl && ok; l++) {
      const NeuralLayer& ly = d.layers[l]; SgLayer& sl = sg.layers[l];
      ok = sg_add_layer(W.data() + ly.w_off, B.data() + ly.b_off, ly.n_out, ly.n_in, sl, SW, TAU);
      sl.act = ly.act; sl.last = l == nl - 1; sl.norm = ly.norm; sl.b_off = ly.b_off;
    }
    sg.enabled = ok ? 1 : 0; sg.n_layers = ok ? nl : 0;
    sg.inputs_identity = 1;
    for (size_t i = 0; i < inputs.size(); i++) if (inputs[i].kind != NI_FIELDREF || inputs[i].slot != (int)i || inputs[i].has_map_missing) sg.inputs_identity = 0;
    if (!ok) { SW.clear(); TAU.clear(); }
  }
  b.put(&d.inputs, inputs); b.put(&d.pts, pts); b.put(&d.weights, W); b.put(&d.weights_t, WT); b.put(&d.biases, B); b.put(&d.out_neuron, out_neuron);
  b.put(&d.sg.weights, SW); b.put(&d.sg.tau_exp, TAU);
  b.set_labels(labels, false); finish_target(b, classify);
}

}  // namespace

CompiledModel compile_pmml(const char* data, size_t len) {
  XmlDoc doc;
  try { xml_parse(data, len, doc); } catch (XmlError& e) { throw CompileError(e.what()); }
  const XNode& root = *doc.root;
  if (root.name != "PMML") malformed("root element is not PMML");
  Builder b; memset(&b.cm.dev, 0, sizeof(b.cm.dev));
  read_dictionary(root, b);
  const XNode* me = nullptr;
  for (const XNode* k = root.first; k && !me; k = k->next) {
    sv n = k->name;
    if (n == "ClusteringModel" || n == "TreeModel" || n == "MiningModel" || n == "RegressionModel" || n == "NeuralNetwork" || n == "GeneralRegressionModel") me = k;
    else if (n.size() > 5 && n.substr(n.size() - 5) == "Model") unsupported("model element " + std::string(n));
    else if (n == "Scorecard" || n == "RuleSetModel" || n == "NaiveBayesModel") unsupported("model element " + std::string(n));
  }
  if (!me) malformed("no model element");
  b.cm.model_name = std::string(me->attr("modelName"));
  read_schema(*me, b);
  read_derived(root, *me, b);
  if (me->name == "ClusteringModel") compile_clustering(*me, b);
  else if (me->name == "TreeModel") compile_tree(*me, b);
  else if (me->name == "MiningModel") compile_mining(*me, b);
  else if (me->name == "RegressionModel") compile_regression(*me, b);
  else if (me->name == "GeneralRegressionModel") compile_general_regression(*me, b);
  else compile_neural(*me, b);
  return std::move(b.cm);
}

}  // namespace fjp
