// oracle.cc -- CPU ORACLE.  TEST INFRASTRUCTURE ONLY: nothing on the product path may link, import or call this.
//
// A literal, per-event, double-precision interpreter of PMML that restates, in plain C++17:
//   (1) the reference's own per-event pipeline around JPMML
//         flink-jpmml-scala/src/main/scala/io/radicalbit/flink/pmml/scala/api/PmmlModel.scala:109-174
//         .../api/converter/VectorConverter.scala:58-103, .../api/pipeline/Pipeline.scala:49-98,
//         .../models/prediction/Prediction.scala:37-62
//   (2) the arithmetic of the third-party dependency the reference delegates to and does NOT vendor:
//         org.jpmml:pmml-evaluator:1.3.9  (project/Dependencies.scala:66-70)
//       restated from the DMG PMML 4.3 specification read the way JPMML-Evaluator 1.3.x implements it
//       (items marked [R] there are recollections isolated in single functions here).
//
// PARITY PINNING: ClusteringModel is pinned by the reference's own known-answer tests
//   (api/PmmlModelSpec.scala:50-83, RichConnectedStreamSpec.scala) -- see tests/test_oracle_golden.py.
//   TreeModel / MiningModel / RegressionModel / GeneralRegressionModel / NeuralNetwork: PARITY UNPINNED --
//   the reference holds no fixture or test for them and JPMML cannot run here (no JVM); they are checked
//   against hand-derived answers, numpy restatements and scikit-learn instead.
//
// XML is read with expat (the product uses its own parser: two independent front-ends).
// Build: oracle/Makefile ->  oracle/liboracle.so
#include <expat.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <limits>
#include <map>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

namespace orc {

// ------------------------------------------------------------------------------------------ XML DOM
struct Xml {
  std::string name;
  std::vector<std::pair<std::string, std::string>> attrs;
  std::vector<std::unique_ptr<Xml>> kids;
  std::string text;
  const std::string* attr(const char* k) const {
    for (auto& a : attrs) if (a.first == k) return &a.second;
    return nullptr;
  }
  std::string get(const char* k, const char* dflt = "") const { auto* p = attr(k); return p ? *p : std::string(dflt); }
  const Xml* child(const char* n) const {
    for (auto& c : kids) if (c->name == n) return c.get();
    return nullptr;
  }
  std::vector<const Xml*> children(const char* n) const {
    std::vector<const Xml*> r;
    for (auto& c : kids) if (c->name == n) r.push_back(c.get());
    return r;
  }
};

struct ParseCtx { std::unique_ptr<Xml> root; std::vector<Xml*> stack; };

static std::string local_name(const char* n) { const char* c = strrchr(n, ':'); return c ? c + 1 : n; }

static void XMLCALL on_start(void* ud, const XML_Char* name, const XML_Char** atts) {
  auto* cx = static_cast<ParseCtx*>(ud);
  auto node = std::make_unique<Xml>();
  node->name = local_name(name);
  for (int i = 0; atts[i]; i += 2) node->attrs.emplace_back(local_name(atts[i]), atts[i + 1]);
  Xml* raw = node.get();
  if (cx->stack.empty()) cx->root = std::move(node); else cx->stack.back()->kids.push_back(std::move(node));
  cx->stack.push_back(raw);
}
static void XMLCALL on_end(void* ud, const XML_Char*) { static_cast<ParseCtx*>(ud)->stack.pop_back(); }
static void XMLCALL on_text(void* ud, const XML_Char* s, int len) {
  auto* cx = static_cast<ParseCtx*>(ud);
  if (!cx->stack.empty()) cx->stack.back()->text.append(s, len);
}

static std::unique_ptr<Xml> parse_xml(const char* data, size_t len) {
  XML_Parser p = XML_ParserCreate(nullptr);
  ParseCtx cx;
  XML_SetUserData(p, &cx);
  XML_SetElementHandler(p, on_start, on_end);
  XML_SetCharacterDataHandler(p, on_text);
  if (XML_Parse(p, data, (int)len, 1) == XML_STATUS_ERROR) {
    std::string msg = std::string("XML error: ") + XML_ErrorString(XML_GetErrorCode(p)) + " at line " +
                      std::to_string(XML_GetCurrentLineNumber(p));
    XML_ParserFree(p);
    throw std::runtime_error(msg);
  }
  XML_ParserFree(p);
  if (!cx.root) throw std::runtime_error("XML error: no root element");
  return std::move(cx.root);
}

// ------------------------------------------------------------------------------------------ values
enum DType : uint8_t { DT_STRING, DT_INTEGER, DT_FLOAT, DT_DOUBLE, DT_BOOLEAN, DT_OTHER };
enum OpType : uint8_t { OP_CONTINUOUS, OP_CATEGORICAL, OP_ORDINAL };

static DType parse_dtype(const std::string& s) {
  if (s == "string") return DT_STRING;
  if (s == "integer") return DT_INTEGER;
  if (s == "float") return DT_FLOAT;
  if (s == "double") return DT_DOUBLE;
  if (s == "boolean") return DT_BOOLEAN;
  return DT_OTHER;
}
static OpType parse_optype(const std::string& s) {
  if (s == "categorical") return OP_CATEGORICAL;
  if (s == "ordinal") return OP_ORDINAL;
  return OP_CONTINUOUS;
}

struct EvalError : std::runtime_error { using std::runtime_error::runtime_error; };      // -> status 3
struct PrepError : std::runtime_error { using std::runtime_error::runtime_error; };      // -> status 2
struct Unsupported : std::runtime_error { using std::runtime_error::runtime_error; };    // load-time

static std::string trim(const std::string& s) {
  size_t a = 0, b = s.size();
  while (a < b && (unsigned char)s[a] <= ' ') a++;
  while (b > a && (unsigned char)s[b - 1] <= ' ') b--;
  return s.substr(a, b - a);
}

// java.lang.Double.parseDouble: trims whitespace, accepts optional trailing d/D/f/F, "NaN", "Infinity", hex floats.
static bool java_parse_double(const std::string& raw, double* out) {
  std::string s = trim(raw);
  if (s.empty()) return false;
  std::string body = s;
  size_t sign = (body[0] == '+' || body[0] == '-') ? 1 : 0;
  std::string rest = body.substr(sign);
  if (rest == "NaN") { *out = std::numeric_limits<double>::quiet_NaN(); return true; }
  if (rest == "Infinity") { *out = (sign && body[0] == '-') ? -INFINITY : INFINITY; return true; }
  char last = body.back();
  bool hex = rest.size() > 1 && rest[0] == '0' && (rest[1] == 'x' || rest[1] == 'X');
  if (!hex && (last == 'd' || last == 'D' || last == 'f' || last == 'F')) body.pop_back();  // Java float-type suffix
  if (body.size() == sign) return false;
  // reject forms strtod accepts but Java does not ("inf", "nan", "infinity")
  for (char c : body.substr(sign, 3)) if (c == 'i' || c == 'I' || c == 'n' || c == 'N') return false;
  char* end = nullptr;
  double v = strtod(body.c_str(), &end);
  if (end == body.c_str() || *end != '\0') return false;
  *out = v;
  return true;
}

struct FV {  // JPMML FieldValue: null | number typed by dataType | string
  enum Kind : uint8_t { MISSING, NUM, STR } kind = MISSING;
  DType dtype = DT_DOUBLE;
  double num = 0;
  std::string str;
  static FV missing() { return FV{}; }
  static FV number(DType dt, double v) { FV f; f.kind = NUM; f.dtype = dt; f.num = v; return f; }
  static FV string(const std::string& s) { FV f; f.kind = STR; f.dtype = DT_STRING; f.str = s; return f; }
  bool is_missing() const { return kind == MISSING; }
};

// java.lang.Double.compare total order (NaN greatest, -0.0 < 0.0).  SURVEY App. A [R]: JPMML compares
// through Comparable.compareTo, not IEEE operators.  Isolated here.
static int java_compare(double a, double b) {
  if (a < b) return -1;
  if (a > b) return 1;
  if (a == b) {
    if (a == 0.0) { bool na = std::signbit(a), nb = std::signbit(b); return na == nb ? 0 : (na ? -1 : 1); }
    return 0;
  }
  bool an = std::isnan(a), bn = std::isnan(b);
  return an == bn ? 0 : (an ? 1 : -1);
}

// TypeUtil.parse(dataType, string)
static FV parse_as(DType dt, const std::string& s) {
  switch (dt) {
    case DT_STRING: return FV::string(s);
    case DT_DOUBLE: { double v; if (!java_parse_double(s, &v)) throw EvalError("not a double: " + s); return FV::number(dt, v); }
    case DT_FLOAT: { double v; if (!java_parse_double(s, &v)) throw EvalError("not a float: " + s);
                     // Float.parseFloat rounds the decimal text once to float; strtof does the same
                     std::string t = trim(s); char* e = nullptr; float f = strtof(t.c_str(), &e);
                     if (e && *e == '\0') return FV::number(dt, (double)f);
                     return FV::number(dt, (double)(float)v); }
    case DT_INTEGER: { double v; if (!java_parse_double(s, &v) || v != std::floor(v) || std::fabs(v) > 2147483647.0)
                         throw EvalError("not an integer: " + s);
                       return FV::number(dt, v); }
    case DT_BOOLEAN: { std::string t = trim(s);
                       if (t == "true" || t == "TRUE" || t == "True") return FV::number(dt, 1.0);
                       if (t == "false" || t == "FALSE" || t == "False") return FV::number(dt, 0.0);
                       double v; if (java_parse_double(t, &v) && (v == 0.0 || v == 1.0)) return FV::number(dt, v);
                       throw EvalError("not a boolean: " + s); }
    default: throw EvalError("unsupported dataType");
  }
}

// TypeUtil.cast(dataType, java.lang.Double) -- the only raw input type a Flink Vector can deliver.
static FV cast_double_to(DType dt, double x) {
  switch (dt) {
    case DT_DOUBLE: return FV::number(dt, x);
    case DT_FLOAT: return FV::number(dt, (double)(float)x);
    case DT_INTEGER: if (std::isfinite(x) && x == std::floor(x) && std::fabs(x) <= 2147483647.0) return FV::number(dt, x);
                     throw PrepError("cannot cast to integer");
    case DT_BOOLEAN: if (x == 0.0 || x == 1.0) return FV::number(dt, x);
                     throw PrepError("cannot cast to boolean");
    default: throw PrepError("cannot cast a Double to this dataType");  // string: api/PmmlModelSpec.scala:131-136
  }
}

static bool fv_equals(const FV& a, const FV& b) {  // Object.equals on the boxed values
  if (a.kind != b.kind) return false;
  if (a.kind == FV::STR) return a.str == b.str;
  return java_compare(a.num, b.num) == 0;
}
static bool fv_equals_string(const FV& v, const std::string& s) { return fv_equals(v, parse_as(v.dtype, s)); }

// ------------------------------------------------------------------------------------------ dictionary
struct Interval { bool has_l = false, has_r = false, open_l = false, open_r = false; double l = 0, r = 0; };
struct DataField {
  std::string name; OpType op = OP_CONTINUOUS; DType dt = DT_DOUBLE;
  std::vector<Interval> intervals;
  std::vector<std::string> valid, invalid, missing;  // <Value property=...>
};
enum InvalidTreatment { IVT_RETURN_INVALID, IVT_AS_IS, IVT_AS_MISSING };
enum OutlierTreatment { OT_AS_IS, OT_AS_MISSING, OT_AS_EXTREME };
struct MiningField {
  std::string name; std::string usage = "active";
  InvalidTreatment ivt = IVT_RETURN_INVALID; OutlierTreatment ot = OT_AS_IS;
  std::optional<std::string> missing_repl; double low = 0, high = 0; bool has_low = false, has_high = false;
};

struct TargetSpec {  // <Targets><Target>
  std::string field; double factor = 1, constant = 0; std::optional<double> mn, mx; std::string cast;  // castInteger
};

// Evaluation context: one FV per known field (DataDictionary order, then derived fields appended).
struct Ctx { std::vector<FV> v; };

// ------------------------------------------------------------------------------------------ predicates (3-valued)
enum Tri : int8_t { T_FALSE = 0, T_TRUE = 1, T_UNKNOWN = 2 };

struct Predicate {
  enum Kind { TRUE_, FALSE_, SIMPLE, SET, COMPOUND } kind = TRUE_;
  int field = -1; std::string op, value;           // SIMPLE / SET (op = isIn|isNotIn)
  std::vector<std::string> set;                    // SET
  // operands parsed once at load in the field's dataType (JPMML re-parses the text per evaluation; the
  // outcome is identical, so the baseline is not charged for that); parse failures surface at evaluation
  FV parsed; bool parse_ok = false; std::vector<FV> parsed_set; bool set_ok = false;
  enum Op : uint8_t { EQ, NE, LT, LE, GT, GE, IS_MISSING, IS_NOT_MISSING, IS_IN, IS_NOT_IN, AND, OR, XOR, SURROGATE, BAD } opc = BAD;
  std::string boolop; std::vector<Predicate> kids; // COMPOUND
};

static Tri eval_pred(const Predicate& p, const Ctx& cx) {
  switch (p.kind) {
    case Predicate::TRUE_: return T_TRUE;
    case Predicate::FALSE_: return T_FALSE;
    case Predicate::SIMPLE: {
      const FV& v = cx.v[p.field];
      if (p.opc == Predicate::IS_MISSING) return v.is_missing() ? T_TRUE : T_FALSE;
      if (p.opc == Predicate::IS_NOT_MISSING) return v.is_missing() ? T_FALSE : T_TRUE;
      if (v.is_missing()) return T_UNKNOWN;
      if (!p.parse_ok) throw EvalError("predicate value does not parse in the field's dataType");
      if (p.opc == Predicate::EQ) return fv_equals(v, p.parsed) ? T_TRUE : T_FALSE;
      if (p.opc == Predicate::NE) return fv_equals(v, p.parsed) ? T_FALSE : T_TRUE;
      int c = v.kind == FV::STR ? (v.str.compare(p.parsed.str) < 0 ? -1 : (v.str == p.parsed.str ? 0 : 1)) : java_compare(v.num, p.parsed.num);
      switch (p.opc) {
        case Predicate::LT: return c < 0 ? T_TRUE : T_FALSE;
        case Predicate::LE: return c <= 0 ? T_TRUE : T_FALSE;
        case Predicate::GT: return c > 0 ? T_TRUE : T_FALSE;
        case Predicate::GE: return c >= 0 ? T_TRUE : T_FALSE;
        default: throw EvalError("bad operator " + p.op);
      }
    }
    case Predicate::SET: {
      const FV& v = cx.v[p.field];
      if (v.is_missing()) return T_UNKNOWN;
      if (!p.set_ok) throw EvalError("set member does not parse in the field's dataType");
      bool in = false;
      for (auto& s : p.parsed_set) if (fv_equals(v, s)) { in = true; break; }
      return (p.opc == Predicate::IS_IN) == in ? T_TRUE : T_FALSE;
    }
    case Predicate::COMPOUND: {
      if (p.opc == Predicate::SURROGATE) {
        for (auto& k : p.kids) { Tri t = eval_pred(k, cx); if (t != T_UNKNOWN) return t; }
        return T_UNKNOWN;
      }
      Tri acc = eval_pred(p.kids[0], cx);
      for (size_t i = 1; i < p.kids.size(); i++) {
        Tri t = eval_pred(p.kids[i], cx);
        if (p.opc == Predicate::AND) acc = (acc == T_FALSE || t == T_FALSE) ? T_FALSE : ((acc == T_UNKNOWN || t == T_UNKNOWN) ? T_UNKNOWN : T_TRUE);
        else if (p.opc == Predicate::OR) acc = (acc == T_TRUE || t == T_TRUE) ? T_TRUE : ((acc == T_UNKNOWN || t == T_UNKNOWN) ? T_UNKNOWN : T_FALSE);
        else if (p.opc == Predicate::XOR) acc = (acc == T_UNKNOWN || t == T_UNKNOWN) ? T_UNKNOWN : ((acc != t) ? T_TRUE : T_FALSE);
        else throw EvalError("bad booleanOperator");
      }
      return acc;
    }
  }
  return T_UNKNOWN;
}

// ------------------------------------------------------------------------------------------ derived fields
// TransformationDictionary / LocalTransformations <DerivedField> expressions over numeric fields (PMML 4.3 Transformations):
// Constant, FieldRef, NormContinuous, NormDiscrete, Discretize, MapValues (one numeric column), Apply (built-in functions).
// A value is a double or missing; booleans are 1.0 / 0.0.  Unless a function says otherwise a missing argument gives a
// missing result (then mapMissingTo applies); a missing result with all arguments present gets defaultValue.
struct DExpr {
  enum Kind { CONST, FIELD, NORMCONT, NORMDISC, DISCRETIZE, MAPVALUES, APPLY, PARAM, USERCALL } kind = CONST;
  bool to_float = false;                                             // USERCALL: the DefineFunction's dataType is float
  double constant = 0; int slot = -1; std::string function;
  std::optional<double> map_missing, dflt;
  std::vector<std::pair<double, double>> pts; int outliers = 0;     // NORMCONT
  double match = 0;                                                  // NORMDISC
  struct Bin { bool has_l, has_r, open_l, open_r; double l, r, value; };
  std::vector<Bin> bins;                                             // DISCRETIZE
  std::vector<std::pair<double, double>> pairs;                      // MAPVALUES in -> out
  std::vector<DExpr> args;                                           // APPLY; NORMCONT/NORMDISC/... read `slot`

  bool eval(const Ctx& cx, double* out) const {   // false = missing
    switch (kind) {
      case CONST: *out = constant; return true;
      case FIELD: { const FV& v = cx.v[slot]; if (v.is_missing()) return fill(map_missing, out); if (v.kind != FV::NUM) throw EvalError("non-numeric FieldRef"); *out = v.num; return true; }
      case NORMDISC: { const FV& v = cx.v[slot]; if (v.is_missing()) return fill(map_missing, out); *out = java_compare(v.num, match) == 0 ? 1.0 : 0.0; return true; }
      case NORMCONT: {
        const FV& v = cx.v[slot]; if (v.is_missing()) return fill(map_missing, out);
        double x = v.num; size_t n = pts.size();
        auto interp = [&](size_t i) { return pts[i].second + (x - pts[i].first) / (pts[i + 1].first - pts[i].first) * (pts[i + 1].second - pts[i].second); };
        if (x < pts[0].first) { if (outliers == 1) return false; *out = outliers == 2 ? pts[0].second : interp(0); return true; }
        if (x > pts[n - 1].first) { if (outliers == 1) return false; *out = outliers == 2 ? pts[n - 1].second : interp(n - 2); return true; }
        size_t i = 0; while (i + 2 < n && x > pts[i + 1].first) i++;
        *out = interp(i); return true;
      }
      case DISCRETIZE: {
        const FV& v = cx.v[slot]; if (v.is_missing()) return fill(map_missing, out);
        for (auto& b : bins) {
          bool lo = !b.has_l || (b.open_l ? v.num > b.l : v.num >= b.l), hi = !b.has_r || (b.open_r ? v.num < b.r : v.num <= b.r);
          if (lo && hi) { *out = b.value; return true; }
        }
        return fill(dflt, out);
      }
      case MAPVALUES: {
        const FV& v = cx.v[slot]; if (v.is_missing()) return fill(map_missing, out);
        for (auto& p : pairs) if (java_compare(v.num, p.first) == 0) { *out = p.second; return true; }
        return fill(dflt, out);
      }
      case APPLY: return apply(cx, out);
      // DefineFunction (JPMML FunctionUtil.evaluate(DefineFunction ...)): the body runs in a context that holds only the
      // ParameterFields, bound to the argument values (a missing argument is a missing parameter).  Loaded by substitution:
      // PARAM = a FieldRef to a parameter (args[0] is the argument's expression; the FieldRef's own mapMissingTo applies),
      // USERCALL = the Apply (args = the arguments, then the body): Apply@mapMissingTo if any argument is missing, else the
      // body's value in the function's dataType, Apply@defaultValue if that is missing.
      case PARAM: if (args[0].eval(cx, out)) return true; return fill(map_missing, out);
      case USERCALL: {
        double t;
        if (map_missing) for (size_t i = 0; i + 1 < args.size(); i++) if (!args[i].eval(cx, &t)) return fill(map_missing, out);
        if (!args.back().eval(cx, out)) return fill(dflt, out);
        if (to_float) *out = (double)(float)*out;
        return true;
      }
    }
    return false;
  }
  static bool fill(const std::optional<double>& v, double* out) { if (v) { *out = *v; return true; } return false; }

  bool apply(const Ctx& cx, double* out) const {
    const std::string& f = function; size_t n = args.size();
    if (f == "isMissing" || f == "isNotMissing") { double t; bool present = args.at(0).eval(cx, &t); *out = (present == (f == "isNotMissing")) ? 1.0 : 0.0; return true; }
    if (f == "if") {
      double c; if (!args.at(0).eval(cx, &c)) return fill(map_missing, out);
      if (c != 0.0) { if (args.at(1).eval(cx, out)) return true; return fill(dflt, out); }
      if (n > 2) { if (args[2].eval(cx, out)) return true; }
      return fill(dflt, out);
    }
    std::vector<double> a(n);
    for (size_t i = 0; i < n; i++) if (!args[i].eval(cx, &a[i])) return fill(map_missing, out);
    auto need = [&](size_t k) { if (n != k) throw EvalError("wrong number of arguments for " + f); };
    double r;
    if (f == "+") { need(2); r = a[0] + a[1]; } else if (f == "-") { need(2); r = a[0] - a[1]; }
    else if (f == "*") { need(2); r = a[0] * a[1]; } else if (f == "/") { need(2); r = a[0] / a[1]; }
    else if (f == "min") { r = a.at(0); for (double v : a) if (java_compare(v, r) < 0) r = v; }
    else if (f == "max") { r = a.at(0); for (double v : a) if (java_compare(v, r) > 0) r = v; }
    else if (f == "sum") { r = 0; for (double v : a) r += v; }
    else if (f == "avg") { r = 0; for (double v : a) r += v; r /= (double)n; }
    else if (f == "product") { r = 1; for (double v : a) r *= v; }
    else if (f == "pow") { need(2); r = std::pow(a[0], a[1]); }
    else if (f == "exp") { need(1); r = std::exp(a[0]); } else if (f == "ln") { need(1); r = std::log(a[0]); }
    else if (f == "log10") { need(1); r = std::log10(a[0]); } else if (f == "sqrt") { need(1); r = std::sqrt(a[0]); }
    else if (f == "abs") { need(1); r = std::fabs(a[0]); } else if (f == "floor") { need(1); r = std::floor(a[0]); }
    else if (f == "ceil") { need(1); r = std::ceil(a[0]); } else if (f == "round") { need(1); r = std::floor(a[0] + 0.5); }
    else if (f == "threshold") { need(2); r = a[0] > a[1] ? 1.0 : 0.0; }
    else if (f == "equal") { need(2); r = java_compare(a[0], a[1]) == 0; } else if (f == "notEqual") { need(2); r = java_compare(a[0], a[1]) != 0; }
    else if (f == "lessThan") { need(2); r = java_compare(a[0], a[1]) < 0; } else if (f == "lessOrEqual") { need(2); r = java_compare(a[0], a[1]) <= 0; }
    else if (f == "greaterThan") { need(2); r = java_compare(a[0], a[1]) > 0; } else if (f == "greaterOrEqual") { need(2); r = java_compare(a[0], a[1]) >= 0; }
    else if (f == "and") { r = 1; for (double v : a) if (v == 0.0) r = 0; } else if (f == "or") { r = 0; for (double v : a) if (v != 0.0) r = 1; }
    else if (f == "not") { need(1); r = a[0] == 0.0 ? 1.0 : 0.0; }
    else if (f == "isIn" || f == "isNotIn") { bool in = false; for (size_t i = 1; i < n; i++) if (java_compare(a[0], a[i]) == 0) in = true; r = (in == (f == "isIn")) ? 1.0 : 0.0; }
    else throw EvalError("unsupported Apply function " + f);
    *out = r; return true;
  }
};
struct DerivedField { int slot; DType dt; DExpr expr; };

// ------------------------------------------------------------------------------------------ model results
struct Result {
  bool has_value = false;        // regression value or classification label present
  bool is_class = false;
  double value = 0;              // regression
  std::string label;             // classification / clustering entity id
  int aux = -1;                  // cluster position (1-based) / tree node id / label index
  std::vector<std::pair<std::string, double>> probs;  // classification distribution in insertion order
};

struct Model;
using FieldIndex = std::unordered_map<std::string, int>;

struct Doc {  // shared dictionary
  std::vector<DataField> fields;
  FieldIndex index;
  int slot(const std::string& n) const { auto it = index.find(n); if (it == index.end()) throw Unsupported("unknown field " + n); return it->second; }
};

struct Model {
  std::string model_name, function;
  std::vector<MiningField> mining;
  std::vector<TargetSpec> targets;
  std::vector<std::pair<int, std::string>> outputs;  // <Output> fields: (context slot, feature) -- feed later modelChain segments
  virtual ~Model() = default;
  virtual Result evaluate(const Ctx& cx, std::vector<int>* node_ids) const = 0;
  const MiningField* target_field() const {
    for (auto& m : mining) if (m.usage == "target" || m.usage == "predicted") return &m;
    return nullptr;
  }
};

// Targets processing for regression results (TargetUtil.processValue): min/max clamp, rescale, cast.
static double apply_target(const Model& m, double v) {
  const MiningField* tf = m.target_field();
  for (auto& t : m.targets) {
    if (!t.field.empty() && tf && t.field != tf->name) continue;
    if (t.mn && v < *t.mn) v = *t.mn;
    if (t.mx && v > *t.mx) v = *t.mx;
    v = v * t.factor + t.constant;
    if (t.cast == "round") v = std::floor(v + 0.5);  // java.lang.Math.round: floor(x + 1/2), so -2.5 -> -2
    else if (t.cast == "ceiling") v = std::ceil(v);
    else if (t.cast == "floor") v = std::floor(v);
    break;
  }
  return v;
}

// ------------------------------------------------------------------------------------------ ClusteringModel (J1)
struct ClusteringModel : Model {
  std::string kind, measure, default_compare; double minkowski_p = 2;
  struct CField { int slot; double weight = 1; std::string compare; double sim_scale = 0; };
  std::vector<CField> cfields;
  std::vector<double> missing_weights;  // empty -> all 1
  struct Cluster { std::string id; std::vector<double> center; };
  std::vector<Cluster> clusters;

  Result evaluate(const Ctx& cx, std::vector<int>*) const override {
    if (kind != "distance") throw EvalError("similarity measures are not supported");
    size_t F = cfields.size();
    // MeasureUtil.calculateAdjustment: sum(q) / sum(q over non-missing)
    double sum_q = 0, sum_nm = 0;
    for (size_t i = 0; i < F; i++) {
      double q = missing_weights.empty() ? 1.0 : missing_weights[i];
      sum_q += q;
      sum_nm += cx.v[cfields[i].slot].is_missing() ? 0.0 : q;
    }
    double adj = sum_q / sum_nm;
    double inner, outer;
    if (measure == "euclidean") { inner = 2; outer = 2; }
    else if (measure == "squaredEuclidean") { inner = 2; outer = 1; }
    else if (measure == "chebychev" || measure == "cityBlock") { inner = 1; outer = 1; }
    else if (measure == "minkowski") { inner = outer = minkowski_p; }
    else throw EvalError("unsupported measure " + measure);
    Result r; r.is_class = true;
    int best = -1; double bestd = 0;
    for (size_t c = 0; c < clusters.size(); c++) {
      double acc = 0; bool cheb = measure == "chebychev";
      for (size_t i = 0; i < F; i++) {
        const FV& v = cx.v[cfields[i].slot];
        if (v.is_missing()) continue;  // missing fields are skipped
        if (v.kind != FV::NUM) throw EvalError("non-numeric clustering field");
        double x = v.num, y = clusters[c].center[i], z;
        const std::string& cf = cfields[i].compare;
        if (cf == "absDiff") z = std::fabs(x - y);
        else if (cf == "gaussSim") { double s = cfields[i].sim_scale; double d = x - y; z = std::exp(-std::log(2.0) * d * d / (s * s)); }
        else if (cf == "delta") z = (java_compare(x, y) == 0) ? 0 : 1;
        else if (cf == "equal") z = (java_compare(x, y) == 0) ? 1 : 0;
        else throw EvalError("unsupported compareFunction " + cf);
        double p = inner == 2 ? z * z : (inner == 1 ? z : std::pow(z, inner));
        double term = cfields[i].weight * p;
        if (cheb) acc = std::max(acc, term); else acc += term;
      }
      double d;
      if (cheb) d = acc * adj;
      else { double t = acc * adj; d = outer == 1 ? t : (outer == 2 ? std::sqrt(t) : std::pow(t, 1.0 / outer)); }
      r.probs.emplace_back(clusters[c].id, d);
      // Classification.getWinner for DISTANCE: strictly smaller wins, first entry kept on ties (java compare order)
      if (best < 0 || java_compare(d, bestd) < 0) { best = (int)c; bestd = d; }
    }
    r.has_value = true; r.label = clusters[best].id; r.aux = best + 1; r.value = bestd;
    return r;
  }
};

// ------------------------------------------------------------------------------------------ TreeModel (J2)
struct TreeNode {
  std::string id; Predicate pred; bool has_score = false; std::string score; double score_num = 0; bool score_ok = false; std::string default_child;
  int default_idx = -1;  // position of defaultChild among kids
  int node_number = 0;  // numeric value of id, or implicit 1-based pre-order number
  std::vector<std::pair<std::string, double>> dist;  // ScoreDistribution value -> recordCount / probability
  bool dist_is_prob = false;
  std::vector<TreeNode> kids;
};

struct TreeModel : Model {
  std::string missing_strategy = "none", no_true_child = "returnNullPrediction";
  enum { MVS_NONE, MVS_DEFAULT_CHILD, MVS_LAST, MVS_NULL, MVS_OTHER } mvs = MVS_NONE;
  bool regression = false; bool last_on_no_true = false;
  TreeNode root;

  // returns the selected node or nullptr (null prediction)
  const TreeNode* walk(const Ctx& cx) const {
    Tri rt = eval_pred(root.pred, cx);
    if (rt != T_TRUE) return nullptr;
    const TreeNode* node = &root;
    const TreeNode* last_scored = root.has_score ? &root : nullptr;
    while (!node->kids.empty()) {
      const TreeNode* next = nullptr; bool null_out = false, stop_here = false;
      for (auto& k : node->kids) {
        Tri t = eval_pred(k.pred, cx);
        if (t == T_TRUE) { next = &k; break; }
        if (t == T_UNKNOWN) {
          if (mvs == MVS_NONE) continue;  // treated as FALSE
          if (mvs == MVS_DEFAULT_CHILD) {
            if (node->default_idx < 0) throw EvalError("defaultChild not found");
            next = &node->kids[node->default_idx];
            break;
          }
          if (mvs == MVS_LAST) { stop_here = true; break; }
          if (mvs == MVS_NULL) { null_out = true; break; }
          throw EvalError("unsupported missingValueStrategy " + missing_strategy);  // weightedConfidence, aggregateNodes
        }
      }
      if (null_out) return nullptr;
      if (stop_here) return last_scored;
      if (!next) {  // no child evaluated to TRUE
        if (last_on_no_true) return last_scored;
        return nullptr;
      }
      node = next;
      if (node->has_score) last_scored = node;
    }
    return node;
  }

  Result evaluate(const Ctx& cx, std::vector<int>* node_ids) const override {
    const TreeNode* n = walk(cx);
    Result r;
    if (node_ids) node_ids->push_back(n ? n->node_number : 0);
    if (!n || !n->has_score) return r;  // null prediction
    r.aux = n->node_number;
    if (regression) {
      if (!n->score_ok) throw EvalError("non-numeric score");
      r.has_value = true; r.value = n->score_num;
    } else {
      r.has_value = true; r.is_class = true; r.label = n->score;
      double tot = 0; for (auto& d : n->dist) tot += d.second;
      for (auto& d : n->dist) r.probs.emplace_back(d.first, n->dist_is_prob ? d.second : d.second / tot);
      if (n->dist.empty()) r.probs.emplace_back(n->score, 1.0);
    }
    return r;
  }
};

// ------------------------------------------------------------------------------------------ RegressionModel (J4)
static double apply_link(const std::string& norm, double y) {
  if (norm == "none" || norm.empty()) return y;
  if (norm == "softmax" || norm == "logit") return 1.0 / (1.0 + std::exp(-y));
  if (norm == "exp") return std::exp(y);
  if (norm == "probit") return 0.5 * std::erfc(-y / std::sqrt(2.0));
  if (norm == "cloglog") return 1.0 - std::exp(-std::exp(y));
  if (norm == "loglog") return std::exp(-std::exp(-y));
  if (norm == "cauchit") return 0.5 + std::atan(y) / M_PI;
  throw EvalError("unsupported normalizationMethod " + norm);
}

struct RegressionModel : Model {
  std::string normalization = "none";
  struct Num { int slot; double coef; int exponent; };
  struct Cat { int slot; std::string value; double coef; };
  struct Term { std::vector<int> slots; double coef; };
  struct Table { double intercept = 0; std::string category; std::vector<Num> nums; std::vector<Cat> cats; std::vector<Term> terms; };
  std::vector<Table> tables;

  // returns false when a numeric predictor is missing (whole table -> no result, SURVEY App. A [R])
  bool eval_table(const Table& t, const Ctx& cx, double* out) const {
    double y = t.intercept;
    for (auto& n : t.nums) {
      const FV& v = cx.v[n.slot];
      if (v.is_missing()) return false;
      if (v.kind != FV::NUM) throw EvalError("non-numeric predictor");
      double x = v.num;
      double p = n.exponent == 1 ? x : std::pow(x, (double)n.exponent);
      y += n.coef * p;
    }
    for (auto& c : t.cats) {
      const FV& v = cx.v[c.slot];
      if (v.is_missing()) continue;  // contributes 0
      if (fv_equals_string(v, c.value)) y += c.coef;
    }
    for (auto& pt : t.terms) {
      double p = 1;
      for (int s : pt.slots) { const FV& v = cx.v[s]; if (v.is_missing()) return false; p *= v.num; }
      y += pt.coef * p;
    }
    *out = y;
    return true;
  }

  Result evaluate(const Ctx& cx, std::vector<int>*) const override {
    Result r;
    if (function == "regression") {
      double y; if (!eval_table(tables.at(0), cx, &y)) return r;
      r.has_value = true; r.value = apply_link(normalization, y);
      return r;
    }
    std::vector<double> y(tables.size());
    for (size_t i = 0; i < tables.size(); i++) if (!eval_table(tables[i], cx, &y[i])) return r;
    size_t C = tables.size();
    std::vector<double> p(C);
    if (normalization == "softmax") {
      double s = 0; for (size_t i = 0; i < C; i++) { p[i] = std::exp(y[i]); s += p[i]; }
      for (size_t i = 0; i < C; i++) p[i] /= s;
    } else if (normalization == "simplemax") {
      double s = 0; for (double v : y) s += v;
      for (size_t i = 0; i < C; i++) p[i] = y[i] / s;
    } else if (normalization == "none") {
      double s = 0; for (size_t i = 0; i + 1 < C; i++) { p[i] = y[i]; s += y[i]; }
      p[C - 1] = 1.0 - s;
    } else {  // logit, probit, cloglog, loglog, cauchit: per category link; last = 1 - sum(others)
      double s = 0; for (size_t i = 0; i + 1 < C; i++) { p[i] = apply_link(normalization, y[i]); s += p[i]; }
      p[C - 1] = 1.0 - s;
    }
    int best = 0;  // first maximal entry in RegressionTable order
    for (size_t i = 1; i < C; i++) if (java_compare(p[i], p[best]) > 0) best = (int)i;
    r.has_value = true; r.is_class = true; r.label = tables[best].category; r.aux = best;
    for (size_t i = 0; i < C; i++) r.probs.emplace_back(tables[i].category, p[i]);
    return r;
  }
};

// ------------------------------------------------------------------------------------------ GeneralRegressionModel (J5)
struct GeneralRegressionModel : Model {
  std::string model_type, link, cum_link; double link_param = 0, offset = 0;
  std::vector<std::string> params;          // ParameterList order
  struct PP { int slot; bool factor; std::string value; double power; };  // per parameter: predictors
  std::vector<std::vector<PP>> pp;          // [param] -> cells
  struct Row { std::string category; std::vector<std::pair<int, double>> beta; };  // (param idx, beta)
  std::vector<Row> rows;                    // one per target category (or single)
  std::vector<std::string> categories;      // target categories (DataField values) for multinomial

  bool design(const Ctx& cx, std::vector<double>* x) const {
    x->assign(params.size(), 1.0);
    for (size_t p = 0; p < params.size(); p++) {
      double v = 1.0;
      for (auto& c : pp[p]) {
        const FV& f = cx.v[c.slot];
        if (f.is_missing()) return false;
        if (c.factor) v *= fv_equals_string(f, c.value) ? 1.0 : 0.0;
        else v *= (c.power == 1.0 ? f.num : std::pow(f.num, c.power));
      }
      (*x)[p] = v;
    }
    return true;
  }
  double inv_link(double eta) const {
    if (link.empty() || link == "identity") return eta;
    if (link == "log") return std::exp(eta);
    if (link == "logit") return 1.0 / (1.0 + std::exp(-eta));
    if (link == "probit") return 0.5 * std::erfc(-eta / std::sqrt(2.0));
    if (link == "cloglog") return 1.0 - std::exp(-std::exp(eta));
    if (link == "loglog") return std::exp(-std::exp(-eta));
    if (link == "logc") return 1.0 - std::exp(eta);
    if (link == "cauchit") return 0.5 + std::atan(eta) / M_PI;
    if (link == "power") return link_param == 0 ? std::exp(eta) : std::pow(eta, 1.0 / link_param);
    if (link == "negbin") return 1.0 / (link_param * (std::exp(-eta) - 1.0));
    throw EvalError("unsupported linkFunction " + link);
  }
  Result evaluate(const Ctx& cx, std::vector<int>*) const override {
    Result r; std::vector<double> x;
    if (!design(cx, &x)) return r;
    auto dot = [&](const Row& row) { double s = 0; for (auto& b : row.beta) s += b.second * x[b.first]; return s; };
    if (function == "regression") {
      double eta = dot(rows.at(0)) + offset;
      double v = (model_type == "generalLinear" || model_type == "generalizedLinear") ? inv_link(eta) : eta;
      r.has_value = true; r.value = v; return r;
    }
    if (model_type != "multinomialLogistic" && model_type != "generalizedLinear" && model_type != "generalLinear")
      throw EvalError("unsupported GeneralRegression modelType " + model_type);
    // multinomialLogistic: categories without a PCell row are the reference (eta = 0); softmax over all
    size_t C = categories.size();
    std::vector<double> p(C, 0.0);
    if (model_type == "multinomialLogistic") {
      double s = 0;
      for (size_t i = 0; i < C; i++) {
        double eta = 0; for (auto& row : rows) if (row.category == categories[i]) eta = dot(row);
        p[i] = std::exp(eta); s += p[i];
      }
      for (auto& v : p) v /= s;
    } else {  // binary generalized linear: first row's category gets inv_link(eta)
      double eta = dot(rows.at(0)) + offset;
      double q = inv_link(eta);
      for (size_t i = 0; i < C; i++) p[i] = (categories[i] == rows[0].category) ? q : (C == 2 ? 1.0 - q : 0.0);
    }
    int best = 0; for (size_t i = 1; i < C; i++) if (java_compare(p[i], p[best]) > 0) best = (int)i;
    r.has_value = true; r.is_class = true; r.label = categories[best]; r.aux = best;
    for (size_t i = 0; i < C; i++) r.probs.emplace_back(categories[i], p[i]);
    return r;
  }
};

// ------------------------------------------------------------------------------------------ NeuralNetwork (J6)
struct Expr {  // DerivedField body: FieldRef | NormContinuous | NormDiscrete | Constant
  enum Kind { FIELDREF, NORMCONT, NORMDISC, CONSTANT } kind = FIELDREF;
  int slot = -1; std::string value; double constant = 0;
  std::vector<std::pair<double, double>> pts;  // LinearNorm orig -> norm
  std::string outliers = "asIs"; std::optional<double> map_missing;
  // returns false if missing
  bool eval(const Ctx& cx, double* out) const {
    if (kind == CONSTANT) { *out = constant; return true; }
    const FV& v = cx.v[slot];
    if (v.is_missing()) { if (map_missing) { *out = *map_missing; return true; } return false; }
    switch (kind) {
      case FIELDREF: if (v.kind != FV::NUM) throw EvalError("non-numeric FieldRef"); *out = v.num; return true;
      case NORMDISC: *out = fv_equals_string(v, value) ? 1.0 : 0.0; return true;
      case NORMCONT: {
        double x = v.num; size_t n = pts.size();
        if (x < pts[0].first) {
          if (outliers == "asMissingValues") return false;
          if (outliers == "asExtremeValues") { *out = pts[0].second; return true; }
          *out = interp(0, x); return true;
        }
        if (x > pts[n - 1].first) {
          if (outliers == "asMissingValues") return false;
          if (outliers == "asExtremeValues") { *out = pts[n - 1].second; return true; }
          *out = interp(n - 2, x); return true;
        }
        size_t i = 0; while (i + 2 < n && x > pts[i + 1].first) i++;
        *out = interp(i, x); return true;
      }
      default: return false;
    }
  }
  double interp(size_t i, double x) const {
    double a0 = pts[i].first, b0 = pts[i].second, a1 = pts[i + 1].first, b1 = pts[i + 1].second;
    return b0 + (x - a0) / (a1 - a0) * (b1 - b0);
  }
  double denorm(double y) const {  // inverse of NormContinuous for regression outputs
    size_t n = pts.size(), i = 0;
    while (i + 2 < n && y > pts[i + 1].second) i++;
    double a0 = pts[i].first, b0 = pts[i].second, a1 = pts[i + 1].first, b1 = pts[i + 1].second;
    return (y - b0) / (b1 - b0) * (a1 - a0) + a0;
  }
};

static double activate(const std::string& f, double z, double threshold) {
  if (f == "threshold") return z > threshold ? 1.0 : 0.0;
  if (f == "logistic") return 1.0 / (1.0 + std::exp(-z));
  if (f == "tanh") return std::tanh(z);
  if (f == "identity") return z;
  if (f == "exponential") return std::exp(z);
  if (f == "reciprocal") return 1.0 / z;
  if (f == "square") return z * z;
  if (f == "Gauss") return std::exp(-(z * z));
  if (f == "sine") return std::sin(z);
  if (f == "cosine") return std::cos(z);
  if (f == "Elliott") return z / (1.0 + std::fabs(z));
  if (f == "arctan") return 2.0 * std::atan(z) / M_PI;
  if (f == "rectifier") return std::max(0.0, z);
  throw EvalError("unsupported activationFunction " + f);
}

struct NeuralNetwork : Model {
  std::string activation = "logistic", normalization = "none"; double threshold = 0;
  struct Input { std::string id; Expr expr; };
  struct Neuron { std::string id; double bias = 0, width = 0, altitude = 1; std::vector<std::pair<int, double>> con; };  // (index in id table, weight)
  struct Layer { std::string activation, normalization; double threshold; std::vector<Neuron> neurons; };
  struct Output { int neuron; Expr expr; };
  std::vector<Input> inputs; std::vector<Layer> layers; std::vector<Output> outputs;
  int n_ids = 0;

  Result evaluate(const Ctx& cx, std::vector<int>*) const override {
    Result r;
    std::vector<double> a(n_ids, 0.0);
    int k = 0;
    for (auto& in : inputs) { double v; if (!in.expr.eval(cx, &v)) return r; a[k++] = v; }
    for (auto& L : layers) {
      size_t first = k;
      for (auto& n : L.neurons) {
        if (L.activation == "radialBasis") {
          // PMML 4.3 NeuralNetwork, radial basis function: Z = sum_i (x_i - w_i)^2 / (2 width^2),
          // activation = exp(fanIn * ln(altitude) - Z); the bias takes no part (JPMML NeuralNetworkEvaluator, RADIAL_BASIS)
          double z = 0;
          for (auto& c : n.con) { double dlt = a[c.first] - c.second; z += dlt * dlt; }
          a[k++] = std::exp((double)n.con.size() * std::log(n.altitude) - z / (2.0 * n.width * n.width));
          continue;
        }
        double z = n.bias;
        for (auto& c : n.con) z += c.second * a[c.first];
        a[k++] = activate(L.activation, z, L.threshold);
      }
      if (L.normalization == "softmax") {
        double s = 0; for (size_t i = first; i < (size_t)k; i++) { a[i] = std::exp(a[i]); s += a[i]; }
        for (size_t i = first; i < (size_t)k; i++) a[i] /= s;
      } else if (L.normalization == "simplemax") {
        double s = 0; for (size_t i = first; i < (size_t)k; i++) s += a[i];
        for (size_t i = first; i < (size_t)k; i++) a[i] /= s;
      }
    }
    if (function == "regression") {
      const Output& o = outputs.at(0);
      double y = a[o.neuron];
      r.has_value = true;
      r.value = o.expr.kind == Expr::NORMCONT ? o.expr.denorm(y) : y;
      return r;
    }
    int best = -1;
    for (size_t i = 0; i < outputs.size(); i++) {
      double p = a[outputs[i].neuron];
      r.probs.emplace_back(outputs[i].expr.value, p);
      if (best < 0 || java_compare(p, r.probs[best].second) > 0) best = (int)i;  // first maximal output
    }
    r.has_value = true; r.is_class = true; r.label = r.probs[best].first; r.aux = best;
    return r;
  }
};

// ------------------------------------------------------------------------------------------ MiningModel (J3)
struct MiningModel : Model {
  std::string method;
  struct Segment { Predicate pred; double weight = 1; std::unique_ptr<Model> model; };
  std::vector<Segment> segments;

  Result evaluate(const Ctx& cx, std::vector<int>* node_ids) const override {
    Result r;
    if (method == "modelChain") {
      // segments run in order; a segment's predictedValue output fields become inputs of the later ones; the result is
      // the last evaluated segment's (this is how boosted-tree classifiers are exported: trees -> link function)
      Ctx chain = cx; Result last;
      for (auto& s : segments) {
        if (eval_pred(s.pred, chain) != T_TRUE) continue;
        Result sr = s.model->evaluate(chain, nullptr);
        if (sr.has_value && !sr.is_class) sr.value = apply_target(*s.model, sr.value);
        for (auto& o : s.model->outputs) {
          if (o.second != "predictedValue") continue;
          chain.v[o.first] = !sr.has_value ? FV::missing() : (sr.is_class ? FV::string(sr.label) : FV::number(DT_DOUBLE, sr.value));
        }
        last = std::move(sr);
      }
      if (node_ids) node_ids->push_back(last.aux);
      return last;
    }
    std::vector<std::pair<const Segment*, Result>> res;
    for (auto& s : segments) {
      if (eval_pred(s.pred, cx) != T_TRUE) { if (node_ids) node_ids->push_back(-1); continue; }
      size_t before = node_ids ? node_ids->size() : 0;
      Result sr = s.model->evaluate(cx, node_ids);
      if (node_ids && node_ids->size() == before) node_ids->push_back(sr.aux);
      if (method == "selectFirst") {
        if (sr.has_value && !sr.is_class) sr.value = apply_target(*s.model, sr.value);
        return sr;
      }
      res.emplace_back(&s, std::move(sr));
    }
    if (method == "selectFirst") return r;  // no segment matched
    if (method == "selectAll") throw EvalError("unsupported multipleModelMethod " + method);
    if (res.empty()) return r;
    if (function == "regression") {
      // a segment without a result makes the aggregation fail (SURVEY App. A [R]) -> evaluation error
      for (auto& p : res) if (!p.second.has_value) throw EvalError("segment produced no result");
      double out;
      if (method == "sum" || method == "average" || method == "weightedSum" || method == "weightedAverage") {
        bool w = method[0] == 'w';
        double s = 0, sw = 0;
        for (auto& p : res) {  // sequential, segment order, double
          double v = apply_target(*p.first->model, p.second.value);
          s += w ? p.first->weight * v : v; sw += p.first->weight;
        }
        out = method == "sum" || method == "weightedSum" ? s : (method == "average" ? s / (double)res.size() : s / sw);
      } else if (method == "median" || method == "max") {
        std::vector<double> v; for (auto& p : res) v.push_back(apply_target(*p.first->model, p.second.value));
        std::sort(v.begin(), v.end());
        // median = JPMML's DoubleVector.median(): the 50th percentile of commons-math3's Percentile in its default
        // (LEGACY) estimation: position p = (n + 1) / 2 in the sorted values, lower + frac(p) * (upper - lower)
        if (method == "max") out = v.back();
        else if (v.size() % 2) out = v[v.size() / 2];
        else { double lo = v[v.size() / 2 - 1], hi = v[v.size() / 2]; out = lo + 0.5 * (hi - lo); }
      } else throw EvalError("unsupported multipleModelMethod " + method);
      r.has_value = true; r.value = out;
      return r;
    }
    // classification
    std::vector<std::pair<std::string, double>> tally;  // insertion order
    auto bump = [&](const std::string& k, double w) { for (auto& t : tally) if (t.first == k) { t.second += w; return; } tally.emplace_back(k, w); };
    if (method == "majorityVote" || method == "weightedMajorityVote") {
      double tot = 0;
      for (auto& p : res) {
        if (!p.second.has_value) throw EvalError("segment produced no result");
        double w = method[0] == 'w' ? p.first->weight : 1.0;
        bump(p.second.label, w); tot += w;
      }
      for (auto& t : tally) r.probs.emplace_back(t.first, t.second / tot);
    } else if (method == "average" || method == "weightedAverage") {
      double sw = 0;
      for (auto& p : res) {
        if (!p.second.has_value) throw EvalError("segment produced no result");
        double w = method[0] == 'w' ? p.first->weight : 1.0; sw += w;
        for (auto& q : p.second.probs) bump(q.first, w * q.second);
      }
      for (auto& t : tally) r.probs.emplace_back(t.first, t.second / sw);
    } else throw EvalError("unsupported multipleModelMethod " + method);
    int best = 0;  // first maximal entry in insertion order (SURVEY App. A [R], isolated here)
    for (size_t i = 1; i < tally.size(); i++) if (java_compare(tally[i].second, tally[best].second) > 0) best = (int)i;
    r.has_value = true; r.is_class = true; r.label = tally[best].first; r.aux = best;
    return r;
  }
};

// ------------------------------------------------------------------------------------------ loader
static double to_d(const std::string& s, const char* what) {
  double v; if (!java_parse_double(s, &v)) throw Unsupported(std::string("bad number for ") + what + ": " + s);
  return v;
}

static std::vector<std::string> split_array(const Xml& arr) {  // <Array> content, honours "quoted strings"
  std::vector<std::string> out; const std::string& t = arr.text; size_t i = 0;
  while (i < t.size()) {
    while (i < t.size() && (unsigned char)t[i] <= ' ') i++;
    if (i >= t.size()) break;
    std::string tok;
    if (t[i] == '"') { i++; while (i < t.size() && t[i] != '"') { if (t[i] == '\\' && i + 1 < t.size()) i++; tok += t[i++]; } i++; }
    else while (i < t.size() && (unsigned char)t[i] > ' ') tok += t[i++];
    out.push_back(tok);
  }
  return out;
}

static Predicate::Op opcode(const std::string& s) {
  static const std::map<std::string, Predicate::Op> m = {{"equal", Predicate::EQ}, {"notEqual", Predicate::NE}, {"lessThan", Predicate::LT},
    {"lessOrEqual", Predicate::LE}, {"greaterThan", Predicate::GT}, {"greaterOrEqual", Predicate::GE}, {"isMissing", Predicate::IS_MISSING},
    {"isNotMissing", Predicate::IS_NOT_MISSING}, {"isIn", Predicate::IS_IN}, {"isNotIn", Predicate::IS_NOT_IN}, {"and", Predicate::AND},
    {"or", Predicate::OR}, {"xor", Predicate::XOR}, {"surrogate", Predicate::SURROGATE}};
  auto it = m.find(s); return it == m.end() ? Predicate::BAD : it->second;
}
static Predicate load_pred_(const Xml& e, const Doc& d);
static Predicate load_pred(const Xml& e, const Doc& d) {
  Predicate p = load_pred_(e, d);
  if (p.kind == Predicate::SIMPLE || p.kind == Predicate::SET) p.opc = opcode(p.op);
  if (p.kind == Predicate::COMPOUND) p.opc = opcode(p.boolop);
  return p;
}
static Predicate load_pred_(const Xml& e, const Doc& d) {
  Predicate p;
  if (e.name == "True") p.kind = Predicate::TRUE_;
  else if (e.name == "False") p.kind = Predicate::FALSE_;
  else if (e.name == "SimplePredicate") {
    p.kind = Predicate::SIMPLE; p.field = d.slot(e.get("field")); p.op = e.get("operator"); p.value = e.get("value");
    if (p.op != "isMissing" && p.op != "isNotMissing") { try { p.parsed = parse_as(d.fields[p.field].dt, p.value); p.parse_ok = true; } catch (EvalError&) {} }
  } else if (e.name == "SimpleSetPredicate") {
    p.kind = Predicate::SET; p.field = d.slot(e.get("field")); p.op = e.get("booleanOperator");
    if (auto* a = e.child("Array")) p.set = split_array(*a);
    try { for (auto& t : p.set) p.parsed_set.push_back(parse_as(d.fields[p.field].dt, t)); p.set_ok = true; } catch (EvalError&) {}
  } else if (e.name == "CompoundPredicate") {
    p.kind = Predicate::COMPOUND; p.boolop = e.get("booleanOperator");
    for (auto& k : e.kids) if (k->name != "Extension") p.kids.push_back(load_pred(*k, d));
    if (p.kids.size() < 2) throw Unsupported("CompoundPredicate needs two operands");
  } else throw Unsupported("unknown predicate " + e.name);
  return p;
}
static const Xml* find_pred(const Xml& e) {
  for (auto& k : e.kids)
    if (k->name == "True" || k->name == "False" || k->name == "SimplePredicate" || k->name == "SimpleSetPredicate" || k->name == "CompoundPredicate") return k.get();
  return nullptr;
}

static thread_local const Doc* g_doc = nullptr;  // dictionary of the document being loaded (Output fields resolve against it)
static void load_common(const Xml& e, Model& m) {
  if (auto* out = e.child("Output"))
    for (auto* f : out->children("OutputField")) m.outputs.emplace_back(g_doc->slot(f->get("name")), f->get("feature", "predictedValue"));
  m.model_name = e.get("modelName"); m.function = e.get("functionName");
  if (auto* ms = e.child("MiningSchema"))
    for (auto* f : ms->children("MiningField")) {
      MiningField mf; mf.name = f->get("name"); mf.usage = f->get("usageType", "active");
      std::string ivt = f->get("invalidValueTreatment", "returnInvalid");
      mf.ivt = ivt == "asIs" ? IVT_AS_IS : (ivt == "asMissing" ? IVT_AS_MISSING : IVT_RETURN_INVALID);
      std::string ot = f->get("outliers", "asIs");
      mf.ot = ot == "asMissingValues" ? OT_AS_MISSING : (ot == "asExtremeValues" ? OT_AS_EXTREME : OT_AS_IS);
      if (auto* r = f->attr("missingValueReplacement")) mf.missing_repl = *r;
      if (auto* l = f->attr("lowValue")) { mf.low = to_d(*l, "lowValue"); mf.has_low = true; }
      if (auto* h = f->attr("highValue")) { mf.high = to_d(*h, "highValue"); mf.has_high = true; }
      m.mining.push_back(mf);
    }
  if (auto* ts = e.child("Targets"))
    for (auto* t : ts->children("Target")) {
      TargetSpec s; s.field = t->get("field");
      if (auto* a = t->attr("rescaleFactor")) s.factor = to_d(*a, "rescaleFactor");
      if (auto* a = t->attr("rescaleConstant")) s.constant = to_d(*a, "rescaleConstant");
      if (auto* a = t->attr("min")) s.mn = to_d(*a, "min");
      if (auto* a = t->attr("max")) s.mx = to_d(*a, "max");
      s.cast = t->get("castInteger");
      m.targets.push_back(s);
    }
}

static Expr load_expr(const Xml& derived, const Doc& d) {
  Expr x;
  for (auto& k : derived.kids) {
    if (k->name == "FieldRef") { x.kind = Expr::FIELDREF; x.slot = d.slot(k->get("field")); if (auto* m = k->attr("mapMissingTo")) x.map_missing = to_d(*m, "mapMissingTo"); return x; }
    if (k->name == "NormDiscrete") { x.kind = Expr::NORMDISC; x.slot = d.index.count(k->get("field")) ? d.slot(k->get("field")) : -1; x.value = k->get("value"); return x; }
    if (k->name == "Constant") { x.kind = Expr::CONSTANT; x.constant = to_d(trim(k->text), "Constant"); return x; }
    if (k->name == "NormContinuous") {
      x.kind = Expr::NORMCONT; x.slot = d.index.count(k->get("field")) ? d.slot(k->get("field")) : -1; x.outliers = k->get("outliers", "asIs");
      if (auto* m = k->attr("mapMissingTo")) x.map_missing = to_d(*m, "mapMissingTo");
      for (auto* ln : k->children("LinearNorm")) x.pts.emplace_back(to_d(ln->get("orig"), "orig"), to_d(ln->get("norm"), "norm"));
      if (x.pts.size() < 2) throw Unsupported("NormContinuous needs two LinearNorm");
      return x;
    }
  }
  throw Unsupported("unsupported DerivedField expression");
}

// user functions of the TransformationDictionary + the parameter bindings of the call being expanded
struct FnEnv { std::map<std::string, const Xml*> fns; const std::map<std::string, DExpr>* params = nullptr; int depth = 0; };

static DExpr load_dexpr(const Xml& e, const Doc& d, const FnEnv* env = nullptr) {
  DExpr x;
  if (env && env->params) {   // inside a DefineFunction body only the ParameterFields are visible
    if (e.name == "FieldRef") {
      auto it = env->params->find(e.get("field")); if (it == env->params->end()) throw Unsupported("DefineFunction body refers to a field that is not a ParameterField");
      x.kind = DExpr::PARAM; x.args.push_back(it->second); if (auto* a = e.attr("mapMissingTo")) x.map_missing = to_d(*a, "mapMissingTo"); return x;
    }
    if (e.name != "Constant" && e.name != "Apply") throw Unsupported("DefineFunction body element " + e.name);
  }
  auto opt = [&](const char* k) -> std::optional<double> { if (auto* a = e.attr(k)) return to_d(*a, k); return std::nullopt; };
  auto numeric_slot = [&](const std::string& f) { int sl = d.slot(f); DType dt = d.fields[sl].dt; if (dt == DT_STRING || dt == DT_OTHER) throw Unsupported("transformation of a non-numeric field"); return sl; };
  if (e.name == "Constant") { x.kind = DExpr::CONST; x.constant = to_d(trim(e.text), "Constant"); return x; }
  if (e.name == "FieldRef") { x.kind = DExpr::FIELD; x.slot = numeric_slot(e.get("field")); x.map_missing = opt("mapMissingTo"); return x; }
  if (e.name == "NormDiscrete") { x.kind = DExpr::NORMDISC; x.slot = numeric_slot(e.get("field")); x.match = to_d(e.get("value"), "NormDiscrete value"); x.map_missing = opt("mapMissingTo"); return x; }
  if (e.name == "NormContinuous") {
    x.kind = DExpr::NORMCONT; x.slot = numeric_slot(e.get("field")); x.map_missing = opt("mapMissingTo");
    std::string o = e.get("outliers", "asIs"); x.outliers = o == "asMissingValues" ? 1 : (o == "asExtremeValues" ? 2 : 0);
    for (auto* ln : e.children("LinearNorm")) x.pts.emplace_back(to_d(ln->get("orig"), "orig"), to_d(ln->get("norm"), "norm"));
    if (x.pts.size() < 2) throw Unsupported("NormContinuous needs two LinearNorm");
    return x;
  }
  if (e.name == "Discretize") {
    x.kind = DExpr::DISCRETIZE; x.slot = numeric_slot(e.get("field")); x.map_missing = opt("mapMissingTo"); x.dflt = opt("defaultValue");
    for (auto* b : e.children("DiscretizeBin")) {
      DExpr::Bin bin{false, false, false, false, 0, 0, to_d(b->get("binValue"), "binValue")};
      const Xml* iv = b->child("Interval"); if (!iv) throw Unsupported("DiscretizeBin without Interval");
      std::string c = iv->get("closure"); bin.open_l = c.rfind("open", 0) == 0; bin.open_r = c.size() >= 4 && c.substr(c.size() - 4) == "Open";
      if (auto* a = iv->attr("leftMargin")) { bin.has_l = true; bin.l = to_d(*a, "leftMargin"); }
      if (auto* a = iv->attr("rightMargin")) { bin.has_r = true; bin.r = to_d(*a, "rightMargin"); }
      x.bins.push_back(bin);
    }
    return x;
  }
  if (e.name == "MapValues") {
    x.kind = DExpr::MAPVALUES; x.map_missing = opt("mapMissingTo"); x.dflt = opt("defaultValue");
    auto pairs = e.children("FieldColumnPair"); if (pairs.size() != 1) throw Unsupported("MapValues with several FieldColumnPairs");
    x.slot = numeric_slot(pairs[0]->get("field")); std::string in_col = pairs[0]->get("column"), out_col = e.get("outputColumn");
    const Xml* tbl = e.child("InlineTable"); if (!tbl) throw Unsupported("MapValues without InlineTable");
    for (auto* row : tbl->children("row")) {
      const Xml* i = row->child(in_col.c_str()); const Xml* o = row->child(out_col.c_str());
      if (!i || !o) throw Unsupported("MapValues row without both columns");
      x.pairs.emplace_back(to_d(trim(i->text), "MapValues input"), to_d(trim(o->text), "MapValues output"));
    }
    return x;
  }
  if (e.name == "Apply") {
    x.kind = DExpr::APPLY; x.function = e.get("function"); x.map_missing = opt("mapMissingTo"); x.dflt = opt("defaultValue");
    for (auto& k : e.kids) if (k->name != "Extension") x.args.push_back(load_dexpr(*k, d, env));
    auto fn = env ? env->fns.find(x.function) : std::map<std::string, const Xml*>::const_iterator();
    if (env && fn != env->fns.end()) {
      const Xml& df = *fn->second;
      if (env->depth > 8) throw Unsupported("DefineFunction calls nested deeper than 8");
      DType fdt = parse_dtype(df.get("dataType", "double"));
      if (fdt != DT_DOUBLE && fdt != DT_FLOAT) throw Unsupported("DefineFunction dataType other than double / float");
      std::map<std::string, DExpr> bind; size_t i = 0; const Xml* body = nullptr;
      for (auto& k : df.kids) {
        if (k->name == "Extension") continue;
        if (k->name == "ParameterField") {
          if (k->attr("dataType") && k->get("dataType") != "double") throw Unsupported("ParameterField dataType other than double");
          if (i >= x.args.size()) throw Unsupported("too few arguments for " + x.function);
          bind[k->get("name")] = x.args[i++];
        } else body = k.get();
      }
      if (i != x.args.size() || !body) throw Unsupported("wrong number of arguments for " + x.function);
      FnEnv inner; inner.fns = env->fns; inner.params = &bind; inner.depth = env->depth + 1;
      x.kind = DExpr::USERCALL; x.to_float = fdt == DT_FLOAT; x.args.push_back(load_dexpr(*body, d, &inner));
    }
    return x;
  }
  throw Unsupported("unsupported transformation element " + e.name);
}

static std::unique_ptr<Model> load_model(const Xml& e, const Doc& d);

// LocalTransformations of a model nested in a Segment: its DerivedFields are visible inside that model only (two segments may use
// one name for different things).  Expressions are pure functions of fields known before any model runs, so they are evaluated
// up front with all the others -- each in a context slot of its own -- and the NAME is bound to that slot only while the nested
// model is being loaded (everything after loading works on slots).
struct NestedSink { std::vector<DerivedField>* derived; const FnEnv* env; int first_output_slot, end_output_slot; };
static thread_local NestedSink* g_nested = nullptr;
static bool dexpr_reads_slots(const DExpr& x, int lo, int hi) {
  if (x.kind != DExpr::CONST && x.kind != DExpr::APPLY && x.kind != DExpr::PARAM && x.kind != DExpr::USERCALL && x.slot >= lo && x.slot < hi) return true;
  for (auto& a : x.args) if (dexpr_reads_slots(a, lo, hi)) return true;
  return false;
}
struct NestedScope {
  Doc* doc = nullptr; std::vector<std::tuple<std::string, bool, int>> saved;   // name, had a binding before, that binding
  NestedScope(const Xml& model, const Doc& d) {
    const Xml* lt = model.child("LocalTransformations");
    if (!lt || !g_nested) return;
    doc = const_cast<Doc*>(&d);
    for (auto* f : lt->children("DerivedField")) {
      DType dt = parse_dtype(f->get("dataType", "double"));
      if (dt == DT_STRING || dt == DT_OTHER || dt == DT_BOOLEAN) throw Unsupported("non-numeric DerivedField");
      const Xml* ex = nullptr; for (auto& k : f->kids) if (k->name != "Extension" && k->name != "Value") { ex = k.get(); break; }
      if (!ex) throw Unsupported("DerivedField without expression");
      DExpr expr = load_dexpr(*ex, *doc, g_nested->env);
      if (dexpr_reads_slots(expr, g_nested->first_output_slot, g_nested->end_output_slot)) throw Unsupported("nested DerivedField over another model's output");
      const std::string user = f->get("name"); const int slot = (int)doc->fields.size();
      DataField df; df.name = user + "\x01" + std::to_string(slot); df.op = parse_optype(f->get("optype", "continuous")); df.dt = dt;
      doc->index[df.name] = slot; doc->fields.push_back(df);
      g_nested->derived->push_back({slot, dt, std::move(expr)});
      auto it = doc->index.find(user);
      saved.emplace_back(user, it != doc->index.end(), it != doc->index.end() ? it->second : -1);
      doc->index[user] = slot;
    }
  }
  ~NestedScope() {
    if (!doc) return;
    for (auto it = saved.rbegin(); it != saved.rend(); ++it) { if (std::get<1>(*it)) doc->index[std::get<0>(*it)] = std::get<2>(*it); else doc->index.erase(std::get<0>(*it)); }
  }
};

static void number_tree(TreeNode& n, int& counter) {
  counter++;
  double v;
  n.node_number = (!n.id.empty() && java_parse_double(n.id, &v) && v == std::floor(v)) ? (int)v : counter;
  if (n.id.empty()) n.id = std::to_string(counter);
  for (auto& k : n.kids) number_tree(k, counter);
}
static TreeNode load_node(const Xml& e, const Doc& d) {
  TreeNode n; n.id = e.get("id");
  if (auto* s = e.attr("score")) { n.has_score = true; n.score = *s; n.score_ok = java_parse_double(*s, &n.score_num); }
  n.default_child = e.get("defaultChild");
  const Xml* pe = find_pred(e);
  if (!pe) throw Unsupported("Node without predicate");
  n.pred = load_pred(*pe, d);
  for (auto* sd : e.children("ScoreDistribution")) {
    if (auto* pr = sd->attr("probability")) { n.dist.emplace_back(sd->get("value"), to_d(*pr, "probability")); n.dist_is_prob = true; }
    else n.dist.emplace_back(sd->get("value"), to_d(sd->get("recordCount"), "recordCount"));
  }
  for (auto* k : e.children("Node")) n.kids.push_back(load_node(*k, d));
  for (size_t i = 0; i < n.kids.size(); i++) if (!n.default_child.empty() && n.kids[i].id == n.default_child) n.default_idx = (int)i;
  return n;
}

static std::unique_ptr<Model> load_model(const Xml& e, const Doc& d) {
  if (e.name == "ClusteringModel") {
    auto m = std::make_unique<ClusteringModel>(); load_common(e, *m);
    if (e.get("modelClass") != "centerBased") throw Unsupported("only centerBased ClusteringModel is supported");
    const Xml* cm = e.child("ComparisonMeasure"); if (!cm) throw Unsupported("no ComparisonMeasure");
    m->kind = cm->get("kind"); m->default_compare = cm->get("compareFunction", "absDiff");
    for (auto& k : cm->kids) if (k->name != "Extension") { m->measure = k->name; if (k->name == "minkowski") m->minkowski_p = to_d(k->get("p-parameter"), "p-parameter"); }
    for (auto* cf : e.children("ClusteringField")) {
      if (cf->get("isCenterField", "true") != "true") continue;
      ClusteringModel::CField c; c.slot = d.slot(cf->get("field")); c.weight = to_d(cf->get("fieldWeight", "1"), "fieldWeight");
      c.compare = cf->get("compareFunction", m->default_compare.c_str());
      if (auto* s = cf->attr("similarityScale")) c.sim_scale = to_d(*s, "similarityScale");
      m->cfields.push_back(c);
    }
    if (auto* mw = e.child("MissingValueWeights")) if (auto* a = mw->child("Array")) for (auto& t : split_array(*a)) m->missing_weights.push_back(to_d(t, "MissingValueWeights"));
    int pos = 0;
    for (auto* c : e.children("Cluster")) {
      pos++; ClusteringModel::Cluster cl; cl.id = c->attr("id") ? c->get("id") : std::to_string(pos);  // EntityUtil.getId
      const Xml* a = c->child("Array"); if (!a) throw Unsupported("Cluster without Array");
      for (auto& t : split_array(*a)) cl.center.push_back(to_d(t, "Cluster Array"));
      if (cl.center.size() != m->cfields.size()) throw Unsupported("Cluster Array size mismatch");
      m->clusters.push_back(cl);
    }
    if (m->clusters.empty()) throw Unsupported("no clusters");
    return m;
  }
  if (e.name == "TreeModel") {
    auto m = std::make_unique<TreeModel>(); load_common(e, *m);
    m->missing_strategy = e.get("missingValueStrategy", "none"); m->no_true_child = e.get("noTrueChildStrategy", "returnNullPrediction");
    const Xml* root = e.child("Node"); if (!root) throw Unsupported("TreeModel without Node");
    m->root = load_node(*root, d); int c = 0; number_tree(m->root, c);
    const std::string& ms = m->missing_strategy;
    m->mvs = ms == "none" ? TreeModel::MVS_NONE : ms == "defaultChild" ? TreeModel::MVS_DEFAULT_CHILD : ms == "lastPrediction" ? TreeModel::MVS_LAST : ms == "nullPrediction" ? TreeModel::MVS_NULL : TreeModel::MVS_OTHER;
    m->regression = m->function == "regression"; m->last_on_no_true = m->no_true_child == "returnLastPrediction";
    return m;
  }
  if (e.name == "RegressionModel") {
    auto m = std::make_unique<RegressionModel>(); load_common(e, *m);
    m->normalization = e.get("normalizationMethod", "none");
    for (auto* t : e.children("RegressionTable")) {
      RegressionModel::Table tb; tb.intercept = to_d(t->get("intercept"), "intercept"); tb.category = t->get("targetCategory");
      for (auto* n : t->children("NumericPredictor")) tb.nums.push_back({d.slot(n->get("name")), to_d(n->get("coefficient"), "coefficient"), (int)to_d(n->get("exponent", "1"), "exponent")});
      for (auto* n : t->children("CategoricalPredictor")) tb.cats.push_back({d.slot(n->get("name")), n->get("value"), to_d(n->get("coefficient"), "coefficient")});
      for (auto* n : t->children("PredictorTerm")) { RegressionModel::Term pt; pt.coef = to_d(n->get("coefficient"), "coefficient"); for (auto* fr : n->children("FieldRef")) pt.slots.push_back(d.slot(fr->get("field"))); tb.terms.push_back(pt); }
      m->tables.push_back(tb);
    }
    if (m->tables.empty()) throw Unsupported("no RegressionTable");
    return m;
  }
  if (e.name == "GeneralRegressionModel") {
    auto m = std::make_unique<GeneralRegressionModel>(); load_common(e, *m);
    m->model_type = e.get("modelType"); m->link = e.get("linkFunction"); m->cum_link = e.get("cumulativeLink");
    if (auto* a = e.attr("linkParameter")) m->link_param = to_d(*a, "linkParameter");
    if (auto* a = e.attr("offsetValue")) m->offset = to_d(*a, "offsetValue");
    if (m->model_type == "ordinalMultinomial" || m->model_type == "CoxRegression") throw Unsupported("GeneralRegression modelType " + m->model_type + " is not supported");
    std::unordered_map<std::string, int> pidx;
    if (auto* pl = e.child("ParameterList")) for (auto* p : pl->children("Parameter")) { pidx[p->get("name")] = (int)m->params.size(); m->params.push_back(p->get("name")); }
    m->pp.resize(m->params.size());
    std::unordered_map<std::string, bool> is_factor;
    if (auto* fl = e.child("FactorList")) for (auto* p : fl->children("Predictor")) is_factor[p->get("name")] = true;
    if (auto* cl = e.child("CovariateList")) for (auto* p : cl->children("Predictor")) is_factor[p->get("name")] = false;
    if (auto* ppm = e.child("PPMatrix")) for (auto* c : ppm->children("PPCell")) {
      auto it = pidx.find(c->get("parameterName")); if (it == pidx.end()) throw Unsupported("PPCell for unknown parameter");
      std::string pn = c->get("predictorName");
      bool fac = is_factor.count(pn) ? is_factor[pn] : false;
      GeneralRegressionModel::PP cell{d.slot(pn), fac, c->get("value"), fac ? 1.0 : to_d(c->get("value", "1"), "PPCell value")};
      m->pp[it->second].push_back(cell);
    }
    if (auto* pm = e.child("ParamMatrix")) for (auto* c : pm->children("PCell")) {
      std::string cat = c->get("targetCategory");
      GeneralRegressionModel::Row* row = nullptr;
      for (auto& r : m->rows) if (r.category == cat) row = &r;
      if (!row) { m->rows.push_back({cat, {}}); row = &m->rows.back(); }
      auto it = pidx.find(c->get("parameterName")); if (it == pidx.end()) throw Unsupported("PCell for unknown parameter");
      row->beta.emplace_back(it->second, to_d(c->get("beta"), "beta"));
    }
    if (m->rows.empty()) throw Unsupported("empty ParamMatrix");
    if (m->function == "classification") {
      const MiningField* tf = m->target_field(); if (!tf) throw Unsupported("classification without target field");
      m->categories = d.fields[d.slot(tf->name)].valid;
      if (m->categories.empty()) throw Unsupported("target field declares no categories");
    }
    return m;
  }
  if (e.name == "NeuralNetwork") {
    auto m = std::make_unique<NeuralNetwork>(); load_common(e, *m);
    m->activation = e.get("activationFunction", "logistic"); m->normalization = e.get("normalizationMethod", "none");
    if (auto* t = e.attr("threshold")) m->threshold = to_d(*t, "threshold");
    std::unordered_map<std::string, int> ids;
    const Xml* ni = e.child("NeuralInputs"); if (!ni) throw Unsupported("no NeuralInputs");
    for (auto* in : ni->children("NeuralInput")) {
      const Xml* df = in->child("DerivedField"); if (!df) throw Unsupported("NeuralInput without DerivedField");
      ids[in->get("id")] = (int)ids.size(); m->inputs.push_back({in->get("id"), load_expr(*df, d)});
    }
    for (auto* L : e.children("NeuralLayer")) {
      NeuralNetwork::Layer layer; layer.activation = L->get("activationFunction", m->activation.c_str());
      layer.normalization = L->get("normalizationMethod", m->normalization.c_str());
      layer.threshold = L->attr("threshold") ? to_d(L->get("threshold"), "threshold") : m->threshold;
      for (auto* n : L->children("Neuron")) {
        NeuralNetwork::Neuron nn; nn.id = n->get("id"); nn.bias = to_d(n->get("bias", "0"), "bias");
        if (layer.activation == "radialBasis") {   // width / altitude: the Neuron's, else the layer's, else the network's
          const Xml* src[3] = {n, L, &e}; bool has_w = false;
          for (auto* x : src) if (!has_w && x->attr("width")) { nn.width = to_d(x->get("width"), "width"); has_w = true; }
          for (auto* x : src) if (x->attr("altitude")) { nn.altitude = to_d(x->get("altitude"), "altitude"); break; }
          if (!has_w) throw Unsupported("radialBasis neuron without width");
        }
        for (auto* c : n->children("Con")) { auto it = ids.find(c->get("from")); if (it == ids.end()) throw Unsupported("Con from unknown neuron"); nn.con.emplace_back(it->second, to_d(c->get("weight"), "weight")); }
        layer.neurons.push_back(nn);
      }
      for (auto& n : layer.neurons) ids[n.id] = (int)ids.size();
      m->layers.push_back(std::move(layer));
    }
    m->n_ids = (int)ids.size();
    const Xml* no = e.child("NeuralOutputs"); if (!no) throw Unsupported("no NeuralOutputs");
    for (auto* o : no->children("NeuralOutput")) {
      const Xml* df = o->child("DerivedField"); if (!df) throw Unsupported("NeuralOutput without DerivedField");
      auto it = ids.find(o->get("outputNeuron")); if (it == ids.end()) throw Unsupported("NeuralOutput for unknown neuron");
      m->outputs.push_back({it->second, load_expr(*df, d)});
    }
    if (m->outputs.empty()) throw Unsupported("no NeuralOutput");
    return m;
  }
  if (e.name == "MiningModel") {
    auto m = std::make_unique<MiningModel>(); load_common(e, *m);
    const Xml* sg = e.child("Segmentation"); if (!sg) throw Unsupported("MiningModel without Segmentation");
    m->method = sg->get("multipleModelMethod");
    for (auto* s : sg->children("Segment")) {
      MiningModel::Segment seg; seg.weight = to_d(s->get("weight", "1"), "weight");
      const Xml* pe = find_pred(*s); if (!pe) throw Unsupported("Segment without predicate");
      seg.pred = load_pred(*pe, d);
      for (auto& k : s->kids) {
        const std::string& n = k->name;
        if (n == "TreeModel" || n == "RegressionModel" || n == "MiningModel" || n == "ClusteringModel" || n == "NeuralNetwork" || n == "GeneralRegressionModel") {
          NestedScope scope(*k, d);   // the model's own LocalTransformations
          seg.model = load_model(*k, d);
        }
      }
      if (!seg.model) throw Unsupported("Segment without a supported model");
      m->segments.push_back(std::move(seg));
    }
    return m;
  }
  throw Unsupported("unsupported model element " + e.name);
}

// ------------------------------------------------------------------------------------------ the per-event pipeline
struct Loaded {
  Doc doc;
  std::unique_ptr<Model> model;
  std::vector<int> active_slots;          // vector slot i <-> i-th active MiningField (top-level MiningSchema order)
  std::vector<const MiningField*> active; // same order
  std::vector<std::string> active_names;
  // target extraction (Pipeline.scala:61-98): first *named* target field
  bool has_named_target = false; DType target_dt = DT_DOUBLE;
  int n_segments = 0;
  std::vector<DerivedField> derived;   // TransformationDictionary then LocalTransformations, document order
};

enum Status : uint8_t { ST_OK = 0, ST_SIZE = 1, ST_PREP = 2, ST_EVAL = 3, ST_EXTRACT = 4, ST_NOMODEL = 5 };

// FieldValueUtil.prepareInputValue (JPMML 1.3.x), SURVEY App. A "Input preparation".
static FV prepare(const DataField& df, const MiningField& mf, bool present, double raw) {
  auto missing_path = [&]() -> FV {
    if (mf.missing_repl) {
      try { return parse_as(df.dt, *mf.missing_repl); } catch (EvalError& e) { throw PrepError(e.what()); }
    }
    return FV::missing();
  };
  if (!present) return missing_path();
  FV v; bool castable = true;
  try { v = cast_double_to(df.dt, raw); } catch (PrepError&) { castable = false; }
  enum { VALID, INVALID, MISSING } st = VALID;
  if (!castable) st = INVALID;
  else {
    bool has_valid_space = !df.valid.empty(); bool matched = false;
    try {
      for (auto& s : df.missing) if (fv_equals_string(v, s)) { st = MISSING; matched = true; break; }
      if (!matched) for (auto& s : df.invalid) if (fv_equals_string(v, s)) { st = INVALID; matched = true; break; }
      if (!matched && has_valid_space) { bool ok = false; for (auto& s : df.valid) if (fv_equals_string(v, s)) { ok = true; break; } st = ok ? VALID : INVALID; matched = true; }
    } catch (EvalError&) { st = INVALID; matched = true; }
    if (!matched && df.op == OP_CONTINUOUS && !df.intervals.empty()) {
      bool in = false;
      for (auto& iv : df.intervals) {
        bool lo = !iv.has_l || (iv.open_l ? v.num > iv.l : v.num >= iv.l);
        bool hi = !iv.has_r || (iv.open_r ? v.num < iv.r : v.num <= iv.r);
        if (lo && hi) { in = true; break; }
      }
      st = in ? VALID : INVALID;
    }
  }
  if (st == MISSING) return missing_path();
  if (st == INVALID) {
    switch (mf.ivt) {
      case IVT_RETURN_INVALID: throw PrepError("invalid value");
      case IVT_AS_IS: if (!castable) throw PrepError("invalid value cannot be used as is"); return v;
      case IVT_AS_MISSING: return missing_path();
    }
  }
  // valid value: outlier treatment
  if (mf.ot != OT_AS_IS && v.kind == FV::NUM) {
    if (mf.has_low && v.num < mf.low) { if (mf.ot == OT_AS_MISSING) return missing_path(); v.num = mf.low; }
    else if (mf.has_high && v.num > mf.high) { if (mf.ot == OT_AS_MISSING) return missing_path(); v.num = mf.high; }
  }
  return v;
}

// PmmlModel.predict (api/PmmlModel.scala:109-119) for one event.
static void predict_one(const Loaded& L, const double* x, const uint32_t* present, bool has_replace, double replace,
                        int vec_size, double* score, uint8_t* status, int32_t* aux, std::vector<int>* node_ids,
                        std::vector<double>* probs, Ctx& cx) {
  *score = std::numeric_limits<double>::quiet_NaN(); if (aux) *aux = -1;
  int F = (int)L.active.size();
  if (vec_size != F) { *status = ST_SIZE; return; }                 // validateInput :127-134
  cx.v.assign(L.doc.fields.size(), FV::missing());
  try {                                                              // prepareInput :143-152
    for (int i = 0; i < F; i++) {
      bool pres = present ? ((present[i >> 5] >> (i & 31)) & 1u) : true;
      double raw = x[i];
      if (!pres && has_replace) { pres = true; raw = replace; }      // map.get(name).orElse(replaceNaN).orNull
      cx.v[L.active_slots[i]] = prepare(L.doc.fields[L.active_slots[i]], *L.active[i], pres, raw);
    }
  } catch (PrepError&) { *status = ST_PREP; return; }
  Result r;
  try {
    for (auto& df : L.derived) {   // derived fields see the prepared values and the derived fields before them
      double v;
      if (!df.expr.eval(cx, &v)) { cx.v[df.slot] = FV::missing(); continue; }
      if (df.dt == DT_FLOAT) v = (double)(float)v;
      else if (df.dt == DT_INTEGER && !(std::isfinite(v) && v == std::floor(v))) throw EvalError("derived value is not an integer");
      cx.v[df.slot] = FV::number(df.dt, v);
    }
    r = L.model->evaluate(cx, node_ids);
  }                        // evaluateInput :159-160
  catch (EvalError&) { *status = ST_EVAL; return; }
  if (probs) { for (auto& p : r.probs) probs->push_back(p.second); }
  if (aux) *aux = r.aux;
  // extractTarget :167-174 + Pipeline.extractTargetValue :94-98
  if (!L.has_named_target) { *status = ST_EXTRACT; return; }
  if (!r.has_value) { *status = ST_EVAL; return; }                   // null prediction -> "Target value is null."
  if (r.is_class) {
    // label is delivered in the target field's dataType: String -> toDouble; Double -> itself; else None
    if (L.target_dt == DT_STRING || L.target_dt == DT_DOUBLE) {
      double v; if (!java_parse_double(r.label, &v)) { *status = ST_EXTRACT; return; }
      *score = v; *status = ST_OK; return;
    }
    *status = ST_EXTRACT; return;
  }
  double v = apply_target(*L.model, r.value);
  if (L.target_dt != DT_DOUBLE) { *status = ST_EXTRACT; return; }    // Float / Integer results are not matched
  *score = v; *status = ST_OK;
}

static std::unique_ptr<Loaded> load(const char* data, size_t len) {
  auto root = parse_xml(data, len);
  if (root->name != "PMML") throw Unsupported("root element is not PMML");
  auto L = std::make_unique<Loaded>();
  const Xml* dd = root->child("DataDictionary"); if (!dd) throw Unsupported("no DataDictionary");
  for (auto* f : dd->children("DataField")) {
    DataField df; df.name = f->get("name"); df.op = parse_optype(f->get("optype")); df.dt = parse_dtype(f->get("dataType"));
    for (auto* iv : f->children("Interval")) {
      Interval i; std::string c = iv->get("closure");
      i.open_l = c.rfind("open", 0) == 0; i.open_r = c.size() >= 4 && c.substr(c.size() - 4) == "Open";
      if (auto* a = iv->attr("leftMargin")) { i.has_l = true; i.l = to_d(*a, "leftMargin"); }
      if (auto* a = iv->attr("rightMargin")) { i.has_r = true; i.r = to_d(*a, "rightMargin"); }
      df.intervals.push_back(i);
    }
    for (auto* v : f->children("Value")) {
      std::string p = v->get("property", "valid");
      (p == "invalid" ? df.invalid : (p == "missing" ? df.missing : df.valid)).push_back(v->get("value"));
    }
    L->doc.index[df.name] = (int)L->doc.fields.size(); L->doc.fields.push_back(df);
  }
  std::vector<const Xml*> derived_xml; FnEnv fn_env;
  if (auto* td = root->child("TransformationDictionary")) {
    for (auto* f : td->children("DefineFunction")) fn_env.fns[f->get("name")] = f;
    for (auto* f : td->children("DerivedField")) derived_xml.push_back(f);
  }
  const Xml* me = nullptr;
  for (auto& k : root->kids) {
    const std::string& n = k->name;
    if (n == "ClusteringModel" || n == "TreeModel" || n == "MiningModel" || n == "RegressionModel" || n == "GeneralRegressionModel" || n == "NeuralNetwork") { me = k.get(); break; }
    if (n.size() > 5 && n.substr(n.size() - 5) == "Model" && n != "MiningModel") throw Unsupported("unsupported model element " + n);
  }
  if (!me) throw Unsupported("no supported model element");
  if (auto* lt = me->child("LocalTransformations")) for (auto* f : lt->children("DerivedField")) derived_xml.push_back(f);
  for (auto* f : derived_xml) {   // each derived field gets a context slot; its expression may use the ones before it
    DType dt = parse_dtype(f->get("dataType", "double"));
    if (dt == DT_STRING || dt == DT_OTHER || dt == DT_BOOLEAN) throw Unsupported("non-numeric DerivedField");
    const Xml* ex = nullptr; for (auto& k : f->kids) if (k->name != "Extension" && k->name != "Value") { ex = k.get(); break; }
    if (!ex) throw Unsupported("DerivedField without expression");
    DExpr expr = load_dexpr(*ex, L->doc, &fn_env);
    DataField df; df.name = f->get("name"); df.op = parse_optype(f->get("optype", "continuous")); df.dt = dt;
    L->doc.index[df.name] = (int)L->doc.fields.size(); L->doc.fields.push_back(df);
    L->derived.push_back({(int)L->doc.fields.size() - 1, dt, std::move(expr)});
  }
  const int first_output_slot = (int)L->doc.fields.size();
  {  // Output fields anywhere in the model tree are addressable by later modelChain segments: give them context slots
    std::function<void(const Xml&)> scan = [&](const Xml& x) {
      if (x.name == "OutputField" && !L->doc.index.count(x.get("name"))) {
        DataField df; df.name = x.get("name"); df.op = parse_optype(x.get("optype", "continuous")); df.dt = parse_dtype(x.get("dataType", "double"));
        L->doc.index[df.name] = (int)L->doc.fields.size(); L->doc.fields.push_back(df);
      }
      for (auto& k : x.kids) scan(*k);
    };
    scan(*me);
  }
  g_doc = &L->doc;
  NestedSink sink{&L->derived, &fn_env, first_output_slot, (int)L->doc.fields.size()};
  g_nested = &sink;
  try { L->model = load_model(*me, L->doc); } catch (...) { g_nested = nullptr; throw; }
  g_nested = nullptr;
  for (auto& mf : L->model->mining) {
    if (mf.usage == "active") { L->active.push_back(&mf); L->active_slots.push_back(L->doc.slot(mf.name)); L->active_names.push_back(mf.name); }
  }
  if (const MiningField* tf = L->model->target_field()) {
    L->has_named_target = true;
    // clustering results are entity ids: always strings (ClusterAffinityDistribution)
    L->target_dt = dynamic_cast<ClusteringModel*>(L->model.get()) ? DT_STRING : L->doc.fields[L->doc.slot(tf->name)].dt;
  }
  if (auto* mm = dynamic_cast<MiningModel*>(L->model.get())) L->n_segments = (int)mm->segments.size();
  return L;
}

}  // namespace orc

// ============================================================================================ C ABI (ctypes)
extern "C" {

void* orc_load(const char* pmml, size_t len, char* err, size_t errcap) {
  try { return orc::load(pmml, len).release(); }
  catch (std::exception& e) { if (err && errcap) { strncpy(err, e.what(), errcap - 1); err[errcap - 1] = 0; } return nullptr; }
}
void orc_free(void* m) { delete static_cast<orc::Loaded*>(m); }
int orc_num_active(void* m) { return (int)static_cast<orc::Loaded*>(m)->active.size(); }
const char* orc_active_name(void* m, int i) { return static_cast<orc::Loaded*>(m)->active_names[i].c_str(); }
const char* orc_model_name(void* m) { return static_cast<orc::Loaded*>(m)->model->model_name.c_str(); }
int orc_num_segments(void* m) { return static_cast<orc::Loaded*>(m)->n_segments; }

// Scores n events.  X: n x vec_size doubles row-major; present: n x ceil(vec_size/32) words or NULL (all present);
// replace: n doubles (NaN = none) or NULL.  node_ids: optional n x n_segments (or n x 1); probs: optional n x n_probs.
int orc_predict(void* mp, const double* X, const uint32_t* present, const double* replace, long n, int vec_size,
                double* score, uint8_t* status, int32_t* aux, int32_t* node_ids, int node_stride,
                double* probs, int prob_stride, int nthreads) {
  auto* L = static_cast<orc::Loaded*>(mp);
  int W = (vec_size + 31) / 32;
  auto work = [&](long lo, long hi) {
    std::vector<int> ids; std::vector<double> pr; orc::Ctx cx;
    for (long i = lo; i < hi; i++) {
      ids.clear(); pr.clear();
      bool has_r = replace && !std::isnan(replace[i]);
      orc::predict_one(*L, X + i * (long)vec_size, present ? present + i * (long)W : nullptr, has_r, has_r ? replace[i] : 0.0,
                       vec_size, score + i, status + i, aux ? aux + i : nullptr, node_ids ? &ids : nullptr, probs ? &pr : nullptr, cx);
      if (node_ids) for (int k = 0; k < node_stride; k++) node_ids[i * (long)node_stride + k] = k < (int)ids.size() ? ids[k] : -1;
      if (probs) for (int k = 0; k < prob_stride; k++) probs[i * (long)prob_stride + k] = k < (int)pr.size() ? pr[k] : std::numeric_limits<double>::quiet_NaN();
    }
  };
  if (nthreads <= 1 || n < 2 * nthreads) { work(0, n); return 0; }
  std::vector<std::thread> th; long chunk = (n + nthreads - 1) / nthreads;
  for (int t = 0; t < nthreads; t++) { long lo = t * chunk, hi = std::min(n, lo + chunk); if (lo < hi) th.emplace_back(work, lo, hi); }
  for (auto& t : th) t.join();
  return 0;
}

}  // extern "C"
