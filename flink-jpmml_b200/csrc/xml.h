// xml.h -- minimal non-validating XML reader for PMML documents (tens of MB for large ensembles).
// Replaces the JAXB unmarshalling step of the reference's model load
// (flink-jpmml-scala/src/main/scala/io/radicalbit/flink/pmml/scala/api/PmmlModel.scala:53-61).
// Namespaces are not interpreted: prefixes are dropped, which also makes PMML 3.2/4.0/4.1/4.2 documents readable
// the way the reference's ImportFilter does (it only rewrites the namespace URI to 4.3).
#pragma once
#include <cstdint>
#include <cstring>
#include <deque>
#include <stdexcept>
#include <string>
#include <string_view>
#include <vector>

namespace fjp {

struct XmlError : std::runtime_error { using std::runtime_error::runtime_error; };

struct XAttr { std::string_view key, val; };

struct XNode {
  std::string_view name;
  uint32_t attr_begin = 0, attr_end = 0;  // into XmlDoc::attrs
  XNode* first = nullptr;                 // first child
  XNode* last = nullptr;
  XNode* next = nullptr;                  // next sibling
  std::string_view text;                  // character data (first chunk; later chunks appended in doc.extra)
  const struct XmlDoc* doc = nullptr;
  // scratch for whoever walks the tree (the PMML compiler memoises per-Node results here instead of in hash maps):
  // values are valid for the walk whose epoch matches
  mutable uint32_t scratch_epoch = 0; mutable uint32_t scratch_set = 0; mutable int32_t scratch[4] = {0, 0, 0, 0};

  std::string_view attr(std::string_view k, std::string_view dflt = {}) const;
  bool has(std::string_view k) const;
  const XNode* child(std::string_view n) const {
    for (const XNode* c = first; c; c = c->next) if (c->name == n) return c;
    return nullptr;
  }
  template <class F> void each(std::string_view n, F&& f) const {
    for (const XNode* c = first; c; c = c->next) if (c->name == n) f(*c);
  }
  size_t count(std::string_view n) const { size_t k = 0; each(n, [&](const XNode&) { k++; }); return k; }
};

struct XmlDoc {
  std::string buf;              // mutable copy; entity-decoded in place
  std::deque<XNode> nodes;      // stable addresses
  std::vector<XAttr> attrs;
  std::deque<std::string> extra;  // storage for concatenated multi-chunk text
  XNode* root = nullptr;
};

inline std::string_view XNode::attr(std::string_view k, std::string_view dflt) const {
  for (uint32_t i = attr_begin; i < attr_end; i++) if (doc->attrs[i].key == k) return doc->attrs[i].val;
  return dflt;
}
inline bool XNode::has(std::string_view k) const {
  for (uint32_t i = attr_begin; i < attr_end; i++) if (doc->attrs[i].key == k) return true;
  return false;
}

namespace detail {
inline bool is_space(char c) { return c == ' ' || c == '\n' || c == '\t' || c == '\r'; }
inline bool name_char(char c) { return !(is_space(c) || c == '>' || c == '/' || c == '=' || c == '<' || c == '\0'); }

// decodes entities of [b,e) in place, returns new end
inline char* decode(char* b, char* e) {
  char* amp = static_cast<char*>(memchr(b, '&', e - b));
  if (!amp) return e;
  char* w = amp; char* r = amp;
  while (r < e) {
    if (*r != '&') { *w++ = *r++; continue; }
    char* semi = static_cast<char*>(memchr(r, ';', e - r));
    if (!semi) throw XmlError("unterminated entity");
    std::string_view ent(r + 1, semi - r - 1);
    if (ent == "lt") *w++ = '<';
    else if (ent == "gt") *w++ = '>';
    else if (ent == "amp") *w++ = '&';
    else if (ent == "quot") *w++ = '"';
    else if (ent == "apos") *w++ = '\'';
    else if (!ent.empty() && ent[0] == '#') {
      unsigned long cp = ent.size() > 1 && (ent[1] == 'x' || ent[1] == 'X') ? strtoul(std::string(ent.substr(2)).c_str(), nullptr, 16)
                                                                           : strtoul(std::string(ent.substr(1)).c_str(), nullptr, 10);
      if (cp < 0x80) *w++ = (char)cp;
      else if (cp < 0x800) { *w++ = (char)(0xC0 | (cp >> 6)); *w++ = (char)(0x80 | (cp & 0x3F)); }
      else if (cp < 0x10000) { *w++ = (char)(0xE0 | (cp >> 12)); *w++ = (char)(0x80 | ((cp >> 6) & 0x3F)); *w++ = (char)(0x80 | (cp & 0x3F)); }
      else { *w++ = (char)(0xF0 | (cp >> 18)); *w++ = (char)(0x80 | ((cp >> 12) & 0x3F)); *w++ = (char)(0x80 | ((cp >> 6) & 0x3F)); *w++ = (char)(0x80 | (cp & 0x3F)); }
    } else throw XmlError("unknown entity &" + std::string(ent) + ";");
    r = semi + 1;
  }
  return w;
}
inline std::string_view strip_prefix(std::string_view n) {
  size_t c = n.rfind(':');
  return c == std::string_view::npos ? n : n.substr(c + 1);
}
}  // namespace detail

inline void xml_parse(const char* data, size_t len, XmlDoc& doc) {
  using namespace detail;
  doc.buf.assign(data, len);
  {  // one attribute per '=' at most: reserve once instead of regrowing (and copying) the array for large ensembles
    size_t eq = 0; for (size_t i = 0; i < len; i++) eq += data[i] == '=';   // vectorised by the compiler
    doc.attrs.reserve(eq + 16);
  }
  doc.buf.push_back('\0');
  char* p = doc.buf.data();
  char* end = p + len;
  std::vector<XNode*> stack;
  auto fail = [&](const char* what) {
    size_t line = 1; for (char* q = doc.buf.data(); q < p && q < end; q++) if (*q == '\n') line++;
    throw XmlError(std::string("XML error: ") + what + " at line " + std::to_string(line));
  };
  auto add_text = [&](char* b, char* e) {
    if (stack.empty()) return;
    bool all_space = true;
    for (char* q = b; q < e; q++) if (!is_space(*q)) { all_space = false; break; }
    if (all_space) return;
    XNode* n = stack.back();
    std::string_view chunk(b, e - b);
    if (n->text.empty()) n->text = chunk;
    else { doc.extra.emplace_back(std::string(n->text) + std::string(chunk)); n->text = doc.extra.back(); }
  };
  while (p < end) {
    if (*p != '<') {
      char* lt = static_cast<char*>(memchr(p, '<', end - p));
      char* e = lt ? lt : end;
      char* we = decode(p, e);
      add_text(p, we);
      p = e;
      continue;
    }
    if (p + 3 < end && p[1] == '!' && p[2] == '-' && p[3] == '-') {  // comment
      char* c = strstr(p + 4, "-->");
      if (!c) fail("unterminated comment");
      p = c + 3; continue;
    }
    if (p + 8 < end && strncmp(p, "<![CDATA[", 9) == 0) {
      char* c = strstr(p + 9, "]]>");
      if (!c) fail("unterminated CDATA");
      add_text(p + 9, c); p = c + 3; continue;
    }
    if (p[1] == '?') { char* c = strstr(p + 2, "?>"); if (!c) fail("unterminated processing instruction"); p = c + 2; continue; }
    if (p[1] == '!') {  // DOCTYPE etc.: skip to matching '>'
      int depth = 0; char* q = p + 2;
      for (; q < end; q++) { if (*q == '[') depth++; else if (*q == ']') depth--; else if (*q == '>' && depth <= 0) break; }
      if (q >= end) fail("unterminated declaration");
      p = q + 1; continue;
    }
    if (p[1] == '/') {  // end tag
      char* q = p + 2; while (q < end && name_char(*q)) q++;
      std::string_view nm = strip_prefix(std::string_view(p + 2, q - p - 2));
      if (stack.empty() || stack.back()->name != nm) fail("mismatched end tag");
      while (q < end && *q != '>') q++;
      if (q >= end) fail("unterminated end tag");
      stack.pop_back(); p = q + 1; continue;
    }
    // start tag
    char* q = p + 1; while (q < end && name_char(*q)) q++;
    if (q == p + 1) fail("bad tag");
    doc.nodes.emplace_back();
    XNode* n = &doc.nodes.back();
    n->doc = &doc; n->name = strip_prefix(std::string_view(p + 1, q - p - 1));
    n->attr_begin = (uint32_t)doc.attrs.size();
    bool self_close = false;
    for (;;) {
      while (q < end && is_space(*q)) q++;
      if (q >= end) fail("unterminated start tag");
      if (*q == '>') { q++; break; }
      if (*q == '/') { if (q + 1 < end && q[1] == '>') { self_close = true; q += 2; break; } fail("bad '/' in tag"); }
      char* kb = q; while (q < end && name_char(*q)) q++;
      std::string_view key = strip_prefix(std::string_view(kb, q - kb));
      while (q < end && is_space(*q)) q++;
      if (q >= end || *q != '=') fail("attribute without value");
      q++; while (q < end && is_space(*q)) q++;
      if (q >= end || (*q != '"' && *q != '\'')) fail("unquoted attribute value");
      char quote = *q++; char* vb = q;
      char* ve = static_cast<char*>(memchr(q, quote, end - q));
      if (!ve) fail("unterminated attribute value");
      char* we = decode(vb, ve);
      doc.attrs.push_back({key, std::string_view(vb, we - vb)});
      q = ve + 1;
    }
    n->attr_end = (uint32_t)doc.attrs.size();
    if (stack.empty()) { if (doc.root) fail("multiple root elements"); doc.root = n; }
    else { XNode* par = stack.back(); if (par->last) par->last->next = n; else par->first = n; par->last = n; }
    if (!self_close) stack.push_back(n);
    p = q;
  }
  if (!stack.empty()) { p = end; fail("unexpected end of document"); }
  if (!doc.root) { p = end; fail("no root element"); }
}

}  // namespace fjp
