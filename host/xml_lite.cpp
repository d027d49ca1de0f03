// xml_lite.cpp — see xml_lite.h.
#include "xml_lite.h"

#include <stdlib.h>
#include <string.h>

#include <fstream>
#include <sstream>

namespace rrt_host {

const char* XmlNode::attr(const char* key) const {
  for (const auto& a : attrs)
    if (a.first == key) return a.second.c_str();
  return nullptr;
}

long XmlNode::int_attr(const char* key, long fallback) const {
  const char* v = attr(key);
  return v ? strtol(v, nullptr, 10) : fallback;
}

const XmlNode* XmlNode::child(const char* tag) const {
  for (const auto& c : children)
    if (!tag || c->name == tag) return c.get();
  return nullptr;
}

const XmlNode* XmlNode::next(const char* tag) const {
  if (!parent) return nullptr;
  for (size_t i = index_in_parent + 1; i < parent->children.size(); ++i)
    if (!tag || parent->children[i]->name == tag) return parent->children[i].get();
  return nullptr;
}

const XmlNode* XmlNode::path(const std::string& q) const {
  const XmlNode* e = this;
  size_t pos = 0;
  while (e && pos <= q.size()) {
    size_t slash = q.find('/', pos);
    std::string token = q.substr(pos, slash == std::string::npos ? std::string::npos : slash - pos);
    e = e->child(token.c_str());
    if (slash == std::string::npos) break;
    pos = slash + 1;
  }
  return e;
}

namespace {

struct Parser {
  const std::string& s;
  size_t i = 0;
  std::string err;
  explicit Parser(const std::string& src) : s(src) {}

  bool starts(const char* lit) const { return s.compare(i, strlen(lit), lit) == 0; }
  void skip_ws() { while (i < s.size() && (s[i] == ' ' || s[i] == '\t' || s[i] == '\n' || s[i] == '\r')) ++i; }
  bool skip_until(const char* lit) {
    size_t p = s.find(lit, i);
    if (p == std::string::npos) { err = std::string("unterminated construct, expected ") + lit; return false; }
    i = p + strlen(lit);
    return true;
  }
  static bool name_char(char c) {
    return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || (c >= '0' && c <= '9') || c == '_' || c == '-' || c == '.' || c == ':';
  }
  std::string name() {
    size_t b = i;
    while (i < s.size() && name_char(s[i])) ++i;
    return s.substr(b, i - b);
  }
  static void append_decoded(std::string& out, const std::string& raw) {
    for (size_t k = 0; k < raw.size(); ++k) {
      if (raw[k] != '&') { out += raw[k]; continue; }
      size_t semi = raw.find(';', k);
      if (semi == std::string::npos) { out += raw[k]; continue; }
      std::string ent = raw.substr(k + 1, semi - k - 1);
      if (ent == "amp") out += '&';
      else if (ent == "lt") out += '<';
      else if (ent == "gt") out += '>';
      else if (ent == "quot") out += '"';
      else if (ent == "apos") out += '\'';
      else if (!ent.empty() && ent[0] == '#') {
        long code = ent.size() > 1 && (ent[1] == 'x' || ent[1] == 'X') ? strtol(ent.c_str() + 2, nullptr, 16) : strtol(ent.c_str() + 1, nullptr, 10);
        if (code > 0 && code < 128) out += (char)code;
      } else { out += raw.substr(k, semi - k + 1); }
      k = semi;
    }
  }

  // parses the content of `node` up to and including its end tag (or end of input for the synthetic root)
  // depth: elements open above this one.  Nesting is bounded (COLLADA needs a dozen levels): the parser, the tree's destructor and every walk of
  // the tree recurse per level, and a file of 300 000 nested tags would otherwise end in a stack overflow instead of a message
  bool content(XmlNode* node, bool is_root, int depth = 0) {
    if (depth > 512) { err = "elements nested more than 512 deep"; return false; }
    bool seen_child = false;
    while (i < s.size()) {
      if (s[i] != '<') {
        size_t lt = s.find('<', i);
        if (lt == std::string::npos) lt = s.size();
        if (!seen_child && !is_root) { append_decoded(node->text, s.substr(i, lt - i)); node->has_text = true; }
        i = lt;
        continue;
      }
      if (starts("<!--")) { if (!skip_until("-->")) return false; continue; }
      if (starts("<![CDATA[")) {
        size_t b = i + 9;
        if (!skip_until("]]>")) return false;
        if (!seen_child && !is_root) { node->text += s.substr(b, i - 3 - b); node->has_text = true; }
        continue;
      }
      if (starts("<?")) { if (!skip_until("?>")) return false; continue; }
      if (starts("<!")) { if (!skip_until(">")) return false; continue; }
      if (starts("</")) {
        i += 2;
        std::string n = name();
        skip_ws();
        if (i >= s.size() || s[i] != '>') { err = "malformed end tag </" + n; return false; }
        ++i;
        if (is_root || n != node->name) { err = "mismatched end tag </" + n + ">"; return false; }
        return true;
      }
      // start tag
      ++i;
      std::unique_ptr<XmlNode> c(new XmlNode());
      c->name = name();
      if (c->name.empty()) { err = "malformed start tag"; return false; }
      c->parent = node; c->index_in_parent = node->children.size();
      bool self_closed = false;
      for (;;) {
        skip_ws();
        if (i >= s.size()) { err = "unterminated start tag <" + c->name; return false; }
        if (s[i] == '>') { ++i; break; }
        if (s[i] == '/' && i + 1 < s.size() && s[i + 1] == '>') { i += 2; self_closed = true; break; }
        std::string key = name();
        skip_ws();
        if (key.empty() || i >= s.size() || s[i] != '=') { err = "malformed attribute in <" + c->name; return false; }
        ++i; skip_ws();
        if (i >= s.size() || (s[i] != '"' && s[i] != '\'')) { err = "unquoted attribute value in <" + c->name; return false; }
        char q = s[i++];
        size_t e = s.find(q, i);
        if (e == std::string::npos) { err = "unterminated attribute value in <" + c->name; return false; }
        std::string val;
        append_decoded(val, s.substr(i, e - i));
        c->attrs.emplace_back(key, val);
        i = e + 1;
      }
      XmlNode* raw = c.get();
      node->children.push_back(std::move(c));
      seen_child = true;
      if (!self_closed && !content(raw, false, depth + 1)) return false;
    }
    if (!is_root) { err = "missing end tag </" + node->name + ">"; return false; }
    return true;
  }
};

}  // namespace

bool XmlDocument::parse(const std::string& bytes) {
  root.reset(new XmlNode());
  error.clear();
  Parser p(bytes);
  if (bytes.size() >= 3 && (unsigned char)bytes[0] == 0xEF && (unsigned char)bytes[1] == 0xBB && (unsigned char)bytes[2] == 0xBF) p.i = 3;
  if (!p.content(root.get(), true)) { error = p.err + " (near byte " + std::to_string(p.i) + ")"; return false; }
  if (root->children.empty()) { error = "no element in document"; return false; }
  return true;
}

bool XmlDocument::load(const std::string& filename) {
  std::ifstream in(filename, std::ios::binary);
  if (!in.is_open()) { error = "cannot open " + filename; return false; }
  std::stringstream ss;
  ss << in.rdbuf();
  return parse(ss.str());
}

}  // namespace rrt_host
