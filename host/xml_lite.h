// xml_lite.h — a small read-only XML DOM, enough for COLLADA 1.4 documents.
// Host-side plumbing for host/dae_loader.cpp (the reference reads the same files through tinyxml2,
// reference src/collada/collada.cpp:131-142).  Elements, attributes, character data, comments, processing instructions,
// CDATA and the five predefined entities; no DTDs, no namespaces.
#ifndef RRT_HOST_XML_LITE_H
#define RRT_HOST_XML_LITE_H

#include <memory>
#include <string>
#include <utility>
#include <vector>

namespace rrt_host {

struct XmlNode {
  std::string name;
  std::vector<std::pair<std::string, std::string>> attrs;
  std::string text;                 // character data that precedes the first child element (tinyxml2's GetText())
  bool has_text = false;
  std::vector<std::unique_ptr<XmlNode>> children;
  XmlNode* parent = nullptr;
  size_t index_in_parent = 0;

  const char* attr(const char* key) const;                    // nullptr when absent
  long int_attr(const char* key, long fallback = 0) const;
  const XmlNode* child(const char* tag = nullptr) const;      // first child element (with that tag)
  const XmlNode* next(const char* tag = nullptr) const;       // next sibling element (with that tag)
  const XmlNode* path(const std::string& slash_separated) const;   // child("a")->child("b")->...
};

struct XmlDocument {
  std::unique_ptr<XmlNode> root;    // synthetic node whose children are the top-level elements
  std::string error;                // empty on success
  bool parse(const std::string& bytes);
  bool load(const std::string& filename);
};

}  // namespace rrt_host

#endif
