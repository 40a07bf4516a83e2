// doc.hpp -- host side of the document: the JSON tree, gabs path operations and the value parser.
//
// This is the host logic that sits above the kernels: pointer-chasing edits of <= 200 bytes each
//  that have no device shape.  The bulk passes -- parse of a whole
// document, serialisation, key-prefix classification -- are CUDA (tape.cu, emit.cu, prefix.cu).
//
// Reference behaviour restated (paths relative to the reference root):
//   vendor/github.com/Jeffail/gabs/gabs.go:81-114   Search (array fan-out)
//   vendor/github.com/Jeffail/gabs/gabs.go:180-208  Set/SetP
//   vendor/github.com/Jeffail/gabs/gabs.go:276-301  Delete/DeleteP
//   state/state.go:149-160                          getClusterKeyParts
//   create/node.go:350-380                          getNewHostnames
#pragma once
#include <cstddef>
#include <cstdint>
#include <map>
#include <string>
#include <string_view>
#include <utility>
#include <vector>

namespace tk {

enum class Kind : uint8_t { Bool, Num, Str, Arr, Map, Rec };

// A null pointer is JSON null / Go's nil interface.
struct Node {
    Kind kind;
    explicit Node(Kind k) : kind(k) {}
};
struct BoolNode : Node { bool v; explicit BoolNode(bool b) : Node(Kind::Bool), v(b) {} };
struct NumNode : Node { double v; explicit NumNode(double d) : Node(Kind::Num), v(d) {} };
struct StrNode : Node { std::string s; explicit StrNode(std::string x) : Node(Kind::Str), s(std::move(x)) {} };
struct ArrNode : Node { std::vector<Node *> items; ArrNode() : Node(Kind::Arr) {} };
// Go map[string]interface{}: unique keys; kept in bytewise key order, which is the order
// encoding/json writes them in, so serialisation and prefix queries walk it linearly.
struct MapNode : Node { std::map<std::string, Node *, std::less<>> m; MapNode() : Node(Kind::Map) {} };
// A Go struct value stored with SetP: members in declaration order, opaque to Search.
struct RecNode : Node { std::vector<std::pair<std::string, Node *>> m; RecNode() : Node(Kind::Rec) {} };

void free_node(Node *n);

// err.Error() texts the reference's tests compare against
extern const char *const kErrCollision;  // gabs.ErrPathCollision
extern const char *const kErrNotObj;     // gabs.ErrNotObj

// json.Unmarshal(text, &interface{}) on the host, for VALUE arguments (kilobyte scale).
// record_mode: objects become RecNode (order kept).  Returns false and fills err with the Go
// error text on failure.
bool parse_value(std::string_view text, bool record_mode, Node **out, std::string *err);

// Go's error text for invalid JSON (scanner.go); used by the device parser's error path to
// phrase a failure it has already detected.  Returns false if text is valid.
bool syntax_error_text(std::string_view text, std::string *err);

// encoding/json floatEncoder (64 bit) -- ES6-style shortest text
void format_go_float(double v, std::string *out);
// strtod in the "C" locale, whatever LC_NUMERIC the host program has set
double strtod_c(const char *s);

std::vector<std::string_view> split(std::string_view s, char sep);

struct Document {
    Node *root = nullptr;  // nullptr == nil
    uint64_t version = 1;  // bumped by every successful or partially applied mutation
    // false while the document exists only as the device tape state.New produced; the point
    // operations build the tree from that tape on first use (Engine::materialise)
    bool host_valid = true;
    ~Document() { free_node(root); }

    // State.Get: string at path or empty view
    std::string_view get(std::string_view path) const;
    // SetP; takes ownership of value on success, frees it on failure
    int set(std::string_view path, Node *value, std::string *err);
    int del(std::string_view path, std::string *err);
    // S("module").ChildrenMap(): nullptr if "module" is not a map
    const MapNode *module_children() const;
};

// getClusterKeyParts
bool cluster_key_parts(std::string_view cluster_key, std::string_view *provider, std::string_view *name,
                       std::string *err);

// getNewHostnames
std::vector<std::string> new_hostnames(const std::vector<std::string_view> &existing, std::string_view node_name,
                                       long long nodes_to_add);

}  // namespace tk
