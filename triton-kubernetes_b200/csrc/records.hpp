// records.hpp -- the renderer records: one field table per Go struct the reference hands to
// SetManager / AddCluster / AddNode / SetTerraformBackendConfig (base structs
// create/manager.go:16-27, create/cluster.go:24-43, create/node.go:19-41 and one struct per
// provider in create/{manager,cluster,node}_<provider>.go, backend/local/backend.go:24-26).
// A record is rendered as JSON text in encoding/json's struct order: fields in declaration order,
// an embedded base struct's fields first, `omitempty` dropping "" and empty lists, `json:"-"` gone.
#pragma once
#include <string>
#include <string_view>

#include "doc.hpp"

namespace tk {

enum FieldType { F_STRING, F_STRINGS, F_LABELS };
struct FieldSpec { const char *json; bool omitempty; FieldType type; };
struct RecordSpec { const char *go_name; const FieldSpec *fields; size_t n; };

const RecordSpec *find_record(std::string_view go_name);
const RecordSpec *record_at(size_t i);  // nullptr past the end

// "manager" | "cluster" | "node" x "triton" | "aws" | "gcp" | "azure" | "baremetal" | "vsphere"
// -> the Go struct name (e.g. bareMetalNodeTerraformConfig); empty if there is no such struct
std::string record_struct_name(std::string_view kind, std::string_view provider);

// `<url>//terraform/modules/<dir>?ref=<ref>` (create/manager.go:160-171, cluster.go:299-311,
// node.go:209-221); empty url / ref mean the defaults github.com/joyent/triton-kubernetes, master
bool module_source(std::string_view kind, std::string_view provider, std::string_view url, std::string_view ref, std::string *out);

// settings: a JSON object {json field name: value}.  Absent fields take Go's zero value.
// Returns false with a message for unknown fields or values of the wrong type.
bool render_record(const RecordSpec &spec, const Node *settings, std::string *out, std::string *err);

// the table itself as JSON (for the tests that pin it to the reference's structs)
void record_table_json(const RecordSpec &spec, std::string *out);

}  // namespace tk
