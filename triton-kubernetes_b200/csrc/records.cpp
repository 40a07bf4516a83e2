// records.cpp -- renderer records (see records.hpp).
#include "records.hpp"

#include <cstdio>
#include <cstring>

namespace tk {

#include "records_table.inc"

const RecordSpec *find_record(std::string_view go_name) {
    for (const RecordSpec &r : kRecords)
        if (go_name == r.go_name) return &r;
    return nullptr;
}
const RecordSpec *record_at(size_t i) { return i < sizeof kRecords / sizeof kRecords[0] ? &kRecords[i] : nullptr; }

namespace {
struct Provider { const char *id, *go, *dir; };
const Provider kProviders[] = {{"triton", "triton", "triton"}, {"aws", "aws", "aws"}, {"gcp", "gcp", "gcp"},
                               {"azure", "azure", "azure"}, {"baremetal", "bareMetal", "bare-metal"}, {"vsphere", "vSphere", "vsphere"}};
const Provider *find_provider(std::string_view id) {
    for (const Provider &p : kProviders)
        if (id == p.id) return &p;
    return nullptr;
}

// transport escaping only: the text is parsed again by the value parser
void put_json_string(std::string *out, std::string_view s) {
    out->push_back('"');
    for (unsigned char c : s) {
        if (c == '"' || c == '\\') { out->push_back('\\'); out->push_back((char)c); }
        else if (c < 0x20) { char t[8]; snprintf(t, sizeof t, "\\u%04x", c); out->append(t); }
        else out->push_back((char)c);
    }
    out->push_back('"');
}
}  // namespace

std::string record_struct_name(std::string_view kind, std::string_view provider) {
    const Provider *p = find_provider(provider);
    if (!p) return {};
    std::string n = p->go;
    if (kind == "manager") n += "ManagerTerraformConfig";
    else if (kind == "cluster") n += "ClusterTerraformConfig";
    else if (kind == "node") n += "NodeTerraformConfig";
    else return {};
    return find_record(n) ? n : std::string();
}

bool module_source(std::string_view kind, std::string_view provider, std::string_view url, std::string_view ref, std::string *out) {
    const Provider *p = find_provider(provider);
    if (!p || record_struct_name(kind, provider).empty()) return false;
    out->assign(url.empty() ? std::string_view("github.com/joyent/triton-kubernetes") : url);
    out->append("//terraform/modules/");
    out->append(p->dir);
    out->append(kind == "manager" ? "-rancher" : kind == "cluster" ? "-rancher-k8s" : "-rancher-k8s-host");
    out->append("?ref=");
    out->append(ref.empty() ? std::string_view("master") : ref);
    return true;
}

bool render_record(const RecordSpec &spec, const Node *settings, std::string *out, std::string *err) {
    static const MapNode kEmpty;
    const MapNode *m = &kEmpty;
    if (settings) {
        if (settings->kind != Kind::Map) { *err = std::string(spec.go_name) + ": settings must be a JSON object"; return false; }
        m = static_cast<const MapNode *>(settings);
    }
    for (auto &kv : m->m) {
        bool known = false;
        for (size_t i = 0; i < spec.n && !known; i++) known = kv.first == spec.fields[i].json;
        if (!known) { *err = std::string(spec.go_name) + " has no field \"" + kv.first + "\""; return false; }
    }
    out->push_back('{');
    bool first = true;
    for (size_t i = 0; i < spec.n; i++) {
        const FieldSpec &f = spec.fields[i];
        auto it = m->m.find(std::string_view(f.json));
        const Node *v = it == m->m.end() ? nullptr : it->second;
        std::string text;
        switch (f.type) {
        case F_STRING: {
            if (v && v->kind != Kind::Str) { *err = std::string(spec.go_name) + "." + f.json + " must be a string"; return false; }
            std::string_view s = v ? std::string_view(static_cast<const StrNode *>(v)->s) : std::string_view();
            if (s.empty() && f.omitempty) continue;
            put_json_string(&text, s);
            break;
        }
        case F_STRINGS: {
            if (v && v->kind != Kind::Arr) { *err = std::string(spec.go_name) + "." + f.json + " must be a list of strings"; return false; }
            const ArrNode *a = static_cast<const ArrNode *>(v);
            if ((!a || a->items.empty()) && f.omitempty) continue;
            if (!a) { text = "null"; break; }  // nil slice
            text.push_back('[');
            for (size_t k = 0; k < a->items.size(); k++) {
                const Node *e = a->items[k];
                if (!e || e->kind != Kind::Str) { *err = std::string(spec.go_name) + "." + f.json + " must be a list of strings"; return false; }
                if (k) text.push_back(',');
                put_json_string(&text, static_cast<const StrNode *>(e)->s);
            }
            text.push_back(']');
            break;
        }
        case F_LABELS:  // a nested struct value: never dropped, its own fields are omitempty
            if (!render_record(*find_record("rancherHostLabelsConfig"), v, &text, err)) return false;
            break;
        }
        if (!first) out->push_back(',');
        first = false;
        put_json_string(out, f.json);
        out->push_back(':');
        out->append(text);
    }
    out->push_back('}');
    return true;
}

void record_table_json(const RecordSpec &spec, std::string *out) {
    static const char *const types[] = {"string", "[]string", "rancherHostLabelsConfig"};
    out->append("[");
    for (size_t i = 0; i < spec.n; i++) {
        if (i) out->push_back(',');
        out->append("{\"json\":\"").append(spec.fields[i].json).append("\",\"omitempty\":").append(spec.fields[i].omitempty ? "true" : "false");
        out->append(",\"type\":\"").append(types[spec.fields[i].type]).append("\"}");
    }
    out->append("]");
}

}  // namespace tk
