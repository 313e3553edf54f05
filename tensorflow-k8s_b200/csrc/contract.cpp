// contract.cpp -- TFJob replica specs in, TF_CONFIG per task out (include/b200ps.h "Contract group").
//
// Stands in for what the reference only names (images/tf3.PNG): pkg/apis/tensorflow/v1alpha1/
// {types,defaults}.go, pkg/apis/tensorflow/validation/validation.go and pkg/trainer/replicas.go.
// The role model (cluster / job / task, PS vs WORKER) is k8s-operator.md:6.  The JSON shapes are
// the upstream operator's as recalled in SURVEY.md CS4 -- they are not in the mounted reference.
#include <cctype>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "common.h"

namespace {

// ------------------------------------------------------------------ a small JSON reader
struct JValue;
using JObject = std::vector<std::pair<std::string, JValue>>;
using JArray = std::vector<JValue>;

struct JValue {
    enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
    bool b = false;
    double num = 0;
    std::string str;
    std::shared_ptr<JArray> arr;
    std::shared_ptr<JObject> obj;

    const JValue* get(const char* key) const {
        if (kind != Object) return nullptr;
        for (auto& kv : *obj)
            if (kv.first == key) return &kv.second;
        return nullptr;
    }
};

class JParser {
   public:
    explicit JParser(const char* s) : p_(s), begin_(s) {}
    bool parse(JValue& out, std::string& err) {
        if (!value(out, 0) || (ws(), *p_ != '\0')) {
            if (err_.empty()) err_ = "trailing characters";
            err = err_ + " at byte " + std::to_string(p_ - begin_);
            return false;
        }
        return true;
    }

   private:
    const char* p_;
    const char* begin_;
    std::string err_;

    void ws() {
        while (*p_ == ' ' || *p_ == '\t' || *p_ == '\n' || *p_ == '\r') ++p_;
    }
    bool bad(const char* m) {
        if (err_.empty()) err_ = m;
        return false;
    }
    bool lit(const char* w) {
        size_t n = strlen(w);
        if (strncmp(p_, w, n) != 0) return bad("bad literal");
        p_ += n;
        return true;
    }
    static void utf8(std::string& s, unsigned cp) {
        if (cp < 0x80) {
            s += (char)cp;
        } else if (cp < 0x800) {
            s += (char)(0xC0 | (cp >> 6));
            s += (char)(0x80 | (cp & 0x3F));
        } else if (cp < 0x10000) {
            s += (char)(0xE0 | (cp >> 12));
            s += (char)(0x80 | ((cp >> 6) & 0x3F));
            s += (char)(0x80 | (cp & 0x3F));
        } else {
            s += (char)(0xF0 | (cp >> 18));
            s += (char)(0x80 | ((cp >> 12) & 0x3F));
            s += (char)(0x80 | ((cp >> 6) & 0x3F));
            s += (char)(0x80 | (cp & 0x3F));
        }
    }
    bool hex4(unsigned& v) {
        v = 0;
        for (int i = 0; i < 4; ++i) {
            char c = *p_++;
            v <<= 4;
            if (c >= '0' && c <= '9') v |= (unsigned)(c - '0');
            else if (c >= 'a' && c <= 'f') v |= (unsigned)(c - 'a' + 10);
            else if (c >= 'A' && c <= 'F') v |= (unsigned)(c - 'A' + 10);
            else return bad("bad \\u escape");
        }
        return true;
    }
    bool string(std::string& s) {
        if (*p_ != '"') return bad("expected string");
        ++p_;
        while (*p_ != '"') {
            unsigned char c = (unsigned char)*p_;
            if (c == 0) return bad("unterminated string");
            if (c < 0x20) return bad("control character in string");
            if (c != '\\') {
                s += (char)c;
                ++p_;
                continue;
            }
            ++p_;
            switch (*p_++) {
                case '"': s += '"'; break;
                case '\\': s += '\\'; break;
                case '/': s += '/'; break;
                case 'b': s += '\b'; break;
                case 'f': s += '\f'; break;
                case 'n': s += '\n'; break;
                case 'r': s += '\r'; break;
                case 't': s += '\t'; break;
                case 'u': {
                    unsigned cp, lo;
                    if (!hex4(cp)) return false;
                    if (cp >= 0xD800 && cp < 0xDC00) {  // a high surrogate must be followed by \uDC00..\uDFFF
                        if (!(p_[0] == '\\' && p_[1] == 'u')) return bad("lone surrogate in \\u escape");
                        p_ += 2;
                        if (!hex4(lo)) return false;
                        if (lo < 0xDC00 || lo > 0xDFFF) return bad("mismatched surrogate pair in \\u escape");
                        cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
                    } else if (cp >= 0xDC00 && cp <= 0xDFFF) {
                        return bad("lone surrogate in \\u escape");
                    }
                    utf8(s, cp);
                    break;
                }
                default: return bad("bad escape");
            }
        }
        ++p_;
        return true;
    }
    bool value(JValue& v, int depth) {
        if (depth > 64) return bad("nesting too deep");
        ws();
        switch (*p_) {
            case '{': {
                ++p_;
                v.kind = JValue::Object;
                v.obj = std::make_shared<JObject>();
                ws();
                if (*p_ == '}') return ++p_, true;
                for (;;) {
                    std::string k;
                    ws();
                    if (!string(k)) return false;
                    ws();
                    if (*p_++ != ':') return bad("expected ':'");
                    JValue c;
                    if (!value(c, depth + 1)) return false;
                    v.obj->emplace_back(std::move(k), std::move(c));
                    ws();
                    if (*p_ == ',') { ++p_; continue; }
                    if (*p_ == '}') return ++p_, true;
                    return bad("expected ',' or '}'");
                }
            }
            case '[': {
                ++p_;
                v.kind = JValue::Array;
                v.arr = std::make_shared<JArray>();
                ws();
                if (*p_ == ']') return ++p_, true;
                for (;;) {
                    JValue c;
                    if (!value(c, depth + 1)) return false;
                    v.arr->push_back(std::move(c));
                    ws();
                    if (*p_ == ',') { ++p_; continue; }
                    if (*p_ == ']') return ++p_, true;
                    return bad("expected ',' or ']'");
                }
            }
            case '"': v.kind = JValue::String; return string(v.str);
            case 't': v.kind = JValue::Bool; v.b = true; return lit("true");
            case 'f': v.kind = JValue::Bool; v.b = false; return lit("false");
            case 'n': v.kind = JValue::Null; return lit("null");
            default: {
                // the JSON number grammar first (strtod alone would also take 0x10, inf, nan, a leading '+'):
                //   -? (0 | [1-9][0-9]*) (. [0-9]+)? ([eE] [+-]? [0-9]+)?
                const char* q = p_;
                if (*q == '-') ++q;
                if (*q == '0') {
                    ++q;
                } else if (*q >= '1' && *q <= '9') {
                    while (isdigit((unsigned char)*q)) ++q;
                } else {
                    return bad(p_ == q ? "unexpected character" : "bad number");
                }
                if (*q == '.') {
                    ++q;
                    if (!isdigit((unsigned char)*q)) return bad("bad number");
                    while (isdigit((unsigned char)*q)) ++q;
                }
                if (*q == 'e' || *q == 'E') {
                    ++q;
                    if (*q == '+' || *q == '-') ++q;
                    if (!isdigit((unsigned char)*q)) return bad("bad number");
                    while (isdigit((unsigned char)*q)) ++q;
                }
                char* end = nullptr;
                v.kind = JValue::Number;
                v.num = strtod(p_, &end);
                if (end != q) return bad("bad number");
                p_ = q;
                return true;
            }
        }
    }
};

// Go's encoding/json string escaping (HTML-safe form, which json.Marshal uses by default)
void put_json_string(std::string& out, const std::string& s) {
    static const char* hex = "0123456789abcdef";
    out += '"';
    for (unsigned char c : s) {
        if (c == '"' || c == '\\') {
            out += '\\';
            out += (char)c;
        } else if (c == '\n') {
            out += "\\n";
        } else if (c == '\r') {
            out += "\\r";
        } else if (c == '\t') {
            out += "\\t";
        } else if (c < 0x20 || c == '<' || c == '>' || c == '&') {
            out += "\\u00";
            out += hex[c >> 4];
            out += hex[c & 15];
        } else {
            out += (char)c;
        }
    }
    out += '"';
}

std::string lower(std::string s) {
    for (auto& c : s) c = (char)tolower((unsigned char)c);
    return s;
}

bool is_int(const JValue& v, long lo, long hi, long& out) {
    if (v.kind != JValue::Number) return false;
    if (!(v.num >= (double)lo && v.num <= (double)hi)) return false;  // also rejects inf / nan before the cast (UB otherwise)
    long x = (long)v.num;
    if ((double)x != v.num) return false;
    out = x;
    return true;
}

bool dns_name(const std::string& s) {
    if (s.empty() || s.size() > 253) return false;
    for (unsigned char c : s)
        if (!(islower(c) || isdigit(c) || c == '-' || c == '.')) return false;
    return isalnum((unsigned char)s.front()) && isalnum((unsigned char)s.back());
}

// 4 characters, like the operator's random RuntimeId, but a pure function of the job name so the
// exact-string fixtures are reproducible
std::string derived_runtime_id(const std::string& name) {
    uint32_t h = 2166136261u;
    for (unsigned char c : name) h = (h ^ c) * 16777619u;
    static const char* abc = "abcdefghijklmnopqrstuvwxyz0123456789";
    std::string r;
    for (int i = 0; i < 4; ++i, h /= 36) r += abc[h % 36];
    return r;
}

bool compute_job(const std::string& j) { return j == "worker" || j == "master" || j == "chief"; }

}  // namespace

struct b200ps_cluster {
    // job name ("chief" | "master" | "ps" | "worker") -> task addresses; std::map keeps the keys
    // sorted, which is the order Go's json.Marshal gives a map
    std::map<std::string, std::vector<std::string>> jobs;

    int count(const std::string& j) const {
        auto it = jobs.find(j);
        return it == jobs.end() ? 0 : (int)it->second.size();
    }
};

namespace {

struct ReplicaSpec {
    std::string job;  // lower-cased replica type
    long replicas = 1;
    long port = 2222;
};

int read_replica(const JValue& spec, const std::string& type_hint, ReplicaSpec& r) {
    using namespace b200ps;
    B200PS_REQUIRE(spec.kind == JValue::Object, B200PS_ERR_VALIDATION, "replica spec must be an object");
    std::string type = type_hint;
    if (const JValue* t = spec.get("tfReplicaType")) {
        B200PS_REQUIRE(t->kind == JValue::String, B200PS_ERR_VALIDATION, "tfReplicaType must be a string");
        type = t->str;
    }
    if (type.empty()) type = "MASTER";  // v1alpha1 default replica type
    r.job = lower(type);
    B200PS_REQUIRE(r.job == "master" || r.job == "chief" || r.job == "ps" || r.job == "worker",
                   B200PS_ERR_VALIDATION,
                   "tfReplicaSpec.TFReplicaType is %s but must be one of MASTER, PS, WORKER (or Chief)", type.c_str());
    if (const JValue* n = spec.get("replicas"))
        if (n->kind != JValue::Null)
            B200PS_REQUIRE(is_int(*n, 1, 1 << 20, r.replicas), B200PS_ERR_VALIDATION,
                           "%s: replicas must be an integer >= 1", type.c_str());
    if (const JValue* p = spec.get("tfPort"))
        if (p->kind != JValue::Null)
            B200PS_REQUIRE(is_int(*p, 1, 65535, r.port), B200PS_ERR_VALIDATION,
                           "%s: tfPort must be in 1..65535", type.c_str());
    if (r.job == "master" || r.job == "chief")
        B200PS_REQUIRE(r.replicas == 1, B200PS_ERR_VALIDATION, "the %s must have replicas = 1", type.c_str());
    return B200PS_OK;
}

}  // namespace

extern "C" int b200ps_cluster_from_tfjob(const char* tfjob_json, b200ps_cluster_t** out) {
    using namespace b200ps;
    B200PS_REQUIRE(tfjob_json && out, B200PS_ERR_INVALID, "null argument");
    JValue root;
    std::string err;
    B200PS_REQUIRE(JParser(tfjob_json).parse(root, err), B200PS_ERR_PARSE, "TFJob JSON: %s", err.c_str());
    B200PS_REQUIRE(root.kind == JValue::Object, B200PS_ERR_PARSE, "TFJob JSON: top level must be an object");

    std::string name = "tfjob";
    if (const JValue* md = root.get("metadata"))
        if (const JValue* n = md->get("name"))
            if (n->kind == JValue::String) name = n->str;
    B200PS_REQUIRE(dns_name(name), B200PS_ERR_VALIDATION, "metadata.name '%s' is not a DNS-1123 name", name.c_str());

    const JValue* spec = root.get("spec");
    B200PS_REQUIRE(spec && spec->kind == JValue::Object, B200PS_ERR_VALIDATION, "TFJob has no spec");

    std::vector<ReplicaSpec> specs;
    bool list_form = false;
    if (const JValue* rs = spec->get("replicaSpecs")) {  // v1alpha1
        B200PS_REQUIRE(rs->kind == JValue::Array, B200PS_ERR_VALIDATION, "spec.replicaSpecs must be a list");
        list_form = true;
        for (auto& e : *rs->arr) {
            ReplicaSpec r;
            if (int rc = read_replica(e, "", r)) return rc;
            specs.push_back(r);
        }
    } else if (const JValue* ms = spec->get("tfReplicaSpecs")) {  // later map form
        B200PS_REQUIRE(ms->kind == JValue::Object, B200PS_ERR_VALIDATION, "spec.tfReplicaSpecs must be a map");
        for (auto& kv : *ms->obj) {
            ReplicaSpec r;
            if (int rc = read_replica(kv.second, kv.first, r)) return rc;
            specs.push_back(r);
        }
    } else {
        return fail(B200PS_ERR_VALIDATION, "spec has neither replicaSpecs nor tfReplicaSpecs");
    }
    B200PS_REQUIRE(!specs.empty(), B200PS_ERR_VALIDATION, "TFJob has no replica specs");

    std::string rid;
    if (const JValue* r = spec->get("runtimeId"))
        if (r->kind == JValue::String) rid = r->str;
    if (rid.empty()) rid = derived_runtime_id(name);
    B200PS_REQUIRE(dns_name(rid), B200PS_ERR_VALIDATION, "spec.runtimeId '%s' is not a DNS-1123 label", rid.c_str());

    auto c = std::make_unique<b200ps_cluster>();
    for (auto& r : specs) {
        B200PS_REQUIRE(!c->jobs.count(r.job), B200PS_ERR_VALIDATION, "replica type %s appears twice", r.job.c_str());
        auto& hosts = c->jobs[r.job];
        for (long i = 0; i < r.replicas; ++i) {
            // v1alpha1: "<name, at most 40 chars>-<type>-<runtimeId>-<index>:<port>";
            // the map form has no runtime id: "<name>-<type>-<index>:<port>"
            std::string h = list_form ? name.substr(0, 40) + "-" + r.job + "-" + rid + "-" + std::to_string(i)
                                      : name + "-" + r.job + "-" + std::to_string(i);
            hosts.push_back(h + ":" + std::to_string(r.port));
        }
    }
    B200PS_REQUIRE(!(c->count("master") && c->count("chief")), B200PS_ERR_VALIDATION,
                   "a TFJob has a Master or a Chief, not both");
    B200PS_REQUIRE(c->count("master") + c->count("chief") + c->count("worker") >= 1, B200PS_ERR_VALIDATION,
                   "a TFJob needs at least one MASTER/Chief or WORKER replica");
    *out = c.release();
    return B200PS_OK;
}

extern "C" int b200ps_cluster_from_tf_config(const char* tf_config_json, b200ps_cluster_t** out, char* task_type,
                                             size_t task_type_cap, int* task_index) {
    using namespace b200ps;
    B200PS_REQUIRE(tf_config_json && out, B200PS_ERR_INVALID, "null argument");
    JValue root;
    std::string err;
    B200PS_REQUIRE(JParser(tf_config_json).parse(root, err), B200PS_ERR_PARSE, "TF_CONFIG: %s", err.c_str());
    const JValue* cl = root.get("cluster");
    B200PS_REQUIRE(cl && cl->kind == JValue::Object, B200PS_ERR_PARSE, "TF_CONFIG has no cluster object");
    auto c = std::make_unique<b200ps_cluster>();
    for (auto& kv : *cl->obj) {
        B200PS_REQUIRE(kv.second.kind == JValue::Array, B200PS_ERR_PARSE, "TF_CONFIG cluster.%s must be a list",
                       kv.first.c_str());
        B200PS_REQUIRE(!c->jobs.count(kv.first), B200PS_ERR_PARSE, "TF_CONFIG cluster.%s appears twice",
                       kv.first.c_str());
        auto& hosts = c->jobs[kv.first];
        for (auto& h : *kv.second.arr) {
            B200PS_REQUIRE(h.kind == JValue::String, B200PS_ERR_PARSE, "TF_CONFIG cluster.%s holds a non-string",
                           kv.first.c_str());
            hosts.push_back(h.str);
        }
    }
    if (const JValue* task = root.get("task")) {
        const JValue* ty = task->get("type");
        const JValue* ix = task->get("index");
        long idx = 0;
        B200PS_REQUIRE(ty && ty->kind == JValue::String, B200PS_ERR_PARSE, "TF_CONFIG task.type missing");
        B200PS_REQUIRE(ix && is_int(*ix, 0, 1 << 20, idx), B200PS_ERR_PARSE, "TF_CONFIG task.index missing");
        B200PS_REQUIRE(idx < c->count(ty->str), B200PS_ERR_VALIDATION,
                       "TF_CONFIG task %s:%ld is not in the cluster spec", ty->str.c_str(), idx);
        if (task_type) {
            B200PS_REQUIRE(ty->str.size() < task_type_cap, B200PS_ERR_RANGE, "task_type buffer too small");
            memcpy(task_type, ty->str.c_str(), ty->str.size() + 1);
        }
        if (task_index) *task_index = (int)idx;
    } else {
        B200PS_REQUIRE(!task_type && !task_index, B200PS_ERR_PARSE, "TF_CONFIG has no task object");
    }
    *out = c.release();
    return B200PS_OK;
}

extern "C" void b200ps_cluster_destroy(b200ps_cluster_t* c) { delete c; }

extern "C" int b200ps_cluster_num_tasks(const b200ps_cluster_t* c, const char* job) {
    using namespace b200ps;
    B200PS_REQUIRE(c && job, B200PS_ERR_INVALID, "null argument");
    return c->count(job);
}

static int copy_out(const std::string& s, char* buf, size_t cap) {
    using namespace b200ps;
    B200PS_REQUIRE(buf && s.size() < cap, B200PS_ERR_RANGE, "buffer of %zu bytes is too small for %zu", cap,
                   s.size() + 1);
    memcpy(buf, s.c_str(), s.size() + 1);
    return (int)s.size();
}

extern "C" int b200ps_cluster_task_address(const b200ps_cluster_t* c, const char* job, int index, char* buf,
                                           size_t cap) {
    using namespace b200ps;
    B200PS_REQUIRE(c && job, B200PS_ERR_INVALID, "null argument");
    B200PS_REQUIRE(index >= 0 && index < c->count(job), B200PS_ERR_RANGE, "no task %s:%d", job, index);
    return copy_out(c->jobs.at(job)[(size_t)index], buf, cap);
}

extern "C" int b200ps_tf_config_for(const b200ps_cluster_t* c, const char* job, int index, char* buf, size_t cap) {
    using namespace b200ps;
    B200PS_REQUIRE(c && job, B200PS_ERR_INVALID, "null argument");
    B200PS_REQUIRE(index >= 0 && index < c->count(job), B200PS_ERR_RANGE, "no task %s:%d", job, index);
    std::string s = "{\"cluster\":{";
    bool first = true;
    for (auto& kv : c->jobs) {
        if (!first) s += ',';
        first = false;
        put_json_string(s, kv.first);
        s += ":[";
        for (size_t i = 0; i < kv.second.size(); ++i) {
            if (i) s += ',';
            put_json_string(s, kv.second[i]);
        }
        s += ']';
    }
    s += "},\"task\":{\"type\":";
    put_json_string(s, job);
    s += ",\"index\":" + std::to_string(index) + "},\"environment\":\"cloud\"}";
    return copy_out(s, buf, cap);
}

extern "C" int b200ps_cluster_num_shards(const b200ps_cluster_t* c) {
    using namespace b200ps;
    B200PS_REQUIRE(c, B200PS_ERR_INVALID, "null argument");
    return c->count("ps");
}

extern "C" int b200ps_cluster_num_workers(const b200ps_cluster_t* c) {
    using namespace b200ps;
    B200PS_REQUIRE(c, B200PS_ERR_INVALID, "null argument");
    return c->count("master") + c->count("chief") + c->count("worker");
}

extern "C" int b200ps_cluster_worker_rank(const b200ps_cluster_t* c, const char* job, int index) {
    using namespace b200ps;
    B200PS_REQUIRE(c && job, B200PS_ERR_INVALID, "null argument");
    std::string j = job;
    B200PS_REQUIRE(compute_job(j), B200PS_ERR_INVALID, "%s is not a compute job", job);
    B200PS_REQUIRE(index >= 0 && index < c->count(j), B200PS_ERR_RANGE, "no task %s:%d", job, index);
    if (j == "worker") return c->count("master") + c->count("chief") + index;
    return index;  // the single master / chief is compute task 0
}

extern "C" int b200ps_placement_gpu(int task_rank, int ngpus) {
    using namespace b200ps;
    B200PS_REQUIRE(task_rank >= 0 && ngpus >= 1, B200PS_ERR_INVALID, "placement needs rank >= 0 and ngpus >= 1");
    return task_rank % ngpus;
}
