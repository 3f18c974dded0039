// checkpoint.cpp — native reader AND writer for the reference's checkpoint files, no libtorch.
//
// Writer (second half of this file): VPRNetModel::SaveCheckpoint's `torch::save(model_, path + ".pt")` (vprnet.cpp:58-63) — the
// same archive, emitted from the engine's named tensors, so that a reference process can `torch::load` what this engine holds.
//
// Replaces the read half of VPRNetModel::LoadCheckpoint (muzero-cpp/vprnet.cpp:66-74): `torch::load(model_, path + ".pt")`.
// What `torch::save(model_, ...)` (vprnet.cpp:58-63) writes is a ZIP archive with STORED (uncompressed) members:
//   <name>/data.pkl      a protocol-2 pickle of the module tree: every module is `NEWOBJ(__torch__...Module)` + BUILD with a dict of
//                        its attributes — sub-modules, `training` flags, and tensors as
//                        torch._utils._rebuild_tensor_v2(persistent_id(('storage', torch.FloatStorage, key, device, numel)),
//                                                        storage_offset, size, stride, requires_grad, OrderedDict())
//   <name>/data/<key>    the raw little-endian storage bytes
//   <name>/constants.pkl, <name>/version, <name>/code/...   not needed here
// The reader walks the central directory, runs the small pickle subset those files use on a value stack, and flattens the
// module tree into dotted names ("dynamics_network.dynamics_resnet_body.0.resnet_0_bn1.running_var") — exactly the names
// mzb_load_weights takes.  Float32 tensors only (num_batches_tracked, int64, is skipped: eval-mode inference never reads it).
#include <algorithm>
#include <cstring>
#include <functional>
#include <map>
#include <memory>

#include "common.h"

struct mzb_engine;
namespace mzb {
// engine.cu: the named fp32 tensors the engine was last loaded with (un-folded host copies) and its int64 scalars
void engine_set_long_scalars(mzb_engine* e, const std::vector<std::pair<std::string, int64_t>>& longs);
int engine_named_tensors(mzb_engine* e, std::vector<mzb_tensor_desc>* tensors, std::vector<const char*>* long_names, std::vector<int64_t>* long_values);
}

namespace {

struct Member { size_t offset = 0, size = 0; bool stored = true; };   // offset of the member's DATA in the file; stored = not compressed

// ------------------------------------------------------------------ ZIP (stored members, no zip64)
uint32_t rd32(const uint8_t* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }
uint32_t rd16(const uint8_t* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8); }

int zip_index(const std::vector<uint8_t>& f, std::map<std::string, Member>* out) {
    const size_t n = f.size();
    MZB_REQUIRE(n >= 22, "checkpoint: file too small to be a zip archive");
    size_t eocd = std::string::npos;
    for (size_t back = 22; back <= n && back <= 22 + 65535; ++back)   // end-of-central-directory record, behind a possible comment
        if (rd32(&f[n - back]) == 0x06054b50u) { eocd = n - back; break; }
    MZB_REQUIRE(eocd != std::string::npos, "checkpoint: no zip end-of-central-directory record (not a torch::save archive?)");
    const uint32_t entries = rd16(&f[eocd + 10]), cd_size = rd32(&f[eocd + 12]), cd_off = rd32(&f[eocd + 16]);
    MZB_REQUIRE(entries != 0xffffu && cd_off != 0xffffffffu, "checkpoint: zip64 archives are not supported");
    MZB_REQUIRE((size_t)cd_off + cd_size <= n, "checkpoint: central directory outside the file");
    size_t p = cd_off;
    for (uint32_t i = 0; i < entries; ++i) {
        MZB_REQUIRE(p + 46 <= n && rd32(&f[p]) == 0x02014b50u, "checkpoint: bad central directory entry %u", i);
        const uint32_t method = rd16(&f[p + 10]), csize = rd32(&f[p + 20]), usize = rd32(&f[p + 24]);
        const uint32_t nlen = rd16(&f[p + 28]), xlen = rd16(&f[p + 30]), clen = rd16(&f[p + 32]), lho = rd32(&f[p + 42]);
        MZB_REQUIRE(p + 46 + nlen <= n, "checkpoint: truncated central directory");
        const std::string name((const char*)&f[p + 46], nlen);
        MZB_REQUIRE((size_t)lho + 30 <= n && rd32(&f[lho]) == 0x04034b50u, "checkpoint: bad local header for %s", name.c_str());
        const size_t data = (size_t)lho + 30 + rd16(&f[lho + 26]) + rd16(&f[lho + 28]);
        MZB_REQUIRE(data + csize <= n, "checkpoint: member %s runs past the end of the file", name.c_str());
        (*out)[name] = Member{data, usize, method == 0 && csize == usize};   // only the code/ members are ever deflated
        p += 46 + nlen + xlen + clen;
    }
    return MZB_OK;
}

// ------------------------------------------------------------------ the pickle subset of data.pkl
struct Val;
typedef std::shared_ptr<Val> P;
struct Val {
    enum Kind { NONE, INT, BOOL, FLOAT, STR, GLOBAL, TUPLE, LIST, DICT, OBJ, PERSID, TENSOR, MARK } kind = NONE;
    int64_t i = 0;
    std::string s;                        // STR / GLOBAL ("module name") / OBJ class
    std::vector<P> items;                 // TUPLE / LIST / PERSID tuple
    std::vector<std::pair<P, P>> dict;    // DICT
    P state;                              // OBJ: what BUILD attached
    // TENSOR
    std::string storage_key, storage_type;
    int64_t offset = 0;
    std::vector<int64_t> dims, strides;
};
P make(Val::Kind k) { P v(new Val); v->kind = k; return v; }

struct Unpickler {
    const uint8_t* b; size_t n, pos = 0;
    std::vector<P> stack;
    std::map<uint32_t, P> memo;
    std::string err;

    bool need(size_t k) { if (pos + k > n) { err = "truncated pickle"; return false; } return true; }
    bool pop(P* out) { if (stack.empty() || stack.back()->kind == Val::MARK) { err = "pickle stack underflow"; return false; } *out = stack.back(); stack.pop_back(); return true; }
    bool pop_to_mark(std::vector<P>* out) {
        size_t k = stack.size();
        while (k > 0 && stack[k - 1]->kind != Val::MARK) --k;
        if (k == 0) { err = "pickle: no MARK on the stack"; return false; }
        out->assign(stack.begin() + k, stack.end());
        stack.resize(k - 1);
        return true;
    }
    bool line(std::string* out) {
        size_t e = pos;
        while (e < n && b[e] != '\n') ++e;
        if (e >= n) { err = "truncated pickle"; return false; }
        out->assign((const char*)b + pos, e - pos);
        pos = e + 1;
        return true;
    }
    static bool ints(const P& t, std::vector<int64_t>* out) {
        if (t->kind != Val::TUPLE) return false;
        for (const P& x : t->items) { if (x->kind != Val::INT) return false; out->push_back(x->i); }
        return true;
    }
    // REDUCE: the two callables the archive uses
    bool reduce(const P& fn, const P& args, P* out) {
        if (fn->kind == Val::GLOBAL && fn->s == "torch._utils _rebuild_tensor_v2") {
            if (args->kind != Val::TUPLE || args->items.size() < 4) { err = "_rebuild_tensor_v2: bad argument tuple"; return false; }
            const P& pid = args->items[0];
            if (pid->kind != Val::PERSID || pid->items.size() < 5 || pid->items[0]->kind != Val::STR || pid->items[0]->s != "storage" ||
                pid->items[1]->kind != Val::GLOBAL || pid->items[2]->kind != Val::STR || args->items[1]->kind != Val::INT) {
                err = "_rebuild_tensor_v2: unexpected persistent id";
                return false;
            }
            P t = make(Val::TENSOR);
            t->storage_type = pid->items[1]->s; t->storage_key = pid->items[2]->s; t->offset = args->items[1]->i;
            if (!ints(args->items[2], &t->dims) || !ints(args->items[3], &t->strides) || t->dims.size() != t->strides.size()) {
                err = "_rebuild_tensor_v2: bad size / stride";
                return false;
            }
            *out = t;
            return true;
        }
        if (fn->kind == Val::GLOBAL && fn->s == "collections OrderedDict") { *out = make(Val::DICT); return true; }
        *out = make(Val::NONE);   // anything else carries nothing this reader needs
        return true;
    }

    bool run(P* root) {
        while (true) {
            if (!need(1)) return false;
            const uint8_t op = b[pos++];
            switch (op) {
                case 0x80: if (!need(1)) return false; ++pos; break;                                   // PROTO
                case 'c': { std::string m, nm; if (!line(&m) || !line(&nm)) return false; P g = make(Val::GLOBAL); g->s = m + " " + nm; stack.push_back(g); break; }
                case 'q': { if (!need(1) || stack.empty()) { err = "BINPUT"; return false; } memo[b[pos++]] = stack.back(); break; }
                case 'r': { if (!need(4) || stack.empty()) { err = "LONG_BINPUT"; return false; } memo[rd32(b + pos)] = stack.back(); pos += 4; break; }
                case 'h': { if (!need(1)) return false; auto it = memo.find(b[pos++]); if (it == memo.end()) { err = "BINGET of an unset memo slot"; return false; } stack.push_back(it->second); break; }
                case 'j': { if (!need(4)) return false; auto it = memo.find(rd32(b + pos)); pos += 4; if (it == memo.end()) { err = "LONG_BINGET of an unset memo slot"; return false; } stack.push_back(it->second); break; }
                case ')': stack.push_back(make(Val::TUPLE)); break;
                case ']': stack.push_back(make(Val::LIST)); break;
                case '}': stack.push_back(make(Val::DICT)); break;
                case '(': stack.push_back(make(Val::MARK)); break;
                case 'N': stack.push_back(make(Val::NONE)); break;
                case 0x88: case 0x89: { P v = make(Val::BOOL); v->i = op == 0x88; stack.push_back(v); break; }
                case 'K': { if (!need(1)) return false; P v = make(Val::INT); v->i = b[pos++]; stack.push_back(v); break; }
                case 'M': { if (!need(2)) return false; P v = make(Val::INT); v->i = rd16(b + pos); pos += 2; stack.push_back(v); break; }
                case 'J': { if (!need(4)) return false; P v = make(Val::INT); v->i = (int32_t)rd32(b + pos); pos += 4; stack.push_back(v); break; }
                case 0x8a: {                                                                            // LONG1
                    if (!need(1)) return false;
                    const uint32_t k = b[pos++];
                    if (k > 8 || !need(k)) { err = "LONG1 wider than 8 bytes"; return false; }
                    uint64_t x = 0;
                    for (uint32_t t = 0; t < k; ++t) x |= (uint64_t)b[pos + t] << (8 * t);
                    if (k > 0 && k < 8 && (b[pos + k - 1] & 0x80)) x |= ~0ull << (8 * k);
                    pos += k;
                    P v = make(Val::INT); v->i = (int64_t)x; stack.push_back(v);
                    break;
                }
                case 'G': { if (!need(8)) return false; pos += 8; stack.push_back(make(Val::FLOAT)); break; }
                case 'X': { if (!need(4)) return false; const uint32_t k = rd32(b + pos); pos += 4; if (!need(k)) return false; P v = make(Val::STR); v->s.assign((const char*)b + pos, k); pos += k; stack.push_back(v); break; }
                case 'Q': { P pid; if (!pop(&pid)) return false; P v = make(Val::PERSID); if (pid->kind == Val::TUPLE) v->items = pid->items; stack.push_back(v); break; }
                case 't': { std::vector<P> it; if (!pop_to_mark(&it)) return false; P v = make(Val::TUPLE); v->items.swap(it); stack.push_back(v); break; }
                case 0x85: case 0x86: case 0x87: {
                    const size_t k = op - 0x84;
                    if (stack.size() < k) { err = "TUPLEn underflow"; return false; }
                    P v = make(Val::TUPLE);
                    v->items.assign(stack.end() - k, stack.end());
                    stack.resize(stack.size() - k);
                    stack.push_back(v);
                    break;
                }
                case 0x81: { P args, cls; if (!pop(&args) || !pop(&cls)) return false; P v = make(Val::OBJ); v->s = cls->s; stack.push_back(v); break; }   // NEWOBJ
                case 'R': { P args, fn, res; if (!pop(&args) || !pop(&fn) || !reduce(fn, args, &res)) return false; stack.push_back(res); break; }
                case 'b': { P st; if (!pop(&st)) return false; if (stack.empty()) { err = "BUILD underflow"; return false; } stack.back()->state = st; break; }
                case 's': { P v, k; if (!pop(&v) || !pop(&k)) return false; if (stack.empty()) { err = "SETITEM underflow"; return false; } stack.back()->dict.emplace_back(k, v); break; }
                case 'u': {
                    std::vector<P> it;
                    if (!pop_to_mark(&it)) return false;
                    if (stack.empty() || it.size() % 2) { err = "SETITEMS: odd item count"; return false; }
                    for (size_t k = 0; k + 1 < it.size(); k += 2) stack.back()->dict.emplace_back(it[k], it[k + 1]);
                    break;
                }
                case 'a': { P v; if (!pop(&v)) return false; if (stack.empty()) { err = "APPEND underflow"; return false; } stack.back()->items.push_back(v); break; }
                case 'e': { std::vector<P> it; if (!pop_to_mark(&it)) return false; if (stack.empty()) { err = "APPENDS underflow"; return false; } for (P& v : it) stack.back()->items.push_back(v); break; }
                case '.': return pop(root);                                                             // STOP
                default: { char m[64]; snprintf(m, sizeof(m), "unsupported pickle opcode 0x%02x at byte %zu", op, pos - 1); err = m; return false; }
            }
        }
    }
};

}  // namespace

struct mzb_checkpoint {
    std::vector<uint8_t> file;
    struct Entry { std::string name; std::vector<int64_t> dims; std::vector<float> data; };
    std::vector<Entry> tensors;
    std::vector<std::pair<std::string, int64_t>> longs;   // int64 scalars (BatchNorm num_batches_tracked), kept for a faithful re-save
};

namespace {

// module tree -> dotted names, depth first in attribute order
int flatten(const P& obj, const std::string& prefix, const std::map<std::string, Member>& members, const std::string& root,
            mzb_checkpoint* ck, int depth = 0) {
    MZB_REQUIRE(depth < 64, "checkpoint: module tree deeper than 64 levels (a cycle in the pickle memo?)");
    if (!obj->state || obj->state->kind != Val::DICT) return MZB_OK;
    for (const auto& kv : obj->state->dict) {
        if (kv.first->kind != Val::STR) continue;
        const std::string name = prefix + kv.first->s;
        const P& v = kv.second;
        if (v->kind == Val::OBJ) {
            int rc = flatten(v, name + ".", members, root, ck, depth + 1);
            if (rc) return rc;
        } else if (v->kind == Val::TENSOR && v->storage_type == "torch LongStorage") {
            auto it = members.find(root + "/data/" + v->storage_key);
            int64_t numel = 1;
            for (int64_t d : v->dims) numel *= d;
            if (it != members.end() && it->second.stored && numel == 1 && v->offset >= 0 && (size_t)(v->offset + 1) * 8 <= it->second.size) {
                int64_t x = 0;
                memcpy(&x, ck->file.data() + it->second.offset + (size_t)v->offset * 8, 8);
                ck->longs.emplace_back(name, x);
            }
        } else if (v->kind == Val::TENSOR && v->storage_type == "torch FloatStorage") {
            auto it = members.find(root + "/data/" + v->storage_key);
            MZB_REQUIRE(it != members.end(), "checkpoint: storage %s of %s is missing from the archive", v->storage_key.c_str(), name.c_str());
            MZB_REQUIRE(it->second.stored, "checkpoint: storage %s of %s is compressed (torch::save stores tensor data uncompressed)", v->storage_key.c_str(), name.c_str());
            MZB_REQUIRE(v->dims.size() <= 4, "checkpoint: %s has %zu dimensions", name.c_str(), v->dims.size());
            // sizes and strides come from the file: every product is checked against the storage it indexes before it is used
            const int64_t cap = (int64_t)(it->second.size / 4);   // float32 elements in the storage
            int64_t numel = 1, last = v->offset;
            MZB_REQUIRE(v->offset >= 0 && v->offset <= cap, "checkpoint: %s starts outside its storage", name.c_str());
            for (size_t d = 0; d < v->dims.size(); ++d) {
                MZB_REQUIRE(v->dims[d] >= 0 && v->strides[d] >= 0, "checkpoint: %s has a negative size or stride", name.c_str());
                MZB_REQUIRE(v->dims[d] <= cap && v->strides[d] <= cap, "checkpoint: %s has a size or stride larger than its storage", name.c_str());
                MZB_REQUIRE(v->dims[d] == 0 || numel <= cap / v->dims[d], "checkpoint: %s has more elements than its storage", name.c_str());
                numel *= v->dims[d];
                if (v->dims[d] > 1) {
                    MZB_REQUIRE(v->strides[d] == 0 || (v->dims[d] - 1) <= (cap - last) / v->strides[d], "checkpoint: %s reaches outside its storage", name.c_str());
                    last += (v->dims[d] - 1) * v->strides[d];
                }
            }
            MZB_REQUIRE(numel == 0 || last + 1 <= cap, "checkpoint: %s reaches outside its storage", name.c_str());
            mzb_checkpoint::Entry e;
            e.name = name; e.dims = v->dims; e.data.resize((size_t)numel);
            const uint8_t* base = ck->file.data() + it->second.offset;
            int64_t d4[4] = {1, 1, 1, 1}, s4[4] = {0, 0, 0, 0};
            const size_t nd = v->dims.size();
            for (size_t d = 0; d < nd; ++d) { d4[4 - nd + d] = v->dims[d]; s4[4 - nd + d] = v->strides[d]; }
            size_t o = 0;
            for (int64_t i0 = 0; i0 < d4[0] && numel; ++i0)
                for (int64_t i1 = 0; i1 < d4[1]; ++i1)
                    for (int64_t i2 = 0; i2 < d4[2]; ++i2)
                        for (int64_t i3 = 0; i3 < d4[3]; ++i3) {
                            const int64_t idx = v->offset + i0 * s4[0] + i1 * s4[1] + i2 * s4[2] + i3 * s4[3];
                            memcpy(&e.data[o++], base + (size_t)idx * 4, 4);   // little-endian float32, as stored
                        }
            ck->tensors.push_back(std::move(e));
        }
    }
    return MZB_OK;
}

int open_impl(const char* path, mzb_checkpoint* ck) {
    FILE* f = fopen(path, "rb");
    MZB_REQUIRE(f != nullptr, "checkpoint: cannot open %s", path);
    fseek(f, 0, SEEK_END);
    const long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    ck->file.resize(sz > 0 ? (size_t)sz : 0);
    const size_t got = ck->file.empty() ? 0 : fread(ck->file.data(), 1, ck->file.size(), f);
    fclose(f);
    MZB_REQUIRE(got == ck->file.size() && got > 0, "checkpoint: cannot read %s", path);
    std::map<std::string, Member> members;
    int rc = zip_index(ck->file, &members);
    if (rc) return rc;
    std::string root;
    for (const auto& m : members) {
        const size_t at = m.first.rfind("/data.pkl");
        if (at != std::string::npos && at + 9 == m.first.size()) { root = m.first.substr(0, at); break; }
    }
    MZB_REQUIRE(!root.empty(), "checkpoint: %s holds no data.pkl (not a torch::save module archive)", path);
    const Member pk = members[root + "/data.pkl"];
    MZB_REQUIRE(pk.stored, "checkpoint: %s: data.pkl is compressed (torch::save stores it uncompressed)", path);
    Unpickler u{ck->file.data() + pk.offset, pk.size};
    P top;
    MZB_REQUIRE(u.run(&top), "checkpoint: %s: %s", path, u.err.c_str());
    MZB_REQUIRE(top->kind == Val::OBJ, "checkpoint: %s: the pickled object is not a module", path);
    rc = flatten(top, "", members, root, ck);
    if (rc) return rc;
    MZB_REQUIRE(!ck->tensors.empty(), "checkpoint: %s holds no float32 tensors", path);
    return MZB_OK;
}

}  // namespace

extern "C" {

int mzb_checkpoint_open(const char* path, mzb_checkpoint** out) {
    MZB_REQUIRE(path && out, "null argument");
    try {   // nothing may unwind across the C boundary (a hostile size field -> bad_alloc / length_error)
        std::unique_ptr<mzb_checkpoint> ck(new mzb_checkpoint());
        int rc = open_impl(path, ck.get());
        if (rc) return rc;
        ck->file.clear();
        ck->file.shrink_to_fit();
        *out = ck.release();
        return MZB_OK;
    } catch (const std::exception& ex) {
        mzb::set_error("checkpoint: %s: %s", path, ex.what());
        return MZB_ERR_INVALID;
    }
}

int mzb_checkpoint_long_count(const mzb_checkpoint* ck) { return ck ? (int)ck->longs.size() : 0; }

int mzb_checkpoint_long(const mzb_checkpoint* ck, int index, const char** name, int64_t* value) {
    MZB_REQUIRE(ck && name && value && index >= 0 && index < (int)ck->longs.size(), "bad checkpoint scalar index");
    *name = ck->longs[index].first.c_str(); *value = ck->longs[index].second;
    return MZB_OK;
}

int mzb_checkpoint_count(const mzb_checkpoint* ck) { return ck ? (int)ck->tensors.size() : 0; }

int mzb_checkpoint_tensor(const mzb_checkpoint* ck, int index, mzb_tensor_desc* out) {
    MZB_REQUIRE(ck && out && index >= 0 && index < (int)ck->tensors.size(), "bad checkpoint tensor index");
    const mzb_checkpoint::Entry& e = ck->tensors[index];
    out->name = e.name.c_str(); out->data = e.data.data(); out->ndim = (int32_t)e.dims.size();
    for (int d = 0; d < 4; ++d) out->dims[d] = d < (int)e.dims.size() ? e.dims[d] : 0;
    return MZB_OK;
}

void mzb_checkpoint_close(mzb_checkpoint* ck) { delete ck; }

}  // extern "C"

// ====================================================================================================================
// Writer: the archive torch::save(model_, path) produces for a torch::nn::Module tree (vprnet.cpp:58-63), rebuilt from dotted
// tensor names.  Layout (all members STORED, CRC-32 checked by libtorch's miniz reader):
//   <root>/data.pkl                 protocol-2 pickle: GLOBAL '__torch__[.___torch_mangle_k] Module', EMPTY_TUPLE, NEWOBJ, dict, BUILD
//   <root>/data/<i>                 raw little-endian storages, one per tensor
//   <root>/code/__torch__[/...].py  one class definition per module (parameters, buffers, attribute / sub-module types)
//   <root>/constants.pkl, version ("3"), byteorder ("little")
// What the C++ loader needs beyond the tensors themselves is every registered child and buffer of the reference's modules
// (torch::load reads them by name and fails on a missing one): BatchNorm's int64 `num_batches_tracked` (model_layers.cpp:36-42,
// track_running_stats) and the parameter-less `activation_i` ReLU modules between an MLP's `linear_i` (model_layers.cpp:46-65).
namespace {

struct WTensor { std::string attr; const float* data = nullptr; std::vector<int64_t> dims; bool is_long = false; int64_t long_value = 0; int storage = -1; };
struct WNode {
    std::string attr, cls, file;          // attribute name in the parent, '__torch__.___torch_mangle_k', 'code/__torch__/___torch_mangle_k.py'
    std::vector<WTensor> tensors;
    std::vector<std::unique_ptr<WNode>> kids;
    WNode* kid(const std::string& name) {
        for (auto& k : kids) if (k->attr == name) return k.get();
        kids.emplace_back(new WNode());
        kids.back()->attr = name;
        return kids.back().get();
    }
};

bool is_buffer_name(const std::string& a) { return a == "running_mean" || a == "running_var" || a == "num_batches_tracked"; }

// modules the tensors do not name: BatchNorm's counter, the MLP's ReLU modules
void complete(WNode* n, const std::map<std::string, int64_t>& longs, const std::string& prefix) {
    bool bn = false, counter = false;
    for (const WTensor& t : n->tensors) { bn = bn || t.attr == "running_var"; counter = counter || t.attr == "num_batches_tracked"; }
    if (bn && !counter) {
        WTensor t;
        t.attr = "num_batches_tracked"; t.is_long = true;
        auto it = longs.find(prefix + "num_batches_tracked");
        t.long_value = it == longs.end() ? 0 : it->second;
        n->tensors.push_back(t);
    }
    int n_linear = 0;
    for (auto& k : n->kids) if (k->attr.compare(0, 7, "linear_") == 0) ++n_linear;
    if (n_linear > 0 && n_linear == (int)n->kids.size()) {   // an MLP's Sequential: linear_0, activation_0, linear_1, ..., linear_{n-1}
        std::vector<std::unique_ptr<WNode>> ordered;
        for (int i = 0; i < n_linear; ++i) {
            for (auto& k : n->kids) if (k && k->attr == "linear_" + std::to_string(i)) ordered.push_back(std::move(k));
            if (i < n_linear - 1) { ordered.emplace_back(new WNode()); ordered.back()->attr = "activation_" + std::to_string(i); }
        }
        if ((int)ordered.size() == 2 * n_linear - 1) n->kids.swap(ordered);
    }
    for (auto& k : n->kids) complete(k.get(), longs, prefix + k->attr + ".");
}

void number(WNode* n, int* next_class, int* next_storage, bool root) {
    if (root) { n->cls = "__torch__"; n->file = "code/__torch__.py"; }
    else { n->cls = "__torch__.___torch_mangle_" + std::to_string(*next_class); n->file = "code/__torch__/___torch_mangle_" + std::to_string(*next_class) + ".py"; ++*next_class; }
    for (WTensor& t : n->tensors) t.storage = (*next_storage)++;
    for (auto& k : n->kids) number(k.get(), next_class, next_storage, false);
}

struct Pickle {
    std::string b;
    void op(uint8_t c) { b.push_back((char)c); }
    void u32(uint32_t v) { for (int i = 0; i < 4; ++i) b.push_back((char)((v >> (8 * i)) & 0xff)); }
    void global(const std::string& module, const std::string& name) { op('c'); b += module; b += '\n'; b += name; b += '\n'; }
    void str(const std::string& v) { op('X'); u32((uint32_t)v.size()); b += v; }
    void integer(int64_t v) {
        if (v >= 0 && v < 256) { op('K'); b.push_back((char)v); }
        else if (v >= 0 && v < 65536) { op('M'); b.push_back((char)(v & 0xff)); b.push_back((char)(v >> 8)); }
        else if (v >= INT32_MIN && v <= INT32_MAX) { op('J'); u32((uint32_t)(int32_t)v); }
        else { op(0x8a); b.push_back(8); for (int i = 0; i < 8; ++i) b.push_back((char)(((uint64_t)v >> (8 * i)) & 0xff)); }
    }
    void int_tuple(const std::vector<int64_t>& v) { op('('); for (int64_t x : v) integer(x); op('t'); }
};

void pickle_tensor(Pickle& p, const WTensor& t) {
    int64_t numel = 1;
    for (int64_t d : t.dims) numel *= d;
    std::vector<int64_t> strides(t.dims.size(), 1);
    for (int d = (int)t.dims.size() - 2; d >= 0; --d) strides[d] = strides[d + 1] * t.dims[d + 1];
    p.global("torch._utils", "_rebuild_tensor_v2");
    p.op('(');
    p.op('('); p.str("storage"); p.global("torch", t.is_long ? "LongStorage" : "FloatStorage"); p.str(std::to_string(t.storage)); p.str("cpu"); p.integer(numel); p.op('t');
    p.op('Q');
    p.integer(0);
    p.int_tuple(t.dims);
    p.int_tuple(strides);
    p.op(is_buffer_name(t.attr) ? 0x89 : 0x88);   // requires_grad: parameters yes, buffers no
    p.global("collections", "OrderedDict"); p.op(')'); p.op('R');
    p.op('t');
    p.op('R');
}

void pickle_module(Pickle& p, const WNode& n) {
    p.global(n.cls, "Module");
    p.op(')'); p.op(0x81); p.op('}'); p.op('(');
    for (const WTensor& t : n.tensors) { p.str(t.attr); pickle_tensor(p, t); }
    for (const auto& k : n.kids) { p.str(k->attr); pickle_module(p, *k); }
    p.op('u'); p.op('b');
}

std::string class_source(const WNode& n) {
    std::string s = "class Module(Module):\n  __parameters__ = [";
    for (const WTensor& t : n.tensors) if (!is_buffer_name(t.attr)) s += "\"" + t.attr + "\", ";
    s += "]\n  __buffers__ = [";
    for (const WTensor& t : n.tensors) if (is_buffer_name(t.attr)) s += "\"" + t.attr + "\", ";
    s += "]\n";
    bool plain = true;   // attribute names that are not identifiers (Sequential's "0", "1", ...) go through __annotations__
    for (const auto& k : n.kids) plain = plain && !k->attr.empty() && !(k->attr[0] >= '0' && k->attr[0] <= '9');
    for (const WTensor& t : n.tensors) s += "  " + t.attr + " : Tensor\n";
    if (!plain) s += "  __annotations__ = []\n";
    for (const auto& k : n.kids)
        s += plain ? "  " + k->attr + " : " + k->cls + ".Module\n" : "  __annotations__[\"" + k->attr + "\"] = " + k->cls + ".Module\n";
    return s;
}

uint32_t crc32_of(const uint8_t* d, size_t n) {
    static uint32_t table[256];
    static bool init = false;
    if (!init) {
        for (uint32_t i = 0; i < 256; ++i) { uint32_t c = i; for (int k = 0; k < 8; ++k) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1; table[i] = c; }
        init = true;
    }
    uint32_t c = 0xFFFFFFFFu;
    for (size_t i = 0; i < n; ++i) c = table[(c ^ d[i]) & 0xff] ^ (c >> 8);
    return c ^ 0xFFFFFFFFu;
}

struct ZipWriter {
    std::vector<uint8_t> out, central;
    int entries = 0;
    static void w16(std::vector<uint8_t>& v, uint32_t x) { v.push_back(x & 0xff); v.push_back((x >> 8) & 0xff); }
    static void w32(std::vector<uint8_t>& v, uint32_t x) { for (int i = 0; i < 4; ++i) v.push_back((x >> (8 * i)) & 0xff); }
    void add(const std::string& name, const void* data, size_t n) {
        const uint8_t* d = (const uint8_t*)data;
        const uint32_t crc = crc32_of(d, n), off = (uint32_t)out.size();
        // tensor data starts on a 64-byte boundary, as libtorch's own writer arranges with a padding extra field (mmap-friendly)
        const size_t hdr = 30 + name.size();
        size_t pad = (64 - (off + hdr + 4) % 64) % 64;
        const uint32_t xlen = (uint32_t)(4 + pad);
        w32(out, 0x04034b50u); w16(out, 20); w16(out, 0); w16(out, 0); w16(out, 0); w16(out, 0);
        w32(out, crc); w32(out, (uint32_t)n); w32(out, (uint32_t)n); w16(out, (uint32_t)name.size()); w16(out, xlen);
        out.insert(out.end(), name.begin(), name.end());
        out.push_back('F'); out.push_back('B'); w16(out, (uint32_t)pad);
        out.insert(out.end(), pad, (uint8_t)'Z');
        out.insert(out.end(), d, d + n);
        w32(central, 0x02014b50u); w16(central, 20); w16(central, 20); w16(central, 0); w16(central, 0); w16(central, 0); w16(central, 0);
        w32(central, crc); w32(central, (uint32_t)n); w32(central, (uint32_t)n); w16(central, (uint32_t)name.size()); w16(central, 0); w16(central, 0);
        w16(central, 0); w16(central, 0); w32(central, 0); w32(central, off);
        central.insert(central.end(), name.begin(), name.end());
        ++entries;
    }
    void finish() {
        const uint32_t cd_off = (uint32_t)out.size(), cd_size = (uint32_t)central.size();
        out.insert(out.end(), central.begin(), central.end());
        w32(out, 0x06054b50u); w16(out, 0); w16(out, 0); w16(out, entries); w16(out, entries); w32(out, cd_size); w32(out, cd_off); w16(out, 0);
    }
};

void emit_code(const WNode& n, const std::string& root, ZipWriter* z) {
    const std::string src = class_source(n);
    z->add(root + "/" + n.file, src.data(), src.size());
    for (const auto& k : n.kids) emit_code(*k, root, z);
}
void emit_storages(const WNode& n, const std::string& root, ZipWriter* z) {
    for (const WTensor& t : n.tensors) {
        int64_t numel = 1;
        for (int64_t d : t.dims) numel *= d;
        if (t.is_long) z->add(root + "/data/" + std::to_string(t.storage), &t.long_value, 8);
        else z->add(root + "/data/" + std::to_string(t.storage), t.data, (size_t)numel * 4);
    }
    for (const auto& k : n.kids) emit_storages(*k, root, z);
}

int write_impl(const char* path, const mzb_tensor_desc* tensors, int n, const char* const* long_names, const int64_t* long_values, int n_longs) {
    WNode top;
    std::map<std::string, int64_t> longs;
    for (int i = 0; i < n_longs; ++i) longs[long_names[i]] = long_values[i];
    size_t total = 0;
    for (int i = 0; i < n; ++i) {
        MZB_REQUIRE(tensors[i].name && tensors[i].data && tensors[i].ndim >= 0 && tensors[i].ndim <= 4, "checkpoint writer: bad tensor descriptor %d", i);
        const std::string name = tensors[i].name;
        WNode* at = &top;
        size_t pos = 0;
        while (true) {
            const size_t dot = name.find('.', pos);
            if (dot == std::string::npos) break;
            MZB_REQUIRE(dot > pos, "checkpoint writer: empty component in tensor name %s", name.c_str());
            at = at->kid(name.substr(pos, dot - pos));
            pos = dot + 1;
        }
        WTensor t;
        t.attr = name.substr(pos); t.data = tensors[i].data;
        MZB_REQUIRE(!t.attr.empty(), "checkpoint writer: tensor name %s ends in a dot", name.c_str());
        t.dims.assign(tensors[i].dims, tensors[i].dims + tensors[i].ndim);
        int64_t numel = 1;
        for (int64_t d : t.dims) { MZB_REQUIRE(d >= 0, "checkpoint writer: negative size in %s", name.c_str()); numel *= d; }
        total += (size_t)numel * 4;
        at->tensors.push_back(t);
    }
    MZB_REQUIRE(total < 0xF0000000ull, "checkpoint writer: archives of 4 GB and more need zip64, which is not written");
    // within a module: parameters, then buffers, then children — the order torch::nn::Module serialises its own
    std::function<void(WNode*)> sort_attrs = [&](WNode* nd) {
        std::stable_sort(nd->tensors.begin(), nd->tensors.end(), [](const WTensor& a, const WTensor& b) { return !is_buffer_name(a.attr) && is_buffer_name(b.attr); });
        for (auto& k : nd->kids) sort_attrs(k.get());
    };
    sort_attrs(&top);
    complete(&top, longs, "");
    int next_class = 0, next_storage = 0;
    number(&top, &next_class, &next_storage, true);
    Pickle pk;
    pk.op(0x80); pk.op(2);
    pickle_module(pk, top);
    pk.op('.');
    std::string root = path;
    const size_t slash = root.find_last_of('/');
    if (slash != std::string::npos) root = root.substr(slash + 1);
    const size_t dot = root.rfind('.');
    if (dot != std::string::npos && dot > 0) root = root.substr(0, dot);
    if (root.empty()) root = "archive";
    ZipWriter z;
    z.add(root + "/data.pkl", pk.b.data(), pk.b.size());
    z.add(root + "/byteorder", "little", 6);
    emit_storages(top, root, &z);
    emit_code(top, root, &z);
    const char constants[] = {(char)0x80, 2, ')', '.'};
    z.add(root + "/constants.pkl", constants, 4);
    z.add(root + "/version", "3\n", 2);
    z.finish();
    const std::string tmp = std::string(path) + ".tmp";
    FILE* f = fopen(tmp.c_str(), "wb");
    MZB_REQUIRE(f != nullptr, "checkpoint writer: cannot create %s", tmp.c_str());
    const size_t put = fwrite(z.out.data(), 1, z.out.size(), f);
    const int closed = fclose(f);
    MZB_REQUIRE(put == z.out.size() && closed == 0, "checkpoint writer: short write to %s", tmp.c_str());
    MZB_REQUIRE(rename(tmp.c_str(), path) == 0, "checkpoint writer: cannot move %s into place", tmp.c_str());   // readers never see half a file
    return MZB_OK;
}

}  // namespace

extern "C" {

int mzb_checkpoint_write(const char* path, const mzb_tensor_desc* tensors, int n_tensors, const char* const* long_names,
                         const int64_t* long_values, int n_longs) {
    MZB_REQUIRE(path && tensors && n_tensors > 0 && (n_longs == 0 || (long_names && long_values)), "null argument");
    try {
        return write_impl(path, tensors, n_tensors, long_names, long_values, n_longs);
    } catch (const std::exception& ex) {
        mzb::set_error("checkpoint writer: %s: %s", path, ex.what());
        return MZB_ERR_INVALID;
    }
}



int mzb_load_checkpoint(mzb_engine* e, const char* path) {
    MZB_REQUIRE(e && path, "null argument");
    mzb_checkpoint* ck = nullptr;
    int rc = mzb_checkpoint_open(path, &ck);
    if (rc) return rc;
    std::vector<mzb_tensor_desc> t(ck->tensors.size());
    for (size_t i = 0; i < t.size(); ++i) mzb_checkpoint_tensor(ck, (int)i, &t[i]);
    rc = mzb_load_weights(e, t.data(), (int)t.size());
    if (rc == MZB_OK) mzb::engine_set_long_scalars(e, ck->longs);
    mzb_checkpoint_close(ck);
    return rc;
}

int mzb_save_checkpoint(mzb_engine* e, const char* path) {
    MZB_REQUIRE(e && path, "null argument");
    std::vector<mzb_tensor_desc> t;
    std::vector<const char*> names;
    std::vector<int64_t> values;
    int rc = mzb::engine_named_tensors(e, &t, &names, &values);
    if (rc) return rc;
    return mzb_checkpoint_write(path, t.data(), (int)t.size(), names.data(), values.data(), (int)names.size());
}

}  // extern "C"
