// checkpoint.cpp — native reader for the reference's checkpoint files, no libtorch.
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
#include <cstring>
#include <map>
#include <memory>

#include "common.h"

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
};

namespace {

// module tree -> dotted names, depth first in attribute order
int flatten(const P& obj, const std::string& prefix, const std::map<std::string, Member>& members, const std::string& root,
            mzb_checkpoint* ck) {
    if (!obj->state || obj->state->kind != Val::DICT) return MZB_OK;
    for (const auto& kv : obj->state->dict) {
        if (kv.first->kind != Val::STR) continue;
        const std::string name = prefix + kv.first->s;
        const P& v = kv.second;
        if (v->kind == Val::OBJ) {
            int rc = flatten(v, name + ".", members, root, ck);
            if (rc) return rc;
        } else if (v->kind == Val::TENSOR && v->storage_type == "torch FloatStorage") {
            auto it = members.find(root + "/data/" + v->storage_key);
            MZB_REQUIRE(it != members.end(), "checkpoint: storage %s of %s is missing from the archive", v->storage_key.c_str(), name.c_str());
            MZB_REQUIRE(it->second.stored, "checkpoint: storage %s of %s is compressed (torch::save stores tensor data uncompressed)", v->storage_key.c_str(), name.c_str());
            MZB_REQUIRE(v->dims.size() <= 4, "checkpoint: %s has %zu dimensions", name.c_str(), v->dims.size());
            int64_t numel = 1, last = v->offset;
            for (size_t d = 0; d < v->dims.size(); ++d) {
                MZB_REQUIRE(v->dims[d] >= 0 && v->strides[d] >= 0, "checkpoint: %s has a negative size or stride", name.c_str());
                numel *= v->dims[d];
                if (v->dims[d] > 0) last += (v->dims[d] - 1) * v->strides[d];
            }
            MZB_REQUIRE(numel == 0 || (v->offset >= 0 && (size_t)(last + 1) * 4 <= it->second.size), "checkpoint: %s reaches outside its storage", name.c_str());
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
    std::unique_ptr<mzb_checkpoint> ck(new mzb_checkpoint());
    int rc = open_impl(path, ck.get());
    if (rc) return rc;
    ck->file.clear();
    ck->file.shrink_to_fit();
    *out = ck.release();
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

int mzb_load_checkpoint(mzb_engine* e, const char* path) {
    MZB_REQUIRE(e && path, "null argument");
    mzb_checkpoint* ck = nullptr;
    int rc = mzb_checkpoint_open(path, &ck);
    if (rc) return rc;
    std::vector<mzb_tensor_desc> t(ck->tensors.size());
    for (size_t i = 0; i < t.size(); ++i) mzb_checkpoint_tensor(ck, (int)i, &t[i]);
    rc = mzb_load_weights(e, t.data(), (int)t.size());
    mzb_checkpoint_close(ck);
    return rc;
}

}  // extern "C"
