// mzbt_io.h — TEST INFRASTRUCTURE.  "MZBT": a minimal named-tensor container used to move weights,
// inputs and outputs between the reference driver, the Python tests and golden fixtures.
//   "MZBT" | u32 version=1 | u32 count | count x { u32 name_len | name | u32 dtype | u32 ndim |
//   i64 dims[ndim] | raw little-endian data }      dtype: 0=f32 1=f64 2=i32 3=u8
#ifndef MZB_ORACLE_MZBT_IO_H_
#define MZB_ORACLE_MZBT_IO_H_
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

namespace mzbt {
struct Tensor {
    std::string name;
    uint32_t dtype = 0;
    std::vector<int64_t> dims;
    std::vector<uint8_t> data;
    int64_t numel() const { int64_t n = 1; for (auto d : dims) n *= d; return n; }
    template <class T> const T* as() const { return reinterpret_cast<const T*>(data.data()); }
};
inline size_t dtype_size(uint32_t dt) { return dt == 0 ? 4 : dt == 1 ? 8 : dt == 2 ? 4 : 1; }

struct Writer {
    std::vector<Tensor> tensors;
    template <class T> void add(const std::string& name, uint32_t dtype, const std::vector<int64_t>& dims, const T* p) {
        Tensor t; t.name = name; t.dtype = dtype; t.dims = dims;
        t.data.resize(t.numel() * dtype_size(dtype));
        std::memcpy(t.data.data(), p, t.data.size());
        tensors.push_back(std::move(t));
    }
    bool save(const std::string& path) const {
        FILE* f = fopen(path.c_str(), "wb");
        if (!f) return false;
        uint32_t ver = 1, n = (uint32_t)tensors.size();
        fwrite("MZBT", 1, 4, f); fwrite(&ver, 4, 1, f); fwrite(&n, 4, 1, f);
        for (auto& t : tensors) {
            uint32_t nl = (uint32_t)t.name.size(), nd = (uint32_t)t.dims.size();
            fwrite(&nl, 4, 1, f); fwrite(t.name.data(), 1, nl, f);
            fwrite(&t.dtype, 4, 1, f); fwrite(&nd, 4, 1, f);
            fwrite(t.dims.data(), 8, nd, f);
            fwrite(t.data.data(), 1, t.data.size(), f);
        }
        fclose(f);
        return true;
    }
};

inline bool load(const std::string& path, std::vector<Tensor>& out) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return false;
    char magic[4]; uint32_t ver, n;
    if (fread(magic, 1, 4, f) != 4 || std::memcmp(magic, "MZBT", 4) != 0) { fclose(f); return false; }
    if (fread(&ver, 4, 1, f) != 1 || fread(&n, 4, 1, f) != 1) { fclose(f); return false; }
    for (uint32_t i = 0; i < n; ++i) {
        Tensor t; uint32_t nl, nd;
        if (fread(&nl, 4, 1, f) != 1) { fclose(f); return false; }
        t.name.resize(nl);
        if (nl && fread(&t.name[0], 1, nl, f) != nl) { fclose(f); return false; }
        if (fread(&t.dtype, 4, 1, f) != 1 || fread(&nd, 4, 1, f) != 1) { fclose(f); return false; }
        t.dims.resize(nd);
        if (nd && fread(t.dims.data(), 8, nd, f) != nd) { fclose(f); return false; }
        t.data.resize(t.numel() * dtype_size(t.dtype));
        if (!t.data.empty() && fread(t.data.data(), 1, t.data.size(), f) != t.data.size()) { fclose(f); return false; }
        out.push_back(std::move(t));
    }
    fclose(f);
    return true;
}
}  // namespace mzbt
#endif
