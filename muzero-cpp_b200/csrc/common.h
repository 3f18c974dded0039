// common.h — shared host-side declarations of libmzb200 (no torch, no libtorch).
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

#include "mzb200.h"

namespace mzb {

void set_error(const char* fmt, ...);
const char* get_error();

#define MZB_CUDA(expr)                                                                              \
    do {                                                                                            \
        cudaError_t _e = (expr);                                                                    \
        if (_e != cudaSuccess) {                                                                    \
            mzb::set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return MZB_ERR_CUDA;                                                                    \
        }                                                                                           \
    } while (0)

#define MZB_REQUIRE(cond, ...)           \
    do {                                 \
        if (!(cond)) {                   \
            mzb::set_error(__VA_ARGS__); \
            return MZB_ERR_INVALID;      \
        }                                \
    } while (0)

// RAII device buffer
struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    ~DevBuf() { release(); }
    cudaError_t alloc(size_t n) {
        release();
        bytes = n;
        if (n == 0) return cudaSuccess;
        return cudaMalloc(&p, n);
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        bytes = 0;
    }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

// Shapes derived from the config (VPRNetModel ctor vprnet.cpp:41-46, MuzeroNetworkImpl ctor models.cpp:171-209)
struct NetDims {
    int cin = 0, in_h = 0, in_w = 0;  // initial inference input
    int C = 0, h = 0, w = 0;          // encoded (hidden) state
    int A = 0, ach = 0;
    int Sv = 0, Sr = 0;               // value / reward support sizes
    int hw() const { return h * w; }
    int hidden_size() const { return C * h * w; }
    int initial_flat() const { return cin * in_h * in_w; }
};

// ValueEncoder constants (util.cpp:45-54)
struct EncoderConsts {
    int support = 0;
    double min_value = 0, max_value = 0, step = 0;
    int contractive = 1;
};
EncoderConsts make_encoder(double min_value, double max_value, bool contractive);

}  // namespace mzb
