// tmem_probe.cu — micro-benchmark of the tower kernel's per-layer epilogue: what bounds it?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I muzero-cpp_b200/csrc -o tools/tmem_probe tools/tmem_probe.cu
//
// One CTA per SM (grid 148), 8 "epilogue" warps (2 per TMEM lane quarter), 224 accumulator columns like a 4-game Connect4 tile.
// Variants (cycles per pass, averaged over `iters` passes, max over warps):
//   0  tcgen05.ld only, all 8 warps (each quarter's two warps split the 224 columns)
//   1  tcgen05.ld only, ONLY quarters 0 and 1 (4 warps)      -> half of (0) if the TMEM read port is shared, equal if per-quarter
//   2  tcgen05.ld only, 4 warps, one per quarter, 224 columns each
//   3  the current epilogue body: ld + bias + ReLU + cvt + run-time-masked 2-byte stores into X
//   4  the lean body: ld + bias + cvt.relu + compile-time pixel mask (W + 1 = 8), unpredicated stores
//   5  as 4, ONLY quarters 0 and 1
//   6  as 4 + fp32 tcgen05.st of the result (the PH_STASH layers)
//   7  as 3 with x32 loads
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "umma.cuh"

using namespace mzb::umma;

constexpr int COLS = 224, P = 56, R = 245, HALO = 9;

__device__ __forceinline__ void sts_u16(uint32_t addr, unsigned short v) { asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"(v) : "memory"); }

template <int J>
__device__ __forceinline__ void pair_masked(uint32_t v0, uint32_t v1, float bias, uint32_t m, uint32_t xa) {
    float y0 = fmaxf(__uint_as_float(v0) + bias, 0.f), y1 = fmaxf(__uint_as_float(v1) + bias, 0.f);
    asm volatile(
        "{\n\t.reg .b32 pk, t0, t1;\n\t.reg .b16 lo, hi;\n\t.reg .pred p0, p1;\n\t"
        "cvt.rn.bf16x2.f32 pk, %1, %0;\n\tmov.b32 {lo, hi}, pk;\n\tand.b32 t0, %2, %4;\n\tand.b32 t1, %2, %5;\n\t"
        "setp.ne.u32 p0, t0, 0;\n\tsetp.ne.u32 p1, t1, 0;\n\t@p0 st.shared.u16 [%3 + %6], lo;\n\t@p1 st.shared.u16 [%3 + %7], hi;\n\t}" ::"f"(y0),
        "f"(y1), "r"(m), "r"(xa), "n"(1u << J), "n"(2u << J), "n"(J * 16), "n"(J * 16 + 16)
        : "memory");
}
// compile-time mask: bit J / J+1 of MASK says whether the column is a pixel
template <int J, uint32_t MASK>
__device__ __forceinline__ void pair_static(uint32_t v0, uint32_t v1, float bias, uint32_t xa) {
    float y0 = __uint_as_float(v0) + bias, y1 = __uint_as_float(v1) + bias;
    uint32_t pk;
    asm volatile("cvt.rn.relu.bf16x2.f32 %0, %2, %1;" : "=r"(pk) : "f"(y0), "f"(y1));
    if (MASK & (1u << J)) sts_u16(xa + J * 16, (unsigned short)(pk & 0xffffu));
    if (MASK & (2u << J)) sts_u16(xa + J * 16 + 16, (unsigned short)(pk >> 16));
}
template <uint32_t MASK>
__device__ __forceinline__ void store16_static(const uint32_t* v, float bias, uint32_t xa) {
    pair_static<0, MASK>(v[0], v[1], bias, xa);   pair_static<2, MASK>(v[2], v[3], bias, xa);
    pair_static<4, MASK>(v[4], v[5], bias, xa);   pair_static<6, MASK>(v[6], v[7], bias, xa);
    pair_static<8, MASK>(v[8], v[9], bias, xa);   pair_static<10, MASK>(v[10], v[11], bias, xa);
    pair_static<12, MASK>(v[12], v[13], bias, xa); pair_static<14, MASK>(v[14], v[15], bias, xa);
}
__device__ __forceinline__ void store16_masked(const uint32_t* v, float bias, uint32_t m, uint32_t xa) {
    pair_masked<0>(v[0], v[1], bias, m, xa);   pair_masked<2>(v[2], v[3], bias, m, xa);
    pair_masked<4>(v[4], v[5], bias, m, xa);   pair_masked<6>(v[6], v[7], bias, m, xa);
    pair_masked<8>(v[8], v[9], bias, m, xa);   pair_masked<10>(v[10], v[11], bias, m, xa);
    pair_masked<12>(v[12], v[13], bias, m, xa); pair_masked<14>(v[14], v[15], bias, m, xa);
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* v) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, "
        "%23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]),
          "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
          "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
}

// 16 lanes x 256 bits, 4 repeats = 32 columns: thread t holds lanes (t / 4) and (t / 4) + 8 of the 16-lane group, columns
// 8 k + 2 (t % 4) + {0, 1} in registers 4 k + {0, 1} (lane t/4) and 4 k + {2, 3} (lane t/4 + 8) — the mma C-fragment layout
__device__ __forceinline__ void tmem_ld_16x256b_x4(uint32_t taddr, uint32_t* v) {
    asm volatile(
        "tcgen05.ld.sync.aligned.16x256b.x4.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_16x256b_x2(uint32_t taddr, uint32_t* v) {
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_st_32x32b_x1(uint32_t taddr, uint32_t v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(taddr), "r"(v) : "memory");
}

__global__ void __launch_bounds__(320, 1) probe(int variant_in, int iters, long long* out, float* sink) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint32_t tmem_ptr;
    __shared__ uint32_t pixmask[8];
    __shared__ uint64_t done_bar;
    const bool cadence = variant_in >= 200;   // every pass starts with all 8 warps released together by a barrier, then an idle gap (the tower kernel's per-layer rhythm)
    variant_in %= 200;
    const bool pollers = variant_in >= 100;   // warps 0 (one lane) and 1 (all lanes) poll an mbarrier meanwhile, like the tower kernel's producer / MMA warps
    variant_in %= 100;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) { mbar_init(&done_bar, 8); fence_mbar_init(); }
    const int variant = variant_in % 10;
    const bool fence_each = variant_in >= 10;
    for (int i = tid; i < 64 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    if (tid < 8) {
        uint32_t m = 0;
        for (int j = 0; j < 32; ++j) { const int r = tid * 32 + j, yy = r / 8, xx = r % 8; if (r < P && yy >= 1 && yy <= 6 && xx < 7) m |= 1u << j; }
        pixmask[tid] = m;
    }
    if (warp == 1) tmem_alloc(&tmem_ptr, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_ptr;
    float acc = 0.f;
    long long t0 = 0, t1 = 0;
    if (warp >= 2) {
        const int q = warp & 3, hsel = (warp - 2) >> 2;
        const int ch = 32 * q + lane;
        const uint32_t t_lane = tmem + ((uint32_t)(32 * q) << 16);
        const uint32_t xs = smem_u32(smem) + (uint32_t)(((ch >> 3) * R + HALO) * 16 + (ch & 7) * 2);
        const uint32_t pm = smem_u32(pixmask);
        const float bias = 0.01f * ch;
        const bool only01 = variant == 1 || variant == 5;
        const bool one_per_q = variant == 2;
        const bool active = only01 ? q < 2 : (one_per_q ? hsel == 0 : true);
        asm volatile("bar.sync 1, 256;" ::: "memory");
        t0 = clock64();
        long long spent = 0, tp = 0;
        if (active) {
            for (int it = 0; it < iters; ++it) {
                if (cadence) {
                    if (it > 0) { spent += clock64() - tp; const long long w0 = clock64(); while (clock64() - w0 < 8000) {} }
                    asm volatile("bar.sync 1, 256;" ::: "memory");
                    tp = clock64();
                }
                // the games of this warp: hsel, hsel + 2 (two per warp) — or all four when one warp per quarter
                if (variant == 9) {
                    const int gq = lane >> 2, qq = lane & 3;
                    const bool relu = iters > 0, stash = false, wx = true;
                    for (int half = 0; half < 2; ++half) {
                        const int g16 = 2 * q + half;
                        const float bA = 0.01f * (16 * g16 + gq), bB = 0.01f * (16 * g16 + 8 + gq);
                        const uint32_t t16 = tmem + ((uint32_t)(32 * q + 16 * half) << 16);
                        const uint32_t xb = smem_u32(smem) + (uint32_t)((((2 * g16 + (gq >> 2)) * R + HALO + 2 * qq) * 16) + (gq & 3) * 4);
                        for (int i = 0; i < 2; ++i) {
                            const int g = hsel + 2 * i, c_lo = 8, c_hi = P;
                            const int cbase = g * P;
                            for (int c0 = c_lo; c0 < c_hi; c0 += 32) {
                                const int left = c_hi - c0;
                                uint32_t v[16];
                                if (left > 16) tmem_ld_16x256b_x4(t16 + cbase + c0, v);
                                else tmem_ld_16x256b_x2(t16 + cbase + c0, v);
                                const int mw_i = c0 >> 5, mw_s = c0 & 31;
                                uint32_t m, m2;
                                asm volatile("ld.shared.u32 %0, [%1];" : "=r"(m) : "r"(pm + mw_i * 4) : "memory");
                                m >>= mw_s;
                                if (mw_s && mw_i + 1 < 8) { asm volatile("ld.shared.u32 %0, [%1];" : "=r"(m2) : "r"(pm + (mw_i + 1) * 4) : "memory"); m |= m2 << (32 - mw_s); }
                                if (left < 32) m &= (1u << left) - 1u;
                                m = wx ? (m >> (2 * qq)) : 0u;
                                tmem_ld_wait();
                                const uint32_t xa = xb + (uint32_t)(cbase + c0) * 16;
#pragma unroll
                                for (int k = 0; k < 4; ++k) {
                                    if (8 * k < left) {
                                        const float y00 = __uint_as_float(v[4 * k]) + bA, y01 = __uint_as_float(v[4 * k + 1]) + bA;
                                        const float y10 = __uint_as_float(v[4 * k + 2]) + bB, y11 = __uint_as_float(v[4 * k + 3]) + bB;
                                        uint32_t p0, p1;
                                        asm volatile("cvt.rn.relu.bf16x2.f32 %0, %2, %1;" : "=r"(p0) : "f"(y00), "f"(y10));
                                        asm volatile("cvt.rn.relu.bf16x2.f32 %0, %2, %1;" : "=r"(p1) : "f"(y01), "f"(y11));
                                        if ((m >> (8 * k)) & 1u) asm volatile("st.shared.u32 [%0], %1;" ::"r"(xa + k * 128), "r"(p0) : "memory");
                                        if ((m >> (8 * k + 1)) & 1u) asm volatile("st.shared.u32 [%0], %1;" ::"r"(xa + k * 128 + 16), "r"(p1) : "memory");
                                    }
                                }
                                (void)relu; (void)stash;
                            }
                        }
                    }
                    if (fence_each) { fence_proxy_async_smem(); tc_fence_before(); __syncwarp(); }
                    continue;
                }
                if (variant == 8) {
                    // channel pairs (g, g + 8) of a 16-lane group from ONE thread: 4-byte stores, half as many store instructions.
                    // X positions: channel 16 G + g -> 2 g, channel 16 G + 8 + g -> 2 g + 1 (the weights' K order is permuted to match)
                    const int gq = lane >> 2, qq = lane & 3;
                    for (int g = hsel; g < 4; g += 2) {
                        const int cbase = g * P;
                        for (int half = 0; half < 2; ++half) {
                            const int G16 = 2 * q + half;   // 16-channel group of this quarter
                            const uint32_t t16 = tmem + ((uint32_t)(32 * q + 16 * half) << 16);
                            const float b0 = 0.01f * (16 * G16 + gq), b1 = 0.01f * (16 * G16 + 8 + gq);
                            const uint32_t xbase = smem_u32(smem) + (uint32_t)((((2 * G16 + (gq >> 2)) * R + HALO) * 16) + (gq & 3) * 4);
                            for (int c0 = 8; c0 < P; c0 += 32) {
                                uint32_t v[16];
                                tmem_ld_16x256b_x4(t16 + cbase + c0, v);
                                tmem_ld_wait();
                                const uint32_t xa = xbase + (uint32_t)(cbase + c0 + 2 * qq) * 16;
#pragma unroll
                                for (int k = 0; k < 4; ++k) {
                                    if (c0 + 8 * k >= P) break;
                                    const float y00 = __uint_as_float(v[4 * k]) + b0, y01 = __uint_as_float(v[4 * k + 1]) + b0;
                                    const float y10 = __uint_as_float(v[4 * k + 2]) + b1, y11 = __uint_as_float(v[4 * k + 3]) + b1;
                                    uint32_t p0, p1;
                                    asm volatile("cvt.rn.relu.bf16x2.f32 %0, %2, %1;" : "=r"(p0) : "f"(y00), "f"(y10));
                                    asm volatile("cvt.rn.relu.bf16x2.f32 %0, %2, %1;" : "=r"(p1) : "f"(y01), "f"(y11));
                                    asm volatile("st.shared.u32 [%0], %1;" ::"r"(xa + k * 128), "r"(p0) : "memory");
                                    if (qq != 3) asm volatile("st.shared.u32 [%0], %1;" ::"r"(xa + k * 128 + 16), "r"(p1) : "memory");
                                }
                            }
                        }
                    }
                    if (fence_each) { fence_proxy_async_smem(); tc_fence_before(); __syncwarp(); }
                    continue;
                }
                for (int g = one_per_q ? 0 : hsel; g < 4; g += one_per_q ? 1 : 2) {
                    const int cbase = g * P;
                    for (int c0 = 8; c0 < P; c0 += 32) {
                        const bool two = P - c0 > 16;
                        uint32_t v[32];
                        if (variant == 7) { tmem_ld32(t_lane + cbase + c0, v); }
                        else { tmem_ld16(t_lane + cbase + c0, v); if (two) tmem_ld16(t_lane + cbase + c0 + 16, v + 16); }
                        uint32_t m = 0;
                        if (variant == 3 || variant == 7) {
                            const int mw_i = c0 >> 5, mw_s = c0 & 31;
                            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(m) : "r"(pm + mw_i * 4) : "memory");
                            m >>= mw_s;
                            if (mw_s && mw_i + 1 < 8) { uint32_t m2; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(m2) : "r"(pm + (mw_i + 1) * 4) : "memory"); m |= m2 << (32 - mw_s); }
                        }
                        tmem_ld_wait();
                        const uint32_t xa = xs + (uint32_t)(cbase + c0) * 16;
                        if (variant <= 2) {
#pragma unroll
                            for (int j = 0; j < 32; ++j) acc += __uint_as_float(v[j]);
                        } else if (variant == 3 || variant == 7) {
                            store16_masked(v, bias, m, xa);
                            if (two) store16_masked(v + 16, bias, m >> 16, xa + 256);
                        } else {
                            store16_static<0x7F7Fu>(v, bias, xa);
                            if (two) store16_static<0x7F7Fu>(v + 16, bias, xa + 256);
                            if (variant == 6) {
#pragma unroll
                                for (int k = 0; k < 4; ++k)
                                    if (c0 + 8 * k < P) tmem_st8(t_lane + 256 + cbase + c0 + 8 * k, v + 8 * k);
                            }
                        }
                    }
                }
                if (variant == 6) tmem_st_wait();
                if (fence_each) { fence_proxy_async_smem(); tc_fence_before(); __syncwarp(); }
            }
        }
        t1 = clock64();
        if (cadence) { spent += t1 - tp; t1 = t0 + spent; }
        __syncwarp();
        if (lane == 0) mbar_arrive(&done_bar);
        asm volatile("bar.sync 1, 256;" ::: "memory");
    } else if (pollers && (warp == 1 || lane == 0)) {
        mbar_wait(&done_bar, 0);
    }
    if (acc == 123.456f) sink[0] = acc;
    if (warp >= 2 && lane == 0) out[(long)blockIdx.x * 8 + (warp - 2)] = t1 - t0;
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem, 512);
}

// register mapping of tcgen05.ld.16x256b: lane l, column c of TMEM hold the value 1000 l + c (written with 32x32b stores);
// every thread reports what it received
__global__ void mapping_check(int* out) {
    __shared__ uint32_t tmem_ptr;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) tmem_alloc(&tmem_ptr, 32);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_ptr;
    if (warp == 0) {
        for (int c = 0; c < 32; ++c) tmem_st_32x32b_x1(tmem + c, (uint32_t)(1000 * lane + c));
        tmem_st_wait();
        uint32_t v[16];
        tmem_ld_16x256b_x4(tmem, v);
        tmem_ld_wait();
        for (int i = 0; i < 16; ++i) out[lane * 16 + i] = (int)v[i];
        tmem_ld_16x256b_x4(tmem + (16u << 16), v);
        tmem_ld_wait();
        for (int i = 0; i < 16; ++i) out[512 + lane * 16 + i] = (int)v[i];
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 32);
}

int main() {
    {
        int* d; cudaMalloc(&d, 1024 * sizeof(int));
        mapping_check<<<1, 32>>>(d);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("mapping_check: %s\n", cudaGetErrorString(e)); return 1; }
        std::vector<int> h(1024);
        cudaMemcpy(h.data(), d, 1024 * sizeof(int), cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int half = 0; half < 2; ++half)
            for (int t = 0; t < 32; ++t)
                for (int i = 0; i < 16; ++i) {
                    const int k = i / 4, r = i % 4;
                    const int want = 1000 * (16 * half + t / 4 + (r >= 2 ? 8 : 0)) + 8 * k + 2 * (t % 4) + (r & 1);
                    if (h[half * 512 + t * 16 + i] != want) { if (bad++ < 8) printf("16x256b mapping: half %d thread %d reg %d = %d, expected %d\n", half, t, i, h[half * 512 + t * 16 + i], want); }
                }
        printf("16x256b.x4 register mapping %s\n", bad ? "DIFFERS from the assumed one" : "as assumed (lanes t/4 and t/4+8, columns 8k + 2(t%%4) + {0,1})");
    }
    long long* d_out; float* d_sink;
    cudaMalloc(&d_out, 148 * 8 * sizeof(long long));
    cudaMalloc(&d_sink, 4);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int iters = 50;
    const char* names[10] = {"ld only, 8 warps", "ld only, quarters 0+1 (4 warps)", "ld only, 1 warp per quarter", "current body (masked stores)",
                            "lean body (static mask, cvt.relu)", "lean body, quarters 0+1", "lean body + fp32 stash", "current body, x32 loads", "16x256b loads, channel pairs, 4-byte stores", "as 8, the kernel's run-time masks / x2 tail / order"};
    for (int rep = 0; rep < 2; ++rep)
        for (int vi = 0; vi < 21; ++vi) {
            static const int order[21] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 19, 13, 14, 16, 18, 103, 109, 203, 209, 208, 200};
            const int v = order[vi];
            cudaMemset(d_out, 0, 148 * 8 * sizeof(long long));
            probe<<<148, 320, 200 * 1024>>>(v, iters, d_out, d_sink);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("variant %d: %s\n", v, cudaGetErrorString(e)); return 1; }
            std::vector<long long> h(148 * 8);
            cudaMemcpy(h.data(), d_out, h.size() * sizeof(long long), cudaMemcpyDeviceToHost);
            double mx = 0, sum = 0; int n = 0;
            for (int b = 0; b < 148; ++b) { long long m = 0; for (int w = 0; w < 8; ++w) m = h[b * 8 + w] > m ? h[b * 8 + w] : m; sum += (double)m; ++n; if ((double)m > mx) mx = (double)m; }
            if (rep == 1) printf("variant %2d%s %-44s  %8.0f cycles per pass (mean over CTAs of the slowest warp), max %8.0f\n", v % 10, v >= 200 ? " +cadence   " : (v >= 100 ? " +pollers   " : (v >= 10 ? " +fence/pass" : "            ")), names[v % 10], sum / n / iters, mx / iters);
        }
    return 0;
}
