// mma_probe.cu — micro-benchmark: cycles per tcgen05.mma (M=128, bf16, K=16) for a given N and operand layout.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I muzero-cpp_b200/csrc -o tools/mma_probe tools/mma_probe.cu
// Timing only: operands are whatever lies in shared memory.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "umma.cuh"

using namespace mzb::umma;

struct ProbeArgs {
    int N, iters, layout;      // layout: 0 = no swizzle (LBO/SBO given), 2 = 128B swizzle, 4 = 64B, 6 = 32B
    int a_lbo, a_sbo, b_lbo, b_sbo;
    int b_off;                 // byte offset of the B start (misalignment experiments)
    int a_same;                // 1 = reuse one A tile (always), else walk through 8 stages
    int fill;                  // 1 = a second warp streams 16 KB bulk copies into the ring while the MMAs run
    long long* out;
    const uint8_t* src;
};

__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo, uint32_t sbo, int layout) {
    uint64_t d = 0;
    d |= (uint64_t)((addr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)((addr >> 7) & 7) << 49;   // base offset (only meaningful for swizzled layouts)
    d |= (uint64_t)layout << 61;
    return d;
}

__device__ __forceinline__ void mma_bf16_parts(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi, uint32_t idesc, int acc) {
    if (acc)
        asm volatile("{\n\t.reg .b64 da, db;\n\t.reg .pred p;\n\tmov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\tsetp.eq.u32 p, 1, 1;\n\t"
                     "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}" ::"r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc) : "memory");
    else
        asm volatile("{\n\t.reg .b64 da, db;\n\t.reg .pred p;\n\tmov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\tsetp.eq.u32 p, 1, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}" ::"r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc) : "memory");
}

__global__ void __launch_bounds__(128, 1) probe_kernel(ProbeArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar, fbar[8];
    __shared__ uint32_t tmem_ptr;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < 200 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    if (tid == 0) { mbar_init(&bar, 1); for (int i = 0; i < 8; ++i) mbar_init(&fbar[i], 1); fence_mbar_init(); }
    if (warp == 0) tmem_alloc(&tmem_ptr, 512);
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_ptr;
    uint8_t* A = smem;                  // 8 stages x 16 KB
    uint8_t* B = smem + 128 * 1024;     // activations
    if (warp == 1 && lane == 0 && a.fill) {
        for (int i = 0; i < a.iters / 4; ++i) {
            const int s = i & 7;
            mbar_expect_tx(&fbar[s], 16384);
            bulk_g2s(smem + 64 * 1024 + (s & 3) * 16384, a.src + (size_t)(i & 63) * 16384, 16384, &fbar[s]);
            mbar_wait(&fbar[s], (i >> 3) & 1);
        }
    }
    if (warp == 0 && lane == 0) {
        const uint32_t idesc = idesc_bf16_f32(128u, (uint32_t)a.N);
        const uint32_t a_addr = smem_u32(A), b_addr = smem_u32(B) + a.b_off;
        // lean issue loop: descriptors precomputed, only the 14-bit address field of the low word changes
        uint32_t a_lo[16], b_lo[4], a_hi, b_hi;
        for (int st = 0; st < 4; ++st)
            for (int kk = 0; kk < 4; ++kk) {
                const uint64_t da = a.layout == 0 ? make_desc(a_addr + st * 16384 + kk * 2 * a.a_lbo, a.a_lbo, a.a_sbo, 0)
                                                  : make_desc(a_addr + st * 16384 + kk * 32, a.a_lbo, a.a_sbo, a.layout);
                a_lo[st * 4 + kk] = (uint32_t)da; a_hi = (uint32_t)(da >> 32);
            }
        for (int kk = 0; kk < 4; ++kk) {
            const uint64_t db = a.layout == 0 ? make_desc(b_addr + kk * 2 * a.b_lbo, a.b_lbo, a.b_sbo, 0) : make_desc(b_addr + kk * 32, a.b_lbo, a.b_sbo, a.layout);
            b_lo[kk] = (uint32_t)db; b_hi = (uint32_t)(db >> 32);
        }
        const long long t0 = clock64();
        mma_bf16_parts(tmem, a_lo[0], a_hi, b_lo[0], b_hi, idesc, 0);
        for (int i = 0; i < a.iters / 16; ++i) {
#pragma unroll
            for (int u = 0; u < 16; ++u) mma_bf16_parts(tmem, a_lo[a.a_same ? 0 : u], a_hi, b_lo[u & 3], b_hi, idesc, 1);
        }
        mma_commit(&bar);
        mbar_wait(&bar, 0);
        const long long t1 = clock64();
        a.out[blockIdx.x] = t1 - t0;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 512);
}

int main(int argc, char** argv) {
    int n_blocks = 148;
    long long* d_out;
    uint8_t* d_src;
    cudaMalloc(&d_out, 1024 * sizeof(long long));
    cudaMalloc(&d_src, 64 * 16384);
    cudaMemset(d_src, 0, 64 * 16384);
    cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    struct Case { const char* name; ProbeArgs a; };
    std::vector<Case> cases;
    auto add = [&](const char* name, int N, int layout, int albo, int asbo, int blbo, int bsbo, int boff, int same, int fill) {
        ProbeArgs p{N, 4000, layout, albo, asbo, blbo, bsbo, boff, same, fill, d_out, d_src};
        cases.push_back({name, p});
    };
    const int R = 243;
    add("noswz N=224 (kernel layout)", 224, 0, 2048, 128, R * 16, 128, 9 * 16, 0, 0);
    add("noswz N=224 + fill", 224, 0, 2048, 128, R * 16, 128, 9 * 16, 0, 1);
    add("noswz N=176 (3-game tile)", 176, 0, 2048, 128, R * 16, 128, 9 * 16, 0, 0);
    add("noswz N=176 + fill", 176, 0, 2048, 128, R * 16, 128, 9 * 16, 0, 1);
    add("noswz N=112 (half tile)", 112, 0, 2048, 128, R * 16, 128, 9 * 16, 0, 0);
    add("noswz N=112 + fill", 112, 0, 2048, 128, R * 16, 128, 9 * 16, 0, 1);
    add("noswz N=224 aligned B, LBO=4096", 224, 0, 2048, 128, 4096, 128, 0, 0, 0);
    add("noswz N=256 aligned", 256, 0, 2048, 128, 4096, 128, 0, 0, 0);
    add("noswz N=128 aligned", 128, 0, 2048, 128, 4096, 128, 0, 0, 0);
    add("noswz N=64 aligned", 64, 0, 2048, 128, 4096, 128, 0, 0, 0);
    add("noswz N=224 A same", 224, 0, 2048, 128, R * 16, 128, 9 * 16, 1, 0);
    add("sw128 N=224 aligned", 224, 2, 16, 1024, 16, 1024, 0, 0, 0);
    add("sw128 N=224 shifted 1 row", 224, 2, 16, 1024, 16, 1024, 128, 0, 0);
    add("sw128 N=224 shifted 9 rows", 224, 2, 16, 1024, 16, 1024, 9 * 128, 0, 0);
    add("sw128 N=224 + fill", 224, 2, 16, 1024, 16, 1024, 128, 0, 1);
    add("sw128 N=256", 256, 2, 16, 1024, 16, 1024, 0, 0, 0);
    add("sw128 N=128", 128, 2, 16, 1024, 16, 1024, 0, 0, 0);
    add("sw64 N=224", 224, 4, 16, 512, 16, 512, 0, 0, 0);
    add("sw32 N=224", 224, 6, 16, 256, 16, 256, 0, 0, 0);
    for (auto& c : cases) {
        for (int rep = 0; rep < 2; ++rep) probe_kernel<<<n_blocks, 128, 200 * 1024>>>(c.a);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("%-36s FAILED: %s\n", c.name, cudaGetErrorString(e)); return 1; }
        std::vector<long long> h(n_blocks);
        cudaMemcpy(h.data(), d_out, n_blocks * sizeof(long long), cudaMemcpyDeviceToHost);
        double s = 0;
        for (auto v : h) s += v;
        const double cyc = s / n_blocks / (c.a.iters / 16 * 16 + 1);
        printf("%-36s %7.1f cycles/MMA  (ideal %5.1f)  operand bytes/cycle %.1f\n", c.name, cyc, c.a.N * 128 / 256.0, (128 + c.a.N) * 32 / cyc);
    }
    return 0;
}
