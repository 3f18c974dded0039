// coreside_probe.cu — does a small block get onto an SM beside a CTA that holds (nearly) all of its shared memory / most of its
// registers?  X = one CTA per SM spinning for ~200 us with a given dynamic shared-memory size and register footprint;
// Y = 148 blocks x 128 threads, launched on another stream 20 us later.  Reports when Y finished relative to X's start.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o coreside_probe tools/coreside_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

template <int NREG, int TMEM = 0>
__global__ void __launch_bounds__(320, 1) x_kernel(long long ns, float* out) {
    extern __shared__ float sm[];
    if (TMEM) {   // what the tower does: one warp allocates all 512 tensor-memory columns
        if (threadIdx.x < 32) {
            unsigned a = (unsigned)__cvta_generic_to_shared(sm);
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(a), "r"(512) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
        __syncthreads();
    }
    float r[NREG];
#pragma unroll
    for (int i = 0; i < NREG; ++i) r[i] = threadIdx.x * 0.5f + i;
    long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
#pragma unroll
        for (int i = 0; i < NREG; ++i) r[i] = r[i] * 1.0001f + r[(i + 1) % NREG];
        long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (t - t0 > ns) break;
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NREG; ++i) s += r[i];
    if (TMEM) {
        __syncthreads();
        if (threadIdx.x < 32) { unsigned t = ((volatile unsigned*)sm)[0]; asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(t), "r"(512) : "memory"); }
        __syncthreads();
    }
    sm[threadIdx.x + 32] = s;
    if (s == 12345.678f) out[0] = sm[threadIdx.x + 32];
}
template <int YREG>
__global__ void __launch_bounds__(128) y_kernel(long long* stamp, double* sink) {
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    if (threadIdx.x == 0) stamp[blockIdx.x] = t;
    if (YREG > 1) {   // a register / fp64 footprint like the tree step's
        double r[YREG];
#pragma unroll
        for (int i = 0; i < YREG; ++i) r[i] = threadIdx.x + i * 0.25;
        for (int k = 0; k < 64; ++k) {
#pragma unroll
            for (int i = 0; i < YREG; ++i) r[i] = r[i] * 1.0000001 + r[(i + 1) % YREG];
        }
        double s = 0;
#pragma unroll
        for (int i = 0; i < YREG; ++i) s += r[i];
        if (s == 1.2345) sink[0] = s;
    }
}
__global__ void stamp_kernel(long long* stamp) { long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); *stamp = t; }

template <int NREG, int TMEM, int YREG>
int run(int smem, int carve, int yblocks = 148, int ythreads = 128) {
    cudaStream_t a, b;
    CK(cudaStreamCreateWithFlags(&a, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&b, cudaStreamNonBlocking));
    long long* d; float* o;
    CK(cudaMalloc(&d, 4096 * 8)); CK(cudaMalloc(&o, 4));
    CK(cudaFuncSetAttribute(x_kernel<NREG, TMEM>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    if (carve >= 0) CK(cudaFuncSetAttribute(y_kernel<YREG>, cudaFuncAttributePreferredSharedMemoryCarveout, carve));
    cudaFuncAttributes fa, fy;
    CK(cudaFuncGetAttributes(&fa, x_kernel<NREG, TMEM>));
    CK(cudaFuncGetAttributes(&fy, y_kernel<YREG>));
    for (int rep = 0; rep < 2; ++rep) {
        stamp_kernel<<<1, 1, 0, a>>>(d + 4000);
        x_kernel<NREG, TMEM><<<148, 320, smem, a>>>(200000, o);
        CK(cudaStreamSynchronize(b));
        // ~20 us later on the other stream
        stamp_kernel<<<1, 1, 0, b>>>(d + 4001);
        y_kernel<YREG><<<yblocks, ythreads, 0, b>>>(d, (double*)o);
        CK(cudaDeviceSynchronize());
    }
    static long long h[4096];
    CK(cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost));
    long long mx = 0, mn = 1ll << 62;
    for (int i = 0; i < yblocks; ++i) { if (h[i] > mx) mx = h[i]; if (h[i] < mn) mn = h[i]; }
    int early = 0;
    for (int i = 0; i < yblocks; ++i) early += (h[i] - h[4000]) * 1e-3 < 150.0;
    printf("[%d of %d Y blocks started while X ran] ", early, yblocks);
    printf("X: %3d regs/thread, TMEM %d, %6d B dynamic smem | Y: %d x %d threads, %2d regs, carve-out pref %3d: Y blocks started %7.1f .. %7.1f us after X's start (X spins 200 us)%s\n",
           fa.numRegs, TMEM, smem, yblocks, ythreads, fy.numRegs, carve, (mn - h[4000]) * 1e-3, (mx - h[4000]) * 1e-3, (mx - h[4000]) * 1e-3 < 150.0 ? "  -> CO-RESIDENT" : "  -> waited for X");
    cudaFree(d); cudaFree(o); cudaStreamDestroy(a); cudaStreamDestroy(b);
    return 0;
}
int main() {
    // the register file is per SM sub-partition (16 K registers each): 10 warps of ~160 registers fill two of the four
    run<150, 1, 1>(231136, 100);
    run<150, 1, 24>(231136, 100);
    run<150, 1, 24>(231136, 100, 296, 64);
    run<150, 1, 24>(231136, 100, 592, 32);
    run<150, 1, 24>(231136, 100, 1024, 32);
    return 0;
}
