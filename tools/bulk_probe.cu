// bulk_probe.cu — micro-benchmark: L2 -> shared memory streaming rate of 1-D bulk copies (the fused tower's weight ring),
// unicast versus cluster multicast (each CTA of a cluster fetches 1/csz of every slot and multicasts it to all members).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I muzero-cpp_b200/csrc -o tools/bulk_probe tools/bulk_probe.cu
// Output: bytes per clock per SM that LAND in shared memory, for grids of 1..148 CTAs.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "umma.cuh"

using namespace mzb::umma;

constexpr int NSTAGE = 4;
constexpr int SLOT = 32768;

struct Args {
    const uint8_t* src;      // weights: region_bytes, every CTA streams all of it, slot after slot
    int region_slots;        // slots in the region
    int iters;               // slots to stream per CTA
    int csz;                 // cluster size (1 = unicast)
    int split;               // unicast: bulk copies per slot
    int nstage;              // slots in flight (<= NSTAGE)
    long long* out;          // cycles per CTA
};

__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void bulk_g2s_mc(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar, uint16_t mask) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;"
                 ::"r"(smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(smem_u32(bar)), "h"(mask) : "memory");
}
__device__ __forceinline__ void mbar_arrive_remote(uint64_t* bar, uint32_t cta) {
    uint32_t ra;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(smem_u32(bar)), "r"(cta));
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(ra) : "memory");
}

__global__ void __launch_bounds__(64, 1) probe_kernel(Args a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t full[NSTAGE], empty[NSTAGE];
    const int tid = threadIdx.x;
    const uint32_t rank = a.csz > 1 ? cluster_ctarank() : 0;
    if (tid == 0) {
        for (int s = 0; s < NSTAGE; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], a.csz); }
        fence_mbar_init();
    }
    __syncthreads();
    if (a.csz > 1) cluster_sync();
    const uint32_t part = SLOT / a.csz;
    const uint16_t mask = (uint16_t)((1u << a.csz) - 1);
    long long t0 = 0, t1 = 0;
    if (tid == 0) {   // producer: this CTA's share of every slot, to every CTA of the cluster
        t0 = clock64();
        for (int it = 0; it < a.iters; ++it) {
            const int s = it % a.nstage;
            if (it >= a.nstage) mbar_wait(&empty[s], ((it / a.nstage) - 1) & 1);
            mbar_expect_tx(&full[s], SLOT);
            const uint8_t* src = a.src + (size_t)(it % a.region_slots) * SLOT + rank * part;
            if (a.csz == 1) { const uint32_t pb = SLOT / a.split; for (int k = 0; k < a.split; ++k) bulk_g2s(smem + s * SLOT + k * pb, src + k * pb, pb, &full[s]); }
            else bulk_g2s_mc(smem + s * SLOT + rank * part, src, part, &full[s], mask);
        }
    } else if (tid == 32) {   // consumer: frees a slot in every CTA of the cluster as soon as it has landed here
        for (int it = 0; it < a.iters; ++it) {
            const int s = it % a.nstage;
            mbar_wait(&full[s], (it / a.nstage) & 1);
            if (a.csz == 1) mbar_arrive(&empty[s]);
            else for (uint32_t c = 0; c < (uint32_t)a.csz; ++c) mbar_arrive_remote(&empty[s], c);
        }
        t1 = clock64();
    }
    __shared__ long long st0;
    if (tid == 0) st0 = t0;
    __syncthreads();
    if (tid == 32) a.out[blockIdx.x] = t1 - st0;
    if (a.csz > 1) cluster_sync();
}

int main() {
    const int region_slots = 160;   // 5 MB of weights, as one recurrent pass
    uint8_t* d_src; long long* d_out;
    cudaMalloc(&d_src, (size_t)region_slots * SLOT);
    cudaMemset(d_src, 1, (size_t)region_slots * SLOT);
    cudaMalloc(&d_out, 1024 * sizeof(long long));
    cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, NSTAGE * SLOT);
    cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    const int iters = 1600;
    struct Case { int csz, split, nstage; };
    for (Case c : {Case{1, 1, 4}, Case{1, 2, 4}, Case{1, 4, 4}, Case{1, 8, 4}, Case{1, 16, 4}, Case{1, 1, 2}, Case{1, 4, 2}, Case{1, 1, 1}, Case{1, 4, 1}, Case{1, 16, 1}, Case{2, 1, 4}}) {
        for (int grid : {8, 148}) {
            const int csz = c.csz;
            if (grid % csz) continue;
            Args a{d_src, region_slots, iters, csz, c.split, c.nstage, d_out};
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(grid); cfg.blockDim = dim3(64); cfg.dynamicSmemBytes = NSTAGE * SLOT;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = csz; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            cudaError_t e = cudaSuccess;
            for (int rep = 0; rep < 2 && e == cudaSuccess; ++rep) e = cudaLaunchKernelEx(&cfg, probe_kernel, a);
            if (e == cudaSuccess) e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("csz %d grid %3d FAILED: %s\n", csz, grid, cudaGetErrorString(e)); cudaGetLastError(); continue; }
            std::vector<long long> h(grid);
            cudaMemcpy(h.data(), d_out, grid * sizeof(long long), cudaMemcpyDeviceToHost);
            double s = 0, mx = 0;
            for (auto v : h) { s += v; if (v > mx) mx = (double)v; }
            const double cyc = s / grid;
            printf("csz %d split %2d stages %d grid %3d: %8.0f cycles per 32 KB slot -> %6.1f B/clk/SM landed, %7.0f B/clk chip\n",
                   csz, c.split, c.nstage, grid, cyc / iters, (double)SLOT * iters / cyc, (double)SLOT * iters / cyc * grid);
        }
    }
    return 0;
}
