// conv_first.cuh — the first convolution of ResidualHeadDownsample (model_layers.cpp:140-155: conv3x3, stride 2, padding 1, bias,
// no BN / ReLU; ALE shape 13 planes x 96 x 96 -> 64 x 48 x 48) on tcgen05, reading the caller's NCHW fp32 observation rows
// directly — it replaces a layout-conversion launch plus the CUDA-core direct kernel (0.55 + 3.07 ms per 1024 observations) —
// and the vectorised AvgPool2d(3, 2, 1) of the same head (model_layers.cpp:31-33).
//
// Implicit GEMM without an im2col tile: D[cout][pixel] = sum over the 9 taps of W_tap[cout][16 planes] x X_tap[pixel][16 planes],
// one K = 16 MMA per tap.  With stride 2 the pixels of ONE tap are every second input pixel, so the band of input rows is staged
// DE-INTERLEAVED by row and column parity: class (rp, cp) holds the input pixels with row parity rp and column parity cp as
// 16-byte entries (8 planes, bf16) in row-major order with Wp = Wo + 1 entries per row.  For tap (dy, dx) output pixel
// (oy, ox) reads input (2 oy + dy - 1, 2 ox + dx - 1) = entry (oy + [dy == 2], ox + [dx == 2]) of class ([dy != 1], [dx != 1]):
// consecutive output pixels are consecutive entries, i.e. the K-major no-swizzle B operand of the tap is that class at a
// constant offset — nine descriptors over the resident band, exactly like the row-shift trick of the stride-1 kernels.  Entry 0
// of an odd-column row is input column -1 (zero) and doubles as the zero column that the row wrap of the last pixel lands on.
//   band  = rbo output rows of one image (N = rbo * Wp rounded up to 16 <= 256 accumulator columns)
//   roles = all 8 warps stage the band (coalesced fp32 reads, bf16 scatter), warp 0 issues the 9 MMAs, the warps of the lane
//           quarters that hold output channels read the accumulator, add the bias and stage bf16 NHWC rows, which leave as
//           16-byte stores (the band's output rows are one contiguous run of the NHWC tensor).
#pragma once
#include <cuda_bf16.h>

#include "umma.cuh"

namespace mzb {
namespace first {

using bf16 = __nv_bfloat16;
using namespace umma;

constexpr int C1_THREADS = 256;
constexpr int C1_W_BYTES = 9 * 4096;   // nine tap images [2 K chunks][128 rows][8] bf16

struct FirstArgs {
    const float* obs; long obs_stride;      // [B][cin][H][W] fp32
    bf16* out;                              // [B][Ho][Wo][cout] bf16
    const uint8_t* wimg;                    // C1_W_BYTES
    const float* bias;                      // [cout]
    int cin, cout, H, W, Ho, Wo, relu;
    int pkm;                                // 1: the output is a PKM tensor [cout / 8][Ho + 2][Wo + 1][8] (conv_band.cuh), else NHWC
    int rbo, bands, N, cls;                 // output rows per band, bands per image, MMA columns, entries per parity class
};

inline int first_rows(int Wo) {   // rbo * Wp <= 256 and the epilogue's last 16-column read stays inside the 256 allocated columns
    const int Wp = Wo + 1, w16 = (Wo + 15) / 16 * 16;
    int r = 256 / Wp;
    while (r > 0 && (r - 1) * Wp + w16 > 256) --r;
    return r;
}
inline int first_cls(int N, int Wo) { return N + (Wo + 1) + 8; }
inline size_t first_smem_bytes(int cls, int rbo, int Wo, int cout) {
    size_t in_b = (size_t)2 * 4 * cls * 16, out_b = (size_t)rbo * (Wo + 1) * cout * 2;
    return (((in_b > out_b ? in_b : out_b) + 127) & ~(size_t)127) + C1_W_BYTES + 64;
}

__global__ void __launch_bounds__(C1_THREADS) conv_first_kernel(const __grid_constant__ FirstArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
    const int img = blockIdx.x / a.bands, band = blockIdx.x - img * a.bands;
    const int oy0 = band * a.rbo;
    const int rows_here = min(a.rbo, a.Ho - oy0);
    const int Wp = a.Wo + 1, CLS = a.cls;
    const size_t in_b = (size_t)2 * 4 * CLS * 16, out_b = (size_t)a.rbo * (a.Wo + 1) * a.cout * 2;
    uint8_t* S = smem;                                   // staged band, later the bf16 output rows
    uint8_t* Wt = smem + (((in_b > out_b ? in_b : out_b) + 127) & ~(size_t)127);  // weight images
    uint64_t* w_full = reinterpret_cast<uint64_t*>(Wt + C1_W_BYTES);
    uint64_t* accum_full = w_full + 1;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(accum_full + 1);

    if (tid == 0) {
        mbar_init(w_full, 1);
        mbar_init(accum_full, 1);
        fence_mbar_init();
        mbar_expect_tx(w_full, C1_W_BYTES);
        bulk_g2s(Wt, a.wimg, C1_W_BYTES, w_full);
    }
    // ---- zero the staging area (pads, planes >= cin, rows outside the image), then scatter the band ----
    for (int i = tid; i < (int)(in_b / 16); i += C1_THREADS) reinterpret_cast<uint4*>(S)[i] = make_uint4(0u, 0u, 0u, 0u);
    __syncthreads();
    if (warp == 0) tmem_alloc(tmem_ptr, 256);
    {
        // input rows iy = 2 oy0 - 1 .. 2 (oy0 + rows_here - 1) + 1.  A work item = (plane half, row, column): the thread reads its
        // pixel from up to 8 planes (each read coalesced across the warp: consecutive lanes = consecutive columns), packs them to
        // bf16 and writes ONE 16-byte entry.  (First version: one plane x 4 columns per thread and 2-byte scatter stores — 70 % of
        // the shared-store wavefronts were bank conflicts, ncu.)
        const int iy0 = 2 * oy0 - 1, nrows = 2 * rows_here + 1;
        const int halves = (a.cin + 7) >> 3;
        const int items = halves * nrows * a.W;
        const float* ob = a.obs + (long)img * a.obs_stride;
        const long plane = (long)a.H * a.W;
        // four items per pass: all 32 loads of a pass are in flight before the first is used (one item at a time was a chain of
        // ~8 DRAM round trips per thread — most of the kernel's time)
        constexpr int U = 4;
        for (int e0 = tid; e0 < items; e0 += U * C1_THREADS) {
            float v[U][8];
            uint32_t dst[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int e = e0 + u * C1_THREADS;
                dst[u] = 0xffffffffu;
#pragma unroll
                for (int p = 0; p < 8; ++p) v[u][p] = 0.f;
                if (e >= items) continue;
                const int ix = e % a.W, t = e / a.W, r = t % nrows, h = t / nrows;
                const int iy = iy0 + r;
                if (iy < 0 || iy >= a.H) continue;
                const float* src = ob + (long)(8 * h) * plane + (long)iy * a.W + ix;
                const int np = min(8, a.cin - 8 * h);
#pragma unroll
                for (int p = 0; p < 8; ++p) if (p < np) v[u][p] = __ldg(src + p * plane);
                const int rp = iy & 1, r2 = ((iy + rp) >> 1) - oy0;      // even rows: r2 = iy / 2 - oy0; odd: (iy + 1) / 2 - oy0
                const int cp = ix & 1, c2 = (ix + cp) >> 1;
                dst[u] = (uint32_t)((((size_t)(h * 4 + rp * 2 + cp) * CLS + (size_t)r2 * Wp + c2) * 16));
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (dst[u] == 0xffffffffu) continue;
                uint4 pk;
                __nv_bfloat162* p2 = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
                for (int p = 0; p < 4; ++p) p2[p] = __floats2bfloat162_rn(v[u][2 * p], v[u][2 * p + 1]);
                *reinterpret_cast<uint4*>(S + dst[u]) = pk;
            }
        }
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_ptr;

    if (warp == 0) {
        // ===== nine MMAs, one per tap (K = 16: K chunk 0 = planes 0-7, chunk 1 = planes 8-15) =====
        mbar_wait(w_full, 0);
        tc_fence_after();
        if (elect_one()) {
            const uint32_t idesc = idesc_bf16_f32(128u, (uint32_t)a.N);
            const uint32_t half_bytes = (uint32_t)(4 * CLS * 16);
#pragma unroll
            for (int tap = 0; tap < 9; ++tap) {
                const int dy = tap / 3, dx = tap - dy * 3;
                const int cls_i = (dy != 1 ? 2 : 0) + (dx != 1 ? 1 : 0);
                const uint32_t b_addr = smem_u32(S) + (uint32_t)((cls_i * CLS + (dy == 2 ? Wp : 0) + (dx == 2 ? 1 : 0)) * 16);
                const uint64_t da = smem_desc_kmajor_noswizzle(smem_u32(Wt) + (uint32_t)tap * 4096u, 2048, 128);
                const uint64_t db = smem_desc_kmajor_noswizzle(b_addr, half_bytes, 128);
                mma_bf16(tmem, da, db, idesc, tap != 0 ? 1u : 0u);
            }
            mma_commit(accum_full);
        }
        __syncwarp();
    }
    mbar_wait(accum_full, 0);
    tc_fence_after();
    __syncthreads();   // every warp is past the barrier: the staged band is dead, S becomes the output rows
    // ===== epilogue: lane quarter q = output channels 32 q .. 32 q + 31; the two warps of a quarter split the band's rows =====
    {
        const int q = warp & 3, hsel = warp >> 2;
        const int co = 32 * q + lane;
        if (32 * q < a.cout) {   // warp-uniform
            const bool co_ok = co < a.cout;
            const float bias = co_ok ? __ldg(a.bias + co) : 0.f;
            const uint32_t t_lane = tmem + ((uint32_t)(32 * q) << 16);
            bf16* So = reinterpret_cast<bf16*>(S);
            for (int r = hsel; r < rows_here; r += 2) {
                for (int x0 = 0; x0 < a.Wo; x0 += 16) {
                    uint32_t v[16];
                    tmem_ld16(t_lane + (uint32_t)(r * Wp + x0), v);   // may run past the row: only x < Wo is used
                    tmem_ld_wait();
                    // NHWC staging [row][x][cout], or PKM staging [K chunk][row][x incl. pad column][8]
                    bf16* sp = a.pkm ? So + ((long)((co >> 3) * a.rbo + r) * Wp + x0) * 8 + (co & 7) : So + (long)(r * a.Wo + x0) * a.cout + co;
                    const long sj = a.pkm ? 8 : a.cout;
#pragma unroll
                    for (int j = 0; j < 16; ++j)
                        if (x0 + j < a.Wo && co_ok) {
                            float y = __uint_as_float(v[j]) + bias;
                            if (a.relu) y = fmaxf(y, 0.f);
                            sp[(long)j * sj] = __float2bfloat16(y);
                        }
                    if (a.pkm && co_ok && x0 + 16 >= a.Wo) sp[(long)(a.Wo - x0) * 8] = __float2bfloat16(0.f);   // the pad column
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    {
        // the band's output rows are contiguous in the NHWC tensor: 16-byte coalesced stores
        const uint4* src = reinterpret_cast<const uint4*>(S);
        if (!a.pkm) {
            const long n16 = (long)rows_here * a.Wo * a.cout * 2 / 16;
            uint4* dst = reinterpret_cast<uint4*>(a.out + ((long)img * a.Ho + oy0) * a.Wo * a.cout);
            for (long i = tid; i < n16; i += C1_THREADS) dst[i] = src[i];
        } else {
            // per K chunk the band's rows (pad column included) are one contiguous run of the PKM tensor
            const int KCo = a.cout >> 3, per = rows_here * Wp;
            uint4* dst = reinterpret_cast<uint4*>(a.out + (long)img * KCo * (a.Ho + 2) * Wp * 8);
            for (int i = tid; i < KCo * per; i += C1_THREADS) {
                const int kc = i / per, rem = i - kc * per;
                dst[((long)kc * (a.Ho + 2) + oy0 + 1) * Wp + rem] = src[(long)kc * a.rbo * Wp + rem];
            }
            // zero border rows of the tensor: row 0 by the first band, row Ho + 1 by the last
            if (oy0 == 0)
                for (int i = tid; i < KCo * Wp; i += C1_THREADS) dst[(long)(i / Wp) * (a.Ho + 2) * Wp + i % Wp] = make_uint4(0u, 0u, 0u, 0u);
            if (oy0 + rows_here == a.Ho)
                for (int i = tid; i < KCo * Wp; i += C1_THREADS) dst[((long)(i / Wp) * (a.Ho + 2) + a.Ho + 1) * Wp + i % Wp] = make_uint4(0u, 0u, 0u, 0u);
        }
    }
    if (warp == 0) tmem_dealloc(tmem, 256);
}

// AvgPool2d(3, stride 2, padding 1), count_include_pad = true -> divisor always 9 (model_layers.cpp:31-33), NHWC bf16, a thread =
// 8 channels of one output pixel: nine 16-byte loads, one 16-byte store (HBM-bound: reads the input once, 2.25 taps per element
// from L1/L2)
__global__ void __launch_bounds__(256) avgpool3_vec_kernel(const bf16* __restrict__ in, bf16* __restrict__ out, int B, int C8, int in_h, int in_w,
                                                           int out_h, int out_w, int pkm_in) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const long total = (long)B * out_h * out_w * C8;
    if (i >= total) return;
    const int c8 = (int)(i % C8);
    long r = i / C8;
    const int ox = (int)(r % out_w); r /= out_w;
    const int oy = (int)(r % out_h);
    const int b = (int)(r / out_h);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
        const int iy = oy * 2 + ky - 1;
        if (iy < 0 || iy >= in_h) continue;
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
            const int ix = ox * 2 + kx - 1;
            if (ix < 0 || ix >= in_w) continue;
            const long idx = pkm_in ? (((long)b * C8 + c8) * (in_h + 2) + iy + 1) * (in_w + 1) + ix : (((long)b * in_h + iy) * in_w + ix) * C8 + c8;
            const uint4 v = __ldg(reinterpret_cast<const uint4*>(in + idx * 8));
            const bf16* p = reinterpret_cast<const bf16*>(&v);
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] += __bfloat162float(p[k]);
        }
    }
    uint4 ov;
    bf16* o = reinterpret_cast<bf16*>(&ov);
#pragma unroll
    for (int k = 0; k < 8; ++k) o[k] = __float2bfloat16(acc[k] / 9.0f);
    reinterpret_cast<uint4*>(out)[i] = ov;
}

}  // namespace first
}  // namespace mzb
