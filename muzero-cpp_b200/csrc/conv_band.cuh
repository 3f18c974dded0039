// conv_band.cuh — one 3x3 / stride-1 convolution (+ folded BN, + residual, ReLU) of a board too large for a whole-game tile
// (ResidualBlock stages of ResidualHeadDownsample, model_layers.cpp:116-162: 48x48, 24x24, 12x12), as an implicit GEMM on
// tcgen05.  Same formulation as tower_fused.cuh — zero-padded pixel rows, a tap is a row shift, weights streamed as 16 KB
// images — but the resident tile is a BAND of `rb` output rows of one image plus one halo row above and below:
//   tile row i (0 .. rb+1)  <->  image row y0 - 1 + i;   tile position = i * (W + 1) + x,  x = W is the shared zero column.
// One CTA per band; N = (rb + 2) * (W + 1) <= 256 accumulator columns, of which the inner rb rows are stored.
// Roles: warp 0 weight producer, warp 1 MMA issue (whole warp, elected lane), warps 2-9 epilogue (lane = cout).
#pragma once
#include <cuda_bf16.h>

#include "umma.cuh"

namespace mzb {
namespace band {

using bf16 = __nv_bfloat16;
using namespace umma;

constexpr int B_THREADS = 320;
constexpr int B_NSTAGE = 4;
constexpr int B_KB_BYTES = 16384;        // one 64-wide K stage: [8 K chunks][128 rows][8] bf16
constexpr int B_STAGE_BYTES = 32768;     // two K stages per ring slot
constexpr int B_TMEM_COLS = 256;

struct BandArgs {
    const bf16* in; const int* in_slot; long in_stride;
    bf16* out; const int* out_slot; long out_stride;
    const bf16* res; const int* res_slot; long res_stride;
    const uint8_t* wimg;       // [9 * cin / 64] images of 16 KB
    const float* bias;         // [cout]
    int H, W, cin, cout, relu;
    int stride;                // 1, or 2: the stride-1 result is computed and only the even rows / columns are stored
    int rb, bands;             // output rows per band, bands per image
    int N, R, HALO, x_bytes;
    int nstage;                // ring slots: 2 when that lets two CTAs share an SM (one's input load / epilogue hides behind the other's MMAs), else 4
};

inline int band_rows(int H, int W) {   // largest rb with (rb + 2) * (W + 1) <= 256, at most H
    int rb = 256 / (W + 1) - 2;
    if (rb > H) rb = H;
    return rb;
}
inline size_t band_smem_bytes(int x_bytes, int nstage) { return (size_t)x_bytes + (size_t)nstage * B_STAGE_BYTES + 256; }

__global__ void __launch_bounds__(B_THREADS, 1) conv_band_kernel(const __grid_constant__ BandArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
    const int img = blockIdx.x / a.bands, band = blockIdx.x - img * a.bands;
    const int y0 = band * a.rb;
    const int W1 = a.W + 1, R = a.R, HALO = a.HALO, KC = a.cin >> 3;
    const int rows_tile = (a.rb + 2) * W1;
    uint8_t* X = smem;
    uint8_t* stages = X + a.x_bytes;
    const uint32_t NS = (uint32_t)a.nstage;
    uint64_t* full = reinterpret_cast<uint64_t*>(stages + NS * B_STAGE_BYTES);
    uint64_t* empty = full + B_NSTAGE;
    uint64_t* accum_full = empty + B_NSTAGE;   // (arrays sized for the deepest ring)
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(accum_full + 1);

    if (tid == 0) {
        for (uint32_t s = 0; s < NS; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        mbar_init(accum_full, 1);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_ptr, B_TMEM_COLS);
    // ---- band input -> X (zero padded): 16 bytes = 8 channels of one pixel per request ----
    {
        const long in_base = (long)(a.in_slot ? a.in_slot[img] : img) * a.in_stride;
        for (int e = tid; e < KC * R; e += B_THREADS) {
            const int kc = e % KC, r = e / KC;
            uint8_t* dst = X + ((size_t)kc * R + r) * 16;
            const int q = r - HALO;
            bool ok = false;
            long off = 0;
            if (q >= 0 && q < rows_tile) {
                const int i = q / W1, x = q - i * W1, y = y0 - 1 + i;
                if (x < a.W && y >= 0 && y < a.H) { ok = true; off = in_base + ((long)(y * a.W + x) * a.cin + kc * 8); }
            }
            if (ok) cp_async16(dst, a.in + off);
            else *reinterpret_cast<uint4*>(dst) = make_uint4(0u, 0u, 0u, 0u);
        }
    }
    cp_async_wait_all();
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = __shfl_sync(0xffffffffu, *tmem_ptr, 0);
    const int nkb = 9 * (a.cin >> 6), kpt = a.cin >> 6;

    if (warp == 0) {
        if (lane == 0) {   // ===== weight producer =====
            const uint8_t* src = a.wimg;
            uint32_t it = 0;
            for (int kb = 0; kb < nkb; kb += 2, ++it) {
                const uint32_t s = it % NS;
                const uint32_t bytes = (uint32_t)min(2, nkb - kb) * B_KB_BYTES;
                if (it >= NS) mbar_wait(&empty[s], ((it / NS) - 1) & 1);
                mbar_expect_tx(&full[s], bytes);
                bulk_g2s(stages + s * B_STAGE_BYTES, src, bytes, &full[s]);
                src += bytes;
            }
        }
    } else if (warp == 1) {
        // ===== MMA issue: the whole warp runs the loop (operands stay warp-uniform), one elected lane issues =====
        const bool leader = elect_one();
        const uint32_t idesc = idesc_bf16_f32(128u, (uint32_t)a.N);
        const uint64_t da0 = smem_desc_kmajor_noswizzle(smem_u32(stages), 128 * 16, 128);
        const uint64_t db0 = smem_desc_kmajor_noswizzle(smem_u32(X) + (uint32_t)HALO * 16u, (uint32_t)R * 16u, 128);
        const uint32_t a_hi = (uint32_t)(da0 >> 32), b_hi = (uint32_t)(db0 >> 32);
        const uint32_t a_lo0 = (uint32_t)da0, b_lo0 = (uint32_t)db0;
        const uint32_t kk_b = 2u * (uint32_t)R, cc_b = 8u * (uint32_t)R;
        uint32_t b_cur = b_lo0 - (uint32_t)(a.W + 2);
        int cc = 0, tx = 0;
        uint32_t it = 0;
        for (int kb = 0; kb < nkb; ++it) {
            const uint32_t s = it % NS;
            const uint32_t a_lo = a_lo0 + s * (B_STAGE_BYTES >> 4);
            mbar_wait(&full[s], (it / NS) & 1);
            tc_fence_after();
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if (kb < nkb) {
                    const uint32_t ah = a_lo + h * (B_KB_BYTES >> 4);
                    uint32_t ub;
                    asm volatile("shfl.sync.idx.b32 %0, %1, 0, 0x1f, 0xffffffff;" : "=r"(ub) : "r"(b_cur));
                    if (leader) {
                        mma_bf16_lohi_rt(tmem, ah, a_hi, ub, b_hi, idesc, kb != 0 ? 1u : 0u);
                        mma_bf16_lohi<1>(tmem, ah + 256u, a_hi, ub + kk_b, b_hi, idesc);
                        mma_bf16_lohi<1>(tmem, ah + 512u, a_hi, ub + 2u * kk_b, b_hi, idesc);
                        mma_bf16_lohi<1>(tmem, ah + 768u, a_hi, ub + 3u * kk_b, b_hi, idesc);
                    }
                    ++kb;
                    b_cur += cc_b;
                    if (++cc == kpt) {
                        cc = 0;
                        b_cur += ((tx == 2) ? (uint32_t)(W1 - 2) : 1u) - (uint32_t)kpt * cc_b;
                        tx = (tx == 2) ? 0 : tx + 1;
                    }
                }
            }
            if (leader) mma_commit(&empty[s]);
            __syncwarp();
        }
        if (leader) mma_commit(accum_full);
        __syncwarp();
    } else {
        // ===== epilogue: lane quarter q = output channels 32q..32q+31, the two warps of a quarter split the tile rows =====
        // cout <= 64: the weight images repeat rows 0-63 in rows 64-127 (net_bf16.cu pack_extra), so quarters 2, 3 hold the same
        // channels as 0, 1 and the four warp pairs split the band's rows four ways
        const bool dup = a.cout <= 64;
        const int q = dup ? (warp & 1) : (warp & 3), hsel = dup ? ((warp - 2) >> 2) + 2 * ((warp & 3) >> 1) : (warp - 2) >> 2;
        const int rstep = dup ? 4 : 2;
        const int co = 32 * q + lane;
        const bool co_ok = co < a.cout;
        const uint32_t t_lane = tmem + ((uint32_t)(32 * (warp & 3)) << 16);
        const float bias = co_ok ? __ldg(a.bias + co) : 0.f;
        const long out_base = (long)(a.out_slot ? a.out_slot[img] : img) * a.out_stride;
        const long res_base = a.res ? (long)(a.res_slot ? a.res_slot[img] : img) * a.res_stride : 0;
        mbar_wait(accum_full, 0);
        tc_fence_after();
        // 1. accumulator + bias -> fp32 staging S[local column][cout] in the (now idle) weight ring: a warp writes 128 contiguous
        //    bytes per column.  2. all epilogue threads: 8 channels of one pixel per thread -> + residual, ReLU, one 16-byte
        //    store (the direct lane = channel stores were 2 bytes per thread: 4x slower than the MMAs they followed).
        float* S = reinterpret_cast<float*>(stages);
        if (32 * q < a.cout) {   // warp-uniform: quarters beyond cout hold zero rows
            for (int i = 1 + hsel; i <= a.rb; i += rstep) {
                if (y0 - 1 + i >= a.H) break;   // (warp-uniform) last band of an image whose height is not a multiple of rb
                for (int x0 = 0; x0 < a.W; x0 += 16) {
                    uint32_t v[16];
                    tmem_ld16(t_lane + (uint32_t)(i * W1 + x0), v);   // may run past the row into the next one: only x < W is used
                    tmem_ld_wait();
                    float* sp = S + (long)((i - 1) * a.W + x0) * a.cout + co;
#pragma unroll
                    for (int j = 0; j < 16; ++j)
                        if (x0 + j < a.W && co_ok) sp[(long)j * a.cout] = __uint_as_float(v[j]) + bias;
                }
            }
        }
        asm volatile("bar.sync 1, %0;" ::"n"(B_THREADS - 64) : "memory");
        {
            const int KCo = a.cout >> 3, et = tid - 64, st = a.stride;
            const int rows_here = min(a.rb, a.H - y0);
            // output pixels of this band: image rows y0 .. y0 + rows_here - 1 and columns 0 .. W - 1 that are multiples of the stride
            const int oW = (a.W + st - 1) / st;
            const int orow0 = (y0 + st - 1) / st;
            const int orows = (y0 + rows_here + st - 1) / st - orow0;
            const int items = orows * oW * KCo;
            for (int e = et; e < items; e += B_THREADS - 64) {
                const int kc = e % KCo, op = e / KCo;
                const int oy = op / oW, ox = op - oy * oW;
                const int lp = ((orow0 + oy) * st - y0) * a.W + ox * st;   // position inside the staged band
                const float4 f0 = *reinterpret_cast<const float4*>(S + (long)lp * a.cout + kc * 8);
                const float4 f1 = *reinterpret_cast<const float4*>(S + (long)lp * a.cout + kc * 8 + 4);
                float x[8] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
                const long off = ((long)(orow0 + oy) * oW + ox) * a.cout + kc * 8;
                if (a.res) {
                    const uint4 rv = *reinterpret_cast<const uint4*>(a.res + res_base + off);
                    const bf16* r8 = reinterpret_cast<const bf16*>(&rv);
#pragma unroll
                    for (int k = 0; k < 8; ++k) x[k] += __bfloat162float(r8[k]);
                }
                uint4 ov;
                bf16* o8 = reinterpret_cast<bf16*>(&ov);
#pragma unroll
                for (int k = 0; k < 8; ++k) o8[k] = __float2bfloat16(a.relu ? fmaxf(x[k], 0.f) : x[k]);
                *reinterpret_cast<uint4*>(a.out + out_base + off) = ov;
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem, B_TMEM_COLS);
}


// ------------------------------------------------------------------------------------------------------------------------------
// conv_band2_kernel — the same convolution as a PERSISTENT, software-pipelined kernel over activations kept in a layout made for
// it.  ncu of conv_band_kernel: tensor pipe active 24-45 %; one or two CTAs per SM each doing load -> MMA -> epilogue in sequence,
// and the load itself — 16-byte cp.async requests scattering NHWC pixels into the K-major tile — ran at ~5 B/clk (8.7 k cycles for
// a 45 KB band against 4.6 k cycles of MMAs; measured by switching parts of the work off, profiles/r02_band_breakdown.txt).
//
// PKM layout ("planar K-major, padded") of an activation [C][H][W] in global memory, per image:
//     [C / 8][H + 2][W + 1][8 channels] bf16,   rows 0 and H + 1 and column W are zero
// i.e. exactly the resident tile's layout.  The band of one K chunk (8 channels) — tile rows y0 - 1 .. y0 + rb, pad column
// included — is ONE contiguous run of (rb + 2)(W + 1) 16-byte entries, so a band is loaded by C / 8 bulk copies (UBLKCP) issued
// by one thread, landing on an mbarrier; image borders need no special case (the zero rows are part of the tensor).
//   one CTA per SM walks over bands blockIdx.x, + gridDim.x, ...; two X buffers, two accumulators (2 x 256 TMEM columns)
//   warp 0     weight producer: 16 KB K stages through a ring, running ahead across bands
//   warp 1     MMA issue: band j into accumulator j % 2 once X[j % 2] has landed and the epilogue of band j - 2 has drained it
//   warps 2-9  epilogue of band j: TMEM -> fp32 staging [K chunk][pixel][8] (its own buffer when that fits — the 64-channel
//              stage — else X[j % 2], which band j's finished MMAs no longer read, in row chunks) -> + residual (prefetched under
//              the MMAs), ReLU -> 16-byte PKM stores, pad column and border rows written as zeros; thread 64 issues the bulk
//              copies of band j + 2 as soon as X[j % 2] is free
struct Band2Args {
    BandArgs b;       // in / out / res are PKM tensors of the conv's input / output shape (slot arrays unused)
    int total;        // bands over all images
    int ns;           // ring slots of 16 KB
    int rc;           // band rows per staging chunk
    int sep;          // 1: staging has its own buffer behind the barriers, 0: it aliases X[j % 2]
    int dbg;          // timing experiments (wrong results): 1 = weight stages shrink to 256 B, 2 = no epilogue work, 4 = no band loads
    int ps;           // staging plane stride in floats (pixels of a chunk * 8 + 8: conflict-free lane-quarter stores)
};
constexpr int B2_WORKERS = 256;
inline size_t band2_smem_bytes(int x_bytes, int ns, size_t stage_bytes) { return (size_t)2 * x_bytes + (size_t)ns * B_KB_BYTES + 256 + stage_bytes; }
inline long pkm_elems(int C, int H, int W) { return (long)(C / 8) * (H + 2) * (W + 1) * 8; }

// thread 64 only: band (img, y0) -> X, one bulk copy per K chunk; x_full completes when all bytes have landed
__device__ __forceinline__ void band2_load(const BandArgs& a, uint8_t* X, int img, int y0, uint64_t* x_full, bool skip) {
    const int W1 = a.W + 1, KC = a.cin >> 3;
    const int rows = min(a.rb + 2, a.H + 2 - y0);                 // PKM rows y0 .. (tile rows y0 - 1 ..); the tensor ends at row H + 1
    const uint32_t bytes = skip ? 16u : (uint32_t)(rows * W1 * 16);
    const uint8_t* src = reinterpret_cast<const uint8_t*>(a.in) + ((long)img * a.in_stride + (long)y0 * W1 * 8) * 2;
    const long plane = (long)(a.H + 2) * W1 * 16;
    // the entry in front of the tile (tile row 0, column -1) is read by the top-left tap of the band's first output pixel: it is
    // the previous row's pad column, i.e. zero.  It lies outside the copied run (and under the aliased staging), so it is rewritten
    // for every band; everything else around the tile only feeds outputs that are not stored.
    for (int kc = 0; kc < KC; ++kc) *reinterpret_cast<uint4*>(X + ((size_t)kc * a.R + a.HALO - 1) * 16) = make_uint4(0u, 0u, 0u, 0u);
    fence_proxy_async_smem();
    mbar_expect_tx(x_full, bytes * (uint32_t)KC);
    for (int kc = 0; kc < KC; ++kc) bulk_g2s(X + ((size_t)kc * a.R + a.HALO) * 16, src + kc * plane, bytes, x_full);
}

__global__ void __launch_bounds__(B_THREADS, 1) conv_band2_kernel(const __grid_constant__ Band2Args g) {
    extern __shared__ __align__(128) uint8_t smem[];
    const BandArgs& a = g.b;
    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
    const int W1 = a.W + 1, R = a.R, HALO = a.HALO;
    const uint32_t NS = (uint32_t)g.ns;
    uint8_t* X0 = smem;
    uint8_t* stages = smem + 2 * (size_t)a.x_bytes;
    uint64_t* w_full = reinterpret_cast<uint64_t*>(stages + NS * B_KB_BYTES);
    uint64_t* w_empty = w_full + 8;
    uint64_t* x_full = w_empty + 8;      // [2]
    uint64_t* acc_full = x_full + 2;     // [2]
    uint64_t* acc_free = acc_full + 2;   // [2]
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(acc_free + 2);
    const int n_mine = (g.total - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;   // bands blockIdx.x + j * gridDim.x
    const int nkb = 9 * (a.cin >> 6), kpt = a.cin >> 6;

    if (tid == 0) {
        for (uint32_t s = 0; s < NS; ++s) { mbar_init(&w_full[s], 1); mbar_init(&w_empty[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&x_full[b], 1); mbar_init(&acc_full[b], 1); mbar_init(&acc_free[b], B2_WORKERS / 32); }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_ptr, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_ptr;

    if (warp == 0) {
        if (lane == 0) {   // ===== weight producer =====
            uint32_t it = 0;
            for (int j = 0; j < n_mine; ++j) {
                const uint8_t* src = a.wimg;
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const uint32_t s = it % NS;
                    if (it >= NS) mbar_wait(&w_empty[s], ((it / NS) - 1) & 1);
                    const uint32_t bytes = (g.dbg & 1) ? 256u : (uint32_t)B_KB_BYTES;
                    mbar_expect_tx(&w_full[s], bytes);
                    bulk_g2s(stages + s * B_KB_BYTES, src, bytes, &w_full[s]);
                    src += B_KB_BYTES;
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issue =====
        const bool leader = elect_one();
        const uint32_t idesc = idesc_bf16_f32(128u, (uint32_t)a.N);
        const uint64_t da0 = smem_desc_kmajor_noswizzle(smem_u32(stages), 128 * 16, 128);
        const uint64_t db0 = smem_desc_kmajor_noswizzle(smem_u32(X0) + (uint32_t)HALO * 16u, (uint32_t)R * 16u, 128);
        const uint32_t a_hi = (uint32_t)(da0 >> 32), b_hi = (uint32_t)(db0 >> 32);
        const uint32_t a_lo0 = (uint32_t)da0, b_lo0 = (uint32_t)db0;
        const uint32_t kk_b = 2u * (uint32_t)R, cc_b = 8u * (uint32_t)R;
        uint32_t it = 0;
        for (int j = 0; j < n_mine; ++j) {
            const uint32_t buf = (uint32_t)j & 1u, use = (uint32_t)j >> 1;
            mbar_wait(&x_full[buf], use & 1);
            if (j >= 2) mbar_wait(&acc_free[buf], (use - 1) & 1);
            tc_fence_after();
            const uint32_t d_tmem = tmem + buf * 256u;
            uint32_t b_cur = b_lo0 + buf * ((uint32_t)a.x_bytes >> 4) - (uint32_t)(a.W + 2);
            int cc = 0, tx = 0;
            for (int kb = 0; kb < nkb; ++kb, ++it) {
                const uint32_t s = it % NS;
                const uint32_t ah = a_lo0 + s * (B_KB_BYTES >> 4);
                mbar_wait(&w_full[s], (it / NS) & 1);
                tc_fence_after();
                uint32_t ub;
                asm volatile("shfl.sync.idx.b32 %0, %1, 0, 0x1f, 0xffffffff;" : "=r"(ub) : "r"(b_cur));
                if (leader) {
                    mma_bf16_lohi_rt(d_tmem, ah, a_hi, ub, b_hi, idesc, kb != 0 ? 1u : 0u);
                    mma_bf16_lohi<1>(d_tmem, ah + 256u, a_hi, ub + kk_b, b_hi, idesc);
                    mma_bf16_lohi<1>(d_tmem, ah + 512u, a_hi, ub + 2u * kk_b, b_hi, idesc);
                    mma_bf16_lohi<1>(d_tmem, ah + 768u, a_hi, ub + 3u * kk_b, b_hi, idesc);
                    mma_commit(&w_empty[s]);
                }
                b_cur += cc_b;
                if (++cc == kpt) {
                    cc = 0;
                    b_cur += ((tx == 2) ? (uint32_t)(W1 - 2) : 1u) - (uint32_t)kpt * cc_b;
                    tx = (tx == 2) ? 0 : tx + 1;
                }
                __syncwarp();
            }
            if (leader) mma_commit(&acc_full[buf]);
            __syncwarp();
        }
    } else {
        // ===== epilogue warps =====
        const int wt = tid - 64;
        const bool sep = g.sep != 0, issuer = tid == 64;
        const bool dup = a.cout <= 64;
        const int rq = warp & 3;                                   // TMEM lane quarter of this warp
        const int q = dup ? (rq & 1) : rq;                         // channel quarter
        const int hs = (warp - 2) >> 2;
        const int rsel = dup ? hs + 2 * (rq >> 1) : hs, rstep = dup ? 4 : 2;
        const int co = 32 * q + lane;
        const bool co_ok = co < a.cout;
        const float bias = co_ok ? __ldg(a.bias + co) : 0.f;
        const int KCo = a.cout >> 3, st = a.stride, ps = g.ps;
        const int oW = (a.W + st - 1) / st, oH = (a.H + st - 1) / st, oW1 = oW + 1;
        float* Ssep = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(w_full) + 256);
        if (issuer)
            for (int j = 0; j < 2 && j < n_mine; ++j) {
                const int gb = (int)blockIdx.x + j * (int)gridDim.x, img = gb / a.bands, band = gb - img * a.bands;
                band2_load(a, X0 + (size_t)j * a.x_bytes, img, band * a.rb, &x_full[j], (g.dbg & 4) != 0);
            }
        for (int j = 0; j < n_mine; ++j) {
            const uint32_t buf = (uint32_t)j & 1u, use = (uint32_t)j >> 1;
            const int gb = (int)blockIdx.x + j * (int)gridDim.x, img = gb / a.bands, band = gb - img * a.bands;
            const int y0 = band * a.rb;
            const int rows_here = min(a.rb, a.H - y0);
            const uint32_t t_lane = tmem + buf * 256u + ((uint32_t)(32 * rq) << 16);
            float* S = sep ? Ssep : reinterpret_cast<float*>(X0 + (size_t)buf * a.x_bytes);
            bf16* outp = a.out + (long)img * a.out_stride;
            const bf16* resp = a.res ? a.res + (long)img * a.res_stride : nullptr;
            // item e of a chunk = (K chunk, output row, output column incl. the pad column), column fastest: 16-byte stores of
            // consecutive threads are consecutive in the PKM tensor
            constexpr int PF = 12;
            uint4 rv[PF];
            auto prefetch = [&](int r0, int r1) {   // residual vectors, fetched before they are needed (first chunk: under the MMAs)
                if (!resp) return;
                const int ya = y0 + r0, yb = y0 + r1;
                const int orow0 = (ya + st - 1) / st, orows = (yb + st - 1) / st - orow0;
                const int items = orows * oW1 * KCo;
#pragma unroll
                for (int k = 0; k < PF; ++k) {
                    const int e = wt + k * B2_WORKERS;
                    if (e < items) {
                        const int ox = e % oW1, t = e / oW1, oy = t % orows, kc = t / orows;
                        if (ox < oW) rv[k] = __ldg(reinterpret_cast<const uint4*>(resp + (((long)kc * (oH + 2) + orow0 + oy + 1) * oW1 + ox) * 8));
                    }
                }
            };
            prefetch(0, min(g.rc, rows_here));
            mbar_wait(&acc_full[buf], use & 1);   // all MMAs of the band are complete: X[buf] is free, the accumulator is final
            tc_fence_after();
            auto next_load = [&]() {
                if (issuer && j + 2 < n_mine) {
                    fence_proxy_async_smem();   // (aliased staging) generic-proxy traffic on X[buf] before the bulk copies' writes
                    const int gb2 = (int)blockIdx.x + (j + 2) * (int)gridDim.x, img2 = gb2 / a.bands, band2 = gb2 - img2 * a.bands;
                    band2_load(a, X0 + (size_t)buf * a.x_bytes, img2, band2 * a.rb, &x_full[buf], (g.dbg & 4) != 0);
                }
            };
            if (sep) next_load();
            for (int r0 = 0; r0 < rows_here; r0 += g.rc) {
                const int r1 = min(r0 + g.rc, rows_here);
                if (g.dbg & 2) {   // (timing experiment) release the accumulator, no staging, no stores
                    if (r1 == rows_here) { tc_fence_before(); __syncwarp(); if (lane == 0) mbar_arrive(&acc_free[buf]); }
                    continue;
                }
                if (r0 > 0) prefetch(r0, r1);
                // 1. accumulator rows r0 .. r1 - 1 of the band (+ bias) -> fp32 staging [K chunk][pixel][8]
                if (32 * q < a.cout) {
                    float* sp0 = S + (co >> 3) * ps + (co & 7);
                    for (int r = r0 + rsel; r < r1; r += rstep) {
                        for (int x0 = 0; x0 < a.W; x0 += 16) {
                            uint32_t v[16];
                            tmem_ld16(t_lane + (uint32_t)((r + 1) * W1 + x0), v);
                            tmem_ld_wait();
                            float* sp = sp0 + ((r - r0) * a.W + x0) * 8;
#pragma unroll
                            for (int jj = 0; jj < 16; ++jj)
                                if (x0 + jj < a.W && co_ok) sp[jj * 8] = __uint_as_float(v[jj]) + bias;
                        }
                    }
                }
                if (r1 == rows_here) {   // last TMEM read of this band: the accumulator may be overwritten by band j + 2
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&acc_free[buf]);
                }
                asm volatile("bar.sync 1, %0;" ::"n"(B2_WORKERS) : "memory");
                // 2. + residual, ReLU, 16-byte PKM stores of the output rows that lie in image rows y0 + r0 .. y0 + r1 - 1
                {
                    const int ya = y0 + r0, yb = y0 + r1;
                    const int orow0 = (ya + st - 1) / st, orows = (yb + st - 1) / st - orow0;
                    const int items = orows * oW1 * KCo;
                    auto emit = [&](int e, const uint4& rvec) {
                        const int ox = e % oW1, t = e / oW1, oy = t % orows, kc = t / orows;
                        uint4 ov = make_uint4(0u, 0u, 0u, 0u);
                        if (ox < oW) {
                            const int lp = ((orow0 + oy) * st - ya) * a.W + ox * st;
                            const float4 f0 = *reinterpret_cast<const float4*>(S + kc * ps + lp * 8);
                            const float4 f1 = *reinterpret_cast<const float4*>(S + kc * ps + lp * 8 + 4);
                            float x[8] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
                            if (resp) {
                                const bf16* r8 = reinterpret_cast<const bf16*>(&rvec);
#pragma unroll
                                for (int k = 0; k < 8; ++k) x[k] += __bfloat162float(r8[k]);
                            }
                            bf16* o8 = reinterpret_cast<bf16*>(&ov);
#pragma unroll
                            for (int k = 0; k < 8; ++k) o8[k] = __float2bfloat16(a.relu ? fmaxf(x[k], 0.f) : x[k]);
                        }
                        *reinterpret_cast<uint4*>(outp + (((long)kc * (oH + 2) + orow0 + oy + 1) * oW1 + ox) * 8) = ov;
                    };
#pragma unroll
                    for (int k = 0; k < PF; ++k) {
                        const int e = wt + k * B2_WORKERS;
                        if (e < items) emit(e, rv[k]);
                    }
                    for (int e = wt + PF * B2_WORKERS; e < items; e += B2_WORKERS) {   // (chunks with more items than the prefetch holds)
                        uint4 rvec = make_uint4(0u, 0u, 0u, 0u);
                        if (resp) {
                            const int ox = e % oW1, t = e / oW1, oy = t % orows, kc = t / orows;
                            if (ox < oW) rvec = *reinterpret_cast<const uint4*>(resp + (((long)kc * (oH + 2) + orow0 + oy + 1) * oW1 + ox) * 8);
                        }
                        emit(e, rvec);
                    }
                    // the tensor's zero border rows: row 0 by the image's first band, row oH + 1 by its last
                    if (r0 == 0 && y0 == 0)
                        for (int e = wt; e < KCo * oW1; e += B2_WORKERS)
                            *reinterpret_cast<uint4*>(outp + ((long)(e / oW1) * (oH + 2) * oW1 + e % oW1) * 8) = make_uint4(0u, 0u, 0u, 0u);
                    if (r1 == rows_here && y0 + rows_here == a.H)
                        for (int e = wt; e < KCo * oW1; e += B2_WORKERS)
                            *reinterpret_cast<uint4*>(outp + (((long)(e / oW1) * (oH + 2) + oH + 1) * oW1 + e % oW1) * 8) = make_uint4(0u, 0u, 0u, 0u);
                }
                asm volatile("bar.sync 1, %0;" ::"n"(B2_WORKERS) : "memory");
            }
            if (!sep) next_load();   // (aliased staging) only now is X[buf] free
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem, 512);
}

}  // namespace band
}  // namespace mzb
