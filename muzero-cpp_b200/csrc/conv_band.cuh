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

}  // namespace band
}  // namespace mzb
