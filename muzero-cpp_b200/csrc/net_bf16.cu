// net_bf16.cu — MZB_BF16 backend: the residual towers as implicit-GEMM tcgen05 kernels (bf16 operands, fp32
// accumulation in tensor memory), activations stored in bf16 NHWC.  Replaces conv3x3+BN+ReLU(+residual) of
// ResidualHead / ResidualBlock (model_layers.cpp:70-109) for every layer whose shape fits the tensor path; the few
// that do not (3-channel representation head, strided downsample convs) run on the direct kernels of net_direct.cuh.
//
// Formulation (per CTA = G whole games):   D[cout, pixel] = sum_{tap, cin} Wf[cout, tap, cin] * X[pixel + shift(tap), cin]
//   * A operand  = BN-folded weights, bf16, pre-packed on the host as ready-made shared-memory images of
//                  [64 K][128 cout] blocks; streamed by 1-D bulk async copies (UBLKCP) through a ring of stages.
//   * B operand  = the CTA's activations, loaded ONCE from HBM/L2 into a zero-padded pixel layout: every game owns
//                  P = (H+1)*(W+1) rows: one zero row of W+1 pixels, then H rows of W real pixels + 1 zero pixel.  In this
//                  layout a 3x3 tap is a constant row shift dy*(W+1)+dx, so the nine taps are nine shared-memory
//                  matrix descriptors over the same resident tile: no im2col copy, no re-read of activations.
//                  No-swizzle K-major core matrices ([K chunk of 8][row][8 elems]) make a row shift a 16-byte offset.
//   * D          = 128 TMEM lanes (cout) x N columns (padded pixels), fp32.
//   * epilogue   = tcgen05.ld, + bias (+ per-action table) (+ residual), ReLU, bf16 store of the real pixels only.
// Roles: warp 0 streams weights, warp 1 allocates TMEM and issues the MMAs (one thread), warps 2-5 run the epilogue.
#include <cuda_bf16.h>

#include <algorithm>
#include <cstdlib>
#include <memory>

#include <map>

#include "conv_band.cuh"
#include "conv_first.cuh"
#include "net_direct.cuh"
#include "tower_fused.cuh"
#include "umma.cuh"

namespace mzb {
namespace {

using bf16 = __nv_bfloat16;
using namespace umma;

constexpr int UM = 128;                 // MMA M = cout rows (zero-padded)
constexpr int KB = 64;                  // K elements per weight stage
constexpr int STAGE_BYTES = KB * UM * 2;  // 16 KB
constexpr int NSTAGE = 3;
constexpr int UMMA_THREADS = 192;
constexpr int TMEM_COLS = 256;

struct UmmaGeom {
    int H, W, P, G, N, HALO, R;   // rows per game, games per tile, MMA N, halo rows, total rows of the resident tile
};

struct UmmaArgs {
    const bf16* in; const int* in_slot; long in_stride;
    bf16* out; const int* out_slot; long out_stride;
    const bf16* res; const int* res_slot; long res_stride;
    const bf16* wpack;          // [9 * C/64][8][128][8]
    const float* bias;          // [cout]
    const float* act_table;     // [A][HW][cout] or null
    const int* action;
    int B, C, cout, relu;
    UmmaGeom g;
    long long* dbg;             // optional per-CTA phase clocks {start, tile staged, accumulator ready, done}
};

__host__ __device__ inline size_t umma_x_bytes(int C, int R) { return (((size_t)(C / 8) * R * 16) + 127) / 128 * 128; }
inline size_t umma_smem_bytes(int C, const UmmaGeom& g) {
    return umma_x_bytes(C, g.R) + (size_t)NSTAGE * STAGE_BYTES + 128 /*barriers + tmem ptr*/ + (size_t)(256 /*colinfo*/ + 256 /*pixinfo*/ + 4 * 64 /*per-game offsets*/) * 4 + 128;
}

__global__ void __launch_bounds__(UMMA_THREADS, 2) umma_conv3_kernel(UmmaArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const UmmaGeom g = a.g;
    const int KC = a.C >> 3;                 // 16-byte K chunks per pixel
    const int KB_PER_TAP = a.C / KB;
    const int NKB = 9 * KB_PER_TAP;
    uint8_t* X = smem;
    uint8_t* stages = X + umma_x_bytes(a.C, g.R);
    uint64_t* full = reinterpret_cast<uint64_t*>(stages + NSTAGE * STAGE_BYTES);
    uint64_t* empty = full + NSTAGE;
    uint64_t* accum = empty + NSTAGE;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(accum + 1);
    int* colinfo = reinterpret_cast<int*>(reinterpret_cast<uint8_t*>(full) + 128);   // [256] padded column -> (game << 16 | pixel) or -1
    int* pixinfo = colinfo + 256;                                                    // [256] real pixel (compact) -> same code
    int* g_in = pixinfo + 256;      // [64] per-game pool slot / buffer index of the input
    int* g_res = g_in + 64;         // residual
    int* g_out = g_res + 64;        // output
    int* g_act = g_out + 64;        // action
    const int game0 = blockIdx.x * g.G;
    const int HW = g.H * g.W;
    if (a.dbg && tid == 0) a.dbg[blockIdx.x * 8 + 0] = clock64();
    const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(a.wpack);

    if (tid == 0) {
        for (int s = 0; s < NSTAGE; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        mbar_init(accum, 1);
        fence_mbar_init();
        // first ring fill: weights do not depend on the activations, start streaming immediately
        for (int kb = 0; kb < NSTAGE && kb < NKB; ++kb) {
            mbar_expect_tx(&full[kb], STAGE_BYTES);
            bulk_g2s(stages + kb * STAGE_BYTES, wsrc + (size_t)kb * STAGE_BYTES, STAGE_BYTES, &full[kb]);
        }
    }
    if (warp == 1) tmem_alloc(tmem_ptr, TMEM_COLS);

    // per-game indirections and the column -> (game, pixel) map: loaded once, so the staging and store loops below
    // carry no dependent global loads and no divisions
    if (tid < g.G) {
        const int game = game0 + tid;
        const bool on = game < a.B;
        g_in[tid] = on ? (a.in_slot ? a.in_slot[game] : game) : 0;
        g_res[tid] = on ? (a.res_slot ? a.res_slot[game] : game) : 0;
        g_out[tid] = on ? (a.out_slot ? a.out_slot[game] : game) : 0;
        g_act[tid] = (on && a.action) ? a.action[game] : 0;
    }
    for (int q = tid; q < 256; q += UMMA_THREADS) {
        int info = -1;
        const int gi = q / g.P, rem = q - gi * g.P;
        const int yy = rem / (g.W + 1), xx = rem - yy * (g.W + 1);
        if (q < g.N && gi < g.G && yy >= 1 && xx < g.W && game0 + gi < a.B) info = (gi << 16) | ((yy - 1) * g.W + xx);
        colinfo[q] = info;
        const int pg = q / HW;
        pixinfo[q] = (pg < g.G && game0 + pg < a.B) ? ((pg << 16) | (q - pg * HW)) : -1;
    }
    __syncthreads();
    // resident activation tile: [K chunk][row][8 x bf16], zero padded.  thread <-> fixed K chunk, rows stride by 192/KC
    {
        const int kc = tid % KC, r0 = tid / KC, rstep = UMMA_THREADS / KC;
        for (int r = r0; r < g.R; r += rstep) {
            const int q = r - g.HALO;
            uint8_t* dst = X + ((size_t)kc * g.R + r) * 16;
            const int info = (q >= 0 && q < 256) ? colinfo[q] : -1;
            if (info >= 0)
                cp_async16(dst, a.in + (long)g_in[info >> 16] * a.in_stride + ((long)(info & 0xffff) * a.C + kc * 8));
            else
                *reinterpret_cast<uint4*>(dst) = make_uint4(0u, 0u, 0u, 0u);
        }
    }
    cp_async_wait_all();
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_ptr;
    if (a.dbg && tid == 0) a.dbg[blockIdx.x * 8 + 1] = clock64();

    if (warp == 0) {
        if (lane == 0) {   // ===== weight producer: remaining K blocks =====
            for (int kb = NSTAGE; kb < NKB; ++kb) {
                const int s = kb % NSTAGE;
                mbar_wait(&empty[s], ((kb / NSTAGE) - 1) & 1);
                mbar_expect_tx(&full[s], STAGE_BYTES);
                bulk_g2s(stages + s * STAGE_BYTES, wsrc + (size_t)kb * STAGE_BYTES, STAGE_BYTES, &full[s]);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {   // ===== MMA issuer =====
            const uint32_t idesc = idesc_bf16_f32(UM, (uint32_t)g.N);
            const uint32_t x_addr = smem_u32(X);
            const uint32_t st_addr = smem_u32(stages);
            const uint32_t lbo_b = (uint32_t)g.R * 16u;
            for (int kb = 0; kb < NKB; ++kb) {
                const int s = kb % NSTAGE;
                mbar_wait(&full[s], (kb / NSTAGE) & 1);
                tc_fence_after();
                const int tap = kb / KB_PER_TAP, cc = kb - tap * KB_PER_TAP;
                const int shift = (tap / 3 - 1) * (g.W + 1) + (tap % 3 - 1);
#pragma unroll
                for (int kk = 0; kk < KB / 16; ++kk) {
                    const uint64_t da = smem_desc_kmajor_noswizzle(st_addr + s * STAGE_BYTES + kk * 2 * (UM * 16), UM * 16, 128);
                    const uint64_t db = smem_desc_kmajor_noswizzle(
                        x_addr + (uint32_t)(((cc * 8 + kk * 2) * g.R + g.HALO + shift) * 16), lbo_b, 128);
                    mma_bf16(tmem, da, db, idesc, (kb | kk) != 0 ? 1u : 0u);
                }
                mma_commit(&empty[s]);   // frees the weight stage once these MMAs have read it
            }
            mma_commit(accum);
            if (a.dbg) a.dbg[blockIdx.x * 8 + 6] = clock64();
        }
    } else {
        // ===== epilogue, phase 1: TMEM -> registers -> fp32 staging tile in shared memory =====
        // All MMAs have completed when `accum` fires, so the activation tile and the weight ring are dead: their
        // space becomes S[real pixel][cout] (fp32), written conflict-free (lane <-> consecutive cout).
        mbar_wait(accum, 0);
        tc_fence_after();
        if (a.dbg && tid == 64) a.dbg[blockIdx.x * 8 + 2] = clock64();
        const int lane_base = 32 * (warp & 3);     // a warp may only touch its own TMEM lane quarter
        const int co = lane_base + lane;
        const bool co_ok = co < a.cout;
        const float bias = co_ok ? a.bias[co] : 0.f;
        float* S = reinterpret_cast<float*>(smem);
        for (int c0 = 0; c0 < g.N; c0 += 16) {
            uint32_t v[16];
            tmem_ld16(tmem + ((uint32_t)lane_base << 16) + (uint32_t)c0, v);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const int info = colinfo[c0 + j];
                if (info >= 0 && co_ok) S[((info >> 16) * HW + (info & 0xffff)) * a.cout + co] = __uint_as_float(v[j]) + bias;
            }
        }
        if (a.dbg && tid == 64) a.dbg[blockIdx.x * 8 + 4] = clock64();
    }
    tc_fence_before();
    __syncthreads();
    if (a.dbg && tid == 0) a.dbg[blockIdx.x * 8 + 5] = clock64();
    if (warp == 1) tmem_dealloc(tmem, TMEM_COLS);
    // ===== epilogue, phase 2 (all warps): coalesced 16-byte residual loads / bf16 stores, 8 independent items in flight =====
    {
        const float* S = reinterpret_cast<const float*>(smem);
        const int CH = a.cout >> 3;                       // 16-byte chunks per pixel; 192 % CH == 0 for cout in {64, 128}
        const int ch = tid % CH, p0 = tid / CH, pstep = UMMA_THREADS / CH;
        const int npix = g.G * HW;
        const bf16* __restrict__ res = a.res;
        bf16* __restrict__ out = a.out;
        const float* __restrict__ table = a.act_table;
        constexpr int U = 8;
        for (int pb = p0; pb < npix; pb += pstep * U) {
            uint4 rv[U]; float4 t0[U], t1[U]; int info[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int pixc = pb + u * pstep;
                info[u] = pixc < npix ? pixinfo[pixc] : -1;
                rv[u] = make_uint4(0u, 0u, 0u, 0u);
                t0[u] = make_float4(0.f, 0.f, 0.f, 0.f); t1[u] = t0[u];
                if (info[u] >= 0) {
                    const int gi = info[u] >> 16, pix = info[u] & 0xffff;
                    if (res) rv[u] = *reinterpret_cast<const uint4*>(res + (long)g_res[gi] * a.res_stride + (long)pix * a.cout + ch * 8);
                    if (table) {
                        const float* tp = table + ((long)g_act[gi] * HW + pix) * a.cout + ch * 8;
                        t0[u] = *reinterpret_cast<const float4*>(tp);
                        t1[u] = *reinterpret_cast<const float4*>(tp + 4);
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (info[u] < 0) continue;
                const int pixc = pb + u * pstep;
                const int gi = info[u] >> 16, pix = info[u] & 0xffff;
                const float4 s0 = *reinterpret_cast<const float4*>(S + pixc * a.cout + ch * 8);
                const float4 s1 = *reinterpret_cast<const float4*>(S + pixc * a.cout + ch * 8 + 4);
                float x[8] = {s0.x + t0[u].x, s0.y + t0[u].y, s0.z + t0[u].z, s0.w + t0[u].w,
                              s1.x + t1[u].x, s1.y + t1[u].y, s1.z + t1[u].z, s1.w + t1[u].w};
                const bf16* r8 = reinterpret_cast<const bf16*>(&rv[u]);
                uint4 ov;
                bf16* o8 = reinterpret_cast<bf16*>(&ov);
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    float v = x[e] + __bfloat162float(r8[e]);
                    o8[e] = __float2bfloat16(a.relu ? fmaxf(v, 0.f) : v);
                }
                *reinterpret_cast<uint4*>(out + (long)g_out[gi] * a.out_stride + (long)pix * a.cout + ch * 8) = ov;
            }
        }
    }
    if (a.dbg && tid == 0) a.dbg[blockIdx.x * 8 + 3] = clock64();
}

}  // namespace

namespace {

// Representation head convolution for a thin observation (ResidualHeadImpl, model_layers.cpp:95-109: conv3x3 + folded BN +
// ReLU on cin = 3..16 planes).  One block per game, one thread per output channel: the zero-padded observation and the
// [tap][cin][cout] weights sit in shared memory, reads are a broadcast (x) and a conflict-free row (w).  Reads the caller's
// NCHW fp32 rows directly and writes NHWC bf16 — replaces a layout-conversion launch plus the generic direct kernel (142 us
// for 1024 Connect4 boards) with ~10 us.
__global__ void __launch_bounds__(128) head_conv_small_kernel(const float* __restrict__ obs, long obs_stride, const float* __restrict__ w,
                                                              const float* __restrict__ bias, bf16* __restrict__ out, long out_stride,
                                                              int cin, int cout, int H, int W, int relu) {
    extern __shared__ float hsm[];
    const int PW = W + 2, PH = H + 2;
    float* xs = hsm;                         // [cin][PH][PW], zero border
    float* ws = hsm + cin * PH * PW;         // [9 * cin][cout]
    const int b = blockIdx.x, tid = threadIdx.x;
    const float* ob = obs + (long)b * obs_stride;
    for (int i = tid; i < cin * PH * PW; i += blockDim.x) {
        const int c = i / (PH * PW), r = i - c * PH * PW, y = r / PW - 1, x = r % PW - 1;
        xs[i] = (y >= 0 && y < H && x >= 0 && x < W) ? ob[(c * H + y) * W + x] : 0.f;
    }
    for (int i = tid; i < 9 * cin * cout; i += blockDim.x) ws[i] = w[i];
    __syncthreads();
    if (cin <= 3) {
        // thin observations (board games: 3 planes): the thread's 27 weights stay in registers, so a pixel costs 27 broadcast
        // shared loads instead of 54 loads (the kernel was LSU-bound: ~100 us per 1024 boards, 1 % of a 50-simulation search)
        for (int co = tid; co < cout; co += blockDim.x) {
            float wr[27];
#pragma unroll
            for (int i = 0; i < 27; ++i) { const int tap = i / 3, c = i - tap * 3; wr[i] = c < cin ? ws[(tap * cin + c) * cout + co] : 0.f; }
            const float bv = bias[co];
            for (int y = 0; y < H; ++y)
                for (int x = 0; x < W; ++x) {
                    float acc = 0.f;
#pragma unroll
                    for (int tap = 0; tap < 9; ++tap) {
                        const float* xp = xs + (y + tap / 3) * PW + (x + tap % 3);
#pragma unroll
                        for (int c = 0; c < 3; ++c) if (c < cin) acc += xp[c * PH * PW] * wr[tap * 3 + c];
                    }
                    acc += bv;
                    if (relu) acc = fmaxf(acc, 0.f);
                    out[(long)b * out_stride + (long)(y * W + x) * cout + co] = __float2bfloat16(acc);
                }
        }
        return;
    }
    for (int co = tid; co < cout; co += blockDim.x) {
        const float bv = bias[co];
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x) {
                float acc = 0.f;
                for (int tap = 0; tap < 9; ++tap) {
                    const float* xp = xs + (y + tap / 3) * PW + (x + tap % 3);
                    const float* wp = ws + (long)tap * cin * cout + co;
                    for (int c = 0; c < cin; ++c) acc += xp[c * PH * PW] * wp[c * cout];
                }
                acc += bv;
                if (relu) acc = fmaxf(acc, 0.f);
                out[(long)b * out_stride + (long)(y * W + x) * cout + co] = __float2bfloat16(acc);
            }
    }
}

}  // namespace

namespace {

// geometry of the padded resident tile for an H x W board; returns false when the shape does not fit the tensor path
bool make_geom(int H, int W, int C, UmmaGeom* g) {
    g->H = H; g->W = W;
    g->P = (H + 1) * (W + 1);
    if (g->P > 256) return false;
    g->G = 256 / g->P;
    g->N = (g->G * g->P + 15) / 16 * 16;
    g->HALO = W + 2;
    g->R = g->HALO + g->N + g->HALO;
    if ((g->R & 1) == 0) g->R += 1;   // odd row count: the 16-byte staging stores of one pixel hit 8 distinct bank groups
    // two CTAs per SM must fit (their prologue / epilogue overlap each other's MMA phase)
    return umma_smem_bytes(C, *g) <= 113 * 1024;
}

class Bf16Backend : public direct::DirectBackend<bf16> {
public:
    const char* name() const override { return "bf16-tcgen05"; }

protected:
    struct Packed { bool ok = false; bool image = false; size_t off = 0; UmmaGeom g; };   // ok: the per-layer kernel takes it; image: packed weights exist (enough for the fused chain)

    bool eligible(const Conv3Host& c, UmmaGeom* g) const {
        if (c.stride != 1 || c.cin % KB != 0 || c.cout > UM || c.cout % 8 != 0) return false;
        if (UMMA_THREADS % (c.cin / 8) != 0 || UMMA_THREADS % (c.cout / 8) != 0) return false;
        if (c.in_h != c.out_h || c.in_w != c.out_w) return false;
        if (!make_geom(c.in_h, c.in_w, c.cin, g)) return false;
        return (size_t)g->G * c.in_h * c.in_w * c.cout * 4 <= umma_x_bytes(c.cin, g->R) + (size_t)NSTAGE * STAGE_BYTES;
    }

    // bf16 weight images: per 64-wide K block a ready-made smem tile [8 K-chunks][128 cout rows][8 elems]
    int pack_extra(const NetPlan& plan, std::vector<uint8_t>* extra) override {
        const Tower* tw[3] = {&plan.representation, &plan.dynamics, &plan.prediction};
        extra->clear();
        for (int t = 0; t < 3; ++t) {
            packed_[t].assign(tw[t]->convs.size(), Packed());
            for (size_t i = 0; i < tw[t]->convs.size(); ++i) {
                const Conv3Host& c = tw[t]->convs[i];
                Packed p;
                // the weight image only needs a 64-multiple of input channels and <= 128 outputs; the per-layer kernel
                // additionally wants its own tile geometry to fit twice per SM
                // (a stride-2 conv is the stride-1 conv sampled at even positions: same image, taken by the band kernel only)
                const bool same = c.stride == 1 && c.in_h == c.out_h && c.in_w == c.out_w;
                const bool half = c.stride == 2 && c.out_h == (c.in_h + 1) / 2 && c.out_w == (c.in_w + 1) / 2;
                if ((!same && !half) || c.cin % KB != 0 || c.cout > UM || c.cout % 8 != 0) continue;
                p.image = true;
                p.ok = same && eligible(c, &p.g);
                while (extra->size() % 128) extra->push_back(0);
                p.off = extra->size();
                const int kb_per_tap = c.cin / KB;
                const size_t bytes = (size_t)9 * kb_per_tap * STAGE_BYTES;
                extra->resize(extra->size() + bytes, 0);
                bf16* dst = reinterpret_cast<bf16*>(extra->data() + p.off);
                for (int tap = 0; tap < 9; ++tap)
                    for (int cc = 0; cc < kb_per_tap; ++cc)
                        for (int kc = 0; kc < 8; ++kc)
                            for (int row = 0; row < UM; ++row)
                                for (int e = 0; e < 8; ++e) {
                                    const int k = cc * KB + kc * 8 + e;
                                    // cout <= 64: rows 64-127 repeat rows 0-63 (the fused chain's epilogue then uses all four
                                    // TMEM lane quarters, tower_fused.cuh `dup`; the per-layer kernels ignore rows >= cout)
                                    const int wr = (c.cout <= 64 && row >= 64) ? row - 64 : row;
                                    const float v = wr < c.cout ? c.w[((size_t)wr * 9 + tap) * c.cin + k] : 0.f;
                                    dst[(((size_t)(tap * kb_per_tap + cc) * 8 + kc) * UM + row) * 8 + e] = __float2bfloat16(v);
                                }
                packed_[t][i] = p;
            }
        }
        // first convolution of a downsample head (stride 2 on <= 16 observation planes): nine K = 16 tap images
        // [2 K chunks][128 rows][8] for conv_first_kernel, which reads the fp32 NCHW observation itself
        first_ok_ = false;
        if (!plan.representation.convs.empty() && !plan.representation.steps.empty() && plan.representation.steps[0].kind == TowerStep::CONV &&
            plan.representation.steps[0].conv0 == 0 && !getenv("MZB_NO_CONV_FIRST")) {
            const Conv3Host& c = plan.representation.convs[0];
            const int rbo = first::first_rows(c.out_w);
            if (c.stride == 2 && c.cin <= 16 && c.cout <= UM && c.cout % 8 == 0 && c.act_table.empty() && c.out_h == (c.in_h - 1) / 2 + 1 &&
                c.out_w == (c.in_w - 1) / 2 + 1 && rbo >= 1) {
                while (extra->size() % 128) extra->push_back(0);
                first_off_ = extra->size();
                extra->resize(extra->size() + first::C1_W_BYTES, 0);
                bf16* dst = reinterpret_cast<bf16*>(extra->data() + first_off_);
                for (int tap = 0; tap < 9; ++tap)
                    for (int kc = 0; kc < 2; ++kc)
                        for (int row = 0; row < UM; ++row)
                            for (int e = 0; e < 8; ++e) {
                                const int k = kc * 8 + e;
                                const float v = (row < c.cout && k < c.cin) ? c.w[((size_t)row * 9 + tap) * c.cin + k] : 0.f;
                                dst[(((size_t)tap * 2 + kc) * UM + row) * 8 + e] = __float2bfloat16(v);
                            }
                first::FirstArgs& a = first_args_;
                memset(&a, 0, sizeof(a));
                a.cin = c.cin; a.cout = c.cout; a.H = c.in_h; a.W = c.in_w; a.Ho = c.out_h; a.Wo = c.out_w; a.relu = c.relu ? 1 : 0;
                a.rbo = std::min(rbo, c.out_h);
                a.bands = (c.out_h + a.rbo - 1) / a.rbo;
                a.N = (a.rbo * (c.out_w + 1) + 15) / 16 * 16;
                a.cls = first::first_cls(a.N, c.out_w);
                const size_t smem = first::first_smem_bytes(a.cls, a.rbo, a.Wo, a.cout);
                if (a.N <= 256 && smem <= 113 * 1024) {
                    MZB_CUDA(cudaFuncSetAttribute(first::conv_first_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                    first_ok_ = true;
                }
            }
        }
        // PKM segment (conv_band.cuh): first conv -> every convolution up to the first pooling step, when all of them fit
        // conv_band2_kernel.  The activations between those kernels then stay in the padded K-major layout.
        pkm_chain_ = false; pkm_max_elems_ = 0;
        if (first_ok_ && !getenv("MZB_BAND_V1")) {
            const Tower& tw = plan.representation;
            size_t pool_at = 0;
            for (size_t i = 1; i < tw.steps.size(); ++i) if (tw.steps[i].kind == TowerStep::POOL) { pool_at = i; break; }
            bool ok = pool_at > 1;
            size_t mxe = (size_t)band::pkm_elems(tw.convs[0].cout, tw.convs[0].out_h, tw.convs[0].out_w);
            for (size_t i = 1; ok && i < pool_at; ++i) {
                const TowerStep& st = tw.steps[i];
                for (int k : {st.conv0, st.kind == TowerStep::BLOCK ? st.conv1 : st.conv0}) {
                    const Conv3Host& c = tw.convs[k];
                    band::Band2Args g2;
                    memset(&g2, 0, sizeof(g2));
                    ok = ok && packed_[0][k].image && !packed_[0][k].ok && c.act_table.empty() && band2_geometry(c.cin, c.cout, c.stride, c.in_h, c.in_w, &g2);
                    mxe = std::max(mxe, (size_t)band::pkm_elems(c.cout, c.out_h, c.out_w));
                }
            }
            if (getenv("MZB_VERBOSE")) fprintf(stderr, "[bf16] downsample head: conv_first %s, PKM segment (conv_band2_kernel) %s\n", first_ok_ ? "yes" : "no", ok ? "yes" : "no");
            if (ok) { pkm_chain_ = true; pkm_max_elems_ = mxe; pkm_last_conv_ = 0; for (size_t i = 1; i < pool_at; ++i) pkm_last_conv_ = std::max(pkm_last_conv_, std::max(tw.steps[i].conv0, tw.steps[i].kind == TowerStep::BLOCK ? tw.steps[i].conv1 : tw.steps[i].conv0)); }
        }
        if (!attr_set_) {
            MZB_CUDA(cudaFuncSetAttribute(umma_conv3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 113 * 1024));
            attr_set_ = true;
        }
        return pack_fused(plan, extra);
    }

    // ------------------------------------------------------------------ whole-chain fusion (tower_fused.cuh)
    struct FusedChain {
        bool ok = false;
        fused::FusedArgs args;     // template: per-call pointers are filled in at launch
        int gmax = 0;
        size_t smem = 0;
    };
    struct TileTab { DevBuf first, count; std::vector<int> h_count; int n_tiles = 0; };

    static bool tower_fusable(const Tower& tw, bool leading_conv, int C, int H, int W) {
        size_t i = 0;
        if (leading_conv) {
            if (tw.steps.empty() || tw.steps[0].kind != TowerStep::CONV) return false;
            i = 1;
        }
        for (; i < tw.steps.size(); ++i) if (tw.steps[i].kind != TowerStep::BLOCK) return false;
        for (size_t k = 0; k < tw.convs.size(); ++k) {
            const Conv3Host& c = tw.convs[k];
            if (c.stride != 1 || c.cin != C || c.cout != C || c.in_h != H || c.in_w != W || c.out_h != H || c.out_w != W) return false;
        }
        return true;
    }

    // 1x1 head convolutions as single-tap weight stages: [C/64][8 K chunks][128 rows][8], rows = stacked head channels
    size_t pack_conv1(const HeadHost* const* hh, int nh, int C, std::vector<uint8_t>* extra, size_t* bias_rel) {
        while (extra->size() % 128) extra->push_back(0);
        const size_t off = extra->size();
        const int kbs = C / KB;
        extra->resize(off + (size_t)kbs * STAGE_BYTES, 0);
        bf16* dst = reinterpret_cast<bf16*>(extra->data() + off);
        std::vector<float> bias(128, 0.f);
        // <= 32 stacked rows: the same rows are repeated in all four TMEM lane quarters (rows sr, 32 + sr, 64 + sr, 96 + sr), so
        // that every epilogue warp — a warp only reaches its own quarter — can take a share of the columns (tower_fused.cuh, PH_RED)
        int total = 0;
        for (int h = 0; h < nh; ++h) total += hh[h]->red;
        const int rep = total <= 32 ? 4 : 1;
        int row0 = 0;
        for (int h = 0; h < nh; ++h) {
            for (int r = 0; r < hh[h]->red; ++r) {
                const int sr = row0 + r;                         // stacked row
                for (int q = 0; q < rep; ++q) {
                    const int lane = sr + 32 * q;
                    bias[lane] = hh[h]->b1[r];
                    for (int k = 0; k < C; ++k) {
                        const int cc = k / KB, kc = (k % KB) / 8, e = k % 8;
                        dst[(((size_t)cc * 8 + kc) * UM + lane) * 8 + e] = __float2bfloat16(hh[h]->w1[(size_t)r * C + k]);
                    }
                }
            }
            row0 += hh[h]->red;
        }
        // (resize may have moved the buffer: dst is not used past this point)
        *bias_rel = extra->size();
        const uint8_t* bp = reinterpret_cast<const uint8_t*>(bias.data());
        extra->insert(extra->end(), bp, bp + 128 * sizeof(float));
        return off;
    }

    // first MLP layer of 1 or 2 heads as tensor-core weight images: rows = stacked hidden units, K = concatenated head
    // inputs (block structure: a head's rows are zero outside its own K range), K padded to a multiple of 64
    size_t pack_mlp0(const HeadHost* const* hh, int nh, int rp, int P, int H, int W, std::vector<uint8_t>* extra, size_t* bias_rel, int* nkb_out) {
        // K order of Z: k = padded pixel row c * rp + reduced channel r (tower_fused.cuh, PH_RED); the reference's flatten
        // index is r * HW + pixel (models.cpp:114,161-162).  Pad rows and channels >= red get zero weights.
        const int ktot = P * rp;
        const int nkb = (ktot + KB - 1) / KB;
        while (extra->size() % 128) extra->push_back(0);
        const size_t off = extra->size();
        extra->resize(off + (size_t)nkb * STAGE_BYTES, 0);
        bf16* dst = reinterpret_cast<bf16*>(extra->data() + off);
        std::vector<float> bias(128, 0.f);
        int row0 = 0;
        for (int h = 0; h < nh; ++h) {
            const int n0 = hh[h]->sizes[0], n1 = hh[h]->sizes[1], red = hh[h]->red, hw = hh[h]->hw;
            for (int j = 0; j < n1; ++j) {
                bias[row0 + j] = hh[h]->b[0][j];
                for (int c = 0; c < P; ++c) {
                    const int yy = c / (W + 1), xx = c % (W + 1);
                    if (yy < 1 || yy > H || xx >= W) continue;
                    const int pix = (yy - 1) * W + xx;
                    for (int r = 0; r < red; ++r) {
                        const int k = c * rp + r, cc = k / KB, kc = (k % KB) / 8, e = k % 8;
                        dst[(((size_t)cc * 8 + kc) * UM + row0 + j) * 8 + e] = __float2bfloat16(hh[h]->W[0][(size_t)j * n0 + (size_t)r * hw + pix]);
                    }
                }
            }
            row0 += n1;
        }
        *bias_rel = extra->size();
        const uint8_t* bp = reinterpret_cast<const uint8_t*>(bias.data());
        extra->insert(extra->end(), bp, bp + 128 * sizeof(float));
        *nkb_out = nkb;
        return off;
    }

    int pack_fused(const NetPlan& plan, std::vector<uint8_t>* extra) {
        fused_rec_ = FusedChain(); fused_init_ = FusedChain();
        fused_final_ = false;
        tiles_.clear();
        static const bool disabled = getenv("MZB_NO_FUSED") != nullptr;
        if (disabled) return MZB_OK;
        // MZB_VERBOSE=1 names the condition that keeps a network off the fused chain (it then runs the per-layer path)
#define MZB_FUSE_REJECT(why) do { if (getenv("MZB_VERBOSE")) fprintf(stderr, "[bf16] fused chain not used: %s\n", why); return MZB_OK; } while (0)
        const NetDims& d = plan.dims;
        const int C = d.C, H = d.h, W = d.w;
        if (C % KB != 0 || C > UM) MZB_FUSE_REJECT("channels not a multiple of 64 or above 128");
        const int P = ((H + 1) * (W + 1) + 7) / 8 * 8;
        if (P > 256) MZB_FUSE_REJECT("padded board above 256 rows");
        const HeadHost* hr[1] = {&plan.reward_head};
        const HeadHost* hpv[2] = {&plan.policy_head, &plan.value_head};
        if (plan.reward_head.red > UM || plan.policy_head.red + plan.value_head.red > UM) MZB_FUSE_REJECT("reduced head channels above 128");
        for (const HeadHost* h : {&plan.reward_head, &plan.policy_head, &plan.value_head}) {
            if ((int)h->W.size() > fused::MLP_MAX_LAYERS) MZB_FUSE_REJECT("MLP deeper than 6 layers");
            for (size_t i = 1; i < h->sizes.size(); ++i) if (h->sizes[i] > 256) MZB_FUSE_REJECT("MLP layer wider than 256");
        }
        const bool dyn_ok = tower_fusable(plan.dynamics, true, C, H, W) && packed_[1].size() == plan.dynamics.convs.size();
        const bool pred_ok = tower_fusable(plan.prediction, false, C, H, W);
        if (!dyn_ok) MZB_FUSE_REJECT("dynamics tower is not head conv + residual blocks at constant shape");
        if (!pred_ok) MZB_FUSE_REJECT("prediction tower is not residual blocks at constant shape");
        for (int t = 1; t < 3; ++t) for (const Packed& p : packed_[t]) if (!p.image) MZB_FUSE_REJECT("a convolution has no packed tensor-core image (channels / stride)");
        // geometry: as many games per tile as 256 accumulator columns hold; the in-kernel head MLPs (first layer on the tensor
        // core, one launch per inference) need <= 8 games per tile and stacked widths <= 128, and are preferred over wider tiles
        const HeadHost& hR = plan.reward_head; const HeadHost& hP = plan.policy_head; const HeadHost& hV = plan.value_head;
        static const bool align8 = getenv("MZB_X_ALIGN8") != nullptr;   // experiment: 128-byte aligned K-chunk planes and game blocks (what a tensor-map copy of the chain input would need)
        const int halo = align8 ? (W + 2 + 7) / 8 * 8 : W + 2;
        int gmax = 0, R = 0, x_bytes = 0, z_bytes = 0, z_rp = 0;
        auto geometry = [&](int g) {
            gmax = g;
            const int nmax = (gmax * P + 15) / 16 * 16;
            R = halo + nmax + (align8 ? (W + 1 + 7) / 8 * 8 : halo);
            if (!align8 && (R & 1) == 0) ++R;
            x_bytes = (int)((((size_t)(C / 8) * R * 16) + 127) / 128 * 128);
        };
        {
            bool ok = hP.W.size() == hV.W.size() && hR.sizes[1] <= UM;
            for (size_t l = 1; ok && l < hP.sizes.size(); ++l) ok = hP.sizes[l] + hV.sizes[l] <= UM;
            for (size_t l = 1; ok && l < hR.sizes.size(); ++l) ok = hR.sizes[l] <= UM;
            static const bool no_z = getenv("MZB_NO_INKERNEL_MLP") != nullptr;
            if (ok && !no_z) {
                geometry(std::min(std::min(256 / P, fused::F_MAX_G), 8));
                z_rp = (std::max(hR.red, std::max(hP.red, hV.red)) + 7) / 8 * 8;
                z_bytes = (P * z_rp + KB - 1) / KB * KB / 8 * 256;   // [K / 8][16 rows][16 B]
                if (fused::fused_smem_bytes(x_bytes, z_bytes, P) > 227 * 1024) z_bytes = 0;
            }
            if (!z_bytes) {
                z_rp = 0;
                geometry(std::min(256 / P, fused::F_MAX_G));
                if (fused::fused_smem_bytes(x_bytes, 0, P) > 227 * 1024) MZB_FUSE_REJECT("tile does not fit shared memory");
            }
        }
        const size_t smem = fused::fused_smem_bytes(x_bytes, z_bytes, P);
        // extra weights
        conv1_rel_[0] = pack_conv1(hr, 1, C, extra, &conv1_bias_rel_[0]);
        conv1_rel_[1] = pack_conv1(hpv, 2, C, extra, &conv1_bias_rel_[1]);
        const HeadHost* hh[3] = {&plan.reward_head, &plan.policy_head, &plan.value_head};
        for (int h = 0; h < 3; ++h) {   // first MLP layer transposed [in][out] for coalesced reads
            while (extra->size() % 16) extra->push_back(0);
            w0t_rel_[h] = extra->size();
            const int n0 = hh[h]->sizes[0], n1 = hh[h]->sizes[1];
            std::vector<float> wt((size_t)n0 * n1);
            // row i' = p * red + r of the transposed matrix <- input r * HW + p of the reference's flatten order: the fused kernel
            // stores the reduced activations pixel-major so that a warp's channels are one contiguous run
            const int red = hh[h]->red, hw = H * W;
            for (int j = 0; j < n1; ++j)
                for (int r = 0; r < red; ++r)
                    for (int px = 0; px < hw; ++px)
                        wt[((size_t)px * red + r) * n1 + j] = hh[h]->W[0][(size_t)j * n0 + (size_t)r * hw + px];
            const uint8_t* bp = reinterpret_cast<const uint8_t*>(wt.data());
            extra->insert(extra->end(), bp, bp + wt.size() * sizeof(float));
        }
        if (z_bytes) {
            const HeadHost* r1[1] = {&hR};
            const HeadHost* pv2[2] = {&hP, &hV};
            mlp0_rel_[0] = pack_mlp0(r1, 1, z_rp, P, H, W, extra, &mlp0_bias_rel_[0], &mlp0_nkb_[0]);
            mlp0_rel_[1] = pack_mlp0(pv2, 2, z_rp, P, H, W, extra, &mlp0_bias_rel_[1], &mlp0_nkb_[1]);
        }
        fused::FusedArgs base;
        memset(&base, 0, sizeof(base));
        base.z_bytes = z_bytes; base.z_rp = z_rp;
        base.C = C; base.H = H; base.W = W; base.P = P; base.HALO = halo; base.R = R; base.HW = H * W; base.x_bytes = x_bytes;
        base.normalize = plan.normalize ? 1 : 0;
        base.dup = (C <= 64 && !getenv("MZB_FUSED_NO_DUP")) ? 1 : 0;
        base.red_rows[0] = plan.reward_head.red; base.red_rows[1] = plan.policy_head.red; base.red_rows[2] = plan.value_head.red;
        fused_rec_.args = base; fused_rec_.gmax = gmax; fused_rec_.smem = smem;
        fused_rec_.ok = true;    // phases are laid out by finalize_fused() once the blob offsets are known
        // the representation head conv (index 0) has its own input channel count and stays on the direct kernel
        fused_init_.args = base; fused_init_.gmax = gmax; fused_init_.smem = smem;
        fused_init_.ok = rep_init_fusable(plan, C, H, W);
        plan_mid_chain(plan, C);
        MZB_CUDA(cudaFuncSetAttribute(fused::tower_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        MZB_CUDA(cudaFuncSetAttribute(fused::mlp_heads_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        return MZB_OK;
    }

    // The fused chain takes over the representation tower at the first step of its trailing run of residual blocks at the
    // hidden-state shape (sets init_tail_): step 1 behind the thin head conv of a board game (ResidualHeadImpl), the step behind the
    // last pooling of ResidualHeadDownsample.  What lies before runs layer by layer (conv_first / band kernels).
    // A run of residual blocks INSIDE the representation tower (not its tail) at a shape that fits a whole-game tile — the 12x12
    // stage of the ALE-shaped downsample head between the two poolings — runs as one fused chain as well: blocks, then a plain copy
    // phase (PH_NORM with normalize = 0) that writes the result as a compact NHWC activation.  One launch instead of two per block.
    void plan_mid_chain(const NetPlan& plan, int C) {
        fused_mid_ = FusedChain(); mid_first_ = mid_end_ = 0;
        if (getenv("MZB_NO_MID_CHAIN")) return;
        const Tower& tw = plan.representation;
        const size_t tail = init_tail_ ? init_tail_ : tw.steps.size();
        for (size_t i = 1; i < tail; ++i) {
            if (tw.steps[i].kind != TowerStep::BLOCK) continue;
            const Conv3Host& c0 = tw.convs[tw.steps[i].conv0];
            const int h = c0.in_h, w = c0.in_w, P = (h + 1) * (w + 1);
            if (c0.cin != C || P > 256) continue;
            size_t j = i;
            auto ok = [&](const TowerStep& st) {
                if (st.kind != TowerStep::BLOCK) return false;
                for (int k : {st.conv0, st.conv1}) {
                    const Conv3Host& c = tw.convs[k];
                    if (c.stride != 1 || c.cin != C || c.cout != C || c.in_h != h || c.in_w != w || !packed_[0][k].image) return false;
                }
                return true;
            };
            while (j < tail && ok(tw.steps[j])) ++j;
            if (j - i < 1 || 2 * (j - i) + 1 > (size_t)fused::F_MAX_PHASES) { i = j; continue; }
            const int halo = w + 2, gmax = std::min(256 / P, fused::F_MAX_G), nmax = (gmax * P + 15) / 16 * 16;
            int R = halo + nmax + halo;
            if ((R & 1) == 0) ++R;
            const int x_bytes = (int)((((size_t)(C / 8) * R * 16) + 127) / 128 * 128);
            const size_t smem = fused::fused_smem_bytes(x_bytes, 0, P);
            if (smem > 227 * 1024 || (fused_rec_.smem && smem > fused_rec_.smem)) return;   // (the kernel's shared-memory attribute is the main chain's)
            fused::FusedArgs base;
            memset(&base, 0, sizeof(base));
            base.C = C; base.H = h; base.W = w; base.P = P; base.HALO = halo; base.R = R; base.HW = h * w; base.x_bytes = x_bytes;
            base.normalize = 0;
            base.dup = (C <= 64 && !getenv("MZB_FUSED_NO_DUP")) ? 1 : 0;
            fused_mid_.args = base; fused_mid_.gmax = gmax; fused_mid_.smem = smem; fused_mid_.ok = true;
            mid_first_ = i; mid_end_ = j;
            return;
        }
    }

    bool rep_init_fusable(const NetPlan& plan, int C, int H, int W) {
        const Tower& tw = plan.representation;
        init_tail_ = 0;
        if (tw.steps.empty()) return false;
        size_t tail = tw.steps.size();
        auto block_ok = [&](const TowerStep& st) {
            if (st.kind != TowerStep::BLOCK) return false;
            for (int k : {st.conv0, st.conv1}) {
                const Conv3Host& c = tw.convs[k];
                if (c.stride != 1 || c.cin != C || c.cout != C || c.in_h != H || c.in_w != W) return false;
                if (!packed_[0][k].image) return false;
            }
            return true;
        };
        while (tail > 0 && block_ok(tw.steps[tail - 1])) --tail;
        if (tail == 0 || tail == tw.steps.size()) return false;   // something must come before, and a block after (the skip preload)
        // shape after the steps in front of the tail
        int c = tw.in_c, h = tw.in_h, w = tw.in_w;
        for (size_t i = 0; i < tail; ++i) {
            const TowerStep& st = tw.steps[i];
            if (st.kind == TowerStep::CONV) { const Conv3Host& cv = tw.convs[st.conv0]; c = cv.cout; h = cv.out_h; w = cv.out_w; }
            else if (st.kind == TowerStep::POOL) { h = (h + 2 - 3) / 2 + 1; w = (w + 2 - 3) / 2 + 1; }
        }
        if (c != C || h != H || w != W) return false;
        init_tail_ = tail;
        return true;
    }

    void add_conv_phase(fused::FusedArgs* fa, int tower, int index, int flags) {
        const direct::ConvDev& c = convs_[tower][index];
        fused::Phase& ph = fa->ph[fa->n_phases++];
        memset(&ph, 0, sizeof(ph));
        ph.w_off = (uint32_t)(fp32_bytes_ + packed_[tower][index].off);
        ph.taps = 9; ph.nkb = 9 * (c.cin / KB);
        ph.bias_off = (int)(c.bias_off * sizeof(float));
        ph.table_off = -1;
        ph.flags = flags | (c.relu ? fused::PH_RELU : 0);
        if (c.has_table) {
            // dynamics head: the action planes' contribution is a per-action table [A][HW][C].  The epilogue warps stage it in
            // the (still unused) R2 accumulator WHILE the head's MMAs run into R1, and the epilogue adds the two — no table
            // look-ups (L2 latency chains) behind the MMAs.  Only the first phase of a chain may do so (R2 must be free).
            fa->table_off = (int)(c.table_off * sizeof(float));
            ph.flags |= fused::PH_TABLE;
        }
    }
    void add_blocks(fused::FusedArgs* fa, int tower, size_t first_step, size_t end_step = (size_t)-1) {
        for (size_t i = first_step; i < steps_[tower].size() && i < end_step; ++i) {
            add_conv_phase(fa, tower, steps_[tower][i].conv0, fused::PH_WRITE_X);
            add_conv_phase(fa, tower, steps_[tower][i].conv1, fused::PH_WRITE_X | fused::PH_ACC_R2 | fused::PH_STASH);
        }
    }
    void add_conv1_phase(fused::FusedArgs* fa, int which, int flags, int C) {
        fused::Phase& ph = fa->ph[fa->n_phases++];
        memset(&ph, 0, sizeof(ph));
        ph.bias_off = -1; ph.table_off = -1;
        ph.flags = flags;
        if (which >= 0) {
            ph.w_off = (uint32_t)(fp32_bytes_ + conv1_rel_[which]);
            ph.taps = 1; ph.nkb = C / KB;
            ph.bias_off = (int)(fp32_bytes_ + conv1_bias_rel_[which]);
            ph.flags |= fused::PH_RED;
            if (which == 0) { ph.red_head0 = 0; ph.red_n0 = fa->red_rows[0]; ph.red_head1 = 0; ph.red_n1 = 0; }
            else { ph.red_head0 = 1; ph.red_n0 = fa->red_rows[1]; ph.red_head1 = 2; ph.red_n1 = fa->red_rows[2]; }
            ph.z_k0 = 0; ph.z_k1 = heads_[ph.red_head0].sizes[0];
        }
    }
    void add_mlp_phase(fused::FusedArgs* fa, int which) {
        if (!fa->z_bytes) return;
        fused::Phase& ph = fa->ph[fa->n_phases++];
        memset(&ph, 0, sizeof(ph));
        ph.table_off = -1;
        ph.flags = fused::PH_MLP;
        ph.w_off = (uint32_t)(fp32_bytes_ + mlp0_rel_[which]);
        ph.taps = 1; ph.nkb = mlp0_nkb_[which];
        ph.bias_off = (int)(fp32_bytes_ + mlp0_bias_rel_[which]);
        if (which == 0) { ph.red_head0 = 0; ph.red_n0 = heads_[0].sizes[1]; ph.red_head1 = 0; ph.red_n1 = 0; }
        else { ph.red_head0 = 1; ph.red_n0 = heads_[1].sizes[1]; ph.red_head1 = 2; ph.red_n1 = heads_[2].sizes[1]; }
    }

    // Staggered half tiles (tower_fused.cuh): a plain 3x3 layer (result -> X, no table / reduction / normalisation) ends per half
    // tile, and the 3x3 layer behind a plain one starts per half tile.  MZB_TOWER_SPLIT="nh,nt" sets the ring slots issued per half
    // at the head / tail of a layer ("0,0" switches the scheme off).
    void mark_split(fused::FusedArgs* fa) {
        static const char* env = getenv("MZB_TOWER_SPLIT");
        int nh = 1, nt = 1;   // measured best (c4d, 4-game tiles): 1,1 0.139 ms | 2,2 0.140 | 3,3 0.141 | off 0.143 per 592 states
        if (env) sscanf(env, "%d,%d", &nh, &nt);
        nh = std::max(0, std::min(nh, 3)); nt = std::max(0, std::min(nt, 3));   // a half's slots stay in the ring (4 slots) until the other half has used them
        fa->split_nh = nh; fa->split_nt = nt;
        auto plain = [&](const fused::Phase& ph) {
            return ph.taps == 9 && (ph.flags & fused::PH_WRITE_X) && !(ph.flags & (fused::PH_RED | fused::PH_NORM | fused::PH_MLP | fused::PH_TABLE));
        };
        for (int p = 0; p < fa->n_phases; ++p) {
            fused::Phase& ph = fa->ph[p];
            ph.flags &= ~(fused::PH_SPLIT_HEAD | fused::PH_SPLIT_TAIL);
            const int nslots = (ph.nkb + 1) / 2;
            if (nt == 0 || ph.taps != 9 || nslots < nh + nt + 1) continue;
            if (plain(ph)) ph.flags |= fused::PH_SPLIT_TAIL;
            if (nh > 0 && p > 0 && plain(fa->ph[p - 1]) && plain(ph)) ph.flags |= fused::PH_SPLIT_HEAD;
        }
    }

    int finalize_fused() {
        if (fused_final_) return MZB_OK;
        fused_final_ = true;
        if (!fused_rec_.ok) return MZB_OK;
        const int C = plan_dims_.C, hw = plan_dims_.hw();
        int dev = 0;
        MZB_CUDA(cudaGetDevice(&dev));
        MZB_CUDA(cudaDeviceGetAttribute(&n_sm_, cudaDevAttrMultiProcessorCount, dev));
        size_t phases_rec = 1 + 2 * (steps_[1].size() - 1) + 2 + 2 * steps_[2].size() + 1 + 2;
        if (phases_rec > (size_t)fused::F_MAX_PHASES) { fused_rec_.ok = false; fused_init_.ok = false; return MZB_OK; }
        {   // recurrent: dynamics head, dynamics blocks, reward conv1x1 + normalise, prediction blocks, policy/value conv1x1
            fused::FusedArgs* fa = &fused_rec_.args;
            fa->n_phases = 0; fa->preload = 0;
            add_conv_phase(fa, 1, steps_[1][0].conv0, fused::PH_WRITE_X | fused::PH_STASH);
            add_blocks(fa, 1, 1);
            if (fa->z_bytes) {
                // reward 1x1 conv -> Z, then normalisation as its own MMA-less phase that releases the MMA warp BEFORE it starts:
                // the reward MLP's first layer (reads Z only) runs on the tensor core while the epilogue warps normalise
                add_conv1_phase(fa, 0, 0, C);
                add_conv1_phase(fa, -1, fused::PH_NORM | fused::PH_EARLY, C);
            } else {
                add_conv1_phase(fa, 0, fused::PH_NORM, C);
            }
            add_mlp_phase(fa, 0);
            add_blocks(fa, 2, 0);
            add_conv1_phase(fa, 1, 0, C);
            add_mlp_phase(fa, 1);
        }
        size_t phases_init = 2 * (steps_[0].size() - init_tail_) + 1 + 2 * steps_[2].size() + 1 + 1;
        if (fused_init_.ok && phases_init > (size_t)fused::F_MAX_PHASES) fused_init_.ok = false;
        if (fused_init_.ok) {   // initial: (head conv on the direct kernel), representation blocks, normalise, prediction, heads
            fused::FusedArgs* fa = &fused_init_.args;
            fa->n_phases = 0; fa->preload = 1;
            add_blocks(fa, 0, init_tail_);
            add_conv1_phase(fa, -1, fused::PH_NORM, C);
            add_blocks(fa, 2, 0);
            add_conv1_phase(fa, 1, 0, C);
            add_mlp_phase(fa, 1);
        }
        if (fused_mid_.ok) {
            fused::FusedArgs* fa = &fused_mid_.args;
            fa->n_phases = 0; fa->preload = 1;
            add_blocks(fa, 0, mid_first_, mid_end_);
            add_conv1_phase(fa, -1, fused::PH_NORM, C);
            for (int p = 0; p < fa->n_phases; ++p) fused::derive_issue_fields(fa->ph[p], fa->R, fa->W);
        }
        for (fused::FusedArgs* fa : {&fused_rec_.args, &fused_init_.args})
            for (int p = 0; p < fa->n_phases; ++p) fused::derive_issue_fields(fa->ph[p], fa->R, fa->W);
        for (fused::FusedArgs* fa : {&fused_rec_.args, &fused_init_.args, &fused_mid_.args}) mark_split(fa);
        // reduced head inputs and MLP descriptors
        red_sets_.clear();   // reduced head inputs: one set of buffers per stream (red_for), so that the parts of a pipelined loop do not share them
        const uint8_t* blob = reinterpret_cast<const uint8_t*>(blob_.p);
        memset(&mlp_, 0, sizeof(mlp_));
        for (int h = 0; h < 3; ++h) {
            fused::MlpHead& m = mlp_.h[h];
            m.x = nullptr;   // per launch: the calling stream's buffers
            m.w0t = reinterpret_cast<const float*>(blob + fp32_bytes_ + w0t_rel_[h]);
            m.n_layers = heads_[h].n_layers;
            for (int l = 0; l <= m.n_layers; ++l) m.sizes[l] = heads_[h].sizes[l];
            for (int l = 0; l < m.n_layers; ++l) { m.W[l] = heads_[h].W[l]; m.b[l] = heads_[h].b[l]; }
            m.out_kind = heads_[h].out_kind; m.step = heads_[h].step; m.min_value = heads_[h].min_value; m.contractive = heads_[h].contractive;
        }
        int n0max = 0;
        for (int h = 0; h < 3; ++h) n0max = std::max(n0max, heads_[h].sizes[0]);
        mlp_mg_ = 16;
        (void)n0max;
        return MZB_OK;
    }

    // balanced tiles: every SM gets ceil-or-floor(B / SMs) games, cut into as few tiles of <= gmax games as possible,
    // emitted wave by wave (the hardware scheduler hands out CTAs in index order)
    int tiles_for(int B, int gmax, const TileTab** out) {
        const long key = ((long)B * 2 + (pipelined_ ? 1 : 0)) * 32 + gmax;
        auto it = tiles_.find(key);
        if (it == tiles_.end()) {
            std::vector<int> first, count;
            const int S = n_sm_ > 0 ? n_sm_ : 148;
            if (pipelined_) {
                // part of a pipelined loop: full tiles (a tile costs nearly the same whatever it holds — the weight stream and the
                // fixed phases do not shrink with it), the remainder spread over the last tiles
                // (MZB_PIPELINE_SM_CUT=1, experiment: smaller tiles on every SM instead — measured slower)
                int nt = (B + gmax - 1) / gmax;
                if (nt < S && getenv("MZB_PIPELINE_SM_CUT")) nt = std::min(S, B);
                int g = 0;
                // the larger tiles FIRST: CTAs are handed out in index order, so the SMs that come free first take them and the
                // SMs that run late get the smaller ones — a late SM catches up instead of falling further behind every simulation
                // (the next launch of this part waits for the LAST tile, then for the tree step)
                for (int k = 0; k < nt; ++k) { const int part = B / nt + (k < B % nt ? 1 : 0); first.push_back(g); count.push_back(part); g += part; }
            } else {
            const int per = (B + S - 1) / S;
            const int waves = (per + gmax - 1) / gmax;
            std::vector<int> left(S), next(S);
            int g = 0;
            for (int k = 0; k < S; ++k) { left[k] = B / S + (k < B % S ? 1 : 0); next[k] = g; g += left[k]; }
            for (int w = 0; w < waves; ++w)
                for (int k = 0; k < S; ++k) {
                    if (left[k] <= 0) continue;
                    const int part = std::min(gmax, (left[k] + (waves - w) - 1) / (waves - w));
                    first.push_back(next[k]); count.push_back(part);
                    next[k] += part; left[k] -= part;
                }
            }
            std::unique_ptr<TileTab> tt(new TileTab());
            tt->n_tiles = (int)first.size();
            tt->h_count = count;
            MZB_CUDA(tt->first.alloc(first.size() * sizeof(int)));
            MZB_CUDA(tt->count.alloc(count.size() * sizeof(int)));
            MZB_CUDA(cudaMemcpy(tt->first.p, first.data(), first.size() * sizeof(int), cudaMemcpyHostToDevice));
            MZB_CUDA(cudaMemcpy(tt->count.p, count.data(), count.size() * sizeof(int), cudaMemcpyHostToDevice));
            it = tiles_.emplace(key, std::move(tt)).first;
        }
        *out = it->second.get();
        return MZB_OK;
    }

    int launch_fused(FusedChain& ch, int B, const bf16* in, const int* in_slot, long in_stride, const int* out_slot,
                     const int* action, const NetOutputs& out, cudaStream_t s, bf16* out_buf = nullptr, long out_stride = 0) {
        const TileTab* tt = nullptr;
        int rc = tiles_for(B, ch.gmax, &tt);
        if (rc) return rc;
        fused::FusedArgs a = ch.args;
        a.in = in; a.in_slot = in_slot; a.in_stride = in_stride;
        a.pool = out_buf ? out_buf : pool_.as<bf16>(); a.out_slot = out_slot; a.pool_stride = out_buf ? out_stride : (long)slot_elems_;
        a.action = action;
        a.blob = reinterpret_cast<const uint8_t*>(blob_.p);
        RedSet* rs = nullptr;
        rc = red_for(s, &rs);
        if (rc) return rc;
        for (int h = 0; h < 3; ++h) a.red[h] = rs->r[h].as<float>();
        if (a.z_bytes) {
            double* outs[3] = {out.reward, out.policy, out.value};
            const int strides[3] = {1, plan_dims_.A, 1};
            for (int h = 0; h < 3; ++h) {
                fused::HeadTail& t = a.head[h];
                t.n_layers = heads_[h].n_layers;
                for (int l = 0; l <= t.n_layers; ++l) t.sizes[l] = heads_[h].sizes[l];
                for (int l = 0; l < t.n_layers; ++l) { t.W[l] = heads_[h].W[l]; t.b[l] = heads_[h].b[l]; }
                t.out_kind = heads_[h].out_kind; t.step = heads_[h].step; t.min_value = heads_[h].min_value; t.contractive = heads_[h].contractive;
                t.out = outs[h]; t.out_stride = strides[h];
            }
        }
        a.tile_first = tt->first.as<int>(); a.tile_count = tt->count.as<int>();
        if (!dbg_host_) {
            MZB_CUDA(cudaHostAlloc((void**)&dbg_host_, 16 * sizeof(int), cudaHostAllocMapped));
            memset(dbg_host_, 0, 16 * sizeof(int));
            MZB_CUDA(cudaHostGetDevicePointer((void**)&dbg_dev_, dbg_host_, 0));
        }
        a.dbg = dbg_dev_;
        static const int dbg_flags = getenv("MZB_FUSED_DBGFLAGS") ? atoi(getenv("MZB_FUSED_DBGFLAGS")) : 0;
        a.dbg_flags = dbg_flags;
        static const bool timing = getenv("MZB_FUSED_TIMING") != nullptr;
        a.tim = nullptr;
        if (timing) {
            const size_t need = (size_t)tt->n_tiles * fused::F_MAX_PHASES * 4 * sizeof(long long);
            if (tim_.bytes < need) MZB_CUDA(tim_.alloc(need));
            MZB_CUDA(cudaMemsetAsync(tim_.p, 0, need, s));
            a.tim = tim_.as<long long>();
        }
        // MZB_TILE_TRACE=1 (diagnostic): every launch's tiles stamp their start / end; dump_tile_trace() prints them per launch
        static const bool tile_trace = getenv("MZB_TILE_TRACE") != nullptr;
        a.tile_trace = nullptr;
        if (tile_trace && trace_launches_.size() < 256) {
            if (!trace_.p) MZB_CUDA(trace_.alloc((size_t)256 * 1024 * 4 * sizeof(long long)));
            a.tile_trace = trace_.as<long long>() + trace_launches_.size() * 1024 * 4;
            trace_launches_.push_back(std::min(tt->n_tiles, 1024));
        }
        fused::tower_fused_kernel<<<tt->n_tiles, fused::F_THREADS, ch.smem, s>>>(a);
        MZB_CUDA(cudaGetLastError());
        if (timing) {
            MZB_CUDA(cudaStreamSynchronize(s));
            std::vector<long long> h((size_t)tt->n_tiles * fused::F_MAX_PHASES * 4);
            MZB_CUDA(cudaMemcpy(h.data(), tim_.p, h.size() * sizeof(long long), cudaMemcpyDeviceToHost));
            fprintf(stderr, "[fused] %d tiles, %d phases; per phase avg cycles over tiles: wait(prev epi done -> mma start) | mma issue | mma start -> accum full | epilogue\n", tt->n_tiles, a.n_phases);
            const int gsel = getenv("MZB_FUSED_TIMING_G") ? atoi(getenv("MZB_FUSED_TIMING_G")) : 0;   // only tiles of this many games (0 = all)
            for (int p = 0; p < a.n_phases; ++p) {
                double issue = 0, full = 0, epi = 0, gap = 0;
                int n = 0;
                for (int t = 0; t < tt->n_tiles; ++t) {
                    if (gsel && tt->h_count[t] != gsel) continue;
                    const long long* d = &h[((size_t)t * fused::F_MAX_PHASES + p) * 4];
                    issue += d[1] - d[0]; full += d[2] - d[0]; epi += d[3] - d[2];
                    if (p > 0) gap += d[0] - h[((size_t)t * fused::F_MAX_PHASES + p - 1) * 4 + 3];
                    ++n;
                }
                if (n == 0) n = 1;
                fprintf(stderr, "[fused]  phase %2d nkb %3d flags %3d: gap %7.0f | issue %7.0f | full %7.0f | epi %7.0f\n", p, a.ph[p].nkb, a.ph[p].flags, gap / n, issue / n, full / n, epi / n);
            }
            // whole-CTA view, by tile size: start -> first MMA (prologue + input load), first MMA -> last epilogue end, -> CTA end
            double pro[fused::F_MAX_G + 1] = {0}, body[fused::F_MAX_G + 1] = {0}, tail[fused::F_MAX_G + 1] = {0}, setup[fused::F_MAX_G + 1] = {0}, inp[fused::F_MAX_G + 1] = {0};
            int cnt[fused::F_MAX_G + 1] = {0};
            for (int t = 0; t < tt->n_tiles; ++t) {
                const long long* row = &h[(size_t)t * fused::F_MAX_PHASES * 4];
                const long long* ends = row + (size_t)(fused::F_MAX_PHASES - 1) * 4;
                const int g = tt->h_count[t];
                setup[g] += ends[2] - ends[0]; inp[g] += ends[3] - ends[2];
                pro[g] += row[0] - ends[0]; body[g] += row[(size_t)(a.n_phases - 1) * 4 + 3] - row[0]; tail[g] += ends[1] - row[(size_t)(a.n_phases - 1) * 4 + 3];
                ++cnt[g];
            }
            {   // last MLP phase, relative to its accumulator-ready time: Hs ready | layers done | outputs written | phase end
                double m0 = 0, m1 = 0, m2 = 0, m3 = 0;
                for (int t = 0; t < tt->n_tiles; ++t) {
                    const long long* row = &h[(size_t)t * fused::F_MAX_PHASES * 4];
                    const long long* pr = row + (size_t)(fused::F_MAX_PHASES - 2) * 4;
                    const long long* last = row + (size_t)(a.n_phases - 1) * 4;
                    m0 += pr[0] - last[2]; m1 += pr[1] - last[2]; m2 += pr[2] - last[2]; m3 += last[3] - last[2];
                }
                const int n = tt->n_tiles;
                fprintf(stderr, "[fused]  last MLP tail: Hs ready %6.0f | layers done %6.0f | outputs written %6.0f | end %6.0f\n", m0 / n, m1 / n, m2 / n, m3 / n);
            }
            for (int g = 1; g <= fused::F_MAX_G; ++g)
                if (cnt[g]) fprintf(stderr, "[fused]  tiles of %d games: %d; prologue %7.0f (set-up %5.0f, input %5.0f) | phases %7.0f | tail %7.0f cycles\n", g, cnt[g], pro[g] / cnt[g], setup[g] / cnt[g], inp[g] / cnt[g], body[g] / cnt[g], tail[g] / cnt[g]);
        }
        static const bool sync_debug = getenv("MZB_FUSED_SYNC") != nullptr;
        if (sync_debug) {
            cudaError_t e = cudaStreamSynchronize(s);
            if (e != cudaSuccess) {
                set_error("tower_fused_kernel failed: %s; barrier diagnostics {code %d (1 empty, 2 act_ready, 3 full, 4 accum_full), block %d, parity %d, aux %d}",
                          cudaGetErrorString(e), dbg_host_[0], dbg_host_[1], dbg_host_[2], dbg_host_[3]);
                return MZB_ERR_CUDA;
            }
        }
        return MZB_OK;
    }
    int launch_mlp(int B, int first_head, const NetOutputs& out, cudaStream_t s) {
        fused::MlpArgs m;
        memset(&m, 0, sizeof(m));
        double* outs[3] = {out.reward, out.policy, out.value};
        const int strides[3] = {1, plan_dims_.A, 1};
        int n = 0, n0max = 0;
        RedSet* rs = nullptr;
        int rrc = red_for(s, &rs);
        if (rrc) return rrc;
        for (int h = first_head; h < 3; ++h) {
            m.h[n] = mlp_.h[h]; m.h[n].x = rs->r[h].as<float>(); m.h[n].out = outs[h]; m.h[n].out_stride = strides[h];
            n0max = std::max(n0max, m.h[n].sizes[0]);
            ++n;
        }
        m.n_heads = n; m.B = B; m.mg = mlp_mg_;
        // split the first layer's K range over several CTAs when it is long (ALE shape: 2304 inputs = 72 K tiles, a latency
        // chain of ~1.3 us per tile on one CTA); a fixed function of the layer size, so results do not depend on the batch
        static const int split_env = getenv("MZB_MLP_SPLITS") ? atoi(getenv("MZB_MLP_SPLITS")) : 0;
        m.splits = split_env > 0 ? std::min(split_env, MLP_SPLITS_MAX) : (n0max >= 1024 ? 8 : 1);   // 2 / 4 / 6 / 8 splits: 0.328 / 0.329 / 0.322 / 0.322 ms per ALE-shaped recurrent pass of 1024 states
        m.part = rs->part.as<float>(); m.part_head_stride = (long)MLP_SPLITS_MAX * max_batch_ * 256; m.counter = rs->counter.as<int>();
        const size_t smem = ((size_t)mlp_mg_ * 512 + 16 * (fused::MLP_KT + 4) + fused::MLP_KT * 64 + fused::MLP_WT) * sizeof(float);
        (void)n0max;
        dim3 grid((B + mlp_mg_ - 1) / mlp_mg_, n, m.splits);
        fused::mlp_heads_kernel<<<grid, fused::MLP_THREADS, smem, s>>>(m);
        MZB_CUDA(cudaGetLastError());
        return MZB_OK;
    }

public:
    // the one-launch fused recurrent pass keeps everything in shared memory / TMEM: no global scratch shared between calls
    void set_pipelined(bool on) override { pipelined_ = on; if (!on) { ev_wait_ = nullptr; ev_record_ = nullptr; dump_tile_trace(); } }
    void set_tower_events(cudaEvent_t wait_before, cudaEvent_t record_after) override { ev_wait_ = wait_before; ev_record_ = record_after; }
    bool recurrent_has_followup() const override { return fused_rec_.ok && fused_rec_.args.z_bytes == 0; }
    void dump_tile_trace() {
        if (trace_launches_.empty()) return;
        cudaDeviceSynchronize();
        std::vector<long long> h((size_t)trace_launches_.size() * 1024 * 4);
        cudaMemcpy(h.data(), trace_.p, h.size() * sizeof(long long), cudaMemcpyDeviceToHost);
        long long t_ref = h[0];
        for (size_t l = 0; l < trace_launches_.size(); ++l) {
            const long long* r = &h[l * 1024 * 4];
            long long s0 = r[0], s1 = r[0], e0 = r[1], e1 = r[1];
            double cyc = 0, ns = 0;
            for (int t = 0; t < trace_launches_[l]; ++t) {
                s0 = std::min(s0, r[t * 4]); s1 = std::max(s1, r[t * 4]); e0 = std::min(e0, r[t * 4 + 1]); e1 = std::max(e1, r[t * 4 + 1]);
                cyc += (double)(r[t * 4 + 3] - r[t * 4 + 2]); ns += (double)(r[t * 4 + 1] - r[t * 4]);
            }
            const int n = trace_launches_[l];
            fprintf(stderr, "[tiles] launch %3zu: %3d tiles; CTA starts %8.1f .. %8.1f us, ends %8.1f .. %8.1f us; per tile %7.0f cycles in %6.1f us (%.0f MHz)\n", l, n,
                    (s0 - t_ref) * 1e-3, (s1 - t_ref) * 1e-3, (e0 - t_ref) * 1e-3, (e1 - t_ref) * 1e-3, cyc / n, ns / n * 1e-3, cyc / ns * 1e3);
        }
        trace_launches_.clear();
    }
    // EXPERIMENT (MZB_PIPELINE_SM_CUT=1; default = equal parts): two parts whose launches are whole waves over the SMs.
    // q = games per SM and simulation.  One wave's worth or less (q <= gmax): the parts take ceil(q/2) and floor(q/2) games
    // per SM.  More: part 0 takes floor(waves/2) waves of full tiles, part 1 the rest (C4d, T = 1024, 148 SMs: q = 7 -> 592
    // trees as 148 tiles of 4, 432 trees as 148 tiles of 3).  Measured SLOWER than equal halves of full tiles (2 x 128 tiles
    // of 4): with 256 tiles over 148 SMs the equal cut already rolls (a part's CTAs start wherever an SM comes free; 231 us
    // per simulation against 249 for two tile times), while the SM-filling cut runs in lock step and every launch waits for
    // its part's last tile plus the tree step (240 us against an ideal of 231).  DESIGN §4.5.
    void pipeline_cut(int T, int K, int* t0) const override {
        const int S = n_sm_ > 0 ? n_sm_ : 148, gmax = fused_rec_.gmax;
        if (K != 2 || gmax <= 0 || T < 2 || !getenv("MZB_PIPELINE_SM_CUT")) { NetBackend::pipeline_cut(T, K, t0); return; }
        const int q = (T + S - 1) / S, waves = (q + gmax - 1) / gmax;
        long T1;
        if (waves <= 1) T1 = ((long)T * ((q + 1) / 2) + q - 1) / q;
        else T1 = (long)S * gmax * (waves / 2);
        if (const char* ov = getenv("MZB_PIPELINE_CUT")) T1 = atol(ov);   // experiments
        if (T1 < 1) T1 = 1;
        if (T1 > T - 1) T1 = T - 1;
        t0[0] = 0; t0[1] = (int)T1; t0[2] = T;
    }
    bool recurrent_concurrent() const override {
        // (wide heads run a second kernel, mlp_heads_kernel, on the reduced activations: those buffers are per stream)
        return uploaded_ && fused_final_ && fused_rec_.ok && (fused_rec_.args.z_bytes > 0 || !getenv("MZB_NO_PIPELINE_WIDE")) && !getenv("MZB_FUSED_TIMING");
    }
    int recurrent(int B, const int* d_in_slot, const int* d_action, const int* d_out_slot, const NetOutputs& out, cudaStream_t s) override {
        MZB_REQUIRE(uploaded_, "weights not loaded (mzb_load_weights)");
        int rc = finalize_fused();
        if (rc) return rc;
        if (!fused_rec_.ok) return direct::DirectBackend<bf16>::recurrent(B, d_in_slot, d_action, d_out_slot, out, s);
        MZB_REQUIRE(B <= max_batch_, "batch %d above engine capacity %d", B, max_batch_);
        if (ev_wait_) MZB_CUDA(cudaStreamWaitEvent(s, ev_wait_, 0));
        rc = launch_fused(fused_rec_, B, pool_.as<bf16>(), d_in_slot, (long)slot_elems_, d_out_slot, d_action, out, s);
        if (rc) return rc;
        if (ev_record_) MZB_CUDA(cudaEventRecord(ev_record_, s));
        ev_wait_ = nullptr; ev_record_ = nullptr;
        launches_recurrent_ = 1;
        if (!fused_rec_.args.z_bytes) {
            rc = launch_mlp(B, 0, out, s);
            if (rc) return rc;
            launches_recurrent_ = 2;
        }
        return MZB_OK;
    }
    int initial(const float* d_obs, int B, const int* d_out_slot, const NetOutputs& out, cudaStream_t s) override {
        MZB_REQUIRE(uploaded_, "weights not loaded (mzb_load_weights)");
        int rc = finalize_fused();
        if (rc) return rc;
        if (!fused_init_.ok) return direct::DirectBackend<bf16>::initial(d_obs, B, d_out_slot, out, s);
        MZB_REQUIRE(B <= max_batch_, "batch %d above engine capacity %d", B, max_batch_);
        const NetDims& d = plan_dims_;
        const direct::ConvDev& c0 = convs_[0][0];
        const size_t head_smem = ((size_t)c0.cin * (c0.in_h + 2) * (c0.in_w + 2) + (size_t)9 * c0.cin * c0.cout) * sizeof(float);
        const bf16* chain_in = bufB_.as<bf16>();
        long chain_stride = (long)slot_elems_;
        if (init_tail_ == 1 && c0.stride == 1 && c0.cin <= 16 && !c0.has_table && c0.in_h == c0.out_h && c0.in_w == c0.out_w && head_smem <= 48 * 1024) {
            const float* base = blob_.as<float>();
            head_conv_small_kernel<<<B, 128, head_smem, s>>>(d_obs, (long)d.initial_flat(), base + c0.w_off, base + c0.bias_off, bufB_.as<bf16>(),
                                                            (long)slot_elems_, c0.cin, c0.cout, c0.in_h, c0.in_w, c0.relu);
            MZB_CUDA(cudaGetLastError());
            launches_ = 1;
        } else {
            // the steps in front of the fused chain, layer by layer: first conv (conv_first_kernel from the fp32 observation when
            // it is a stride-2 head, else layout conversion + the generic path), then band / per-layer tensor kernels and pools
            bf16* cur = nullptr;
            launches_ = 0;
            if (first_ok_) {
                first::FirstArgs a = first_args_;
                a.obs = d_obs; a.obs_stride = (long)d.initial_flat();
                a.out = bufB_.as<bf16>();
                a.wimg = reinterpret_cast<const uint8_t*>(blob_.p) + fp32_bytes_ + first_off_;
                a.bias = blob_.as<float>() + c0.bias_off;
                a.pkm = pkm_chain_ ? 1 : 0;
                cur_pkm_ = pkm_chain_;
                first::conv_first_kernel<<<B * a.bands, first::C1_THREADS, first::first_smem_bytes(a.cls, a.rbo, a.Wo, a.cout), s>>>(a);
                MZB_CUDA(cudaGetLastError());
                ++launches_;
                rc = run_tower(0, B, bufB_.as<bf16>(), nullptr, 0, nullptr, s, &cur, 1, init_tail_);
            } else {
                long total = (long)B * d.initial_flat();
                direct::nchw_to_nhwc_kernel<bf16><<<(unsigned)((total + 255) / 256), 256, 0, s>>>(d_obs, bufA_.as<bf16>(), nullptr, (long)d.initial_flat(), B, d.cin, d.in_h * d.in_w);
                ++launches_;
                rc = run_tower(0, B, bufA_.as<bf16>(), nullptr, 0, nullptr, s, &cur, 0, init_tail_);
            }
            if (rc) return rc;
            chain_in = cur;
            chain_stride = (long)d.C * d.hw();
        }
        rc = launch_fused(fused_init_, B, chain_in, nullptr, chain_stride, d_out_slot, nullptr, out, s);
        if (rc) return rc;
        if (!fused_init_.args.z_bytes) {
            rc = launch_mlp(B, 1, out, s);
            if (rc) return rc;
        }
        initial_reward_kernel<<<(B + 127) / 128, 128, 0, s>>>(out.reward, B, renc_.support, (float)renc_.step, (float)renc_.min_value, renc_.contractive);
        MZB_CUDA(cudaGetLastError());
        launches_initial_ = launches_ + 2 + (fused_init_.args.z_bytes ? 0 : 1);   // + fused chain, (+ MLP kernel), + initial reward
        return MZB_OK;
    }

protected:
    FusedChain fused_rec_, fused_init_, fused_mid_;
    size_t mid_first_ = 0, mid_end_ = 0;   // representation steps [mid_first_, mid_end_) run as fused_mid_
    bool fused_final_ = false;
    bool pipelined_ = false;
    cudaEvent_t ev_wait_ = nullptr, ev_record_ = nullptr;
    size_t init_tail_ = 0;            // first representation step of the fused initial chain
    bool first_ok_ = false; size_t first_off_ = 0; first::FirstArgs first_args_;   // conv_first.cuh
    bool pkm_chain_ = false, cur_pkm_ = false; size_t pkm_max_elems_ = 0; int pkm_last_conv_ = 0;   // PKM segment of the downsample head
    size_t conv1_rel_[2] = {0, 0}, conv1_bias_rel_[2] = {0, 0}, w0t_rel_[3] = {0, 0, 0};
    size_t mlp0_rel_[2] = {0, 0}, mlp0_bias_rel_[2] = {0, 0};
    int mlp0_nkb_[2] = {0, 0};
    static constexpr int MLP_SPLITS_MAX = 8;
    struct RedSet { DevBuf r[3], part, counter; };   // + split-K scratch of mlp_heads_kernel
    std::map<cudaStream_t, std::unique_ptr<RedSet>> red_sets_;
    int red_for(cudaStream_t s, RedSet** out) {
        auto it = red_sets_.find(s);
        if (it == red_sets_.end()) {
            std::unique_ptr<RedSet> rs(new RedSet());
            const int hw = plan_dims_.hw();
            for (int h = 0; h < 3; ++h) MZB_CUDA(rs->r[h].alloc((size_t)max_batch_ * heads_[h].red * hw * sizeof(float)));
            MZB_CUDA(rs->part.alloc((size_t)3 * MLP_SPLITS_MAX * max_batch_ * 256 * sizeof(float)));
            MZB_CUDA(rs->counter.alloc((size_t)3 * (max_batch_ + 1) * sizeof(int)));
            MZB_CUDA(cudaMemset(rs->counter.p, 0, rs->counter.bytes));
            it = red_sets_.emplace(s, std::move(rs)).first;
        }
        *out = it->second.get();
        return MZB_OK;
    }
    fused::MlpArgs mlp_;
    int mlp_mg_ = 16, n_sm_ = 0;
    DevBuf trace_;
    std::vector<int> trace_launches_;
    int* dbg_host_ = nullptr; int* dbg_dev_ = nullptr;
    DevBuf tim_;
    std::map<long, std::unique_ptr<TileTab>> tiles_;

    int fast_conv(const direct::ConvCall<bf16>& call, cudaStream_t s) override {
        const direct::ConvDev& c = *call.conv;
        const Packed& p = packed_[c.tower][c.index];
        if (!p.ok) return p.image && !c.has_table ? band_conv(call, p, s) : 0;
        const float* base = blob_.as<float>();
        UmmaArgs a;
        a.in = call.in; a.in_slot = call.in_slot; a.in_stride = call.in_stride;
        a.out = call.out; a.out_slot = call.out_slot; a.out_stride = call.out_stride;
        a.res = call.res; a.res_slot = call.res_slot; a.res_stride = call.res_stride;
        a.wpack = reinterpret_cast<const bf16*>(reinterpret_cast<const uint8_t*>(blob_.p) + fp32_bytes_ + p.off);
        a.bias = base + c.bias_off;
        a.act_table = c.has_table ? base + c.table_off : nullptr;
        a.action = call.action;
        a.B = call.B; a.C = c.cin; a.cout = c.cout; a.relu = c.relu;
        a.g = p.g;
        const int tiles = (call.B + p.g.G - 1) / p.g.G;
        static const bool debug = getenv("MZB_UMMA_DEBUG") != nullptr;
        a.dbg = nullptr;
        if (debug) {
            if (dbg_.bytes < (size_t)tiles * 8 * sizeof(long long)) MZB_CUDA(dbg_.alloc((size_t)tiles * 8 * sizeof(long long)));
            a.dbg = dbg_.as<long long>();
        }
        umma_conv3_kernel<<<tiles, UMMA_THREADS, umma_smem_bytes(c.cin, p.g), s>>>(a);
        if (debug) {
            MZB_CUDA(cudaStreamSynchronize(s));
            std::vector<long long> h((size_t)tiles * 8);
            MZB_CUDA(cudaMemcpy(h.data(), dbg_.p, h.size() * sizeof(long long), cudaMemcpyDeviceToHost));
            double st = 0, issue = 0, mma = 0, p1 = 0, sync = 0, p2 = 0;
            long long t0 = h[0], t1 = h[3];
            for (int i = 0; i < tiles; ++i) {
                const long long* d = &h[(size_t)i * 8];
                st += d[1] - d[0]; issue += d[6] - d[1]; mma += d[2] - d[1]; p1 += d[4] - d[2]; sync += d[5] - d[4]; p2 += d[3] - d[5];
                if (d[0] < t0) t0 = d[0];
                if (d[3] > t1) t1 = d[3];
            }
            fprintf(stderr, "[umma] tower %d conv %d tiles %d N %d (cycles/CTA avg): stage %.0f | mma issue %.0f, accum ready %.0f | tmem->smem %.0f | barrier %.0f | store %.0f | kernel span %lld\n",
                    c.tower, c.index, tiles, p.g.N, st / tiles, issue / tiles, mma / tiles, p1 / tiles, sync / tiles, p2 / tiles, t1 - t0);
        }
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { set_error("umma_conv3_kernel launch failed: %s", cudaGetErrorString(e)); return MZB_ERR_CUDA; }
        return 1;
    }

    // geometry of conv_band2_kernel for one convolution; false when it does not fit
    static bool band2_geometry(int cin, int cout, int stride, int H, int W, band::Band2Args* g2) {
        if (cin % 64 != 0 || cout > 128 || cout % 8 != 0 || (stride != 1 && stride != 2)) return false;
        band::BandArgs& a = g2->b;
        a.stride = stride;
        a.rb = band::band_rows(H, W);
        if (stride == 2) a.rb = a.rb / 2 * 2;
        if (a.rb < 1) return false;
        a.bands = (H + a.rb - 1) / a.rb;
        a.N = ((a.rb + 2) * (W + 1) + 15) / 16 * 16;
        a.HALO = W + 2;
        a.R = a.HALO + a.N + a.HALO;
        a.x_bytes = (int)((((size_t)(cin / 8) * a.R * 16) + 127) / 128 * 128);
        a.H = H; a.W = W; a.cin = cin; a.cout = cout;
        // fp32 staging [K chunk][pixel][8] with 8 floats of padding per plane; its own buffer for the whole band when that fits
        // beside two X buffers and >= 4 ring slots (the 64-channel stage), else aliased on X[j % 2] in row chunks
        const size_t kco = cout / 8;
        auto stage_bytes = [&](int rows) { return kco * ((size_t)rows * W * 8 + 8) * sizeof(float); };
        long avail = 227L * 1024 - 2L * a.x_bytes - 256;
        g2->sep = (avail - (long)stage_bytes(a.rb) >= 4L * band::B_KB_BYTES && !getenv("MZB_BAND_NO_SEP")) ? 1 : 0;
        if (g2->sep) { avail -= (long)stage_bytes(a.rb); g2->rc = a.rb; }
        else { g2->rc = a.rb; while (g2->rc >= 1 && stage_bytes(g2->rc) > (size_t)a.x_bytes) --g2->rc; }
        g2->ns = (int)std::min<long>(8, avail / band::B_KB_BYTES);
        g2->ps = g2->rc * W * 8 + 8;
        return a.N <= 256 && g2->ns >= 2 && g2->rc >= 1;
    }
    size_t act_elems(size_t mx) const override { return std::max(mx, pkm_max_elems_); }

    int fast_steps(int tower, size_t si, int B, const bf16* in, bf16* out, cudaStream_t s, size_t* used) override {
        *used = 0;
        if (tower != 0 || !fused_mid_.ok || si != mid_first_ || cur_pkm_ || !fused_final_ || fused_mid_.args.n_phases == 0) return 0;
        const long stride = (long)fused_mid_.args.C * fused_mid_.args.HW;
        NetOutputs none;
        int rc = launch_fused(fused_mid_, B, in, nullptr, stride, nullptr, nullptr, none, s, out, stride);
        if (rc) return rc;
        *used = mid_end_ - mid_first_;
        return 1;
    }

    int fast_pool(const bf16* in, bf16* out, int B, int C, int in_h, int in_w, int out_h, int out_w, cudaStream_t s) override {
        if (C % 8 != 0) return 0;
        const long total = (long)B * out_h * out_w * (C / 8);
        const int pkm_in = cur_pkm_ ? 1 : 0;
        cur_pkm_ = false;   // pools write NHWC
        first::avgpool3_vec_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(in, out, B, C / 8, in_h, in_w, out_h, out_w, pkm_in);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { set_error("avgpool3_vec_kernel launch failed: %s", cudaGetErrorString(e)); return MZB_ERR_CUDA; }
        return 1;
    }

    // boards too large for a whole-game tile (downsample representation head): one CTA per band of rows (conv_band.cuh)
    int band_conv(const direct::ConvCall<bf16>& call, const Packed& p, cudaStream_t s) {
        static const bool off = getenv("MZB_NO_BAND_CONV") != nullptr;
        if (off) return 0;
        const direct::ConvDev& c = *call.conv;
        band::BandArgs a;
        a.stride = c.stride;
        a.rb = band::band_rows(c.in_h, c.in_w);
        if (c.stride == 2) a.rb = a.rb / 2 * 2;   // bands start on even rows
        if (a.rb < 1) return 0;
        a.bands = (c.in_h + a.rb - 1) / a.rb;
        a.N = ((a.rb + 2) * (c.in_w + 1) + 15) / 16 * 16;
        a.HALO = c.in_w + 2;
        a.R = a.HALO + a.N + a.HALO;
        a.x_bytes = (int)((((size_t)(c.cin / 8) * a.R * 16) + 127) / 128 * 128);
        // fp32 output staging reuses the ring after the MMAs: it must hold rb rows x W pixels x cout floats
        const size_t stage_need = (size_t)a.rb * c.in_w * c.cout * sizeof(float);
        a.nstage = (band::band_smem_bytes(a.x_bytes, 2) <= 113 * 1024 && stage_need <= 2 * (size_t)band::B_STAGE_BYTES) ? 2 : 4;
        const size_t smem = band::band_smem_bytes(a.x_bytes, a.nstage);
        a.in = call.in; a.in_slot = call.in_slot; a.in_stride = call.in_stride;
        a.out = call.out; a.out_slot = call.out_slot; a.out_stride = call.out_stride;
        a.res = call.res; a.res_slot = call.res_slot; a.res_stride = call.res_stride;
        a.wimg = reinterpret_cast<const uint8_t*>(blob_.p) + fp32_bytes_ + p.off;
        a.bias = blob_.as<float>() + c.bias_off;
        a.H = c.in_h; a.W = c.in_w; a.cin = c.cin; a.cout = c.cout; a.relu = c.relu;
        if (cur_pkm_) {
            // inside the PKM segment: persistent, double-buffered kernel over padded K-major activations (conv_band2_kernel)
            if (!(pkm_chain_ && c.tower == 0 && c.index <= pkm_last_conv_)) { set_error("convolution %d of tower %d got a PKM activation outside the PKM segment", c.index, c.tower); return MZB_ERR_STATE; }
            band::Band2Args g2;
            memset(&g2, 0, sizeof(g2));
            if (!band2_geometry(c.cin, c.cout, c.stride, c.in_h, c.in_w, &g2)) { set_error("conv_band2_kernel geometry"); return MZB_ERR_STATE; }
            band::BandArgs& b = g2.b;
            b.in = call.in; b.out = call.out; b.res = call.res;
            b.in_slot = nullptr; b.out_slot = nullptr; b.res_slot = nullptr;
            b.in_stride = band::pkm_elems(c.cin, c.in_h, c.in_w);
            b.out_stride = band::pkm_elems(c.cout, c.out_h, c.out_w);
            b.res_stride = b.out_stride;
            b.wimg = a.wimg; b.bias = a.bias; b.relu = c.relu;
            static const int band_dbg = getenv("MZB_BAND_DBG") ? atoi(getenv("MZB_BAND_DBG")) : 0;
            g2.dbg = band_dbg;
            g2.total = call.B * b.bands;
            if (n_sm_ <= 0) { int dev = 0; MZB_CUDA(cudaGetDevice(&dev)); MZB_CUDA(cudaDeviceGetAttribute(&n_sm_, cudaDevAttrMultiProcessorCount, dev)); }
            if (!band2_attr_set_) {
                MZB_CUDA(cudaFuncSetAttribute(band::conv_band2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
                band2_attr_set_ = true;
            }
            const size_t stage = g2.sep ? (size_t)(c.cout / 8) * g2.ps * sizeof(float) : 0;
            band::conv_band2_kernel<<<std::min(n_sm_, g2.total), band::B_THREADS, band::band2_smem_bytes(b.x_bytes, g2.ns, stage), s>>>(g2);
            cudaError_t e2 = cudaGetLastError();
            if (e2 != cudaSuccess) { set_error("conv_band2_kernel launch failed: %s", cudaGetErrorString(e2)); return MZB_ERR_CUDA; }
            return 1;
        }
        if (a.N > 256 || smem > 227 * 1024 || stage_need > (size_t)a.nstage * band::B_STAGE_BYTES) return 0;
        if (!band_attr_set_) {
            MZB_CUDA(cudaFuncSetAttribute(band::conv_band_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
            band_attr_set_ = true;
        }
        a.in = call.in; a.in_slot = call.in_slot; a.in_stride = call.in_stride;
        a.out = call.out; a.out_slot = call.out_slot; a.out_stride = call.out_stride;
        a.res = call.res; a.res_slot = call.res_slot; a.res_stride = call.res_stride;
        a.wimg = reinterpret_cast<const uint8_t*>(blob_.p) + fp32_bytes_ + p.off;
        a.bias = blob_.as<float>() + c.bias_off;
        a.H = c.in_h; a.W = c.in_w; a.cin = c.cin; a.cout = c.cout; a.relu = c.relu;
        band::conv_band_kernel<<<call.B * a.bands, band::B_THREADS, smem, s>>>(a);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { set_error("conv_band_kernel launch failed: %s", cudaGetErrorString(e)); return MZB_ERR_CUDA; }
        return 1;
    }

    std::vector<Packed> packed_[3];
    bool band_attr_set_ = false, band2_attr_set_ = false;
    DevBuf dbg_;
    bool attr_set_ = false;
};

}  // namespace

NetBackend* make_bf16_backend() { return new Bf16Backend(); }

}  // namespace mzb
