// tower_fused.cuh — the whole residual tower chain of one inference call as ONE tcgen05 kernel.
//
// Replaces, in MZB_BF16 mode, the per-layer launches of MuzeroNetworkImpl::recurrent_inference / initial_inference
// (models.cpp:222-265): DynamicsNetwork head + blocks (models.cpp:105-117), reward conv1x1, normalize_hidden_state
// (models.cpp:16-26), PredictionNetwork blocks + policy/value conv1x1 (models.cpp:153-166); ResidualBlock =
// relu(bn2(conv2(relu(bn1(conv1 x)))) + x) (model_layers.cpp:70-88), eval-mode BN folded on the host.
//
// One CTA owns a tile of G whole games for the entire chain; nothing but the parent hidden state (in) and the child
// hidden state + the reduced head inputs (out) touches HBM:
//   * shared memory holds ONE activation buffer X, bf16, zero-padded pixel rows (every game = P rows: a zero row of
//     W+1 pixels, then H rows of W pixels + 1 zero pixel, padded to a multiple of 8) in K-major no-swizzle core-matrix
//     order [K chunk of 8 channels][row][16 B].  A 3x3 tap is a constant row shift, so the nine taps are nine
//     descriptors over the same resident tile (no im2col, no re-read).  Layer outputs overwrite X in place.
//   * tensor memory holds two fp32 accumulator regions of N <= 256 columns (column = padded pixel row, lane = output
//     channel): R1 for convolutions that start from zero, R2 for the second convolution of a residual block, which
//     accumulates ON TOP OF the block input kept there in fp32 — the skip connection never leaves TMEM and is never
//     rounded to bf16.
//   * BN-folded bf16 weights stream from L2 through a ring of 16 KB stages (64 K x 128 cout images pre-packed on the
//     host) with 1-D bulk async copies; the producer runs ahead across layer boundaries.
//   * per phase: [MMA issuer: 9 taps x C/64 stages x 4 tcgen05.mma] -> accumulator-full barrier -> [8 epilogue warps:
//     tcgen05.ld, + bias (+ per-action table), ReLU, bf16 store into X, fp32 tcgen05.st of the block input into R2]
//     -> activations-ready barrier -> next phase.  The 1x1 head convolutions are extra single-tap phases whose rows go
//     to a small fp32 buffer (channel-major flatten, models.cpp:114,161-162); the per-(game, channel) min-max
//     normalisation is a pure per-thread reduction because a thread owns one channel and its columns are pixels.
#pragma once
#include <cuda_bf16.h>

#include "umma.cuh"

namespace mzb {
namespace fused {

using bf16 = __nv_bfloat16;
using namespace umma;

constexpr int F_EPI_WARPS = 8;        // 2 per TMEM lane quarter (16 measured no faster: the epilogue is issue-bound); a warp owns games g = e (mod 2)
constexpr int F_EPI_THREADS = 32 * F_EPI_WARPS;
constexpr int F_EPI_SPLIT = F_EPI_WARPS / 4;
// Three warpgroups: warps 0-3 = service (0: weight producer, 1: MMA issuer + TMEM owner, 2-3 idle), warps 4-11 = epilogue.  The CTA is
// launched at 168 registers per thread (384 x 168 = the whole register file) and redistributes them with setmaxnreg: the service
// group shrinks to F_SVC_REGS, the two epilogue groups grow to F_EPI_REGS (per SM sub-partition: 72 + 2 x 216 = 504 <= 512).
// With 10 warps (3 on one sub-partition) every warp was held at 168 registers, which every deeper epilogue variant overflowed.
constexpr int F_SVC_THREADS = 128;
#ifndef MZB_SVC_REGS
#define MZB_SVC_REGS 72
#endif
constexpr int F_SVC_REGS = MZB_SVC_REGS, F_EPI_REGS = (512 - 8 - MZB_SVC_REGS) / 2 / 8 * 8;
constexpr int F_THREADS = F_SVC_THREADS + F_EPI_THREADS;
constexpr int F_NSTAGE = 4;                   // ring slots of two 64-K weight images each
constexpr int F_KB_BYTES = 64 * 128 * 2;      // one weight image: 64 K x 128 cout, bf16 (16 KB)
constexpr int F_STAGE_BYTES = 2 * F_KB_BYTES; // a ring slot = 128 K: halves the issuer's wait/commit overhead per MMA
constexpr int F_TMEM_COLS = 512;
constexpr int F_R2_COL = 256;
constexpr int F_MAX_PHASES = 40;
constexpr int F_MAX_G = 16;

enum : int {
    PH_RELU = 1,      // ReLU in the epilogue
    PH_ACC_R2 = 2,    // accumulate into R2 on top of the block input stored there (second conv of a residual block)
    PH_STASH = 4,     // keep the epilogue result in R2 (fp32) as the next block's skip input
    PH_TABLE = 8,     // dynamics head: stage the per-action table in R2 during the MMAs, add it in the epilogue
    PH_WRITE_X = 16,  // store the epilogue result into X (bf16)
    PH_RED = 32,      // 1x1 head convolution: rows of R1 -> red[] buffers (no activation)
    PH_NORM = 64,     // normalize_hidden_state on R2 -> X, node pool slot, R2
    PH_EARLY = 256,   // (with PH_NORM) the next phase does not read this one's output: release the MMA warp before the epilogue work
    PH_SPLIT_HEAD = 512,   // (host-derived) the phase's first K stages may run per half tile: the previous phase's epilogue hands over games [0, G/2) early
    PH_SPLIT_TAIL = 1024,  // (host-derived) the phase's last K stages run per half tile, so that the epilogue of games [0, G/2) starts under the other half's MMAs
    PH_MLP = 128,     // head MLPs: first layer on the tensor core (A = stacked W0 images, B = Z), later layers + output transform on CUDA cores
};

struct Phase {
    uint32_t w_off;        // byte offset of this phase's weight stages in the packed blob
    int nkb;               // 64-wide K stages (taps * C / 64); 0 = no tensor work
    int taps;              // 9 (3x3) or 1 (1x1)
    int bias_off;          // byte offset of the fp32 bias vector, -1 = none
    int table_off;         // unused (kept for layout stability)
    int flags;
    int red_head0, red_n0, red_head1, red_n1;   // PH_RED: rows [0,n0) -> red[head0], rows [n0, n0+n1) -> red[head1]
                                                // PH_MLP: hidden rows [0,n0) belong to head0, [n0, n0+n1) to head1 (n1 = 0: one head)
    int z_k0, z_k1;                             // PH_RED with Z: first K index of head0 / head1 inside Z
    // MMA issue parameters, derived on the host (derive_issue_fields) so that the issuing warp reads them as plain uniform
    // constant-bank loads: no division, no select on the device
    int kpt;               // K stages per tap
    int tap_off0;          // B row offset of the first tap: -(W+2) for 3x3, 0 for 1x1 / MLP
    uint32_t kk_b, cc_b;   // B descriptor step per MMA (16 K elements) / per K stage, in 16-byte units
    uint32_t is_mlp, onto; // 0 / 1
};
inline void derive_issue_fields(Phase& ph, int R, int W) {
    const bool mlp = (ph.flags & 128) != 0;   // PH_MLP
    ph.kpt = ph.taps > 0 && ph.nkb > 0 ? ph.nkb / ph.taps : 1;
    ph.tap_off0 = (ph.taps == 9) ? -(W + 2) : 0;
    ph.kk_b = mlp ? 32u : 2u * (uint32_t)R;
    ph.cc_b = mlp ? 128u : 8u * (uint32_t)R;
    ph.is_mlp = mlp ? 1u : 0u;
    ph.onto = (ph.flags & 2) ? 1u : 0u;       // PH_ACC_R2
}

// one head's MLP tail (layers >= 1) and output transform, for the in-kernel path
struct HeadTail {
    const float* W[6];     // [out][in] per layer (W[0] unused: layer 0 runs on the tensor core)
    const float* b[6];
    int sizes[7];
    int n_layers;
    int out_kind;          // 0 = softmax policy, 1 = support -> scalar
    float step, min_value; int contractive;
    double* out; int out_stride;
};

struct FusedArgs {
    const bf16* in; const int* in_slot; long in_stride;       // chain input, NHWC bf16 per game
    bf16* pool; const int* out_slot; long pool_stride;        // PH_NORM destination (node pool)
    const int* action;                                        // PH_TABLE
    const uint8_t* blob;                                      // weights: fp32 section + packed bf16 stages
    float* red[3];                                            // [B][red_h * HW] fp32 per head (reward, policy, value)
    int red_rows[3];
    const int* tile_first; const int* tile_count;             // tile -> first game, number of games
    int C, H, W, P, HALO, R, HW, x_bytes;
    int n_phases, preload, normalize;
    int split_nh, split_nt;                                   // staggered half tiles: ring slots issued per half at the head / tail of a 3x3 phase (0 = off)
    int dup;                                                  // C <= 64: weight-image rows 64-127 repeat rows 0-63, so lane quarters 2, 3 hold the same channels as 0, 1 and the four warp pairs split the GAMES four ways
    int dbg_flags;                                            // experiments: bit 0 = all taps unshifted (timing only, wrong results)
    int* dbg;                                                 // host-mapped diagnostics (barrier time-outs), may be null
    long long* tim;                                           // optional [tile][phase][4] clocks: mma start, mma issued, accum full, epilogue done
    long long* tile_trace;                                    // optional [tile][4]: globaltimer ns at CTA start / end, clock64 at start / end (MZB_TILE_TRACE)
    int table_off;                                            // PH_TABLE: byte offset of the fp32 per-action table [A][HW][C]
    int z_bytes;                                              // > 0: head inputs stay in shared memory (Z) and the MLPs run in-kernel
    int z_rp;                                                 // channels per padded pixel row in Z's K order (multiple of 8)
    HeadTail head[3];
    Phase ph[F_MAX_PHASES];
};

constexpr int F_HS_BYTES = 2 * 8 * 128 * 4;   // two ping-pong buffers [8 games][128 stacked units] fp32 for the MLP tails
// pixel table entries: look-ups run in 8-row chunks up to P + 7.  Sized by P (not a flat 256) so that the CTA stays >= 1 KB under
// the SM's shared memory: a tree-step block (no shared memory of its own, 1 KB driver reserve) must fit BESIDE a tower CTA
// for the pipelined simulation loop to overlap the two.
__host__ __device__ inline int fused_ptab_entries(int P) { return (P + 8 + 63) & ~63; }
inline size_t fused_smem_bytes(int x_bytes, int z_bytes, int P) {
    return (size_t)x_bytes + (size_t)F_NSTAGE * F_STAGE_BYTES + 256 + (size_t)fused_ptab_entries(P) * 4 + 3 * F_MAX_G * 4 + 32 + 128 + (z_bytes ? (size_t)z_bytes + F_HS_BYTES + 128 : 0);
}

// explicit shared-space accesses (32-bit shared addresses: no generic-pointer conversion per access)
__device__ __forceinline__ void sts_u16(uint32_t addr, unsigned short v) { asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"(v) : "memory"); }
__device__ __forceinline__ uint32_t lds_u16(uint32_t addr) {
    unsigned short v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ uint4 lds_v4u(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void lds_v4(uint32_t addr, int* p) {
    asm volatile("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(p[0]), "=r"(p[1]), "=r"(p[2]), "=r"(p[3]) : "r"(addr) : "memory");
}

// Two adjacent accumulator columns (rows J, J+1 of the chunk) of one lane: + bias, optional ReLU, one packed bf16x2
// conversion, predicated 2-byte stores into X (bit J / J+1 of m = the row is a real pixel).  Hand-scheduled PTX: the
// epilogue is instruction-issue bound (measured: 16 warps no faster than 8), so instructions per element is the metric.
template <int J, bool RELU>
__device__ __forceinline__ void epi_pair(uint32_t& v0, uint32_t& v1, float bias, uint32_t m, uint32_t xa) {
    float y0 = __uint_as_float(v0) + bias, y1 = __uint_as_float(v1) + bias;
    if (RELU) { y0 = fmaxf(y0, 0.f); y1 = fmaxf(y1, 0.f); }
    v0 = __float_as_uint(y0); v1 = __float_as_uint(y1);
    asm volatile(
        "{\n\t"
        ".reg .b32 pk, t0, t1;\n\t"
        ".reg .b16 lo, hi;\n\t"
        ".reg .pred p0, p1;\n\t"
        "cvt.rn.bf16x2.f32 pk, %1, %0;\n\t"
        "mov.b32 {lo, hi}, pk;\n\t"
        "and.b32 t0, %2, %4;\n\t"
        "and.b32 t1, %2, %5;\n\t"
        "setp.ne.u32 p0, t0, 0;\n\t"
        "setp.ne.u32 p1, t1, 0;\n\t"
        "@p0 st.shared.u16 [%3 + %6], lo;\n\t"
        "@p1 st.shared.u16 [%3 + %7], hi;\n\t"
        "}" ::"f"(y0), "f"(y1), "r"(m), "r"(xa), "n"(1u << J), "n"(2u << J), "n"(J * 16), "n"(J * 16 + 16)
        : "memory");
}
template <bool RELU>
__device__ __forceinline__ void epi_store16(uint32_t* v, float bias, uint32_t m, uint32_t xa) {
    epi_pair<0, RELU>(v[0], v[1], bias, m, xa);   epi_pair<2, RELU>(v[2], v[3], bias, m, xa);
    epi_pair<4, RELU>(v[4], v[5], bias, m, xa);   epi_pair<6, RELU>(v[6], v[7], bias, m, xa);
    epi_pair<8, RELU>(v[8], v[9], bias, m, xa);   epi_pair<10, RELU>(v[10], v[11], bias, m, xa);
    epi_pair<12, RELU>(v[12], v[13], bias, m, xa); epi_pair<14, RELU>(v[14], v[15], bias, m, xa);
}

// Per-action table of the dynamics head -> R2 (fp32), for the games of one epilogue warp.  All of a game's loads are issued
// before the first is used: the L2 latency is paid once per 64 columns, not once per 8-column chunk.  Kept out of line so that its
// 64 staging registers do not weigh on the register allocation of the per-layer epilogue.
#ifdef MZB_TABLE_INLINE
__device__ __forceinline__
#else
__device__ __noinline__
#endif
void stage_table(const float* table_ch, const int* g_act, uint32_t pt, uint32_t t_r2, int P, int G, int sel, int nsplit, long a_stride, int C, bool ch_ok) {
    for (int g = sel; g < G; g += nsplit) {
        const float* tg = table_ch + (long)g_act[g] * a_stride;
        for (int cb = 0; cb < P; cb += 64) {
            uint32_t tv[64];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                if (cb + 8 * k < P) {
                    int pix[8];
                    lds_v4(pt + (cb + 8 * k) * 4, pix); lds_v4(pt + (cb + 8 * k) * 4 + 16, pix + 4);
#pragma unroll
                    for (int j = 0; j < 8; ++j) tv[8 * k + j] = (ch_ok && pix[j] >= 0) ? __float_as_uint(__ldg(tg + (long)pix[j] * C)) : 0u;
                }
            }
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if (cb + 8 * k < P) tmem_st8(t_r2 + (uint32_t)(g * P + cb + 8 * k), tv + 8 * k);
        }
    }
    tmem_st_wait();
}

// Register budget: 384 threads x 168 registers = the SM's whole register file (the launch allocation; setmaxnreg then moves
// registers from the service warpgroup to the epilogue warpgroups).  A tree-step block therefore never fits beside a tower CTA
// (tools/coreside_probe.cu): it runs on the SMs a half-cohort launch leaves free.
// __maxnreg__ and __launch_bounds__ exclude each other.
#ifndef MZB_TOWER_MAXNREG
#define MZB_TOWER_MAXNREG 168
#endif
__global__ void __maxnreg__(MZB_TOWER_MAXNREG) tower_fused_kernel(const __grid_constant__ FusedArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    // warp index / tile size through a lane-0 shuffle: the compiler's uniformity analysis then knows they are warp-uniform,
    // so the role branches below are uniform branches and the MMA issuer's operands can live in uniform registers
    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
    const int game0 = __shfl_sync(0xffffffffu, a.tile_first[blockIdx.x], 0), G = __shfl_sync(0xffffffffu, a.tile_count[blockIdx.x], 0);
    if (a.tim && tid == 0) a.tim[((long)blockIdx.x * F_MAX_PHASES + F_MAX_PHASES - 1) * 4 + 0] = clock64();   // CTA start (last row is free: n_phases < F_MAX_PHASES)
    if (a.tile_trace && tid == 0) {
        long long g;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g));
        a.tile_trace[(long)blockIdx.x * 4 + 0] = g; a.tile_trace[(long)blockIdx.x * 4 + 2] = clock64();
    }
    const int P = a.P, R = a.R, HALO = a.HALO, C = a.C, KC = C >> 3, HW = a.HW, W1 = a.W + 1;
    const int N = (G * P + 15) & ~15;
    uint8_t* X = smem;
    uint8_t* stages = X + a.x_bytes;
    uint64_t* full = reinterpret_cast<uint64_t*>(stages + F_NSTAGE * F_STAGE_BYTES);
    uint64_t* empty = full + F_NSTAGE;
    uint64_t* accum_full = empty + F_NSTAGE;
    uint64_t* act_ready = accum_full + 1;
    uint64_t* acc_half = act_ready + 1;
    uint64_t* act_half = acc_half + 1;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(act_half + 1);
    // Staggered half tiles (4-game tiles): the epilogue of a 3x3 layer is strictly behind its MMAs and in front of the next layer's —
    // unless the tile is treated as two halves at the layer boundaries.  The last split_nt ring slots of a layer are issued twice, as
    // N/2-column MMAs over games [0, G/2) and then over [G/2, G), and the first split_nh slots of the next layer likewise: the first
    // half's epilogue runs under the second half's last MMAs, the second half's under the first half's first MMAs of the next layer.
    // A half's MMAs read rows of the other half only for output columns that are padding (masked in every epilogue).
    const int Nh = (G >> 1) * P;
    // (not for duplicated-row towers, C <= 64: there a warp owns ONE game, a half's epilogue is as long as the whole tile's, and the
    //  split only adds MMA cycles — measured 0.076 -> 0.078 ms per 592 states on c4r)
    const bool split = a.split_nt > 0 && !a.dup && (G & 3) == 0 && (Nh & 15) == 0 && Nh >= 112;
    int* pixtab = reinterpret_cast<int*>(reinterpret_cast<uint8_t*>(full) + 256);   // [P] row within a game -> pixel or -1
    const int PT = fused_ptab_entries(P);
    int* g_in = pixtab + PT;
    int* g_out = g_in + F_MAX_G;
    int* g_act = g_out + F_MAX_G;
    uint32_t* pixmask = reinterpret_cast<uint32_t*>(g_act + F_MAX_G);   // [8] bit r = row r of a game is a real pixel
    uint8_t* Z = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(pixmask + 8) + 127) & ~(uintptr_t)127);   // [K/8][8 games][16 B] bf16 head inputs
    float* Hs = reinterpret_cast<float*>(Z + a.z_bytes);

    if (tid == 0) {
        for (int s = 0; s < F_NSTAGE; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        mbar_init(accum_full, 1);
        mbar_init(act_ready, F_EPI_WARPS);
        mbar_init(acc_half, 1);
        mbar_init(act_half, F_EPI_WARPS);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_ptr, F_TMEM_COLS);
    if (tid < G) {
        const int game = game0 + tid;
        g_in[tid] = a.in_slot ? a.in_slot[game] : game;
        g_out[tid] = a.out_slot ? a.out_slot[game] : game;
        g_act[tid] = a.action ? a.action[game] : 0;
    }
    for (int r = tid; r < PT; r += F_THREADS) {
        const int yy = r / W1, xx = r - yy * W1;
        pixtab[r] = (r < P && yy >= 1 && yy <= a.H && xx < a.W) ? (yy - 1) * a.W + xx : -1;
    }
    if (tid < 256) {   // one row per thread, a ballot per warp (was: 8 threads x 32 rows with two divisions each, ~1.3 k cycles in front of the first barrier)
        const int r = tid, yy = r / W1, xx = r - yy * W1;
        const uint32_t m = __ballot_sync(0xffffffffu, r < P && yy >= 1 && yy <= a.H && xx < a.W);
        if (lane == 0) pixmask[tid >> 5] = m;
    }
    for (int i = tid; i < a.z_bytes / 16; i += F_THREADS) reinterpret_cast<uint4*>(Z)[i] = make_uint4(0u, 0u, 0u, 0u);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (a.tim && tid == 0) a.tim[((long)blockIdx.x * F_MAX_PHASES + F_MAX_PHASES - 1) * 4 + 2] = clock64();   // set-up done
    // ---- chain input -> X (zero padded), 16 bytes = 8 channels of one pixel per request ----
    {
        const int kc = tid % KC, r0 = tid / KC, rstep = F_THREADS / KC;
        const int rows = G * P;
        for (int r = r0; r < R; r += rstep) {
            uint8_t* dst = X + ((size_t)kc * R + r) * 16;
            const int q = r - HALO;
            int pix = -1, gi = 0;
            if (q >= 0 && q < rows) { gi = q / P; pix = pixtab[q - gi * P]; }
            if (pix >= 0)
                cp_async16(dst, a.in + (long)g_in[gi] * a.in_stride + ((long)pix * C + kc * 8));
            else
                *reinterpret_cast<uint4*>(dst) = make_uint4(0u, 0u, 0u, 0u);
        }
    }
    const uint32_t tmem = __shfl_sync(0xffffffffu, *tmem_ptr, 0);
    cp_async_wait_all();
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (a.tim && tid == 0) a.tim[((long)blockIdx.x * F_MAX_PHASES + F_MAX_PHASES - 1) * 4 + 3] = clock64();   // chain input (and table) resident

    // (each role's code is dominated by its own setmaxnreg: ptxas allocates registers per region only then)
    if (warp < F_SVC_THREADS / 32) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(F_SVC_REGS));
    if (warp == 0) {
        if (lane == 0) {   // ===== weight producer: every stage of every phase, in order, as far ahead as the ring allows =====
            uint32_t it = 0;
            for (int p = 0; p < a.n_phases; ++p) {
                const uint8_t* src = a.blob + a.ph[p].w_off;
                const int nkb = a.ph[p].nkb;
                for (int kb = 0; kb < nkb; kb += 2, ++it) {
                    const uint32_t s = it % F_NSTAGE;
                    uint32_t bytes = (uint32_t)min(2, nkb - kb) * F_KB_BYTES;
                    if (a.dbg_flags & 2) bytes = 256;   // timing experiment only (wrong results): take the weight stream out of the picture
                    if (it >= F_NSTAGE) mbar_wait_dbg(&empty[s], ((it / F_NSTAGE) - 1) & 1, a.dbg, 1, (int)it);
                    mbar_expect_tx(&full[s], bytes);
                    bulk_g2s(stages + s * F_STAGE_BYTES, src, bytes, &full[s]);
                    src += bytes;
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        // The WHOLE warp runs this loop and only the tcgen05 instructions are predicated on lane 0: every descriptor word is
        // then provably warp-uniform and lives in uniform registers.  With a single-lane branch around the loop the compiler
        // moved each operand vector->uniform per MMA (R2UR + ELECT loops): ~117 cycles per MMA whatever N (measured: 3-game
        // tiles, N = 176, floor 88, ran no faster than 4-game tiles).
        {
            const bool leader = elect_one();   // warp 1 is converged here and stays so: the same lane is elected every time
            const uint32_t idescN = idesc_bf16_f32(128u, (uint32_t)N), idesc16 = idesc_bf16_f32(128u, 16u);
            // descriptors: everything but the 14-bit start-address field is constant, so the loop only adds to the low words
            const uint64_t da0 = smem_desc_kmajor_noswizzle(smem_u32(stages), 128 * 16, 128);
            const uint64_t db0 = smem_desc_kmajor_noswizzle(smem_u32(X) + (uint32_t)HALO * 16u, (uint32_t)R * 16u, 128);
            const uint64_t dz0 = smem_desc_kmajor_noswizzle((smem_u32(pixmask + 8) + 127u) & ~127u, 256, 128);   // = Z, in 32-bit shared-space arithmetic (uniform datapath)
            const uint32_t a_hi = (uint32_t)(da0 >> 32), x_hi = (uint32_t)(db0 >> 32), z_hi = (uint32_t)(dz0 >> 32);
            const uint32_t a_lo0 = (uint32_t)da0, b_lo0 = (uint32_t)db0, z_lo0 = (uint32_t)dz0;
            const uint32_t idescH = idesc_bf16_f32(128u, (uint32_t)(split ? Nh : N));
            uint32_t it = 0, ar = 0, hr = 0;
            for (int p = 0; p < a.n_phases; ++p) {
                const bool head = split && (a.ph[p].flags & PH_SPLIT_HEAD), tail = split && (a.ph[p].flags & PH_SPLIT_TAIL);
                if (head) { mbar_wait_dbg(act_half, hr & 1, a.dbg, 5, p); ++hr; tc_fence_after(); }
                else if (p > 0 || a.preload) { mbar_wait_dbg(act_ready, ar & 1, a.dbg, 2, p); ++ar; tc_fence_after(); }
                const long long t_ph = a.tim ? clock64() : 0;
                long long w_full = 0;   // dbg_flags bit 0: the "issue" column reports the cycles spent waiting for weight slots
                if (a.tim && leader) a.tim[((long)blockIdx.x * F_MAX_PHASES + p) * 4 + 0] = t_ph;
                const int nkb = a.ph[p].nkb, kpt = a.ph[p].kpt;
                const uint32_t is_mlp = a.ph[p].is_mlp, onto = a.ph[p].onto;   // 0 / 1: arithmetic instead of selects keeps the chain uniform
                const uint32_t d_tmem = tmem + onto * (uint32_t)F_R2_COL;
                const uint32_t idesc = idescN + is_mlp * (idesc16 - idescN);      // MLP: B = Z, 16 rows per K chunk (head0's games, head1's games), N = 16
                const uint32_t b_hi = x_hi + is_mlp * (z_hi - x_hi);
                const uint32_t kk_b = a.ph[p].kk_b;    // two 8-element K chunks per MMA, in 16-byte units
                const uint32_t cc_b = a.ph[p].cc_b;    // eight K chunks per weight image
                uint32_t b_cur = b_lo0 + is_mlp * (z_lo0 - b_lo0) + (uint32_t)a.ph[p].tap_off0;   // B start of the current K stage
                int cc = 0, tx = 0, kb = 0;
                const int nslots = (nkb + 1) >> 1, nh = head ? a.split_nh : 0, nt = tail ? a.split_nt : 0;
                int s_cc = 0, s_tx = 0, s_kb = 0;
                uint32_t s_b = 0, s_it = 0;
                // segments: 0 / 1 = the first nh slots for the first / second half tile, 2 = full width, 3 / 4 = the last nt slots per half
                for (int seg = 0; seg < 5; ++seg) {
                const int nseg = seg < 2 ? nh : (seg == 2 ? nslots - nh - nt : nt);
                if (nseg == 0) continue;
                const bool first = (seg == 0) | (seg == 3), second = (seg == 1) | (seg == 4);
                if (first) { s_cc = cc; s_tx = tx; s_kb = kb; s_b = b_cur; s_it = it; }
                if (second) { cc = s_cc; tx = s_tx; kb = s_kb; b_cur = s_b; it = s_it; }
                if (seg == 1) { mbar_wait_dbg(act_ready, ar & 1, a.dbg, 2, p); ++ar; tc_fence_after(); }
                const uint32_t half_off = second ? (uint32_t)Nh : 0u;
                const uint32_t d_seg = d_tmem + half_off;
                const uint32_t idesc_seg = (seg == 2) ? idesc : idescH;
                for (int js = 0; js < nseg; ++js, ++it) {
                    const uint32_t s = it % F_NSTAGE;
                    const uint32_t a_lo = a_lo0 + s * (F_STAGE_BYTES >> 4);
                    if (a.tim && (a.dbg_flags & 1)) { const long long t = clock64(); mbar_wait_dbg(&full[s], (it / F_NSTAGE) & 1, a.dbg, 3, (int)it); w_full += clock64() - t; }
                    else mbar_wait_dbg(&full[s], (it / F_NSTAGE) & 1, a.dbg, 3, (int)it);
                    tc_fence_after();
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        if (kb < nkb) {
                            const uint32_t ah = a_lo + h * (F_KB_BYTES >> 4);
                            const uint32_t acc0 = (onto | (uint32_t)(kb != 0)) ? 1u : 0u;
                            // ptxas keeps the loop-carried b_cur in a vector register; an opaque lane-0 shuffle (which it knows to be
                            // warp-uniform) moves it to the uniform file ONCE per K stage instead of once per MMA
                            uint32_t ub;
                            asm volatile("shfl.sync.idx.b32 %0, %1, 0, 0x1f, 0xffffffff;" : "=r"(ub) : "r"(b_cur + half_off));
                            if (leader) {
                                mma_bf16_lohi_rt(d_seg, ah, a_hi, ub, b_hi, idesc_seg, acc0);
                                mma_bf16_lohi<1>(d_seg, ah + 256u, a_hi, ub + kk_b, b_hi, idesc_seg);
                                mma_bf16_lohi<1>(d_seg, ah + 512u, a_hi, ub + 2u * kk_b, b_hi, idesc_seg);
                                mma_bf16_lohi<1>(d_seg, ah + 768u, a_hi, ub + 3u * kk_b, b_hi, idesc_seg);
                            }
                            ++kb;
                            b_cur += cc_b;
                            if (++cc == kpt) {   // next tap: back to channel 0, one pixel to the right or to the start of the next row
                                cc = 0;
                                b_cur += ((tx == 2) ? (uint32_t)(W1 - 2) : 1u) - (uint32_t)kpt * cc_b;
                                tx = (tx == 2) ? 0 : tx + 1;
                            }
                        }
                    }
                    if (leader && !first) mma_commit(&empty[s]);
                    __syncwarp();
                }
                if (seg == 3 && leader) mma_commit(acc_half);
                }
                if (leader) {
                    if (nkb > 0) mma_commit(accum_full); else mbar_arrive(accum_full);
                    if (a.tim) a.tim[((long)blockIdx.x * F_MAX_PHASES + p) * 4 + 1] = (a.dbg_flags & 1) ? t_ph + w_full : clock64();
                }
                __syncwarp();
            }
        }
    }
    } else {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(F_EPI_REGS));
        // ===== epilogue warps: lane quarter q (hardware rule: warp id % 4), games of parity hsel =====
        // Every TMEM access below is a warp-collective: all control flow around it is warp-uniform, lanes are only
        // predicated on their stores.  The pixel map of a chunk is fetched BEFORE its stores (two 16-byte shared loads),
        // so the eight bf16 stores of a chunk issue back to back instead of alternating with dependent table look-ups.
        const int q = warp & 3, hsel = (warp - F_SVC_THREADS / 32) >> 2;   // hsel in [0, F_EPI_SPLIT)
        // `row` = accumulator lane = weight-image row.  For the 3x3 phases it is the output channel, except with duplicated rows
        // (a.dup, C <= 64): there quarters 2, 3 repeat the channels of quarters 0, 1 and take other games — all eight warps
        // work on a 64-channel layer instead of four (its epilogue halves).  `sel` / `nsplit` = which games this warp owns.
        const int row = 32 * q + lane;
        const int ch = a.dup ? 32 * (q & 1) + lane : row;
        const int sel = a.dup ? hsel + 2 * (q >> 1) : hsel, nsplit = a.dup ? 4 : F_EPI_SPLIT;
        const bool ch_ok = ch < C;
        const uint32_t t_lane = tmem + ((uint32_t)(32 * q) << 16);
        const uint32_t xs = smem_u32(X) + (uint32_t)(((ch >> 3) * R + HALO) * 16 + (ch & 7) * 2);   // + row * 16
        const uint32_t pt = smem_u32(pixtab), pm = smem_u32(pixmask);
        if (a.preload) {   // the chain starts with a residual block: its skip input = the chain input, fp32 into R2
            for (int g = sel; g < G; g += nsplit) {
                const int cbase = g * P;
                for (int c0 = 0; c0 < P; c0 += 8) {
                    int pix[8];
                    lds_v4(pt + c0 * 4, pix); lds_v4(pt + c0 * 4 + 16, pix + 4);
                    uint32_t v[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        uint32_t raw = 0;
                        if (pix[j] >= 0 && ch_ok) raw = lds_u16(xs + (uint32_t)(cbase + c0 + j) * 16);
                        v[j] = raw << 16;   // bf16 -> fp32 bit pattern
                    }
                    tmem_st8(t_lane + F_R2_COL + cbase + c0, v);
                }
            }
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(act_ready);
        }
        bool pool_pending = false;
        uint32_t qh = 0;
        for (int p = 0; p < a.n_phases; ++p) {
            const Phase& ph = a.ph[p];
            const int flags = ph.flags;
            // the per-channel bias is fetched BEFORE waiting for the accumulator: its L2 latency hides behind the MMAs
            const float bias_pf = (ph.bias_off >= 0) ? __ldg(reinterpret_cast<const float*>(a.blob + ph.bias_off) + ((flags & (PH_RED | PH_MLP)) ? row : ch)) : 0.f;
            if (flags & PH_TABLE)   // dynamics head: per-action table -> R2 (fp32) while the MMAs of this phase run into R1
                stage_table(reinterpret_cast<const float*>(a.blob + a.table_off) + ch, g_act, pt, t_lane + (uint32_t)F_R2_COL, P, G, sel, nsplit, (long)HW * C, C, ch_ok);
            const bool tail = split && (flags & PH_SPLIT_TAIL);
            const bool sig_half = split && p + 1 < a.n_phases && (a.ph[p + 1].flags & PH_SPLIT_HEAD);
            if (tail) { mbar_wait_dbg(acc_half, qh & 1, a.dbg, 6, p); ++qh; }
            else mbar_wait_dbg(accum_full, p & 1, a.dbg, 4, p);
            tc_fence_after();
            if (a.tim && tid == F_SVC_THREADS) a.tim[((long)blockIdx.x * F_MAX_PHASES + p) * 4 + 2] = clock64();
            if (flags & PH_EARLY) {   // nothing the next phase reads is produced here (its inputs were fenced by the previous arrive)
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(act_ready);
            }
            if (flags & PH_RED) {
                // 1x1 head convolution: lane = stacked reduced channel (head0's rows, then head1's), column = padded pixel row
                const int nred = ph.red_n0 + ph.red_n1;
                if (a.z_bytes && nred <= 32) {
                    // the stacked rows are repeated in every lane quarter (pack_conv1): quarter q takes the 16-column chunks
                    // k = q, q + 4, ... of its games, so all eight warps share the work instead of the two of quarter 0
                    const bool row_ok = lane < nred;
                    const bool second = lane >= ph.red_n0;
                    const int rr = second ? lane - ph.red_n0 : lane;
                    const float bias = row_ok ? bias_pf : 0.f;
                    const uint32_t zl = smem_u32(Z) + (second ? 128u : 0u) + (uint32_t)((rr >> 3) * 256 + (rr & 7) * 2);
                    const uint32_t cstride = (uint32_t)(a.z_rp >> 3) * 256u;
                    for (int g = hsel; g < G; g += F_EPI_SPLIT) {
                        const int cbase = g * P;
                        for (int c0 = 16 * q; c0 < P; c0 += 64) {
                            uint32_t v[16];
                            tmem_ld16(t_lane + cbase + c0, v);
                            tmem_ld_wait();
                            const uint32_t za = zl + (uint32_t)g * 16u + (uint32_t)c0 * cstride;
#pragma unroll
                            for (int j = 0; j < 16; ++j)
                                if (row_ok && c0 + j < P)
                                    sts_u16(za + (uint32_t)j * cstride, __bfloat16_as_ushort(__float2bfloat16(__uint_as_float(v[j]) + bias)));
                        }
                    }
                } else if (32 * q < nred) {   // warp-uniform; lanes beyond nred only skip the stores
                    const bool row_ok = row < nred;
                    const bool second = row >= ph.red_n0;
                    const int rr = second ? row - ph.red_n0 : row;
                    const float bias = row_ok ? bias_pf : 0.f;
                    if (a.z_bytes) {
                        // in-kernel heads.  Z is the B operand of the first MLP layer: [K / 8][16 rows][8] bf16, rows 0-7 = head0's
                        // games, rows 8-15 = head1's; K index = column * z_rp + channel (pad columns are stored too: their
                        // weights are zero), so a lane's stores are an affine function of the column — no pixel look-up
                        const uint32_t zl = smem_u32(Z) + (second ? 128u : 0u) + (uint32_t)((rr >> 3) * 256 + (rr & 7) * 2);
                        const uint32_t cstride = (uint32_t)(a.z_rp >> 3) * 256u;
                        for (int g = hsel; g < G; g += F_EPI_SPLIT) {
                            const int cbase = g * P;
                            for (int c0 = 0; c0 < P; c0 += 16) {
                                uint32_t v[16];
                                tmem_ld16(t_lane + cbase + c0, v);
                                tmem_ld_wait();
                                const uint32_t za = zl + (uint32_t)g * 16u + (uint32_t)c0 * cstride;
#pragma unroll
                                for (int j = 0; j < 16; ++j)
                                    if (row_ok && c0 + j < P)
                                        sts_u16(za + (uint32_t)j * cstride, __bfloat16_as_ushort(__float2bfloat16(__uint_as_float(v[j]) + bias)));
                            }
                        }
                    } else {
                        const int head = second ? ph.red_head1 : ph.red_head0;
                        // [game][pixel][reduced channel]: the lanes of a warp (consecutive channels) store one contiguous run per
                        // pixel; the first MLP layer's rows are permuted to this order on the host (w0t, net_bf16.cu)
                        const int rows = a.red_rows[head];
                        float* dst0 = a.red[head] + rr;
                        const long gstride = (long)rows * HW;
                        for (int g = hsel; g < G; g += F_EPI_SPLIT) {
                            float* dst = dst0 + (long)(game0 + g) * gstride;
                            const int cbase = g * P;
                            // 16 columns per pass, the pixel map fetched while the TMEM load is in flight; 32-bit offsets (the 8-column
                            // version was a chain of TMEM round trips with 64-bit address arithmetic per store: 8.3 k cycles per phase on
                            // the ALE shape)
                            const int rows4 = rows * 4;
                            char* dstb = reinterpret_cast<char*>(dst);
                            for (int c0 = 0; c0 < P; c0 += 16) {
                                uint32_t v[16];
                                tmem_ld16(t_lane + cbase + c0, v);
                                int pix[16];
                                lds_v4(pt + c0 * 4, pix); lds_v4(pt + c0 * 4 + 16, pix + 4); lds_v4(pt + c0 * 4 + 32, pix + 8); lds_v4(pt + c0 * 4 + 48, pix + 12);
                                tmem_ld_wait();
#pragma unroll
                                for (int j = 0; j < 16; ++j)
                                    if (pix[j] >= 0 && row_ok) *reinterpret_cast<float*>(dstb + pix[j] * rows4) = __uint_as_float(v[j]) + bias;
                            }
                        }
                    }
                }
            }
            if (flags & PH_NORM) {
                // normalize_hidden_state (models.cpp:16-26): a thread owns one channel, its columns are the pixels
                const int nskip = (W1 >> 3) << 3;   // the pad row of a game: no real pixel, nothing to normalise (its R2 columns are never read)
                for (int g = sel; g < G; g += nsplit) {
                    const int cbase = g * P;
                    float mn = 3.402823466e38f, mx = -3.402823466e38f;
                    if (a.normalize) {
                        for (int c0 = nskip; c0 < P; c0 += 32) {
                            const bool two = P - c0 > 16;
                            uint32_t v[32];
                            tmem_ld16(t_lane + F_R2_COL + cbase + c0, v);
                            if (two) tmem_ld16(t_lane + F_R2_COL + cbase + c0 + 16, v + 16);
                            const int mw_i = c0 >> 5, mw_s = c0 & 31;
                            uint32_t m = lds_u32(pm + mw_i * 4) >> mw_s;
                            if (mw_s && mw_i + 1 < 8) m |= lds_u32(pm + (mw_i + 1) * 4) << (32 - mw_s);
                            tmem_ld_wait();
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if ((j < 16 || two) && ((m >> j) & 1u)) { const float x = __uint_as_float(v[j]); mn = fminf(mn, x); mx = fmaxf(mx, x); }
                        }
                    }
                    float scale = mx - mn;
                    if (scale < 1e-5f) scale = 1e-5f;
                    if (!a.normalize) { mn = 0.f; scale = 1.f; }
                    const float inv = 1.0f / scale;
                    for (int c0 = nskip; c0 < P; c0 += 32) {
                        const bool two = P - c0 > 16;
                        uint32_t v[32];
                        tmem_ld16(t_lane + F_R2_COL + cbase + c0, v);
                        if (two) tmem_ld16(t_lane + F_R2_COL + cbase + c0 + 16, v + 16);
                        const int mw_i = c0 >> 5, mw_s = c0 & 31;
                        uint32_t m = lds_u32(pm + mw_i * 4) >> mw_s;
                        if (mw_s && mw_i + 1 < 8) m |= lds_u32(pm + (mw_i + 1) * 4) << (32 - mw_s);
                        tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            if (j < 16 || two) {
                                const float y = (__uint_as_float(v[j]) - mn) * inv;   // reciprocal multiply: IEEE division costs 10x here; 1 ulp, far inside the bf16 tolerance
                                v[j] = __float_as_uint(y);
                                if (ch_ok && ((m >> j) & 1u)) sts_u16(xs + (uint32_t)(cbase + c0 + j) * 16, __bfloat16_as_ushort(__float2bfloat16(y)));
                            }
                        }
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            if (c0 + 8 * k < P) tmem_st8(t_lane + F_R2_COL + cbase + c0 + 8 * k, v + 8 * k);
                    }
                }
                tmem_st_wait();
            }
            if (!(flags & (PH_RED | PH_NORM | PH_MLP))) {
                const uint32_t src_col = (flags & PH_ACC_R2) ? (uint32_t)F_R2_COL : 0u;
                const float bias = ch_ok ? bias_pf : 0.f;
                const bool relu = (flags & PH_RELU) != 0, stash = (flags & PH_STASH) != 0, wx = (flags & PH_WRITE_X) != 0 && ch_ok && !(a.dbg_flags & 8);   // bit 3: timing experiment, no stores into X (wrong results)
                if (flags & PH_TABLE) {   // accumulator (R1) + staged per-action table (R2), 16 columns at a time
                    for (int g = sel; g < G; g += nsplit) {
                        const int cbase = g * P;
                        for (int c0 = (W1 >> 3) << 3; c0 < P; c0 += 16) {   // (from the first real row on, as below)
                            uint32_t v[16], w[16];
                            tmem_ld16(t_lane + src_col + cbase + c0, v);
                            tmem_ld16(t_lane + F_R2_COL + cbase + c0, w);
                            const int mw_i = c0 >> 5, mw_s = c0 & 31;
                            uint32_t m = lds_u32(pm + mw_i * 4) >> mw_s;
                            if (mw_s > 16 && mw_i + 1 < 8) m |= lds_u32(pm + (mw_i + 1) * 4) << (32 - mw_s);
                            tmem_ld_wait();
#pragma unroll
                            for (int j = 0; j < 16; ++j) v[j] = __float_as_uint(__uint_as_float(v[j]) + __uint_as_float(w[j]));
                            const uint32_t xa = xs + (uint32_t)(cbase + c0) * 16;
                            if (relu) epi_store16<true>(v, bias, wx ? m : 0u, xa); else epi_store16<false>(v, bias, wx ? m : 0u, xa);
                            if (stash) {
                                tmem_st8(t_lane + F_R2_COL + cbase + c0, v);
                                if (c0 + 8 < P) tmem_st8(t_lane + F_R2_COL + cbase + c0 + 8, v + 8);
                            }
                        }
                    }
                } else {
                    // a game's first W + 1 rows are its zero pad row: nothing to read or store there (whole 8-column groups only,
                    // to keep the TMEM column addresses aligned) — C4: 48 instead of 64 columns per game through the TMEM read port
                    const int skip = (W1 >> 3) << 3;
                    // Games go to the two warp halves alternately; the odd game out of a 1- or 3-game tile is cut at a column
                    // boundary and shared, so that a 3-game tile's epilogue costs 1.5 games per warp, not 2.
                    static_assert(F_EPI_SPLIT == 2, "the column split below assumes two warp halves");
                    const int G2 = G & ~1;
                    // ... and nothing behind a game's last real pixel either (6x6 board: 7 x 7 = 49 rows padded to P = 56, last pixel at
                    // row 47 -> 48 of 56 columns)
                    const int c_end = min(P, (a.H * W1 + a.W + 7) & ~7);
                    const int mid = skip + ((((c_end - skip) >> 1) + 7) & ~7);
                    const int n_mine = a.dup ? (G > sel ? (G - sel + 3) >> 2 : 0) : (G2 > hsel ? (G2 - hsel + 1) >> 1 : 0) + (G & 1);
                    // staggered half tiles: a warp takes its games in ascending order; before its first game of the second half
                    // (or after its last game, if it has none there) it hands the first half to the next layer's MMAs, then waits
                    // for the second half's accumulator
                    bool crossed = !(tail || sig_half);
                    const int g_half = G >> 1;
                    for (int i = 0; i <= n_mine; ++i) {
                        int g = a.dup ? sel + 4 * i : hsel + 2 * i, c_lo = skip, c_hi = c_end;
                        if (!crossed && (i == n_mine || g >= g_half)) {
                            crossed = true;
                            if (a.tim && tid == F_SVC_THREADS && (a.dbg_flags & 4)) a.tim[((long)blockIdx.x * F_MAX_PHASES + p) * 4 + 3] = clock64();   // diagnostic: first half done
                            if (sig_half) {
                                fence_proxy_async_smem();
                                tc_fence_before();
                                __syncwarp();
                                if (lane == 0) mbar_arrive(act_half);
                            }
                            // (also a warp with no game in the second half: every warp observes every phase of accum_full, or a later
                            // parity wait could be satisfied by the phase before)
                            if (tail) { mbar_wait_dbg(accum_full, p & 1, a.dbg, 4, p); tc_fence_after(); }
                        }
                        if (i == n_mine) break;
                        if (!a.dup && g >= G2) { g = G - 1; if (hsel == 0) c_hi = mid; else c_lo = mid; }
                        const int cbase = g * P;
                        for (int c0 = c_lo; c0 < c_hi; c0 += 32) {
                            const int left = c_hi - c0;
                            const bool two = left > 16;
                            uint32_t v[32];
                            tmem_ld16(t_lane + src_col + cbase + c0, v);
                            if (two) tmem_ld16(t_lane + src_col + cbase + c0 + 16, v + 16);
                            const int mw_i = c0 >> 5, mw_s = c0 & 31;
                            uint32_t m = lds_u32(pm + mw_i * 4) >> mw_s;
                            if (mw_s && mw_i + 1 < 8) m |= lds_u32(pm + (mw_i + 1) * 4) << (32 - mw_s);
                            if (left < 32) m &= (1u << left) - 1u;
                            tmem_ld_wait();
                            const uint32_t mw = wx ? m : 0u;
                            const uint32_t xa = xs + (uint32_t)(cbase + c0) * 16;
                            if (relu) { epi_store16<true>(v, bias, mw, xa); if (two) epi_store16<true>(v + 16, bias, mw >> 16, xa + 256); }
                            else { epi_store16<false>(v, bias, mw, xa); if (two) epi_store16<false>(v + 16, bias, mw >> 16, xa + 256); }
                            if (stash) {
#pragma unroll
                                for (int k = 0; k < 4; ++k)
                                    if (c0 + 8 * k < c_hi) tmem_st8(t_lane + F_R2_COL + cbase + c0 + 8 * k, v + 8 * k);
                            }
                        }
                    }
                }
                if (stash) tmem_st_wait();
            }
            if (flags & PH_MLP) {
                // first MLP layer came off the tensor core: lane = stacked hidden unit, column = game.  bias + ReLU -> Hs
                if (hsel == 0) {
                    uint32_t v[16];   // columns 0-7: head0's games, 8-15: head1's games (Z rows 8-15)
                    tmem_ld16(t_lane, v);
                    tmem_ld_wait();
                    if (row < ph.red_n0 + ph.red_n1) {
                        const bool second = row >= ph.red_n0;
                        const HeadTail& hd = a.head[second ? ph.red_head1 : ph.red_head0];
                        const float b0 = bias_pf;
                        const bool relu0 = hd.n_layers > 1;
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            float y = __uint_as_float(second ? v[8 + j] : v[j]) + b0;
                            if (relu0) y = fmaxf(y, 0.f);
                            Hs[j * 128 + row] = y;
                        }
                    }
                }
            }
            fence_proxy_async_smem();   // generic-proxy stores into X / Z -> visible to the tensor core's operand reads
            tc_fence_before();
            __syncwarp();
            if (lane == 0 && !(flags & PH_EARLY)) mbar_arrive(act_ready);
            if (flags & PH_MLP) {
                // remaining layers (model_layers.cpp:46-65) and the output transform on CUDA cores, overlapped with the next
                // phase's MMAs: policy softmax (vprnet.cpp:118-119,178-179) / ValueEncoder::decode (util.cpp:80-87)
                asm volatile("bar.sync 1, %0;" ::"n"(F_EPI_THREADS) : "memory");
                if (a.tim && tid == F_SVC_THREADS) a.tim[((long)blockIdx.x * F_MAX_PHASES + F_MAX_PHASES - 2) * 4 + 0] = clock64();
                const int et = tid - F_SVC_THREADS;
                const int nh = ph.red_n1 > 0 ? 2 : 1;
                float* cur = Hs;
                float* nxt = Hs + 8 * 128;
                const int n_layers = a.head[ph.red_head0].n_layers;
                for (int l = 1; l < n_layers; ++l) {
                    // both heads in ONE pass (their weight rows come from L2: one exposed latency per layer, not one per head);
                    // 4 threads per output when that fits the 256 epilogue threads, else 2
                    const HeadTail& h0 = a.head[ph.red_head0];
                    const HeadTail& h1 = a.head[ph.red_head1];
                    const int nin0 = h0.sizes[l], nout0 = h0.sizes[l + 1], nout1 = nh > 1 ? h1.sizes[l + 1] : 0;
                    const int tot = nout0 + nout1;
                    const int tpo = (G * tot * 4 <= F_EPI_THREADS) ? 4 : 2;   // warp-uniform
                    for (int base = 0; base < G * tot; base += F_EPI_THREADS / tpo) {   // warp-uniform trip count
                        const int i = base + et / tpo, part = et % tpo;
                        const bool on = i < G * tot;
                        const int g = on ? i / tot : 0, o = on ? i - g * tot : 0;
                        const bool second = o >= nout0;
                        const HeadTail& hd = second ? h1 : h0;
                        const int oo = second ? o - nout0 : o;
                        const int nin = hd.sizes[l];
                        const float* wr = hd.W[l] + (long)oo * nin;
                        const float* xr = cur + g * 128 + (second ? nin0 : 0);
                        const float bl = (on && part == 0) ? __ldg(hd.b[l] + oo) : 0.f;   // in flight together with the weights
                        float acc = 0.f;
                        if (on) {
        #pragma unroll 16
                            for (int k = part; k < nin; k += tpo) acc += __ldg(wr + k) * xr[k];
                        }
                        acc += __shfl_xor_sync(0xffffffffu, acc, 1);
                        if (tpo == 4) acc += __shfl_xor_sync(0xffffffffu, acc, 2);
                        if (on && part == 0) {
                            acc += bl;
                            if (l < n_layers - 1) acc = fmaxf(acc, 0.f);
                            nxt[g * 128 + (second ? nout0 : 0) + oo] = acc;
                        }
                    }
                    asm volatile("bar.sync 1, %0;" ::"n"(F_EPI_THREADS) : "memory");
                    float* t = cur; cur = nxt; nxt = t;
                }
                if (a.tim && tid == F_SVC_THREADS) a.tim[((long)blockIdx.x * F_MAX_PHASES + F_MAX_PHASES - 2) * 4 + 1] = clock64();
                // output transform, one warp per (game, head), lane = logit: softmax by warp reductions (vprnet.cpp:118-119,
                // 178-179) / expected support value + inverse contractive map (util.cpp:80-87).  Heads wider than a warp take the
                // serial form on one lane.
                for (int item = (tid - F_SVC_THREADS) >> 5; item < G * nh; item += F_EPI_WARPS) {   // warp-uniform
                    const int g = item / nh, hh = item - g * nh;
                    const HeadTail& hd = a.head[hh ? ph.red_head1 : ph.red_head0];
                    const int n = hd.sizes[n_layers];
                    const float* lg = cur + g * 128 + (hh ? a.head[ph.red_head0].sizes[n_layers] : 0);
                    const long b = game0 + g;
                    if (n <= 32) {
                        const float x = lane < n ? lg[lane] : -3.402823466e38f;
                        float mx = x;
        #pragma unroll
                        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
                        const float e = lane < n ? expf(x - mx) : 0.f;
                        float sum = e;
        #pragma unroll
                        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
                        const float pr = e / sum;
                        if (hd.out_kind == 0) {
                            if (lane < n) hd.out[b * hd.out_stride + lane] = (double)pr;
                        } else {
                            // sum_i i * p_i in index order, as the reference's torch::sum over the support does on one row
                            float ev = (float)lane * pr;   // lanes >= n hold 0
        #pragma unroll
                            for (int o = 16; o > 0; o >>= 1) ev += __shfl_xor_sync(0xffffffffu, ev, o);
                            if (lane == 0) {
                                float v = ev * hd.step + hd.min_value;
                                if (hd.contractive) {
                                    const float eps = 0.001f;
                                    const float sg = (v > 0.f) ? 1.f : ((v < 0.f) ? -1.f : 0.f);
                                    const float inner = sqrtf(1.f + (4.f * eps) * (fabsf(v) + 1.f + eps)) - 1.f;
                                    const float qq = inner / (2.f * eps);
                                    v = sg * (qq * qq - 1.f);
                                }
                                hd.out[b * hd.out_stride] = (double)v;
                            }
                        }
                    } else if (lane == 0) {
                        float mx = lg[0];
                        for (int i = 1; i < n; ++i) mx = fmaxf(mx, lg[i]);
                        float sum = 0.f;
                        for (int i = 0; i < n; ++i) sum += expf(lg[i] - mx);
                        if (hd.out_kind == 0) {
                            for (int i = 0; i < n; ++i) hd.out[b * hd.out_stride + i] = (double)(expf(lg[i] - mx) / sum);
                        } else {
                            float ev = 0.f;
                            for (int i = 0; i < n; ++i) ev += (float)i * (expf(lg[i] - mx) / sum);
                            float v = ev * hd.step + hd.min_value;
                            if (hd.contractive) {
                                const float eps = 0.001f;
                                const float sg = (v > 0.f) ? 1.f : ((v < 0.f) ? -1.f : 0.f);
                                const float inner = sqrtf(1.f + (4.f * eps) * (fabsf(v) + 1.f + eps)) - 1.f;
                                const float qq = inner / (2.f * eps);
                                v = sg * (qq * qq - 1.f);
                            }
                            hd.out[b * hd.out_stride] = (double)v;
                        }
                    }
                }
                if (a.tim && tid == F_SVC_THREADS) a.tim[((long)blockIdx.x * F_MAX_PHASES + F_MAX_PHASES - 2) * 4 + 2] = clock64();
                asm volatile("bar.sync 1, %0;" ::"n"(F_EPI_THREADS) : "memory");   // Hs is free again
            }
            if (flags & PH_EARLY) pool_pending = true;   // the copy below would sit on the critical path: do it after the next arrive
            if (((flags & PH_NORM) && !(flags & PH_EARLY)) || (pool_pending && !(flags & PH_EARLY))) {
                pool_pending = false;
                // child hidden state -> node pool, straight from X (bf16, already normalised): 16-byte coalesced stores,
                // overlapped with the next phase's MMAs (which only read X)
                asm volatile("bar.sync 1, %0;" ::"n"(F_EPI_THREADS) : "memory");   // every epilogue warp has finished writing X
                const int et = tid - F_SVC_THREADS;
                const int kc = et % KC, rsub = et / KC, rstep = F_EPI_THREADS / KC;   // a thread keeps its 8-channel chunk; rows stride
                const uint32_t xk = smem_u32(X) + (uint32_t)((kc * R + HALO) * 16);
                for (int g = 0; g < G; ++g) {
                    bf16* pg = a.pool + (long)g_out[g] * a.pool_stride + kc * 8;
                    for (int r = rsub; r < P; r += rstep) {
                        const int pix = pixtab[r];
                        if (pix >= 0) *reinterpret_cast<uint4*>(pg + (long)pix * C) = lds_v4u(xk + (uint32_t)(g * P + r) * 16);
                    }
                }
                asm volatile("bar.sync 1, %0;" ::"n"(F_EPI_THREADS) : "memory");   // nobody overwrites X (next epilogue) before the copy is done
            }
            if (a.tim && tid == F_SVC_THREADS && !((a.dbg_flags & 4) && (tail || sig_half))) a.tim[((long)blockIdx.x * F_MAX_PHASES + p) * 4 + 3] = clock64();
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem, F_TMEM_COLS);
    if (a.tim && tid == 0) a.tim[((long)blockIdx.x * F_MAX_PHASES + F_MAX_PHASES - 1) * 4 + 1] = clock64();   // CTA end
    if (a.tile_trace && tid == 0) {
        long long g;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g));
        a.tile_trace[(long)blockIdx.x * 4 + 1] = g; a.tile_trace[(long)blockIdx.x * 4 + 3] = clock64();
    }
}

// ------------------------------------------------------------------ head MLPs on the reduced activations
// red[head] : [B][n0] fp32 in [pixel][reduced channel] order (the reference flattens r*HW + p, models.cpp:114,161-162; w0t's rows
// are permuted to match)  ->  MLP (model_layers.cpp:46-65)  ->
// policy softmax (vprnet.cpp:118-119,178-179) or support-to-scalar (ValueEncoder::decode, util.cpp:80-87).
constexpr int MLP_THREADS = 256;
constexpr int MLP_MAX_LAYERS = 6;
struct MlpHead {
    const float* x;            // [B][n0]
    const float* w0t;          // first layer transposed: [n0][n1]
    const float* W[MLP_MAX_LAYERS];   // later layers [out][in] (W[0] unused)
    const float* b[MLP_MAX_LAYERS];
    int sizes[MLP_MAX_LAYERS + 1];
    int n_layers;
    int out_kind;              // 0 = softmax policy, 1 = support -> scalar
    float step, min_value; int contractive;
    double* out; int out_stride;
};
struct MlpArgs {
    MlpHead h[3]; int n_heads, B, mg;
    // split-K of the first layer (gridDim.z = splits > 1): partial sums [head][split][B][n1] in global memory, one arrival
    // counter per (head, game group); the last CTA to arrive adds the partials in split order (deterministic) and carries on
    int splits; float* part; long part_head_stride; int* counter;
};

constexpr int MLP_KT = 32;    // K tile of the first layer
constexpr int MLP_WT = 64 * 65;   // floats of the transposed tail-layer weight buffer

// One CTA = mg (<= 16) games of one head.  First layer (the only sizeable one: n0 = red*HW inputs) as a shared-memory
// tiled SGEMM: x tile [mg][KT] and transposed-weight tile [KT][n1] staged per K step, thread = (output j, 4 games).
__global__ void __launch_bounds__(MLP_THREADS) mlp_heads_kernel(const __grid_constant__ MlpArgs a) {
    extern __shared__ float msm[];
    const MlpHead& hd = a.h[blockIdx.y];
    const int tid = threadIdx.x;
    const int b0 = blockIdx.x * a.mg;
    const int nb = min(a.mg, a.B - b0);
    const int n0 = hd.sizes[0], n1 = hd.sizes[1];
    const int S = a.splits > 1 ? a.splits : 1, split = blockIdx.z;
    const int n_kt = (n0 + MLP_KT - 1) / MLP_KT;
    const int k_lo = (int)((long)n_kt * split / S) * MLP_KT, k_hi = min(n0, (int)((long)n_kt * (split + 1) / S) * MLP_KT);
    float* part = a.part + blockIdx.y * a.part_head_stride;
    float* bufA = msm;                    // [mg][256]
    float* bufB = bufA + a.mg * 256;      // [mg][256]
    float* xs = bufB + a.mg * 256;        // [16][KT + 4]
    float* ws = xs + 16 * (MLP_KT + 4);   // [KT][n1]
    float* wt = ws + MLP_KT * 64;         // [nin][nout + 1]: one tail layer's weights, transposed (layers up to 64 x 64)
    {
        const int groups = MLP_THREADS / 64;                 // 4 groups of 64 threads: group <-> 4 games
        const int tx = tid & 63, ty = tid >> 6;
        for (int j0 = 0; j0 < n1; j0 += 64) {
            const int j = j0 + tx;
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
            // the next K tile's global loads are issued before this tile is multiplied: their latency (the kernel's main
            // stall, ncu source page) hides behind the FMAs instead of sitting in front of the shared-memory stores
            constexpr int XR = 16 * MLP_KT / MLP_THREADS, WR = MLP_KT * 64 / MLP_THREADS;
            float xq[XR], wq[WR];
            auto gload = [&](int k0) {
                const int kt = min(MLP_KT, n0 - k0);
#pragma unroll
                for (int r = 0; r < XR; ++r) {
                    const int i = tid + r * MLP_THREADS, g = i / MLP_KT, k = i - g * MLP_KT;
                    xq[r] = (g < nb && k < kt) ? __ldg(hd.x + (long)(b0 + g) * n0 + k0 + k) : 0.f;
                }
#pragma unroll
                for (int r = 0; r < WR; ++r) {
                    const int i = tid + r * MLP_THREADS, k = i >> 6, jj = i & 63;
                    wq[r] = (k < kt && j0 + jj < n1) ? __ldg(hd.w0t + (long)(k0 + k) * n1 + j0 + jj) : 0.f;
                }
            };
            gload(k_lo);
            for (int k0 = k_lo; k0 < k_hi; k0 += MLP_KT) {
                __syncthreads();
#pragma unroll
                for (int r = 0; r < XR; ++r) {
                    const int i = tid + r * MLP_THREADS, g = i / MLP_KT, k = i - g * MLP_KT;
                    xs[g * (MLP_KT + 4) + k] = xq[r];
                }
#pragma unroll
                for (int r = 0; r < WR; ++r) ws[tid + r * MLP_THREADS] = wq[r];
                __syncthreads();
                if (k0 + MLP_KT < k_hi) gload(k0 + MLP_KT);
                const float* xr = xs + (ty * 4) * (MLP_KT + 4);
#pragma unroll
                for (int k = 0; k < MLP_KT; k += 4) {
                    const float w0 = ws[(k + 0) * 64 + tx], w1 = ws[(k + 1) * 64 + tx], w2 = ws[(k + 2) * 64 + tx], w3 = ws[(k + 3) * 64 + tx];
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        const float4 x = *reinterpret_cast<const float4*>(xr + g * (MLP_KT + 4) + k);
                        acc[g] += w0 * x.x + w1 * x.y + w2 * x.z + w3 * x.w;
                    }
                }
            }
            if (j < n1) {
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    const int gg = ty * 4 + g;
                    if (gg < nb) {
                        if (S > 1) { part[((long)split * a.B + b0 + gg) * n1 + j] = acc[g]; continue; }
                        float v = acc[g] + hd.b[0][j];
                        if (hd.n_layers > 1) v = fmaxf(v, 0.f);
                        bufA[gg * 256 + j] = v;
                    }
                }
            }
        }
        (void)groups;
    }
    if (S > 1) {
        __shared__ int s_last;
        __threadfence();
        __syncthreads();
        int* cnt = a.counter + blockIdx.y * gridDim.x + blockIdx.x;
        if (tid == 0) {
            const int old = atomicAdd(cnt, 1);
            s_last = (old == S - 1);
            if (s_last) *cnt = 0;   // ready for the next launch on this stream
        }
        __syncthreads();
        if (!s_last) return;
        __threadfence();
        for (int i = tid; i < nb * n1; i += MLP_THREADS) {
            const int g = i / n1, j = i - g * n1;
            float v = 0.f;
            for (int sp = 0; sp < S; ++sp) v += __ldcg(part + ((long)sp * a.B + b0 + g) * n1 + j);   // split order: run-to-run identical
            v += hd.b[0][j];
            if (hd.n_layers > 1) v = fmaxf(v, 0.f);
            bufA[g * 256 + j] = v;
        }
    }
    __syncthreads();
    float* cur = bufA; float* nxt = bufB;
    for (int l = 1; l < hd.n_layers; ++l) {
        const int nin = hd.sizes[l], nout = hd.sizes[l + 1];
        if (nin <= 64 && nout <= 64 && a.mg <= 16) {
            // the layer's weights go through shared memory once, transposed and padded (conflict-free both ways); thread =
            // (output j, 4 games): one weight load feeds four FMAs, the activations are warp-wide broadcasts
            const int ld = nout + 1;
            for (int i = tid; i < nin * nout; i += MLP_THREADS) {
                const int j = i / nin, k = i - j * nin;
                wt[k * ld + j] = __ldg(hd.W[l] + i);
            }
            __syncthreads();
            const int j = tid & 63, g0 = (tid >> 6) * 4;
            if (j < nout) {
                float acc[4] = {0.f, 0.f, 0.f, 0.f};
                const float* xr = cur + g0 * 256;
                for (int k = 0; k < nin; ++k) {
                    const float w = wt[k * ld + j];
#pragma unroll
                    for (int q = 0; q < 4; ++q) acc[q] += w * xr[q * 256 + k];
                }
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (g0 + q < nb) {
                        float v = acc[q] + hd.b[l][j];
                        if (l < hd.n_layers - 1) v = fmaxf(v, 0.f);
                        nxt[(g0 + q) * 256 + j] = v;
                    }
            }
            __syncthreads();
            float* t = cur; cur = nxt; nxt = t;
            continue;
        }
        // wider layers: warp per output, the lanes walk the weight row together (one coalesced request per load)
        const int warp = tid >> 5, lane = tid & 31;
        for (int o = warp; o < nb * nout; o += MLP_THREADS / 32) {
            const int g = o / nout, j = o - g * nout;
            const float* wr = hd.W[l] + (long)j * nin;
            const float* xr = cur + g * 256;
            float acc = 0.f;
            for (int i = lane; i < nin; i += 32) acc += __ldg(wr + i) * xr[i];
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
            if (lane == 0) {
                acc += hd.b[l][j];
                if (l < hd.n_layers - 1) acc = fmaxf(acc, 0.f);
                nxt[g * 256 + j] = acc;
            }
        }
        __syncthreads();
        float* t = cur; cur = nxt; nxt = t;
    }
    if (tid < nb) {
        const int b = b0 + tid;
        const float* lg = cur + tid * 256;
        float* ex = nxt + tid * 256;
        const int n = hd.sizes[hd.n_layers];
        float mx = lg[0];
        for (int i = 1; i < n; ++i) mx = fmaxf(mx, lg[i]);
        float sum = 0.f;
        for (int i = 0; i < n; ++i) { const float e = expf(lg[i] - mx); ex[i] = e; sum += e; }
        if (hd.out_kind == 0) {
            for (int i = 0; i < n; ++i) hd.out[(long)b * hd.out_stride + i] = (double)(ex[i] / sum);
        } else {
            float ev = 0.f;
            for (int i = 0; i < n; ++i) ev += (float)i * (ex[i] / sum);
            float v = ev * hd.step + hd.min_value;
            if (hd.contractive) {
                const float eps = 0.001f;
                const float s = (v > 0.f) ? 1.f : ((v < 0.f) ? -1.f : 0.f);
                const float inner = sqrtf(1.f + (4.f * eps) * (fabsf(v) + 1.f + eps)) - 1.f;
                const float qq = inner / (2.f * eps);
                v = s * (qq * qq - 1.f);
            }
            hd.out[(long)b * hd.out_stride] = (double)v;
        }
    }
}

}  // namespace fused
}  // namespace mzb
