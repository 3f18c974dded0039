// learn_step.cu — the whole training step of VPRNetModel::Learn (muzero-cpp/vprnet.cpp:192-311): K + 1 unrolled
// train-mode network calls, MuzeroNetworkImpl::loss (models.cpp:268-276), the backward pass with Learn's gradient hooks (0.5 on every
// recurrent hidden state, gradient_scale on the losses of steps >= 1, the [B,1] x [B] importance-weight broadcast) and
// torch::optim::Adam with L2 weight decay (vprnet.cpp:35-36,301-303) — fp32, on the GPU, on the module's own tensors.
//
// Layout: activations live as [C][N][H][W] ("CNHW"): a channel is one contiguous run of N*H*W floats, so
//   * a 3x3 convolution is ONE GEMM  Y[Cout][N*HW] = W[Cout][Cin*9] x col[Cin*9][N*HW]  (W is the module's tensor as it lies),
//     its weight gradient  dW += dY x col^T  and its data gradient  dcol = W^T x dY  (folded back by a gather, no atomics);
//   * BatchNorm statistics are reductions over contiguous memory, one block per channel;
//   * torch::cat({state, action planes}, 1) is an append.
// The GEMM is a hand-written 64x64x16 shared-memory tiled fp32 kernel (sgemm_kernel); every op records a closure on a tape that the
// backward pass replays in reverse, accumulating into gradient buffers.  Parameters, Adam moments and BatchNorm buffers stay
// resident in HBM between calls (LearnStep); mzb_learn_commit hands the trained tensors to the search engine's packed weights.
// A correctness-first fp32 path (the reference trains in fp32): every gradient is checked against the unmodified reference's
// autograd (tests/test_learn.py); it is not tuned for throughput.
#include <cmath>
#include <functional>
#include <map>
#include <string>

#include "engine.h"

namespace mzb {
namespace learn_detail {

struct Tn {   // activation [c][n][h][w] with its gradient (g == nullptr: no gradient wanted)
    float* p = nullptr; float* g = nullptr;
    int c = 0, n = 0, h = 0, w = 0;
    long numel() const { return (long)c * n * h * w; }
    long cols() const { return (long)n * h * w; }
};
struct Mat { float* p = nullptr; float* g = nullptr; int rows = 0, cols = 0; };   // [rows = samples][cols = features]

struct Param {
    DevBuf w, g, m, v;          // value, gradient, Adam first / second moment
    std::vector<int64_t> dims;
    size_t numel = 0;
    bool trainable = false;     // BatchNorm running statistics are buffers: no gradient, no Adam
    bool touched = false;       // took part in this step's graph (torch's Adam skips parameters whose grad is undefined)
    int64_t steps = 0;          // Adam's per-parameter step count
};

// bump allocator over chunks that persist between calls: a training step allocates several hundred activations
struct Arena {
    std::vector<std::unique_ptr<DevBuf>> chunks;
    std::vector<size_t> sizes;
    size_t chunk = 0, off = 0;
    void reset() { chunk = 0; off = 0; }
    float* get(size_t bytes) {
        bytes = (bytes + 255) & ~(size_t)255;
        while (chunk < chunks.size() && off + bytes > sizes[chunk]) { ++chunk; off = 0; }
        if (chunk == chunks.size()) {
            const size_t sz = std::max(bytes, (size_t)256 << 20);
            chunks.emplace_back(new DevBuf());
            if (chunks.back()->alloc(sz) != cudaSuccess) { chunks.pop_back(); return nullptr; }
            sizes.push_back(sz);
            off = 0;
        }
        float* p = (float*)((char*)chunks[chunk]->p + off);
        off += bytes;
        return p;
    }
};

inline unsigned nblk(long n) { return (unsigned)((n + 255) / 256); }

// ---------------------------------------------------------------- GEMM: C[M][N] (+)= A (M x K) x B (K x N), fp32
// A(i,k) = A[i * sAi + k * sAk], B(k,j) = B[k * sBk + j * sBj]: one of each pair of strides is 1 (a matrix or its transpose as it
// lies).  64 x 64 tile of C per block of 256 threads (4 x 4 per thread), K walked 16 at a time through shared memory.
template <bool A_KCONTIG, bool B_JCONTIG>
__global__ void __launch_bounds__(256) sgemm_kernel(const float* __restrict__ A, long sAi, long sAk, const float* __restrict__ B, long sBk, long sBj,
                                                   float* __restrict__ C, long ldc, int M, int N, int K, int accumulate,
                                                   const float* __restrict__ bias_row, const float* __restrict__ bias_col) {
    __shared__ float As[16][64 + 4], Bs[16][64 + 4];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int i0 = blockIdx.y * 64, j0 = blockIdx.x * 64;
    float acc[4][4] = {};
    for (int k0 = 0; k0 < K; k0 += 16) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int e = tid + r * 256;
            int ii, kk;
            if (A_KCONTIG) { kk = e & 15; ii = e >> 4; } else { ii = e & 63; kk = e >> 6; }
            const long gi = i0 + ii, gk = k0 + kk;
            As[kk][ii] = (gi < M && gk < K) ? A[gi * sAi + gk * sAk] : 0.f;
            int jj, kb;
            if (B_JCONTIG) { jj = e & 63; kb = e >> 6; } else { kb = e & 15; jj = e >> 4; }
            const long gj = j0 + jj, gkb = k0 + kb;
            Bs[kb][jj] = (gj < N && gkb < K) ? B[gkb * sBk + gj * sBj] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 16; ++kk) {
            float a[4], b[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) { a[r] = As[kk][ty * 4 + r]; b[r] = Bs[kk][tx * 4 + r]; }
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < 4; ++c) acc[r][c] = fmaf(a[r], b[c], acc[r][c]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const long gi = i0 + ty * 4 + r;
        if (gi >= M) continue;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const long gj = j0 + tx * 4 + c;
            if (gj >= N) continue;
            float v = acc[r][c];
            if (bias_row) v += bias_row[gi];
            if (bias_col) v += bias_col[gj];
            float* out = C + gi * ldc + gj;
            *out = accumulate ? *out + v : v;
        }
    }
}
// host side: the strides pick the variant
inline void sgemm(cudaStream_t s, const float* A, long sAi, long sAk, const float* B, long sBk, long sBj, float* C, long ldc, int M, int N, int K,
                  bool accumulate, const float* bias_row = nullptr, const float* bias_col = nullptr) {
    dim3 grid((N + 63) / 64, (M + 63) / 64);
    const bool ak = sAk == 1, bj = sBj == 1;
    if (ak && bj) sgemm_kernel<true, true><<<grid, 256, 0, s>>>(A, sAi, sAk, B, sBk, sBj, C, ldc, M, N, K, accumulate, bias_row, bias_col);
    else if (ak) sgemm_kernel<true, false><<<grid, 256, 0, s>>>(A, sAi, sAk, B, sBk, sBj, C, ldc, M, N, K, accumulate, bias_row, bias_col);
    else if (bj) sgemm_kernel<false, true><<<grid, 256, 0, s>>>(A, sAi, sAk, B, sBk, sBj, C, ldc, M, N, K, accumulate, bias_row, bias_col);
    else sgemm_kernel<false, false><<<grid, 256, 0, s>>>(A, sAi, sAk, B, sBk, sBj, C, ldc, M, N, K, accumulate, bias_row, bias_col);
}

// ---------------------------------------------------------------- convolution support: im2col / col2im on CNHW tensors
// col[(ci * 9 + ky * 3 + kx)][n, yo, xo] = x[ci][n][yo * stride + ky - 1][xo * stride + kx - 1] (zero outside)
__global__ void k_im2col(const float* __restrict__ x, float* __restrict__ col, int Cin, int N, int H, int W, int Ho, int Wo, int stride) {
    const long cols = (long)N * Ho * Wo;
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)Cin * 9 * cols) return;
    const long j = i % cols;
    const int r = (int)(i / cols), ci = r / 9, ky = (r % 9) / 3, kx = r % 3;
    const int xo = (int)(j % Wo), yo = (int)((j / Wo) % Ho), n = (int)(j / ((long)Wo * Ho));
    const int yy = yo * stride + ky - 1, xx = xo * stride + kx - 1;
    col[i] = (yy >= 0 && yy < H && xx >= 0 && xx < W) ? x[(((long)ci * N + n) * H + yy) * W + xx] : 0.f;
}
// dx[ci][n][y][x] += sum over the <= 9 output positions that read this input element
__global__ void k_col2im_add(const float* __restrict__ dcol, float* __restrict__ dx, int Cin, int N, int H, int W, int Ho, int Wo, int stride) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)Cin * N * H * W) return;
    const int x = (int)(i % W), y = (int)((i / W) % H), n = (int)((i / ((long)W * H)) % N), ci = (int)(i / ((long)W * H * N));
    const long cols = (long)N * Ho * Wo;
    float acc = 0.f;
    for (int ky = 0; ky < 3; ++ky) {
        const int ty = y + 1 - ky;
        if (ty < 0 || ty % stride) continue;
        const int yo = ty / stride;
        if (yo >= Ho) continue;
        for (int kx = 0; kx < 3; ++kx) {
            const int tx = x + 1 - kx;
            if (tx < 0 || tx % stride) continue;
            const int xo = tx / stride;
            if (xo >= Wo) continue;
            acc += dcol[(long)(ci * 9 + ky * 3 + kx) * cols + ((long)n * Ho + yo) * Wo + xo];
        }
    }
    dx[i] += acc;
}

// ---------------------------------------------------------------- BatchNorm2d, training mode (model_layers.cpp:36-42: eps 1e-4, momentum 0.01)
// one block per channel over its contiguous run of `cnt` values.  Saves mean / inv_std for the backward pass.
__global__ void k_bn_fwd(const float* __restrict__ x, float* __restrict__ y, const float* __restrict__ gamma, const float* __restrict__ beta,
                         float* running_mean, float* running_var, const float* __restrict__ residual, float* __restrict__ save_mean,
                         float* __restrict__ save_inv, long cnt, float eps, float momentum, int relu) {
    const int c = blockIdx.x;
    const float* xc = x + (long)c * cnt;
    __shared__ double s_a[256], s_b[256];
    double sum = 0.0, sq = 0.0;
    for (long i = threadIdx.x; i < cnt; i += blockDim.x) { const double v = (double)xc[i]; sum += v; sq += v * v; }
    s_a[threadIdx.x] = sum; s_b[threadIdx.x] = sq;
    __syncthreads();
    for (int o = blockDim.x / 2; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) { s_a[threadIdx.x] += s_a[threadIdx.x + o]; s_b[threadIdx.x] += s_b[threadIdx.x + o]; }
        __syncthreads();
    }
    const double mean = s_a[0] / (double)cnt;
    double var = s_b[0] / (double)cnt - mean * mean;
    if (var < 0.0) var = 0.0;
    const float fmean = (float)mean, inv = (float)(1.0 / sqrt(var + (double)eps));
    const float g = gamma[c], b = beta[c];
    for (long i = threadIdx.x; i < cnt; i += blockDim.x) {
        float v = (xc[i] - fmean) * inv * g + b;
        if (residual) v += residual[(long)c * cnt + i];
        if (relu) v = fmaxf(v, 0.f);
        y[(long)c * cnt + i] = v;
    }
    if (threadIdx.x == 0) {
        save_mean[c] = fmean; save_inv[c] = inv;
        const double unbiased = cnt > 1 ? var * (double)cnt / (double)(cnt - 1) : var;
        running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * fmean;
        running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
    }
}
// dz = dy * [y > 0] (ReLU) ; residual.g += dz ; dgamma += sum dz xhat ; dbeta += sum dz ;
// dx += gamma * inv / cnt * (cnt * dz - sum dz - xhat * sum(dz xhat))
__global__ void k_bn_bwd(const float* __restrict__ x, const float* __restrict__ y, const float* __restrict__ dy, float* __restrict__ dx,
                         float* __restrict__ dres, const float* __restrict__ gamma, float* __restrict__ dgamma, float* __restrict__ dbeta,
                         const float* __restrict__ save_mean, const float* __restrict__ save_inv, long cnt, int relu) {
    const int c = blockIdx.x;
    const long base = (long)c * cnt;
    const float mean = save_mean[c], inv = save_inv[c];
    __shared__ double s_a[256], s_b[256];
    double s1 = 0.0, s2 = 0.0;
    for (long i = threadIdx.x; i < cnt; i += blockDim.x) {
        float dz = dy[base + i];
        if (relu && !(y[base + i] > 0.f)) dz = 0.f;
        s1 += (double)dz;
        s2 += (double)dz * (double)((x[base + i] - mean) * inv);
    }
    s_a[threadIdx.x] = s1; s_b[threadIdx.x] = s2;
    __syncthreads();
    for (int o = blockDim.x / 2; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) { s_a[threadIdx.x] += s_a[threadIdx.x + o]; s_b[threadIdx.x] += s_b[threadIdx.x + o]; }
        __syncthreads();
    }
    const float sum_dz = (float)s_a[0], sum_dzx = (float)s_b[0];
    const float k = gamma[c] * inv / (float)cnt;
    for (long i = threadIdx.x; i < cnt; i += blockDim.x) {
        float dz = dy[base + i];
        if (relu && !(y[base + i] > 0.f)) dz = 0.f;
        if (dres) dres[base + i] += dz;
        const float xhat = (x[base + i] - mean) * inv;
        dx[base + i] += k * ((float)cnt * dz - sum_dz - xhat * sum_dzx);
    }
    if (threadIdx.x == 0) { dgamma[c] += sum_dzx; dbeta[c] += sum_dz; }
}

// ---------------------------------------------------------------- normalize_hidden_state (models.cpp:16-26), one thread per (c, n) row
__global__ void k_norm_fwd(const float* __restrict__ x, float* __restrict__ y, long rows, int HW) {
    const long r = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    const float* row = x + r * HW;
    float mn = row[0], mx = row[0];
    for (int i = 1; i < HW; ++i) { mn = fminf(mn, row[i]); mx = fmaxf(mx, row[i]); }
    float scale = mx - mn;
    if (scale < 1e-5f) scale = 1e-5f;
    for (int i = 0; i < HW; ++i) y[r * HW + i] = (row[i] - mn) / scale;
}
// y = (x - mn) / s, s = mx - mn unless clamped (index_put_ of a constant: no gradient through s then).  min / max send their
// gradient to the FIRST element that attains them (torch's min(dim) / max(dim) indices on ties).
__global__ void k_norm_bwd(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ dx, long rows, int HW, float hook_scale) {
    const long r = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    const float* row = x + r * HW;
    const float* g = dy + r * HW;
    float mn = row[0], mx = row[0];
    int imn = 0, imx = 0;
    for (int i = 1; i < HW; ++i) {
        if (row[i] < mn) { mn = row[i]; imn = i; }
        if (row[i] > mx) { mx = row[i]; imx = i; }
    }
    float s = mx - mn;
    const bool clamped = s < 1e-5f;
    if (clamped) s = 1e-5f;
    float sum_g = 0.f, sum_gy = 0.f;
    for (int i = 0; i < HW; ++i) { const float gi = g[i] * hook_scale; sum_g += gi; sum_gy += gi * (row[i] - mn); }
    const float d_s = clamped ? 0.f : -sum_gy / (s * s);
    const float d_mn = -sum_g / s - d_s, d_mx = d_s;
    for (int i = 0; i < HW; ++i) {
        float v = g[i] * hook_scale / s;
        if (i == imn) v += d_mn;
        if (i == imx) v += d_mx;
        dx[r * HW + i] += v;
    }
}

// ---------------------------------------------------------------- small ops
__global__ void k_scale(float* __restrict__ x, long n, float a) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) x[i] *= a;
}
// NCHW (the caller's batch) -> CNHW
__global__ void k_nchw_to_cnhw(const float* __restrict__ x, float* __restrict__ y, int N, int C, int HW) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)N * C * HW) return;
    const int p = (int)(i % HW), n = (int)((i / HW) % N), c = (int)(i / ((long)HW * N));
    y[i] = x[((long)n * C + c) * HW + p];
}
// AvgPool2d(3, 2, 1), padding counted (model_layers.cpp:31-33); planes = C * N
__global__ void k_pool_fwd(const float* __restrict__ x, float* __restrict__ y, long planes, int H, int W, int Ho, int Wo) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= planes * Ho * Wo) return;
    const int xo = (int)(i % Wo), yo = (int)((i / Wo) % Ho);
    const long pl = i / ((long)Wo * Ho);
    float acc = 0.f;
    for (int ky = 0; ky < 3; ++ky)
        for (int kx = 0; kx < 3; ++kx) {
            const int yy = yo * 2 + ky - 1, xx = xo * 2 + kx - 1;
            if (yy >= 0 && yy < H && xx >= 0 && xx < W) acc += x[(pl * H + yy) * W + xx];
        }
    y[i] = acc / 9.f;
}
__global__ void k_pool_bwd(const float* __restrict__ dy, float* __restrict__ dx, long planes, int H, int W, int Ho, int Wo) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= planes * H * W) return;
    const int x = (int)(i % W), y = (int)((i / W) % H);
    const long pl = i / ((long)W * H);
    float acc = 0.f;
    for (int ky = 0; ky < 3; ++ky) {
        const int ty = y + 1 - ky;
        if (ty < 0 || (ty & 1) || ty / 2 >= Ho) continue;
        for (int kx = 0; kx < 3; ++kx) {
            const int tx = x + 1 - kx;
            if (tx < 0 || (tx & 1) || tx / 2 >= Wo) continue;
            acc += dy[(pl * Ho + ty / 2) * Wo + tx / 2];
        }
    }
    dx[i] += acc / 9.f;
}
// torch::cat({state, action planes}, 1) in CNHW: channels C .. C + ach - 1 = planes[action[n, step]]
__global__ void k_action_planes(const float* __restrict__ planes, const int* __restrict__ actions, int a_stride, int step, float* __restrict__ y,
                                int N, int ach, int HW) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)ach * N * HW) return;
    const int p = (int)(i % HW), n = (int)((i / HW) % N), c = (int)(i / ((long)HW * N));
    y[i] = planes[((long)actions[(long)n * a_stride + step] * ach + c) * HW + p];
}
__global__ void k_add(float* __restrict__ dst, const float* __restrict__ src, long n) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] += src[i];
}
// .view(batch, -1) of an NCHW tensor: rows[n][c * HW + p] = x[c][n][p]; and its adjoint
__global__ void k_cnhw_to_rows(const float* __restrict__ x, float* __restrict__ rows, int C, int N, int HW) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)C * N * HW) return;
    const int p = (int)(i % HW), n = (int)((i / HW) % N), c = (int)(i / ((long)HW * N));
    rows[((long)n * C + c) * HW + p] = x[i];
}
__global__ void k_rows_to_cnhw_add(const float* __restrict__ drows, float* __restrict__ dx, int C, int N, int HW) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)C * N * HW) return;
    const int p = (int)(i % HW), n = (int)((i / HW) % N), c = (int)(i / ((long)HW * N));
    dx[i] += drows[((long)n * C + c) * HW + p];
}
__global__ void k_relu(float* __restrict__ x, long n) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) x[i] = fmaxf(x[i], 0.f);
}
__global__ void k_relu_bwd(const float* __restrict__ y, float* __restrict__ dy, long n) {   // in place: dy *= [y > 0]
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && !(y[i] > 0.f)) dy[i] = 0.f;
}
// out[j] += sum_i m[i * ld_i + j * ld_j]  (bias gradients: column sums of [N][Out], row sums of [C][cols])
__global__ void k_sum_add(const float* __restrict__ m, long ld_i, long ld_j, long count, int outs, float* __restrict__ out) {
    const int j = blockIdx.x;
    if (j >= outs) return;
    __shared__ double s[256];
    double acc = 0.0;
    for (long i = threadIdx.x; i < count; i += blockDim.x) acc += (double)m[i * ld_i + j * ld_j];
    s[threadIdx.x] = acc;
    __syncthreads();
    for (int o = blockDim.x / 2; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[j] += (float)s[0];
}
// models.cpp:271-273: loss[n] += sum_k -target[n, k] * log_softmax(logits[n])[k]
__global__ void k_ce_fwd(const float* __restrict__ logits, const float* __restrict__ target, int t_stride, double* loss, int N, int S) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const float* lg = logits + (long)n * S;
    const float* tg = target + (long)n * t_stride;
    float mx = lg[0];
    for (int k = 1; k < S; ++k) mx = fmaxf(mx, lg[k]);
    float sum = 0.f;
    for (int k = 0; k < S; ++k) sum += expf(lg[k] - mx);
    const float lse = mx + logf(sum);
    float acc = 0.f;
    for (int k = 0; k < S; ++k) acc += -tg[k] * (lg[k] - lse);
    loss[n] += (double)acc;
}
// d logits[n, k] += up * (row_scale ? row_scale[n] : 1) * (softmax_k * sum(target) - target_k)
__global__ void k_ce_bwd(const float* __restrict__ logits, const float* __restrict__ target, int t_stride, float* __restrict__ dlogits, int N, int S,
                         float up, const float* __restrict__ row_scale) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const float* lg = logits + (long)n * S;
    const float* tg = target + (long)n * t_stride;
    float mx = lg[0];
    for (int k = 1; k < S; ++k) mx = fmaxf(mx, lg[k]);
    float sum = 0.f, tsum = 0.f;
    for (int k = 0; k < S; ++k) { sum += expf(lg[k] - mx); tsum += tg[k]; }
    const float u = up * (row_scale ? row_scale[n] : 1.f);
    for (int k = 0; k < S; ++k) dlogits[(long)n * S + k] += u * (expf(lg[k] - mx) / sum * tsum - tg[k]);
}
// torch::optim::Adam (betas 0.9 / 0.999, eps 1e-8, amsgrad off) with weight_decay folded into the gradient (L2)
__global__ void k_adam(float* __restrict__ w, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, long n, float lr, float wd,
                       float bc1, float bc2_sqrt) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float b1 = 0.9f, b2 = 0.999f, eps = 1e-8f;
    const float grad = g[i] + wd * w[i];
    const float mi = m[i] + (grad - m[i]) * (1.f - b1);           // exp_avg.lerp_(grad, 1 - beta1)
    const float vi = v[i] * b2 + (1.f - b2) * grad * grad;
    m[i] = mi; v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    w[i] = w[i] - (lr / bc1) * (mi / denom);
}

// ---------------------------------------------------------------- the network in train mode with a tape
struct LearnStep {
    mzb_engine* e = nullptr;
    cudaStream_t s = nullptr;
    std::map<std::string, Param> P;
    Arena arena;
    DevBuf col, dcol;                        // im2col scratch (forward + recompute in backward) and its gradient
    std::vector<std::function<void()>> tape;
    std::string err;
    int64_t weights_serial = -1;
    int64_t steps_done = 0;
    bool device_dirty = false;               // the device holds values the host training state has not seen

    void fail(const std::string& m) { if (err.empty()) err = m; }
    Param* param(const std::string& name) {
        auto it = P.find(name);
        if (it == P.end()) { fail("missing tensor " + name); return nullptr; }
        it->second.touched = true;
        return &it->second;
    }
    float* alloc(size_t floats, bool zero) {
        float* p = arena.get(std::max<size_t>(floats, 1) * sizeof(float));
        if (!p) { fail("out of device memory"); return nullptr; }
        if (zero) cudaMemsetAsync(p, 0, floats * sizeof(float), s);
        return p;
    }
    Tn tensor(int c, int n, int h, int w, bool grad = true) {
        Tn t; t.c = c; t.n = n; t.h = h; t.w = w;
        t.p = alloc((size_t)t.numel(), false);
        if (grad) t.g = alloc((size_t)t.numel(), true);
        return t;
    }
    Mat matrix(int rows, int cols) {
        Mat m; m.rows = rows; m.cols = cols;
        m.p = alloc((size_t)rows * cols, false);
        m.g = alloc((size_t)rows * cols, true);
        return m;
    }
    float* scratch(DevBuf& b, size_t floats) {
        if (b.bytes < floats * sizeof(float)) {
            cudaStreamSynchronize(s);   // the buffer may still be read by queued kernels
            if (b.alloc(floats * sizeof(float) * 5 / 4) != cudaSuccess) { fail("out of device memory"); return nullptr; }
        }
        return b.as<float>();
    }

    // Conv2d(3x3, padding 1, no bias): model_layers.cpp:9-21
    Tn conv3(const Tn& x, const std::string& name, int cout, int stride) {
        Tn y = tensor(cout, x.n, (x.h + 2 - 3) / stride + 1, (x.w + 2 - 3) / stride + 1);
        Param* W = param(name + ".weight");
        const int K9 = x.c * 9;
        const long cols = y.cols();
        float* cl = scratch(col, (size_t)K9 * cols);
        if (!W || !y.p || !y.g || !cl) return y;
        k_im2col<<<nblk((long)K9 * cols), 256, 0, s>>>(x.p, cl, x.c, x.n, x.h, x.w, y.h, y.w, stride);
        sgemm(s, W->w.as<float>(), K9, 1, cl, cols, 1, y.p, cols, cout, (int)cols, K9, false);
        tape.push_back([=]() {
            float* c2 = scratch(col, (size_t)K9 * cols);
            if (!c2) return;
            k_im2col<<<nblk((long)K9 * cols), 256, 0, s>>>(x.p, c2, x.c, x.n, x.h, x.w, y.h, y.w, stride);
            sgemm(s, y.g, cols, 1, c2, 1, cols, W->g.as<float>(), K9, cout, K9, (int)cols, true);                 // dW += dY col^T
            if (x.g) {
                float* dc = scratch(dcol, (size_t)K9 * cols);
                if (!dc) return;
                sgemm(s, W->w.as<float>(), 1, K9, y.g, cols, 1, dc, cols, K9, (int)cols, cout, false);             // dcol = W^T dY
                k_col2im_add<<<nblk(x.numel()), 256, 0, s>>>(dc, x.g, x.c, x.n, x.h, x.w, y.h, y.w, stride);
            }
        });
        return y;
    }
    // BatchNorm2d (train) [+ residual] + ReLU
    Tn bn(const Tn& x, const std::string& name, const Tn* residual, bool relu) {
        Tn y = tensor(x.c, x.n, x.h, x.w);
        Param *g = param(name + ".weight"), *b = param(name + ".bias"), *rm = param(name + ".running_mean"), *rv = param(name + ".running_var");
        float* save = alloc((size_t)2 * x.c, false);
        if (!g || !b || !rm || !rv || !y.p || !y.g || !save) return y;
        const long cnt = x.cols();
        const Tn res = residual ? *residual : Tn();
        k_bn_fwd<<<x.c, 256, 0, s>>>(x.p, y.p, g->w.as<float>(), b->w.as<float>(), rm->w.as<float>(), rv->w.as<float>(), res.p, save, save + x.c, cnt,
                                    1e-4f, 0.01f, relu ? 1 : 0);
        tape.push_back([=]() {
            k_bn_bwd<<<x.c, 256, 0, s>>>(x.p, y.p, y.g, x.g, res.g, g->w.as<float>(), g->g.as<float>(), b->g.as<float>(), save, save + x.c, cnt, relu ? 1 : 0);
        });
        return y;
    }
    // ResidualBlock (model_layers.cpp:70-88)
    Tn block(const Tn& x, const std::string& prefix, int idx) {
        const std::string p = prefix + ".resnet_" + std::to_string(idx) + "_";
        Tn a = bn(conv3(x, p + "conv1", x.c, 1), p + "bn1", nullptr, true);
        return bn(conv3(a, p + "conv2", x.c, 1), p + "bn2", &x, true);
    }
    // normalize_hidden_state; `hook` = the factor Learn's register_hook puts on the gradient of the returned state (vprnet.cpp:272)
    Tn normalize(const Tn& x, float hook) {
        if (!e->cfg.normalize_hidden_states) {
            if (hook != 1.f) tape.push_back([=]() { k_scale<<<nblk(x.numel()), 256, 0, s>>>(x.g, x.numel(), hook); });
            return x;
        }
        Tn y = tensor(x.c, x.n, x.h, x.w);
        if (!y.p || !y.g) return y;
        const long rows = (long)x.c * x.n;
        const int HW = x.h * x.w;
        k_norm_fwd<<<nblk(rows), 256, 0, s>>>(x.p, y.p, rows, HW);
        tape.push_back([=]() { k_norm_bwd<<<nblk(rows), 256, 0, s>>>(x.p, y.g, x.g, rows, HW, hook); });
        return y;
    }
    Tn pool(const Tn& x) {
        Tn y = tensor(x.c, x.n, (x.h + 2 - 3) / 2 + 1, (x.w + 2 - 3) / 2 + 1);
        if (!y.p || !y.g) return y;
        const long planes = (long)x.c * x.n;
        k_pool_fwd<<<nblk(y.numel()), 256, 0, s>>>(x.p, y.p, planes, x.h, x.w, y.h, y.w);
        tape.push_back([=]() { k_pool_bwd<<<nblk(x.numel()), 256, 0, s>>>(y.g, x.g, planes, x.h, x.w, y.h, y.w); });
        return y;
    }
    // Conv2d(1x1, bias) -> .view(batch, -1) -> MLP (model_layers.cpp:46-65): logits [N][out]
    Mat head(const Tn& x, const std::string& conv, const std::string& mlp, int red, const int32_t* layers, int n_layers, int out) {
        Tn r = tensor(red, x.n, x.h, x.w);
        Param *cw = param(conv + ".weight"), *cb = param(conv + ".bias");
        const long cols = x.cols();
        const int Cin = x.c, HW = x.h * x.w, N = x.n;
        Mat cur = matrix(N, red * HW);
        if (!cw || !cb || !r.p || !r.g || !cur.p || !cur.g) return cur;
        sgemm(s, cw->w.as<float>(), Cin, 1, x.p, cols, 1, r.p, cols, red, (int)cols, Cin, false, cb->w.as<float>(), nullptr);
        tape.push_back([=]() {
            sgemm(s, r.g, cols, 1, x.p, 1, cols, cw->g.as<float>(), Cin, red, Cin, (int)cols, true);                  // dW += dR X^T
            k_sum_add<<<red, 256, 0, s>>>(r.g, 1, cols, cols, red, cb->g.as<float>());                              // db += row sums
            sgemm(s, cw->w.as<float>(), 1, Cin, r.g, cols, 1, x.g, cols, Cin, (int)cols, red, true);                  // dX += W^T dR
        });
        k_cnhw_to_rows<<<nblk(r.numel()), 256, 0, s>>>(r.p, cur.p, red, N, HW);
        tape.push_back([=]() { k_rows_to_cnhw_add<<<nblk(r.numel()), 256, 0, s>>>(cur.g, r.g, red, N, HW); });
        int in = red * HW;
        for (int l = 0; l <= n_layers; ++l) {
            const int o = l < n_layers ? layers[l] : out;
            const bool relu = l < n_layers;
            Mat y = matrix(N, o);
            Param *lw = param(mlp + ".linear_" + std::to_string(l) + ".weight"), *lb = param(mlp + ".linear_" + std::to_string(l) + ".bias");
            if (!lw || !lb || !y.p || !y.g) return y;
            const Mat xin = cur;
            sgemm(s, xin.p, in, 1, lw->w.as<float>(), 1, in, y.p, o, N, o, in, false, nullptr, lb->w.as<float>());     // Y = X W^T + b
            if (relu) k_relu<<<nblk((long)N * o), 256, 0, s>>>(y.p, (long)N * o);
            tape.push_back([=]() {
                if (relu) k_relu_bwd<<<nblk((long)N * o), 256, 0, s>>>(y.p, y.g, (long)N * o);
                sgemm(s, y.g, 1, o, xin.p, in, 1, lw->g.as<float>(), in, o, in, N, true);                             // dW += dY^T X
                k_sum_add<<<o, 256, 0, s>>>(y.g, o, 1, N, o, lb->g.as<float>());                                    // db += column sums
                sgemm(s, y.g, o, 1, lw->w.as<float>(), in, 1, xin.g, in, N, in, o, true);                             // dX += dY W
            });
            cur = y; in = o;
        }
        return cur;
    }
    // loss term: forward adds to the per-sample sums, backward seeds the logits' gradient
    void cross_entropy(const Mat& logits, const float* target, int t_stride, double* loss, float up, const float* row_scale) {
        const int N = logits.rows, S = logits.cols;
        k_ce_fwd<<<(N + 127) / 128, 128, 0, s>>>(logits.p, target, t_stride, loss, N, S);
        tape.push_back([=]() { k_ce_bwd<<<(N + 127) / 128, 128, 0, s>>>(logits.p, target, t_stride, logits.g, N, S, up, row_scale); });
    }
    // RepresentationNetworkImpl::forward (models.cpp:31-65) + normalize (models.cpp:227-229)
    Tn representation(const Tn& obs) {
        const mzb_config& c = e->cfg;
        const int C = c.resnet_channels;
        const std::string rh = "representation_network.representation_head.0";
        Tn x;
        if (c.downsample) {   // model_layers.cpp:140-155
            x = conv3(obs, rh + ".resnet_head_downsample_conv1", C / 2, 2);
            for (int i = 0; i < 2; ++i) x = block(x, rh + ".resnet_head_downsample_resblock1." + std::to_string(i), i);
            x = conv3(x, rh + ".resnet_head_downsample_conv2", C, 2);
            for (int i = 2; i < 5; ++i) x = block(x, rh + ".resnet_head_downsample_resblock2." + std::to_string(i - 2), i);
            x = pool(x);
            for (int i = 5; i < 8; ++i) x = block(x, rh + ".resnet_head_downsample_resblock3." + std::to_string(i - 5), i);
            x = pool(x);
        } else {
            x = bn(conv3(obs, rh + ".representation_resnet_head_conv", C, 1), rh + ".representation_resnet_head_bn", nullptr, true);
        }
        for (int i = 0; i < c.representation_blocks; ++i) x = block(x, "representation_network.representation_layers." + std::to_string(i), i);
        return normalize(x, 1.f);
    }
    // PredictionNetworkImpl::forward (models.cpp:153-166)
    void prediction(const Tn& state, Mat* policy, Mat* value) {
        const mzb_config& c = e->cfg;
        Tn x = state;
        for (int i = 0; i < c.prediction_blocks; ++i) x = block(x, "prediction_network.prediction_resnet_body." + std::to_string(i), i);
        *policy = head(x, "prediction_network.prediction_policy_conv", "prediction_network.prediction_policy_mlp.prediction_policy_mlp",
                       c.policy_reduced_channels, c.policy_head_layers, c.n_policy_head_layers, e->dims.A);
        *value = head(x, "prediction_network.prediction_value_conv", "prediction_network.prediction_value_mlp.prediction_value_mlp",
                      c.value_reduced_channels, c.value_head_layers, c.n_value_head_layers, e->dims.Sv);
    }
    // DynamicsNetworkImpl::forward (models.cpp:105-117): reward from the UN-normalised output; next state normalised, its gradient halved
    Tn dynamics(const Tn& state, const float* planes, const int* actions, int a_stride, int step, Mat* reward) {
        const mzb_config& c = e->cfg;
        const int C = c.resnet_channels, ach = c.action_channels;
        Tn in = tensor(C + ach, state.n, state.h, state.w);
        if (!in.p || !in.g) return in;
        const long cols = state.cols();
        cudaMemcpyAsync(in.p, state.p, (size_t)C * cols * sizeof(float), cudaMemcpyDeviceToDevice, s);
        k_action_planes<<<nblk((long)ach * cols), 256, 0, s>>>(planes, actions, a_stride, step, in.p + (size_t)C * cols, state.n, ach, state.h * state.w);
        tape.push_back([=]() { k_add<<<nblk((long)C * cols), 256, 0, s>>>(state.g, in.g, (long)C * cols); });
        const std::string dh = "dynamics_network.dynamics_resnet_head";
        Tn x = bn(conv3(in, dh + ".resnet_head_conv", C, 1), dh + ".resnet_head_bn", nullptr, true);
        for (int i = 0; i < c.dynamics_blocks; ++i) x = block(x, "dynamics_network.dynamics_resnet_body." + std::to_string(i), i);
        *reward = head(x, "dynamics_network.dynamics_reward_conv", "dynamics_network.dynamics_reward_mlp.dynamics_reward_mlp",
                       c.reward_reduced_channels, c.reward_head_layers, c.n_reward_head_layers, e->dims.Sr);
        return normalize(x, 0.5f);
    }
};

void learn_step_free(void* p) { delete static_cast<LearnStep*>(p); }

}  // namespace learn_detail
}  // namespace mzb

using namespace mzb;
using namespace mzb::learn_detail;

namespace mzb {
// the device-resident training state -> the host map mzb_learn_get_tensor / mzb_learn_forward / mzb_save_checkpoint read
int learn_sync_host(mzb_engine* e) {
    LearnStep* st = static_cast<LearnStep*>(e->learn_step);
    if (!st || !st->device_dirty || st->weights_serial != e->weights_serial) return MZB_OK;
    MZB_CUDA(cudaSetDevice(e->device));
    for (auto& kv : st->P) {
        std::vector<float>& h = e->learn_state[kv.first];
        h.resize(kv.second.numel);
        MZB_CUDA(cudaMemcpyAsync(h.data(), kv.second.w.p, kv.second.numel * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    }
    MZB_CUDA(cudaStreamSynchronize(e->stream));
    e->learn_state_serial = e->weights_serial;
    st->device_dirty = false;
    return MZB_OK;
}
}  // namespace mzb

static bool is_running_stat(const std::string& n) {
    return (n.size() > 12 && n.compare(n.size() - 12, 12, "running_mean") == 0) || (n.size() > 11 && n.compare(n.size() - 11, 11, "running_var") == 0);
}

// (re)build the device state: from the loaded tensors after a load (Adam starts over, like a fresh VPRNetModel), or refresh the values
// from the host map when mzb_learn_forward moved BatchNorm statistics there since the last step
static int learn_step_prepare(mzb_engine* e, LearnStep** out) {
    LearnStep* st = static_cast<LearnStep*>(e->learn_step);
    if (!st) {
        st = new LearnStep();
        e->learn_step = st;
        e->learn_step_free = &learn_step_free;
    }
    st->e = e; st->s = e->stream;
    if (st->weights_serial != e->weights_serial) {
        st->P.clear();
        st->steps_done = 0;
        const bool from_host = e->learn_state_serial == e->weights_serial && !e->learn_state.empty();
        for (const mzb_engine::RawTensor& r : e->raw_weights) {
            Param& p = st->P[r.name];
            p.dims = r.dims; p.numel = r.data.size();
            p.trainable = !is_running_stat(r.name);
            const std::vector<float>& src = from_host ? e->learn_state[r.name] : r.data;
            MZB_REQUIRE(src.size() == p.numel, "training state of %s has the wrong size", r.name.c_str());
            MZB_CUDA(p.w.alloc(std::max<size_t>(p.numel, 1) * sizeof(float)));
            MZB_CUDA(cudaMemcpyAsync(p.w.p, src.data(), p.numel * sizeof(float), cudaMemcpyHostToDevice, st->s));
            if (p.trainable) {
                MZB_CUDA(p.g.alloc(std::max<size_t>(p.numel, 1) * sizeof(float)));
                MZB_CUDA(p.m.alloc(std::max<size_t>(p.numel, 1) * sizeof(float)));
                MZB_CUDA(p.v.alloc(std::max<size_t>(p.numel, 1) * sizeof(float)));
                MZB_CUDA(cudaMemsetAsync(p.m.p, 0, p.numel * sizeof(float), st->s));
                MZB_CUDA(cudaMemsetAsync(p.v.p, 0, p.numel * sizeof(float), st->s));
            }
        }
        MZB_CUDA(cudaStreamSynchronize(st->s));
        st->weights_serial = e->weights_serial;
        e->learn_host_dirty = false;
        st->device_dirty = false;
    } else if (e->learn_host_dirty) {
        for (auto& kv : st->P) {
            auto it = e->learn_state.find(kv.first);
            if (it == e->learn_state.end() || it->second.size() != kv.second.numel) continue;
            MZB_CUDA(cudaMemcpyAsync(kv.second.w.p, it->second.data(), kv.second.numel * sizeof(float), cudaMemcpyHostToDevice, st->s));
        }
        MZB_CUDA(cudaStreamSynchronize(st->s));
        e->learn_host_dirty = false;
    }
    *out = st;
    return MZB_OK;
}

extern "C" {

int mzb_learn_step(mzb_engine* e, const mzb_batch* batch, int num_unroll_steps, double value_loss_weight, double learning_rate, double l2_weight_decay,
                   mzb_loss_info* out) {
    MZB_REQUIRE(e && batch && out, "null argument");
    MZB_REQUIRE(e->cfg.evaluator == MZB_EVAL_NETWORK && e->weights_loaded && !e->raw_weights.empty(), "no weights loaded (mzb_load_weights / mzb_load_checkpoint)");
    const int B = batch->num_samples, K1 = num_unroll_steps + 1, A = e->dims.A;
    MZB_REQUIRE(B >= 1 && num_unroll_steps >= 0, "empty batch or negative unroll");
    MZB_REQUIRE(batch->actions && batch->stacked_observations && batch->target_values && batch->target_rewards && batch->target_policies && batch->priorities,
                "null batch field");
    MZB_REQUIRE(num_unroll_steps == 0 || batch->gradient_scale, "gradient_scale is needed when unrolling");
    for (int i = 0; i < B * K1; ++i) MZB_REQUIRE(batch->actions[i] >= 0 && batch->actions[i] < A, "batch action %d outside the action space", batch->actions[i]);
    MZB_CUDA(cudaSetDevice(e->device));
    LearnStep* st = nullptr;
    int rc = learn_step_prepare(e, &st);
    if (rc) return rc;
    try {
        LearnStep& net = *st;
        cudaStream_t s = net.s;
        net.arena.reset();
        net.tape.clear();
        net.err.clear();
        for (auto& kv : net.P) {   // model_->zero_grad()
            kv.second.touched = false;
            if (kv.second.trainable) MZB_CUDA(cudaMemsetAsync(kv.second.g.p, 0, kv.second.numel * sizeof(float), s));
        }
        const NetDims& d = e->dims;
        const int hw_in = d.in_h * d.in_w, hw = d.hw(), Sv = d.Sv, Sr = d.Sr;
        // inputs: the batch as Learn builds its tensors (vprnet.cpp:208-233), doubles -> float on the host
        float* stage = net.alloc((size_t)B * d.cin * hw_in, false);
        Tn obs = net.tensor(d.cin, B, d.in_h, d.in_w, false);
        MZB_REQUIRE(stage && obs.p, "out of device memory");
        MZB_CUDA(cudaMemcpyAsync(stage, batch->stacked_observations, (size_t)obs.numel() * sizeof(float), cudaMemcpyHostToDevice, s));
        k_nchw_to_cnhw<<<nblk(obs.numel()), 256, 0, s>>>(stage, obs.p, B, d.cin, hw_in);
        std::vector<float> planes((size_t)A * std::max(1, d.ach) * hw);
        for (int a = 0; a < A; ++a)
            for (int i = 0; i < d.ach * hw; ++i)
                planes[(size_t)a * d.ach * hw + i] = e->cfg.action_planes ? e->cfg.action_planes[(size_t)a * d.ach * hw + i] : (float)((double)a / A);
        std::vector<float> tv((size_t)B * K1), tr((size_t)B * K1), tp((size_t)B * K1 * A), tv_enc((size_t)B * K1 * Sv), tr_enc((size_t)B * K1 * Sr), gs(B, 1.f);
        for (size_t i = 0; i < tv.size(); ++i) { tv[i] = (float)batch->target_values[i]; tr[i] = (float)batch->target_rewards[i]; }
        for (size_t i = 0; i < tp.size(); ++i) tp[i] = (float)batch->target_policies[i];
        float psum = 0.f;
        for (int b = 0; b < B; ++b) { psum += (float)batch->priorities[b]; if (batch->gradient_scale) gs[b] = (float)batch->gradient_scale[b]; }
        MZB_CUDA(cudaStreamSynchronize(s));
        rc = mzb_value_encode(e, 0, tv.data(), B * K1, tv_enc.data());
        if (rc) return rc;
        rc = mzb_value_encode(e, 1, tr.data(), B * K1, tr_enc.data());
        if (rc) return rc;
        int* d_actions = (int*)net.alloc((size_t)B * K1, false);
        float* d_planes = net.alloc(planes.size(), false);
        float* d_tv = net.alloc(tv_enc.size(), false);
        float* d_tr = net.alloc(tr_enc.size(), false);
        float* d_tp = net.alloc(tp.size(), false);
        float* d_gs = net.alloc(B, false);
        double* d_loss = (double*)net.alloc((size_t)6 * B, true);   // value | reward | policy per sample, as doubles
        MZB_REQUIRE(d_actions && d_planes && d_tv && d_tr && d_tp && d_gs && d_loss, "out of device memory");
        MZB_CUDA(cudaMemcpyAsync(d_actions, batch->actions, (size_t)B * K1 * sizeof(int), cudaMemcpyHostToDevice, s));
        MZB_CUDA(cudaMemcpyAsync(d_planes, planes.data(), planes.size() * sizeof(float), cudaMemcpyHostToDevice, s));
        MZB_CUDA(cudaMemcpyAsync(d_tv, tv_enc.data(), tv_enc.size() * sizeof(float), cudaMemcpyHostToDevice, s));
        MZB_CUDA(cudaMemcpyAsync(d_tr, tr_enc.data(), tr_enc.size() * sizeof(float), cudaMemcpyHostToDevice, s));
        MZB_CUDA(cudaMemcpyAsync(d_tp, tp.data(), tp.size() * sizeof(float), cudaMemcpyHostToDevice, s));
        MZB_CUDA(cudaMemcpyAsync(d_gs, gs.data(), (size_t)B * sizeof(float), cudaMemcpyHostToDevice, s));
        double *lv = d_loss, *lr = lv + B, *lp = lr + B;
        // d total / d (per-sample loss sum) = mean(priorities) / B: the [B,1] x [B] broadcast of vprnet.cpp:295 under .mean()
        const float up = psum / (float)B / (float)B;
        const float wv = (float)value_loss_weight;

        // ---- forward (records the tape)
        Mat policy, value, reward;
        Tn state = net.representation(obs);
        net.prediction(state, &policy, &value);
        MZB_REQUIRE(net.err.empty(), "%s", net.err.c_str());
        net.cross_entropy(value, d_tv, K1 * Sv, lv, up * wv, nullptr);
        net.cross_entropy(policy, d_tp, K1 * A, lp, up, nullptr);
        std::vector<float> v0((size_t)B * Sv);
        MZB_CUDA(cudaMemcpyAsync(v0.data(), value.p, v0.size() * sizeof(float), cudaMemcpyDeviceToHost, s));
        for (int step = 1; step < K1; ++step) {
            state = net.dynamics(state, d_planes, d_actions, K1, step, &reward);
            net.prediction(state, &policy, &value);
            MZB_REQUIRE(net.err.empty(), "%s", net.err.c_str());
            // the step losses' gradients are scaled per sample by gradient_scale (vprnet.cpp:278-286)
            net.cross_entropy(value, d_tv + (size_t)step * Sv, K1 * Sv, lv, up * wv, d_gs);
            net.cross_entropy(reward, d_tr + (size_t)step * Sr, K1 * Sr, lr, up, d_gs);
            net.cross_entropy(policy, d_tp + (size_t)step * A, K1 * A, lp, up, d_gs);
        }
        std::vector<double> loss((size_t)3 * B);
        MZB_CUDA(cudaMemcpyAsync(loss.data(), d_loss, loss.size() * sizeof(double), cudaMemcpyDeviceToHost, s));
        // ---- backward
        for (auto it = net.tape.rbegin(); it != net.tape.rend(); ++it) (*it)();
        MZB_REQUIRE(net.err.empty(), "%s", net.err.c_str());
        MZB_CUDA(cudaGetLastError());
        // ---- optimizer
        for (auto& kv : net.P) {
            Param& p = kv.second;
            if (!p.trainable || !p.touched) continue;
            ++p.steps;
            const double bc1 = 1.0 - std::pow(0.9, (double)p.steps), bc2 = 1.0 - std::pow(0.999, (double)p.steps);
            k_adam<<<nblk((long)p.numel), 256, 0, s>>>(p.w.as<float>(), p.g.as<float>(), p.m.as<float>(), p.v.as<float>(), (long)p.numel,
                                                     (float)learning_rate, (float)l2_weight_decay, (float)bc1, (float)std::sqrt(bc2));
        }
        MZB_CUDA(cudaGetLastError());
        MZB_CUDA(cudaStreamSynchronize(s));
        net.tape.clear();
        ++net.steps_done;
        net.device_dirty = true;
        // LossInfo (vprnet.cpp:293-310)
        double sv = 0, sr = 0, sp = 0, stot = 0, spri = 0;
        for (int b = 0; b < B; ++b) {
            sv += loss[b]; sr += loss[B + b]; sp += loss[2 * B + b];
            stot += loss[b] * value_loss_weight + loss[B + b] + loss[2 * B + b];
            spri += batch->priorities[b];
        }
        out->value = sv / B; out->reward = sr / B; out->policy = sp / B;
        out->total_loss = (spri / B) * (stot / B);
        if (out->errors) {
            std::vector<float> dec(B);
            rc = mzb_value_decode(e, 0, v0.data(), B, 1, dec.data());
            if (rc) return rc;
            for (int b = 0; b < B; ++b) out->errors[b] = (double)dec[b] - (double)tv[(size_t)b * K1];
        }
        return MZB_OK;
    } catch (const std::exception& ex) {
        set_error("mzb_learn_step: %s", ex.what());
        return MZB_ERR_INVALID;
    }
}

// the gradient the last mzb_learn_step left on a parameter (before the optimizer used it)
int mzb_learn_get_gradient(mzb_engine* e, const char* name, float* out, int64_t numel) {
    MZB_REQUIRE(e && name && out, "null argument");
    LearnStep* st = static_cast<LearnStep*>(e->learn_step);
    MZB_REQUIRE(st && st->steps_done > 0, "mzb_learn_step has not run");
    auto it = st->P.find(name);
    MZB_REQUIRE(it != st->P.end() && it->second.trainable, "%s is not a trainable parameter", name);
    MZB_REQUIRE((int64_t)it->second.numel == numel, "parameter %s has %zu elements, not %lld", name, it->second.numel, (long long)numel);
    MZB_CUDA(cudaSetDevice(e->device));
    MZB_CUDA(cudaMemcpy(out, it->second.g.p, it->second.numel * sizeof(float), cudaMemcpyDeviceToHost));
    return MZB_OK;
}

// learner.cpp:93-101 on one GPU: the trained tensors become the weights the engine searches with (and SaveCheckpoint writes)
int mzb_learn_commit(mzb_engine* e) {
    MZB_REQUIRE(e, "null argument");
    LearnStep* st = static_cast<LearnStep*>(e->learn_step);
    MZB_REQUIRE(st && st->weights_serial == e->weights_serial, "no training state (mzb_learn_step has not run since the last load)");
    int rc = learn_sync_host(e);
    if (rc) return rc;
    std::vector<mzb_engine::RawTensor> keep = e->raw_weights;
    std::vector<std::pair<std::string, int64_t>> longs = e->raw_longs;
    std::vector<mzb_tensor_desc> descs;
    for (mzb_engine::RawTensor& r : keep) {
        auto it = e->learn_state.find(r.name);
        if (it != e->learn_state.end() && it->second.size() == r.data.size()) r.data = it->second;
        mzb_tensor_desc d;
        memset(&d, 0, sizeof(d));
        d.name = r.name.c_str(); d.data = r.data.data(); d.ndim = (int32_t)r.dims.size();
        for (size_t k = 0; k < r.dims.size(); ++k) d.dims[k] = r.dims[k];
        descs.push_back(d);
    }
    rc = mzb_load_weights(e, descs.data(), (int)descs.size());
    if (rc) return rc;
    e->raw_longs = longs;
    // the load bumped the serial; the training state (Adam moments included) stays the current one
    st->weights_serial = e->weights_serial;
    e->learn_state_serial = e->weights_serial;
    return MZB_OK;
}

}  // extern "C"
