// ahk_train.cu — the backward pass and the optimizer step of `model.fit` (ai/reinforcement/rl_agents.py:189,276;
// compiled at rl_agents.py:111,170 and ai/supervised/keras/models.py:119), hand-written: no autograd, no cuDNN, no cuBLAS.
//
// Forward (training mode): ahk_policy_forward_train (ahk_policy.cu) — the tcgen05 forward kernels with each BatchNorm's
// BATCH statistics folded into the next layer.  It leaves r1 = relu(conv1), r2, r3, h1, h2 (bf16) in the policy handle.
// Backward, from d loss / d logits (ahk_rl_ddqn_targets / ahk_rl_pg_targets), layer by layer:
//   * dense2 / dense1 data gradients: the TMA-fed tcgen05 GEMM (ahk_gemm.cuh) on bf16 copies, ReLU mask in its epilogue;
//   * every other GEMM-shaped gradient (weight gradients, whose reduction runs over batch x pixels; the transposed
//     convolutions, split by the parity of the input pixel so that no tap is a zero) on `simt_gemm`, a register-tiled
//     float32 CUDA-core GEMM whose operands are FUNCTORS: im2col, the BatchNorm affine map, the Keras layouts and the
//     (x - 128) / 128 of the uint8 frames are index arithmetic inside the loads, nothing is materialised;
//   * conv1's weight gradient (47 % of the network's MACs, uint8 input, 5 output channels): `conv1_wgrad_kernel`;
//   * BatchNorm backward with batch statistics (two passes: per-channel sums, then the elementwise map + ReLU mask);
//   * `sgd_step_kernel` / `adam_step_kernel`: one launch over the flat parameter vector — L2 kernel regulariser
//     (models.py:17), clipnorm on the global norm, learning-rate decay, momentum / Adam moments, BatchNorm moving statistics.
// Layouts are the Keras ones (conv [kh][kw][cin][cout], dense [in][out]) so the flat vector IS model.get_weights().
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/ahk.h"
#include "ahk_gemm.cuh"
#include "ahk_kernels.cuh"

namespace ahk {
namespace train {

constexpr int kP1 = 41 * 41, kP2 = 19 * 19, kP3 = 64;
constexpr float kBnEps = 1e-3f; // Keras BatchNormalization default (models.py:26)

__device__ __forceinline__ float bf(const __nv_bfloat16 *p, long long i) { return __bfloat162float(p[i]); }

// ------------------------------------------------------------------------------------------------ functor GEMM
// C[m][n] (+)= sum_k A(m, k) * B(k, n).  F provides  float a(int m, int k), float b(int k, int n), void store(int m, int n, float v, bool atomic)
// and the flags kAAlongM / kBAlongN: which index consecutive threads walk when a tile is staged (the contiguous one of the source).
template <class F, int TM, int TN, int RM, int RN>
__global__ void __launch_bounds__((TM / RM) * (TN / RN)) simt_gemm(F f, int M, int N, int K, int k_per_split)
{
    constexpr int TK = 16, kThreads = (TM / RM) * (TN / RN);
    __shared__ float s_a[TK][TM + 4], s_b[TK][TN + 4];
    const int tid = threadIdx.x, tn = tid % (TN / RN), tm = tid / (TN / RN);
    const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;
    const int k_begin = blockIdx.z * k_per_split, k_end = min(K, k_begin + k_per_split);
    float acc[RM][RN];
#pragma unroll
    for (int i = 0; i < RM; ++i)
#pragma unroll
        for (int j = 0; j < RN; ++j) acc[i][j] = 0.0f;
    for (int k0 = k_begin; k0 < k_end; k0 += TK) {
#pragma unroll
        for (int e = tid; e < TM * TK; e += kThreads) {
            const int mm = F::kAAlongM ? e % TM : e / TK, kk = F::kAAlongM ? e / TM : e % TK;
            s_a[kk][mm] = (m0 + mm < M && k0 + kk < k_end) ? f.a(m0 + mm, k0 + kk) : 0.0f;
        }
#pragma unroll
        for (int e = tid; e < TN * TK; e += kThreads) {
            const int nn = F::kBAlongN ? e % TN : e / TK, kk = F::kBAlongN ? e / TN : e % TK;
            s_b[kk][nn] = (n0 + nn < N && k0 + kk < k_end) ? f.b(k0 + kk, n0 + nn) : 0.0f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < TK; ++kk) {
            float av[RM], bv[RN];
#pragma unroll
            for (int i = 0; i < RM; ++i) av[i] = s_a[kk][tm * RM + i];
#pragma unroll
            for (int j = 0; j < RN; ++j) bv[j] = s_b[kk][tn * RN + j];
#pragma unroll
            for (int i = 0; i < RM; ++i)
#pragma unroll
                for (int j = 0; j < RN; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < RM; ++i)
#pragma unroll
        for (int j = 0; j < RN; ++j) {
            const int m = m0 + tm * RM + i, n = n0 + tn * RN + j;
            if (m < M && n < N) f.store(m, n, acc[i][j], gridDim.z > 1);
        }
}

template <class F, int TM = 64, int TN = 64, int RM = 4, int RN = 4>
cudaError_t run_gemm(const F &f, int M, int N, int K, int want_ctas, cudaStream_t s)
{
    if (M <= 0 || N <= 0 || K <= 0) return cudaSuccess;
    const int gm = (M + TM - 1) / TM, gn = (N + TN - 1) / TN;
    int splits = want_ctas / (gm * gn);
    if (splits < 1) splits = 1;
    int per = ((K + splits - 1) / splits + 15) / 16 * 16;
    if (per < 64) per = 64;
    splits = (K + per - 1) / per;
    simt_gemm<F, TM, TN, RM, RN><<<dim3(gn, gm, splits), (TM / RM) * (TN / RN), 0, s>>>(f, M, N, K, per);
    return cudaGetLastError();
}

// ---- weight gradients: C = dW (Keras layout, float32, zeroed before), reduction over rows
// dense: dW[in][out] = sum_b act(b, in) * dz[b][out];  act(b, in) = scale[in % cdiv] * x[b][in'] + shift[...]   (BatchNorm output, or plain)
struct DenseWgrad {
    static constexpr bool kAAlongM = true, kBAlongN = true;
    const __nv_bfloat16 *x; // [B][M]
    const float *dz;        // [B][N]
    float *dw;              // [rows][N]
    const float *s, *t;     // BatchNorm affine map per channel (nullptr: identity); channel = m % 64 (dense1: input index pixel * 64 + channel)
    int M, N, keras_perm;   // keras_perm: dW row of input (pixel, channel) is channel * 64 + pixel (Flatten of (C, H, W), models.py:44)
    __device__ float a(int m, int k) const
    {
        const float v = bf(x, static_cast<long long>(k) * M + m);
        return s ? fmaf(s[m & 63], v, t[m & 63]) : v;
    }
    __device__ float b(int k, int n) const { return dz[static_cast<long long>(k) * N + n]; }
    __device__ void store(int m, int n, float v, bool atomic) const
    {
        float *p = dw + static_cast<long long>(keras_perm ? (m & 63) * 64 + (m >> 6) : m) * N + n;
        if (atomic) atomicAdd(p, v); else *p = v;
    }
};
// conv (4x4, stride 2): dW[(tap, c)][o] = sum_{s, y, x} bn(r[s][2y + ky][2x + kx][c]) * dz[(s, y, x)][o]
// (geometry as template constants: every division below is by a constant)
template <int CIN, int CPAD, int IN, int OUT, int N>
struct ConvWgrad {
    static constexpr bool kAAlongM = true, kBAlongN = true;
    const __nv_bfloat16 *r; // [B][IN^2][CPAD]
    const float *dz;        // [B * OUT^2][N]
    float *dw;              // [16][CIN][N]
    const float *s, *t;
    __device__ float a(int m, int k) const
    {
        const int tap = m / CIN, c = m - tap * CIN, ky = tap >> 2, kx = tap & 3;
        const int smp = k / (OUT * OUT), p = k - smp * (OUT * OUT), y = p / OUT, x = p - y * OUT;
        const float v = bf(r, (static_cast<long long>(smp) * (IN * IN) + (2 * y + ky) * IN + 2 * x + kx) * CPAD + c);
        return fmaf(s[c], v, t[c]);
    }
    __device__ float b(int k, int n) const { return dz[static_cast<long long>(k) * N + n]; }
    __device__ void store(int m, int n, float v, bool atomic) const
    {
        float *p = dw + static_cast<long long>(m) * N + n;
        if (atomic) atomicAdd(p, v); else *p = v;
    }
};
// ---- data gradients
// dense: dx[b][in] = sum_out dz[b][out] * W[keras(in)][out]
struct DenseDgrad {
    static constexpr bool kAAlongM = false, kBAlongN = false;
    const float *dz;  // [B][K]
    const float *w;   // Keras [in][K]
    float *dx;        // [B][N]
    const __nv_bfloat16 *mask; // ReLU output of the layer below ([B][N]; nullptr: no mask)
    __nv_bfloat16 *dx_bf16;    // optional bf16 copy
    int K, N, keras_perm;
    __device__ float a(int m, int k) const { return dz[static_cast<long long>(m) * K + k]; }
    __device__ float b(int k, int n) const { return w[static_cast<long long>(keras_perm ? (n & 63) * 64 + (n >> 6) : n) * K + k]; }
    __device__ void store(int m, int n, float v, bool) const
    {
        const long long at = static_cast<long long>(m) * N + n;
        if (mask && !(bf(mask, at) > 0.0f)) v = 0.0f;
        dx[at] = v;
        if (dx_bf16) dx_bf16[at] = __float2bfloat16_rn(v);
    }
};
// transposed 4x4 stride-2 convolution, one PARITY CLASS (py, px) of the input pixels per launch: input pixel (Y, X) =
// (2 y' + py, 2 x' + px) receives taps ky = py + 2 jy, kx = px + 2 jx from output pixel (y' - jy, x' - jx):
// da[s][Y][X][c] = sum_{jy, jx, o} dz[s][y' - jy][x' - jx][o] * W[ky][kx][c][o]      (K = 4 * cout, no zero taps)
template <int CIN, int CPAD, int COUT, int IN, int OUT>
struct ConvDgrad {
    static constexpr bool kAAlongM = false, kBAlongN = false;
    const float *dz; // [B][OUT^2][COUT]
    const float *w;  // Keras [16][CIN][COUT]
    float *da;       // [B][IN^2][CPAD]
    int py, px, ny, nx; // parity class; ny x nx input pixels of this parity per sample
    __device__ float a(int m, int k) const
    {
        const int per = ny * nx, smp = m / per, p = m - smp * per, yq = p / nx, xq = p - yq * nx;
        const int j = k / COUT, o = k - j * COUT, y = yq - (j >> 1), x = xq - (j & 1);
        if (y < 0 || x < 0 || y >= OUT || x >= OUT) return 0.0f;
        return dz[(static_cast<long long>(smp) * (OUT * OUT) + y * OUT + x) * COUT + o];
    }
    __device__ float b(int k, int n) const
    {
        const int j = k / COUT, o = k - j * COUT, ky = py + 2 * (j >> 1), kx = px + 2 * (j & 1);
        return w[(static_cast<long long>(ky * 4 + kx) * CIN + n) * COUT + o];
    }
    __device__ void store(int m, int n, float v, bool) const
    {
        const int per = ny * nx, smp = m / per, p = m - smp * per, yq = p / nx, xq = p - yq * nx;
        da[(static_cast<long long>(smp) * (IN * IN) + (2 * yq + py) * IN + 2 * xq + px) * CPAD + n] = v;
    }
};

// conv2's transposed convolution (da1, 5 channels) as a DIRECT kernel: the 4 x 4 x 5 x 32 kernel (10 KB) sits in constant
// memory, one launch per parity class with the class as template constants, so every weight is an immediate constant
// operand of its FFMA and the only loads are the 128-byte gradient rows: 640 FMAs per thread against 32 16-byte loads.
// (As a functor GEMM with N = 5 it was bound by the shared-memory return path: 3.4 ms per 4096 samples.)
__constant__ float c_w2[16 * 5 * 32];
template <int PY, int PX>
__global__ void __launch_bounds__(128) conv2_dgrad_kernel(const float *__restrict__ dz2, int n, float *__restrict__ da1)
{
    constexpr int NY = PY ? 20 : 21, NX = PX ? 20 : 21;
    const int i = blockIdx.x * 128 + threadIdx.x;
    if (i >= n * NY * NX) return;
    const int smp = i / (NY * NX), p = i - smp * (NY * NX), yq = p / NX, xq = p - yq * NX;
    float acc[5] = {0, 0, 0, 0, 0};
#pragma unroll
    for (int jy = 0; jy < 2; ++jy)
#pragma unroll
        for (int jx = 0; jx < 2; ++jx) {
            const int y = yq - jy, x = xq - jx;
            if (y < 0 || x < 0 || y >= 19 || x >= 19) continue;
            constexpr int kDummy = 0; (void)kDummy;
            const float4 *src = reinterpret_cast<const float4 *>(dz2 + (static_cast<size_t>(smp) * kP2 + y * 19 + x) * 32);
            const int tap = (PY + 2 * jy) * 4 + PX + 2 * jx; // compile-time after unrolling
#pragma unroll
            for (int o4 = 0; o4 < 8; ++o4) {
                const float4 d = __ldg(src + o4);
#pragma unroll
                for (int c = 0; c < 5; ++c) {
                    const float *w = c_w2 + (tap * 5 + c) * 32 + o4 * 4;
                    acc[c] = fmaf(d.x, w[0], fmaf(d.y, w[1], fmaf(d.z, w[2], fmaf(d.w, w[3], acc[c]))));
                }
            }
        }
    float *out = da1 + (static_cast<size_t>(smp) * kP1 + (2 * yq + PY) * 41 + 2 * xq + PX) * 8;
    *reinterpret_cast<float4 *>(out) = make_float4(acc[0], acc[1], acc[2], acc[3]);
    out[4] = acc[4];
}

// ------------------------------------------------------------------------------------------------ small kernels
// column sums: out[c] = sum_rows x[row][pitch * .. + c]  (bias gradients)
__global__ void __launch_bounds__(256) colsum_kernel(const float *__restrict__ x, long long rows, int cols, int pitch, float *__restrict__ out)
{
    // grid.x blocks stride over rows; thread = (row lane, column)
    const int c = threadIdx.x % cols, lanes = 256 / cols, rl = threadIdx.x / cols;
    if (rl >= lanes) return;
    float acc = 0.0f;
    for (long long row = static_cast<long long>(blockIdx.x) * lanes + rl; row < rows; row += static_cast<long long>(gridDim.x) * lanes) acc += x[row * pitch + c];
    atomicAdd(out + c, acc);
}

// output layer (512 -> 10, linear): dh2 = dq W^T masked by h2 > 0 -> dz5;  one thread per (b, k)
__global__ void __launch_bounds__(256) out_dgrad_kernel(const float *__restrict__ dq, const float *__restrict__ w_out, const __nv_bfloat16 *__restrict__ h2,
                                                        int batch, float *__restrict__ dz, __nv_bfloat16 *__restrict__ dz_bf16)
{
    const long long i = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x;
    if (i >= static_cast<long long>(batch) * 512) return;
    const int b = static_cast<int>(i >> 9), k = static_cast<int>(i & 511);
    float acc = 0.0f;
#pragma unroll
    for (int o = 0; o < 10; ++o) acc = fmaf(dq[b * 10 + o], w_out[k * 10 + o], acc);
    if (!(bf(h2, i) > 0.0f)) acc = 0.0f;
    dz[i] = acc;
    dz_bf16[i] = __float2bfloat16_rn(acc);
}

// BatchNorm backward, training mode (batch statistics), for y = gamma * (r - mean) / sigma + beta applied to r = relu(z):
//   pass 1: sums[c] = sum da, sums[cpad + c] = sum da * xhat          (= d beta, d gamma)
//   pass 2: dr = gamma / sigma * (da - dbeta / n - xhat * dgamma / n);  dz = dr where r > 0
template <int kCPad>
__global__ void __launch_bounds__(256) bn_bwd_sums_kernel(const float *__restrict__ da, const __nv_bfloat16 *__restrict__ r, long long rows,
                                                          const float *__restrict__ mean, const float *__restrict__ var, int da_pitch, double *__restrict__ sums)
{
    constexpr int kVecs = kCPad / 8;
    __shared__ float s_part[256][17];
    const int tid = threadIdx.x, grp = tid % kVecs;
    float mu[8], inv[8], sb[8], sg[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { mu[j] = mean[grp * 8 + j]; inv[j] = rsqrtf(var[grp * 8 + j] + kBnEps); sb[j] = 0.0f; sg[j] = 0.0f; }
    const long long stride = static_cast<long long>(gridDim.x) * (256 / kVecs);
    for (long long row = static_cast<long long>(blockIdx.x) * (256 / kVecs) + tid / kVecs; row < rows; row += stride) {
        const long long at = row * kCPad + grp * 8;
        const uint4 rv = __ldg(reinterpret_cast<const uint4 *>(r + at));
        const float4 d0 = __ldg(reinterpret_cast<const float4 *>(da + at)), d1 = __ldg(reinterpret_cast<const float4 *>(da + at + 4));
        const float dv[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
        const uint32_t rw[4] = {rv.x, rv.y, rv.z, rv.w};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float rr = __uint_as_float((j & 1) ? (rw[j >> 1] & 0xFFFF0000u) : (rw[j >> 1] << 16));
            sb[j] += dv[j];
            sg[j] = fmaf(dv[j], (rr - mu[j]) * inv[j], sg[j]);
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) { s_part[tid][j] = sb[j]; s_part[tid][8 + j] = sg[j]; }
    __syncthreads();
    if (tid < kCPad * 2) {
        const int c = tid % kCPad, which = tid / kCPad, g = c / 8, j = c % 8;
        double acc = 0.0;
        for (int t = g; t < 256; t += kVecs) acc += static_cast<double>(s_part[t][which * 8 + j]);
        atomicAdd(sums + which * kCPad + c, acc);
    }
}
template <int kCPad>
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const float *__restrict__ da, const __nv_bfloat16 *__restrict__ r, long long rows,
                                                           const float *__restrict__ gamma, const float *__restrict__ mean, const float *__restrict__ var,
                                                           int c_real, int da_pitch, const double *__restrict__ sums, float *__restrict__ dz,
                                                           float *__restrict__ dgamma, float *__restrict__ dbeta, __nv_bfloat16 *__restrict__ dz_bf16 = nullptr,
                                                           float *__restrict__ dz_colsum = nullptr)
{
    // dz_colsum (nullable, zeroed by the caller): += column sums of dz — the bias gradient of the layer below, formed here
    // instead of a second pass over dz.
    // per channel: dz = k[c] * (da - db[c] - (r - mu[c]) * inv[c] * dg[c]) where r > 0   (the 8 channels of a thread's vector from shared memory)
    __shared__ float s_k[kCPad], s_db[kCPad], s_mu[kCPad], s_inv[kCPad], s_dg[kCPad];
    if (threadIdx.x < kCPad) {
        const int c = threadIdx.x;
        const float inv = rsqrtf(var[c] + kBnEps);
        s_inv[c] = inv;
        s_mu[c] = mean[c];
        s_k[c] = c < c_real ? gamma[c] * inv : 0.0f;
        s_db[c] = static_cast<float>(sums[c] / static_cast<double>(rows));
        s_dg[c] = static_cast<float>(sums[kCPad + c] / static_cast<double>(rows));
    }
    __syncthreads();
    constexpr int kVecs = kCPad / 8; // a thread handles 8 consecutive channels of one row (always the same 8: 256 % kVecs == 0)
    const long long total = rows * kVecs;
    float colsum[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (long long i = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * 256) {
        const int c0 = static_cast<int>(i % kVecs) * 8;
        const long long at = i * 8;
        const uint4 rv = __ldg(reinterpret_cast<const uint4 *>(r + at));
        const float4 d0 = __ldg(reinterpret_cast<const float4 *>(da + at)), d1 = __ldg(reinterpret_cast<const float4 *>(da + at + 4));
        const float dv[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
        const uint32_t rw[4] = {rv.x, rv.y, rv.z, rv.w};
        float out[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float rr = __uint_as_float((j & 1) ? (rw[j >> 1] & 0xFFFF0000u) : (rw[j >> 1] << 16));
            const int c = c0 + j;
            out[j] = rr > 0.0f ? s_k[c] * (dv[j] - s_db[c] - (rr - s_mu[c]) * s_inv[c] * s_dg[c]) : 0.0f;
            colsum[j] += out[j];
        }
        *reinterpret_cast<float4 *>(dz + at) = make_float4(out[0], out[1], out[2], out[3]);
        *reinterpret_cast<float4 *>(dz + at + 4) = make_float4(out[4], out[5], out[6], out[7]);
        if (dz_bf16) {
            uint32_t pk[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const __nv_bfloat162 h = __floats2bfloat162_rn(out[2 * j], out[2 * j + 1]);
                pk[j] = *reinterpret_cast<const uint32_t *>(&h);
            }
            *reinterpret_cast<uint4 *>(dz_bf16 + at) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        }
    }
    if (dz_colsum) {
        __shared__ float s_cs[256][9];
#pragma unroll
        for (int j = 0; j < 8; ++j) s_cs[threadIdx.x][j] = colsum[j];
        __syncthreads();
        if (threadIdx.x < kCPad) {
            const int c = threadIdx.x, g = c / 8, j = c % 8;
            float acc = 0.0f;
            for (int t = g; t < 256; t += kVecs) acc += s_cs[t][j];
            if (c < c_real) atomicAdd(dz_colsum + c, acc);
        }
    }
    if (blockIdx.x == 0 && threadIdx.x < c_real) {
        dbeta[threadIdx.x] = static_cast<float>(sums[threadIdx.x]);
        dgamma[threadIdx.x] = static_cast<float>(sums[kCPad + threadIdx.x]);
    }
}
// per-channel affine map of a BatchNorm on batch statistics: s = gamma / sigma, t = beta - mean * s
__global__ void bn_affine_kernel(const float *gamma, const float *beta, const float *mean, const float *var, int c, float *s, float *t)
{
    const int i = threadIdx.x;
    if (i >= c) return;
    const float sc = gamma[i] * rsqrtf(var[i] + kBnEps);
    s[i] = sc;
    t[i] = beta[i] - mean[i] * sc;
}

// conv1 weight gradient: dW1[ky][kx][c][o] = sum_{s, y, x} ((pix[s][c][3y + ky][3x + kx] - 128) / 128) * dz1[s][y][x][o]
// (c = stacked plane: age * 3 + colour, newest frame first, processor.py:47-49).  A unit = (sample, 4 output rows): the 17
// input rows of the 9 planes are staged in shared memory as bf16 (0..255 is exact; a pixel is converted ONCE per unit, and
// bf16 -> float32 in the inner loop is a shift, not a conversion-pipe instruction), the 4 x 41 x 5 gradients as float32.
// Thread = (output row r of the unit, plane c, ky parity, kx): FOUR taps ky = kyp, kyp + 2, kyp + 4, kyp + 6 against the same
// broadcast gradients, 20 sums in registers; three CTAs per SM overlap one CTA's staging with the others' arithmetic.
// Partial sums are added to dW1 with atomics at the end.  History (per 4096 samples, profiles/fit_launches_r2*.csv): one tap
// per thread, one output row per unit 4.2 ms (staging latency, then the shared-memory return path: every thread fetched the
// pixel's 20 bytes of gradient for 5 FMAs); four taps per thread on uint8 pixels 3.2 ms (one I2F per 5 FMAs on the quarter-rate pipe).
namespace c1w {
constexpr int kRowsOut = 4, kRowsIn = 3 * (kRowsOut - 1) + 8, kPitch = 144, kUnitsPerSample = (41 + kRowsOut - 1) / kRowsOut;
constexpr int kPlane = kRowsIn * kPitch + 16;                 // bf16 elements; + 16: planes fall into different bank groups
constexpr int kImgBytes = 9 * kPlane * 2, kDzBytes = kRowsOut * 41 * 8 * 4, kSmem = kImgBytes + kDzBytes; // 44 352 + 5 248
}
__global__ void __launch_bounds__(576) conv1_wgrad_kernel(const uint8_t *__restrict__ frames, int n, int head, const float *__restrict__ dz1,
                                                          float *__restrict__ dw)
{
    using namespace c1w;
    extern __shared__ __align__(16) uint8_t smem[];
    uint16_t *img = reinterpret_cast<uint16_t *>(smem);
    float *dzs = reinterpret_cast<float *>(smem + kImgBytes);
    const int tid = threadIdx.x, r = tid / 144, rem = tid - r * 144, c = rem >> 4, kyp = (rem >> 3) & 1, kx = rem & 7;
    float acc[4][5];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int o = 0; o < 5; ++o) acc[i][o] = 0.0f;
    const int units = n * kUnitsPerSample;
    for (int u = blockIdx.x; u < units; u += gridDim.x) {
        const int smp = u / kUnitsPerSample, y0 = (u - smp * kUnitsPerSample) * kRowsOut;
        __syncthreads(); // the previous unit's reads are over
        for (int i = tid; i < 9 * kRowsIn * 8; i += 576) { // 16 pixels per step: plane, row, 8 vectors per row
            const int pl = i / (kRowsIn * 8), rm = i - pl * (kRowsIn * 8), row = rm >> 3, v = rm & 7, age = pl / 3, col = pl - age * 3;
            const int yy = 3 * y0 + row;
            uint4 px = make_uint4(0, 0, 0, 0);
            if (yy < 128) px = __ldg(reinterpret_cast<const uint4 *>(frames + ((static_cast<size_t>((head + 3 - age) % 3) * n + smp) * 3 + col) * (128 * 128) + yy * 128 + v * 16));
            const uint32_t w[4] = {px.x, px.y, px.z, px.w};
            uint32_t h[8]; // 16 bf16: the top halves of the floats 0..255
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float f0 = static_cast<float>(w[j] & 0xFFu), f1 = static_cast<float>((w[j] >> 8) & 0xFFu), f2 = static_cast<float>((w[j] >> 16) & 0xFFu),
                            f3 = static_cast<float>(w[j] >> 24);
                h[2 * j] = __byte_perm(__float_as_uint(f0), __float_as_uint(f1), 0x7632);
                h[2 * j + 1] = __byte_perm(__float_as_uint(f2), __float_as_uint(f3), 0x7632);
            }
            uint4 *dst = reinterpret_cast<uint4 *>(img + pl * kPlane + row * kPitch + v * 16);
            dst[0] = make_uint4(h[0], h[1], h[2], h[3]);
            dst[1] = make_uint4(h[4], h[5], h[6], h[7]);
        }
        for (int i = tid; i < kRowsOut * 41 * 2; i += 576) { // gradients: 32 bytes per output pixel; rows past 40 are zeros
            const int rr = (i >> 1) / 41;
            float4 d = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            if (y0 + rr < 41) d = __ldg(reinterpret_cast<const float4 *>(dz1 + (static_cast<size_t>(smp) * kP1 + y0 * 41) * 8 + i * 4));
            *reinterpret_cast<float4 *>(dzs + i * 4) = d;
        }
        __syncthreads();
        const float *dr = dzs + r * 41 * 8;
        const uint16_t *rowp = img + c * kPlane + (3 * r + kyp) * kPitch + kx;
#pragma unroll 2
        for (int x = 0; x < 41; ++x) {
            const float4 d = *reinterpret_cast<const float4 *>(dr + x * 8);
            const float d4 = dr[x * 8 + 4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float v = __uint_as_float(static_cast<uint32_t>(rowp[2 * i * kPitch + 3 * x]) << 16);
                acc[i][0] = fmaf(v, d.x, acc[i][0]); acc[i][1] = fmaf(v, d.y, acc[i][1]); acc[i][2] = fmaf(v, d.z, acc[i][2]);
                acc[i][3] = fmaf(v, d.w, acc[i][3]); acc[i][4] = fmaf(v, d4, acc[i][4]);
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float *out = dw + (((kyp + 2 * i) * 8 + kx) * 9 + c) * 5;
#pragma unroll
        for (int o = 0; o < 5; ++o) atomicAdd(out + o, acc[i][o] * (1.0f / 128.0f));
    }
}
// dW1 -= sum dz1 (the "- 128 / 128" of the normalisation): every tap sees every output pixel once
__global__ void conv1_wgrad_shift_kernel(float *__restrict__ dw, const float *__restrict__ db)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 576 * 5) dw[i] -= db[i % 5];
}

// float32 Keras weights -> the bf16 [N = in][K = out] operand of the data-gradient GEMMs; keras_perm: row n = pixel * 64 + channel
// of the copy is Keras row channel * 64 + pixel (dense1's input order here vs Flatten of (C, H, W))
__global__ void __launch_bounds__(256) cast_weights_kernel(const float *__restrict__ w, __nv_bfloat16 *__restrict__ out, int rows, int cols, int keras_perm)
{
    const size_t total = static_cast<size_t>(rows) * cols;
    for (size_t i = blockIdx.x * 256ull + threadIdx.x; i < total; i += gridDim.x * 256ull) {
        const int n = static_cast<int>(i / cols), k = static_cast<int>(i - static_cast<size_t>(n) * cols);
        const int src = keras_perm ? (n & 63) * 64 + (n >> 6) : n;
        out[i] = __float2bfloat16_rn(w[static_cast<size_t>(src) * cols + k]);
    }
}

// dense1's weight gradient from the raw product G = r3^T dz4 (rows in (pixel, channel) order): the BatchNorm affine map
// a3 = s3[c] r3 + t3[c] and Keras' row order (channel * 64 + pixel): dW[c * 64 + px][o] = s3[c] G[px * 64 + c][o] + t3[c] db[o]
__global__ void __launch_bounds__(256) dense1_wgrad_fix_kernel(const float *__restrict__ g, const float *__restrict__ s3, const float *__restrict__ t3,
                                                               const float *__restrict__ db, float *__restrict__ dw)
{
    for (int i = blockIdx.x * 256 + threadIdx.x; i < 4096 * 512; i += gridDim.x * 256) {
        const int row = i >> 9, o = i & 511, c = row >> 6, px = row & 63;
        dw[i] = fmaf(s3[c], g[(px * 64 + c) * 512 + o], t3[c] * db[o]);
    }
}

// conv2's im2col, materialised once per fit as the MN-major A operand of its weight-gradient GEMM: col[row = (s, y2, x2)][tap][8]
// = r1[s][2 y2 + ky][2 x2 + kx][8]; one 16-byte copy per (row, tap).  (As a tensor map the 64-byte windows would need a
// 64-byte-swizzle MN-major variant of the GEMM; 378 MB of extra traffic per 4096 samples is 0.15 ms.)
__global__ void __launch_bounds__(256) conv2_im2col_kernel(const __nv_bfloat16 *__restrict__ r1, int n, uint4 *__restrict__ col)
{
    const long long total = static_cast<long long>(n) * kP2 * 16;
    for (long long i = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * 256) {
        const int tap = static_cast<int>(i & 15);
        const long long row = i >> 4;
        const int smp = static_cast<int>(row / kP2), p = static_cast<int>(row - static_cast<long long>(smp) * kP2), y = p / 19, x = p - y * 19;
        col[i] = __ldg(reinterpret_cast<const uint4 *>(r1) + static_cast<long long>(smp) * kP1 + (2 * y + (tap >> 2)) * 41 + 2 * x + (tap & 3));
    }
}
// dW2[(tap, c)][o] = s1[c] G[tap * 8 + c][o] + t1[c] db[o]  (Keras [4][4][5][32])
__global__ void __launch_bounds__(256) conv2_wgrad_fix_kernel(const float *__restrict__ g, const float *__restrict__ s1, const float *__restrict__ t1,
                                                              const float *__restrict__ db, float *__restrict__ dw)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= 80 * 32) return;
    const int o = i & 31, tc = i >> 5, tap = tc / 5, c = tc - tap * 5;
    dw[i] = fmaf(s1[c], g[(tap * 8 + c) * 32 + o], t1[c] * db[o]);
}

// conv3's weight gradient from the raw product G = im2col(r2)^T dz3: dW[(tap, c)][o] = s2[c] G[(tap, c)][o] + t2[c] db[o]
__global__ void __launch_bounds__(256) conv3_wgrad_fix_kernel(const float *__restrict__ g, const float *__restrict__ s2, const float *__restrict__ t2,
                                                              const float *__restrict__ db, float *__restrict__ dw)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i < 512 * 64) dw[i] = fmaf(s2[(i >> 6) & 31], g[i], t2[(i >> 6) & 31] * db[i & 63]);
}

// ------------------------------------------------------------------------------------------------ optimizer
struct Segments { ahk_param_segment seg[24]; int count; };

// sum over the trainable elements of (g + 2 l2 p [regularised])^2 -> norm2 (double); sum of l2 * p^2 -> penalty (double)
__global__ void __launch_bounds__(256) grad_norm_kernel(const float *__restrict__ p, const float *__restrict__ g, Segments segs, float l2,
                                                        double *__restrict__ norm2, double *__restrict__ penalty)
{
    double acc = 0.0, pen = 0.0;
    for (int si = 0; si < segs.count; ++si) {
        const ahk_param_segment sg = segs.seg[si];
        if (!(sg.flags & AHK_SEG_TRAINABLE)) continue;
        const bool reg = (sg.flags & AHK_SEG_L2) != 0;
        for (int i = blockIdx.x * 256 + threadIdx.x; i < sg.count; i += gridDim.x * 256) {
            const float w = p[sg.offset + i];
            const float gv = g[sg.offset + i] + (reg ? 2.0f * l2 * w : 0.0f);
            acc += static_cast<double>(gv) * gv;
            if (reg) pen += static_cast<double>(l2) * w * w;
        }
    }
    __shared__ double s_acc[256], s_pen[256];
    s_acc[threadIdx.x] = acc;
    s_pen[threadIdx.x] = pen;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
        if (threadIdx.x < w) { s_acc[threadIdx.x] += s_acc[threadIdx.x + w]; s_pen[threadIdx.x] += s_pen[threadIdx.x + w]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { atomicAdd(norm2, s_acc[0]); atomicAdd(penalty, s_pen[0]); }
}

// keras.optimizers.SGD (Keras 2.x): g *= clipnorm / max(norm, clipnorm); v = momentum * v - lr_t * g; p += v
// keras.optimizers.Adam:            m = b1 m + (1 - b1) g; v = b2 v + (1 - b2) g^2; p -= lr_t * m / (sqrt(v) + eps),
//                                   lr_t = lr * sqrt(1 - b2^t) / (1 - b1^t) passed in by the host
// BatchNorm moving statistics (momentum 0.99): moving = 0.99 * moving + 0.01 * batch
template <bool kAdam>
__global__ void __launch_bounds__(256) step_kernel(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m1, float *__restrict__ m2,
                                                   Segments segs, float lr_t, float momentum, float clipnorm, float l2, float beta1, float beta2, float eps,
                                                   const double *__restrict__ norm2, const float *__restrict__ batch_stats, float bn_momentum,
                                                   float *__restrict__ loss, const double *__restrict__ penalty)
{
    float scale = 1.0f;
    if (clipnorm > 0.0f) {
        const float norm = static_cast<float>(sqrt(*norm2));
        if (norm >= clipnorm) scale = clipnorm / norm; // K.clip_norm
    }
    if (loss && blockIdx.x == 0 && threadIdx.x == 0) *loss += static_cast<float>(*penalty);
    for (int si = 0; si < segs.count; ++si) {
        const ahk_param_segment sg = segs.seg[si];
        if (sg.flags & (AHK_SEG_BN_MEAN | AHK_SEG_BN_VAR)) {
            if (!batch_stats) continue;
            const float *src = batch_stats + ((sg.flags & AHK_SEG_BN_VAR) ? AHK_BN_STATS : 0) + sg.stat_offset;
            for (int i = blockIdx.x * 256 + threadIdx.x; i < sg.count; i += gridDim.x * 256)
                p[sg.offset + i] = bn_momentum * p[sg.offset + i] + (1.0f - bn_momentum) * src[i];
            continue;
        }
        if (!(sg.flags & AHK_SEG_TRAINABLE)) continue;
        const bool reg = (sg.flags & AHK_SEG_L2) != 0;
        for (int i = blockIdx.x * 256 + threadIdx.x; i < sg.count; i += gridDim.x * 256) {
            const int at = sg.offset + i;
            const float w = p[at];
            const float gv = (g[at] + (reg ? 2.0f * l2 * w : 0.0f)) * scale;
            if (kAdam) {
                const float a = beta1 * m1[at] + (1.0f - beta1) * gv, b = beta2 * m2[at] + (1.0f - beta2) * gv * gv;
                m1[at] = a;
                m2[at] = b;
                p[at] = w - lr_t * a / (sqrtf(b) + eps);
            } else {
                const float v = momentum * m1[at] - lr_t * gv;
                m1[at] = v;
                p[at] = w + v;
            }
        }
    }
}

} // namespace train
} // namespace ahk

// ------------------------------------------------------------------------------------------------ C ABI
using namespace ahk::train;

struct ahk_trainer {
    int device = 0, max_batch = 0, sms = 148;
    // gradients with respect to layer inputs / pre-activations (float32) and bf16 copies where a tensor-core GEMM reads them
    float *dz5 = nullptr, *dz4 = nullptr, *da3 = nullptr, *dz3 = nullptr, *da2 = nullptr, *dz2 = nullptr, *da1 = nullptr, *dz1 = nullptr;
    __nv_bfloat16 *dz5h = nullptr, *dz4h = nullptr;
    __nv_bfloat16 *w5h = nullptr, *w4p = nullptr; // bf16 [in][out] copies of dense2 / dense1 (rows in this file's input order) for the data-gradient GEMMs
    CUtensorMap tm_dz5h, tm_dz4h, tm_w5h, tm_w4p;
    float *g4 = nullptr;      // [4096][512] raw dense1 weight gradient (tensor-core path); its head doubles as conv3's [512][64]
    __nv_bfloat16 *dz2h = nullptr, *col2 = nullptr; // bf16 dz2 [n * 361][32] and conv2's im2col [n * 361][128] for its weight-gradient GEMM
    __nv_bfloat16 *dz3h = nullptr, *w3h = nullptr; // bf16 dz3 [n * 64][64] and Keras conv3 kernel [16 * 32][64] for the conv3 GEMMs
    CUtensorMap tm_w3h;
    bool simt_dense = false;  // A/B switch AHK_TRAIN_DENSE=simt: dense data gradients on the functor GEMM
    float *affine = nullptr;  // s1[8] t1[8] s2[32] t2[32] s3[64] t3[64]
    double *sums = nullptr;   // [2][64] BatchNorm backward sums; [128..129] gradient norm^2, penalty
    uint64_t launches = 0;
};

namespace {

int tfail(int code, const char *fmt, const char *detail)
{
    char text[512];
    snprintf(text, sizeof text, fmt, detail);
    ahk::set_error_text(text);
    return code;
}
#define TR_CUDA(expr)                                                                        \
    do {                                                                                     \
        cudaError_t err__ = (expr);                                                          \
        if (err__ != cudaSuccess) return tfail(AHK_ERR_CUDA, #expr ": %s", cudaGetErrorString(err__)); \
    } while (0)

struct DeviceGuard {
    int prev = -1, cur;
    explicit DeviceGuard(int device) : cur(device) { cudaGetDevice(&prev); if (prev != cur) cudaSetDevice(cur); }
    ~DeviceGuard() { if (prev != cur && prev >= 0) cudaSetDevice(prev); }
};

} // namespace

extern "C" {

int ahk_trainer_create(int device, int max_batch, ahk_trainer **out)
{
    if (!out || max_batch <= 0 || max_batch > 16384) return tfail(AHK_ERR_ARG, "%s", "ahk_trainer_create: bad argument (max_batch must be in [1, 16384])");
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return tfail(AHK_ERR_NO_DEVICE, "%s", "no CUDA device: the trainer has no CPU fallback");
    DeviceGuard guard(device);
    ahk_trainer *t = new ahk_trainer();
    t->device = device;
    t->max_batch = max_batch;
    cudaDeviceGetAttribute(&t->sms, cudaDevAttrMultiProcessorCount, device);
    const size_t nb = static_cast<size_t>(max_batch);
    cudaError_t e = cudaMalloc(&t->dz5, nb * 512 * 4);
    if (e == cudaSuccess) e = cudaMalloc(&t->dz4, nb * 512 * 4);
    if (e == cudaSuccess) e = cudaMalloc(&t->da3, nb * 4096 * 4);
    if (e == cudaSuccess) e = cudaMalloc(&t->dz3, nb * 4096 * 4);
    if (e == cudaSuccess) e = cudaMalloc(&t->da2, nb * kP2 * 32 * 4);
    if (e == cudaSuccess) e = cudaMalloc(&t->dz2, nb * kP2 * 32 * 4);
    if (e == cudaSuccess) e = cudaMalloc(&t->da1, nb * kP1 * 8 * 4);
    if (e == cudaSuccess) e = cudaMalloc(&t->dz1, nb * kP1 * 8 * 4);
    if (e == cudaSuccess) e = cudaMalloc(&t->dz5h, nb * 512 * 2);
    if (e == cudaSuccess) e = cudaMalloc(&t->dz4h, nb * 512 * 2);
    if (e == cudaSuccess) e = cudaMalloc(&t->g4, 4096 * 512 * 4);
    if (e == cudaSuccess) e = cudaMalloc(&t->dz3h, nb * 4096 * 2);
    if (e == cudaSuccess) e = cudaMalloc(&t->dz2h, nb * kP2 * 32 * 2);
    if (e == cudaSuccess) e = cudaMalloc(&t->col2, nb * kP2 * 128 * 2);
    if (e == cudaSuccess) e = cudaMalloc(&t->w3h, 512 * 64 * 2);
    if (e == cudaSuccess) e = ahk::tg::matrix_map(&t->tm_w3h, t->w3h, 512, 64, 64, 32);
    if (e == cudaSuccess) e = cudaMalloc(&t->w5h, 512 * 512 * 2);
    if (e == cudaSuccess) e = cudaMalloc(&t->w4p, 4096 * 512 * 2);
    if (e == cudaSuccess) e = ahk::tg::matrix_map(&t->tm_dz5h, t->dz5h, nb, 512, 64, 128);
    if (e == cudaSuccess) e = ahk::tg::matrix_map(&t->tm_dz4h, t->dz4h, nb, 512, 64, 128);
    if (e == cudaSuccess) e = ahk::tg::matrix_map(&t->tm_w5h, t->w5h, 512, 512, 64, 256);
    if (e == cudaSuccess) e = ahk::tg::matrix_map(&t->tm_w4p, t->w4p, 4096, 512, 64, 256);
    if (const char *sw = getenv("AHK_TRAIN_DENSE")) t->simt_dense = strcmp(sw, "simt") == 0;
    if (e == cudaSuccess) e = cudaMalloc(&t->affine, 256 * 4);
    if (e == cudaSuccess) e = cudaMalloc(&t->sums, 160 * 8);
    if (e == cudaSuccess) e = cudaMemset(t->sums, 0, 160 * 8);
    if (e == cudaSuccess) e = cudaMemset(t->da1, 0, nb * kP1 * 8 * 4); // channels 5..7 are never written
    if (e != cudaSuccess) {
        ahk_trainer_destroy(t);
        return tfail(AHK_ERR_CUDA, "ahk_trainer_create: %s", cudaGetErrorString(e));
    }
    *out = t;
    return AHK_OK;
}

void ahk_trainer_destroy(ahk_trainer *t)
{
    if (!t) return;
    DeviceGuard guard(t->device);
    cudaFree(t->dz5); cudaFree(t->dz4); cudaFree(t->da3); cudaFree(t->dz3); cudaFree(t->da2); cudaFree(t->dz2); cudaFree(t->da1); cudaFree(t->dz1);
    cudaFree(t->dz5h); cudaFree(t->dz4h); cudaFree(t->w5h); cudaFree(t->w4p); cudaFree(t->g4); cudaFree(t->dz3h); cudaFree(t->w3h); cudaFree(t->dz2h); cudaFree(t->col2); cudaFree(t->affine); cudaFree(t->sums);
    delete t;
}

int ahk_train_backward(ahk_trainer *t, ahk_policy *p, const ahk_policy_weights *w_dev, const uint8_t *frames_dev, int n, int head,
                       const float *batch_stats_dev, const float *dlogits_dev, const ahk_policy_weights *grads_dev, void *stream)
{
    if (!t || !p || !w_dev || !frames_dev || !batch_stats_dev || !dlogits_dev || !grads_dev) return tfail(AHK_ERR_ARG, "%s", "ahk_train_backward: NULL argument");
    if (n <= 0 || n > t->max_batch || head < 0 || head > 2) return tfail(AHK_ERR_ARG, "%s", "ahk_train_backward: n must be in [1, max_batch], head in [0, 3)");
    ahk_policy_acts acts;
    if (int rc = ahk_policy_get_activations(p, &acts)) return rc;
    if (acts.max_batch < n) return tfail(AHK_ERR_ARG, "%s", "ahk_train_backward: the policy handle's max_batch is smaller than n");
    const float *const *wf = reinterpret_cast<const float *const *>(w_dev);
    float *const *gf = reinterpret_cast<float *const *>(const_cast<ahk_policy_weights *>(grads_dev));
    for (int i = 0; i < 24; ++i)
        if (!wf[i] || !gf[i]) return tfail(AHK_ERR_ARG, "%s", "ahk_train_backward: every weight / gradient pointer must be non-NULL");
    DeviceGuard guard(t->device);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const __nv_bfloat16 *r1 = static_cast<const __nv_bfloat16 *>(acts.act1), *r2 = static_cast<const __nv_bfloat16 *>(acts.act2),
                        *r3 = static_cast<const __nv_bfloat16 *>(acts.act3), *h1 = static_cast<const __nv_bfloat16 *>(acts.h1), *h2 = static_cast<const __nv_bfloat16 *>(acts.h2);
    const float *mean = batch_stats_dev, *var = batch_stats_dev + AHK_BN_STATS;
    // field indices of ahk_policy_weights
    enum { C1W = 0, C1B, G1, B1, M1, V1, C2W, C2B, G2, B2, M2, V2, C3W, C3B, G3, B3, M3, V3, D1W, D1B, D2W, D2B, OW, OB };
    float *s1 = t->affine, *t1 = s1 + 8, *s2 = s1 + 16, *t2 = s1 + 48, *s3 = s1 + 80, *t3 = s1 + 144;
    bn_affine_kernel<<<1, 64, 0, s>>>(wf[G1], wf[B1], mean, var, 5, s1, t1);
    bn_affine_kernel<<<1, 64, 0, s>>>(wf[G2], wf[B2], mean + 8, var + 8, 32, s2, t2);
    bn_affine_kernel<<<1, 64, 0, s>>>(wf[G3], wf[B3], mean + 40, var + 40, 64, s3, t3);
    const int ctas = t->sms * 2;
    auto zero = [&](float *ptr, size_t count) { return cudaMemsetAsync(ptr, 0, count * 4, s); };
    // bias gradients are column sums (atomics): clear them, and the weight gradients that are accumulated with split K
    TR_CUDA(zero(gf[OB], 10)); TR_CUDA(zero(gf[D2B], 512)); TR_CUDA(zero(gf[D1B], 512)); TR_CUDA(zero(gf[C3B], 64)); TR_CUDA(zero(gf[C2B], 32)); TR_CUDA(zero(gf[C1B], 5));
    TR_CUDA(zero(gf[OW], 5120)); TR_CUDA(zero(gf[D2W], 512 * 512)); TR_CUDA(zero(gf[D1W], 4096 * 512)); TR_CUDA(zero(gf[C3W], 512 * 64)); TR_CUDA(zero(gf[C2W], 80 * 32)); TR_CUDA(zero(gf[C1W], 576 * 5));

    // ---- output layer: dW6 = h2^T dq, db6 = sum dq, dz5 = (dq W6^T) . [h2 > 0]
    {
        DenseWgrad f{h2, dlogits_dev, gf[OW], nullptr, nullptr, 512, 10, 0};
        TR_CUDA((run_gemm<DenseWgrad>(f, 512, 10, n, ctas, s)));
        colsum_kernel<<<min(ctas, (n + 24) / 25), 256, 0, s>>>(dlogits_dev, n, 10, 10, gf[OB]);
        out_dgrad_kernel<<<(n * 512 + 255) / 256, 256, 0, s>>>(dlogits_dev, wf[OW], h2, n, t->dz5, t->dz5h);
    }
    // ---- dense2: dW5 = h1^T dz5, db5, dz4 = (dz5 W5^T) . [h1 > 0]
    {
        if (t->simt_dense) {
            DenseWgrad f{h1, t->dz5, gf[D2W], nullptr, nullptr, 512, 512, 0};
            TR_CUDA((run_gemm<DenseWgrad>(f, 512, 512, n, ctas, s)));
        } else { // tcgen05, both operands MN-major (the reduction runs over the batch): dW5 [in][out] += h1^T dz5, split K, atomics
            namespace tg = ahk::tg;
            CUtensorMap ta, tb;
            TR_CUDA(tg::matrix_map(&ta, h1, n, 512, 64, 64));
            TR_CUDA(tg::matrix_map(&tb, t->dz5h, n, 512, 64, 64));
            tg::Args g{};
            const int chunks = (n + 63) / 64;
            g.k_splits = chunks < 16 ? chunks : 16; g.k_chunks = (chunks + g.k_splits - 1) / g.k_splits;
            g.tiles_m = 4; g.blocks_n = 2; g.m_total = 512; g.n_valid = 256; g.out_f32 = gf[D2W]; g.out_pitch = 512; g.scale = 1.0f;
            TR_CUDA((tg::launch<256, 128, tg::kLinearMn, true, tg::kAtomicF32, 4>(g, ta, tb, t->sms, s)));
        }
        colsum_kernel<<<min(ctas, n), 256, 0, s>>>(t->dz5, n, 256, 512, gf[D2B]);
        colsum_kernel<<<min(ctas, n), 256, 0, s>>>(t->dz5 + 256, n, 256, 512, gf[D2B] + 256);
        if (t->simt_dense) {
            DenseDgrad d{t->dz5, wf[D2W], t->dz4, h1, t->dz4h, 512, 512, 0};
            TR_CUDA((run_gemm<DenseDgrad>(d, n, 512, 512, 1, s)));
        } else { // tcgen05: dz5 (bf16) [n][512 out] x W5 (bf16) [512 in][512 out]^T, ReLU mask of h1 in the epilogue
            namespace tg = ahk::tg;
            cast_weights_kernel<<<t->sms * 2, 256, 0, s>>>(wf[D2W], t->w5h, 512, 512, 0);
            tg::Args g{};
            g.tiles_m = (n + 127) / 128; g.blocks_n = 2; g.k_chunks = 8; g.k_splits = 1; g.m_total = n; g.n_valid = 256;
            g.out_bf16 = t->dz4h; g.out_f32 = t->dz4; g.mask_bf16 = h1; g.out_pitch = 512; g.scale = 1.0f;
            TR_CUDA((tg::launch<256, 128, tg::kLinear, false, tg::kMaskBoth, 4>(g, t->tm_dz5h, t->tm_w5h, t->sms, s)));
        }
    }
    // ---- dense1: dW4 = bn3(r3)^T dz4 (Keras row order channel * 64 + pixel), db4, da3 = dz4 W4^T
    {
        colsum_kernel<<<min(ctas, n), 256, 0, s>>>(t->dz4, n, 256, 512, gf[D1B]);
        colsum_kernel<<<min(ctas, n), 256, 0, s>>>(t->dz4 + 256, n, 256, 512, gf[D1B] + 256);
        if (t->simt_dense) {
            DenseWgrad f{r3, t->dz4, gf[D1W], s3, t3, 4096, 512, 1};
            TR_CUDA((run_gemm<DenseWgrad>(f, 4096, 512, n, ctas, s)));
        } else { // tcgen05: G = r3^T dz4 on the raw activations, then the BatchNorm affine map and Keras' row order in one pass
            namespace tg = ahk::tg;
            CUtensorMap ta, tb;
            TR_CUDA(tg::matrix_map(&ta, r3, n, 4096, 64, 64));
            TR_CUDA(tg::matrix_map(&tb, t->dz4h, n, 512, 64, 64));
            TR_CUDA(zero(t->g4, 4096 * 512));
            tg::Args g{};
            const int chunks = (n + 63) / 64;
            g.k_splits = chunks < 4 ? chunks : 4; g.k_chunks = (chunks + g.k_splits - 1) / g.k_splits;
            g.tiles_m = 32; g.blocks_n = 2; g.m_total = 4096; g.n_valid = 256; g.out_f32 = t->g4; g.out_pitch = 512; g.scale = 1.0f;
            TR_CUDA((tg::launch<256, 128, tg::kLinearMn, true, tg::kAtomicF32, 4>(g, ta, tb, t->sms, s)));
            dense1_wgrad_fix_kernel<<<t->sms * 4, 256, 0, s>>>(t->g4, s3, t3, gf[D1B], gf[D1W]);
        }
        if (t->simt_dense) {
            DenseDgrad d{t->dz4, wf[D1W], t->da3, nullptr, nullptr, 512, 4096, 1};
            TR_CUDA((run_gemm<DenseDgrad>(d, n, 4096, 512, 1, s)));
        } else { // tcgen05: dz4 (bf16) [n][512] x W4 (bf16, rows in (pixel, channel) order) [4096][512]^T -> da3 float32
            namespace tg = ahk::tg;
            cast_weights_kernel<<<t->sms * 4, 256, 0, s>>>(wf[D1W], t->w4p, 4096, 512, 1);
            tg::Args g{};
            g.tiles_m = (n + 127) / 128; g.blocks_n = 16; g.k_chunks = 8; g.k_splits = 1; g.m_total = n; g.n_valid = 256;
            g.out_f32 = t->da3; g.out_pitch = 4096; g.scale = 1.0f;
            TR_CUDA((tg::launch<256, 128, tg::kLinear, false, tg::kStoreF32, 4>(g, t->tm_dz4h, t->tm_w4p, t->sms, s)));
        }
    }
    // ---- BatchNorm 3 + ReLU backward: da3 [n * 64 rows][64] -> dz3, d gamma3, d beta3
    {
        const long long rows = static_cast<long long>(n) * 64;
        bn_bwd_sums_kernel<64><<<ctas, 256, 0, s>>>(t->da3, r3, rows, mean + 40, var + 40, 64, t->sums);
        bn_bwd_apply_kernel<64><<<ctas * 2, 256, 0, s>>>(t->da3, r3, rows, wf[G3], mean + 40, var + 40, 64, 64, t->sums, t->dz3, gf[G3], gf[B3], t->dz3h, gf[C3B]);
        TR_CUDA(cudaMemsetAsync(t->sums, 0, 128 * 8, s));
    }
    // ---- conv3: dW3 = im2col(bn2(r2))^T dz3, db3, da2 = transposed conv (4 parity classes)
    {
        if (t->simt_dense) {
            using W3 = ConvWgrad<32, 32, 19, 8, 64>;
            W3 f{r2, t->dz3, gf[C3W], s2, t2};
            TR_CUDA((run_gemm<W3>(f, 512, 64, n * 64, ctas, s)));
            for (int cls = 0; cls < 4; ++cls) {
                const int py = cls >> 1, px = cls & 1, ny = py ? 9 : 10, nx = px ? 9 : 10;
                using D3 = ConvDgrad<32, 32, 64, 19, 8>;
                D3 d{t->dz3, wf[C3W], t->da2, py, px, ny, nx};
                TR_CUDA((run_gemm<D3, 128, 32, 8, 2>(d, n * ny * nx, 32, 4 * 64, 1, s)));
            }
        } else { // both on tcgen05 (ahk_gemm.cuh: kConv3Wgrad, kConv3Dgrad): the tensor maps do the im2col / the transposed gather
            namespace tg = ahk::tg;
            CUtensorMap ta, tb, td;
            { // act2 [sample][19 x 19][32] as conv3's im2col, one sample per box
                const unsigned long long dims[4] = {static_cast<unsigned long long>(kP2) * 32, 8, 8, static_cast<unsigned long long>(n)}, strides[3] = {2 * 32 * 2, 2 * 19 * 32 * 2, static_cast<unsigned long long>(kP2) * 32 * 2};
                const unsigned box[4] = {64, 8, 8, 1};
                TR_CUDA(tg::encode_bf16_map(&ta, r2, 4, dims, strides, box, 128));
            }
            TR_CUDA(tg::matrix_map(&tb, t->dz3h, static_cast<unsigned long long>(n) * 64, 64, 64, 64));
            { // dz3 [sample][8][8][64]: channel, x, y, sample
                const unsigned long long dims[4] = {64, 8, 8, static_cast<unsigned long long>(n)}, strides[3] = {128, 1024, 8192};
                const unsigned box[4] = {64, 10, 10, 1};
                TR_CUDA(tg::encode_bf16_map(&td, t->dz3h, 4, dims, strides, box, 128));
            }
            TR_CUDA(zero(t->g4, 512 * 64));
            tg::Args g{};
            g.k_splits = n < 37 ? n : 37; g.k_chunks = (n + g.k_splits - 1) / g.k_splits;
            g.tiles_m = 4; g.blocks_n = 1; g.m_total = 512; g.n_valid = 64; g.out_f32 = t->g4; g.out_pitch = 64; g.scale = 1.0f;
            TR_CUDA((tg::launch<64, 128, tg::kConv3Wgrad, true, tg::kAtomicF32, 6>(g, ta, tb, t->sms, s)));
            conv3_wgrad_fix_kernel<<<(512 * 64 + 255) / 256, 256, 0, s>>>(t->g4, s2, t2, gf[C3B], gf[C3W]);
            cast_weights_kernel<<<32, 256, 0, s>>>(wf[C3W], t->w3h, 512, 64, 0);
            tg::Args d{};
            d.k_splits = 1; d.k_chunks = 4; d.tiles_m = n; d.blocks_n = 1; d.m_total = n; d.n_valid = 32; d.out_f32 = t->da2; d.out_pitch = 32; d.scale = 1.0f;
            for (int cls = 0; cls < 4; ++cls) {
                d.py = cls >> 1; d.px = cls & 1;
                TR_CUDA((tg::launch<32, 128, tg::kConv3Dgrad, false, tg::kStoreF32, 8>(d, td, t->tm_w3h, t->sms, s)));
            }
        }
    }
    // ---- BatchNorm 2 + ReLU backward
    {
        const long long rows = static_cast<long long>(n) * kP2;
        bn_bwd_sums_kernel<32><<<ctas, 256, 0, s>>>(t->da2, r2, rows, mean + 8, var + 8, 32, t->sums);
        bn_bwd_apply_kernel<32><<<ctas * 2, 256, 0, s>>>(t->da2, r2, rows, wf[G2], mean + 8, var + 8, 32, 32, t->sums, t->dz2, gf[G2], gf[B2], t->dz2h, gf[C2B]);
        TR_CUDA(cudaMemsetAsync(t->sums, 0, 128 * 8, s));
    }
    // ---- conv2: dW2 = im2col(bn1(r1))^T dz2 (80 x 32), db2, da1 (5 channels of 8)
    {
        if (t->simt_dense) {
            using W2 = ConvWgrad<5, 8, 41, 19, 32>;
            W2 f{r1, t->dz2, gf[C2W], s1, t1};
            TR_CUDA((run_gemm<W2, 80, 32, 5, 2>(f, 80, 32, n * kP2, ctas, s))); // the whole 80 x 32 gradient is one tile: split K over the CTAs
        } else { // tcgen05: G[128][32] = col^T dz2 with both operands MN-major, split K over the SMs, then the BatchNorm affine map
            namespace tg = ahk::tg;
            const unsigned long long rows = static_cast<unsigned long long>(n) * kP2;
            conv2_im2col_kernel<<<t->sms * 8, 256, 0, s>>>(r1, n, reinterpret_cast<uint4 *>(t->col2));
            CUtensorMap ta, tb;
            TR_CUDA(tg::matrix_map(&ta, t->col2, rows, 128, 64, 64));
            TR_CUDA(tg::matrix_map(&tb, t->dz2h, rows, 32, 64, 64)); // box wider than the 32 columns: the other 32 are zero fill
            TR_CUDA(zero(t->g4, 128 * 32));
            tg::Args g{};
            const int chunks = static_cast<int>((rows + 63) / 64);
            g.k_splits = chunks < t->sms ? chunks : t->sms; g.k_chunks = (chunks + g.k_splits - 1) / g.k_splits;
            g.tiles_m = 1; g.blocks_n = 1; g.m_total = 128; g.n_valid = 32; g.out_f32 = t->g4; g.out_pitch = 32; g.scale = 1.0f;
            TR_CUDA((tg::launch<64, 128, tg::kLinearMn, true, tg::kAtomicF32, 6>(g, ta, tb, t->sms, s)));
            conv2_wgrad_fix_kernel<<<(80 * 32 + 255) / 256, 256, 0, s>>>(t->g4, s1, t1, gf[C2B], gf[C2W]);
        }
        TR_CUDA(cudaMemcpyToSymbolAsync(c_w2, wf[C2W], sizeof(float) * 16 * 5 * 32, 0, cudaMemcpyDeviceToDevice, s));
        conv2_dgrad_kernel<0, 0><<<(n * 21 * 21 + 127) / 128, 128, 0, s>>>(t->dz2, n, t->da1);
        conv2_dgrad_kernel<0, 1><<<(n * 21 * 20 + 127) / 128, 128, 0, s>>>(t->dz2, n, t->da1);
        conv2_dgrad_kernel<1, 0><<<(n * 20 * 21 + 127) / 128, 128, 0, s>>>(t->dz2, n, t->da1);
        conv2_dgrad_kernel<1, 1><<<(n * 20 * 20 + 127) / 128, 128, 0, s>>>(t->dz2, n, t->da1);
    }
    // ---- BatchNorm 1 + ReLU backward, conv1 weight gradient
    {
        const long long rows = static_cast<long long>(n) * kP1;
        bn_bwd_sums_kernel<8><<<ctas, 256, 0, s>>>(t->da1, r1, rows, mean, var, 8, t->sums);
        bn_bwd_apply_kernel<8><<<ctas * 2, 256, 0, s>>>(t->da1, r1, rows, wf[G1], mean, var, 5, 8, t->sums, t->dz1, gf[G1], gf[B1], nullptr, gf[C1B]);
        TR_CUDA(cudaMemsetAsync(t->sums, 0, 128 * 8, s));
        static bool prepared = false;
        if (!prepared) {
            TR_CUDA(cudaFuncSetAttribute(conv1_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, c1w::kSmem));
            prepared = true;
        }
        conv1_wgrad_kernel<<<min(t->sms * 3, n * c1w::kUnitsPerSample), 576, c1w::kSmem, s>>>(frames_dev, n, head, t->dz1, gf[C1W]);
        conv1_wgrad_shift_kernel<<<(576 * 5 + 255) / 256, 256, 0, s>>>(gf[C1W], gf[C1B]);
    }
    TR_CUDA(cudaGetLastError());
    t->launches += 45;
    return AHK_OK;
}

static int optimizer_step(bool adam, ahk_trainer *t, float *params_dev, const float *grads_dev, float *m1_dev, float *m2_dev, const ahk_param_segment *segs,
                          int n_segs, float lr_t, float momentum, float clipnorm, float l2, float beta1, float beta2, float eps, const float *batch_stats_dev,
                          float bn_momentum, float *loss_dev, void *stream)
{
    if (!t || !params_dev || !grads_dev || !m1_dev || !segs || (adam && !m2_dev)) return tfail(AHK_ERR_ARG, "%s", "ahk_train_*_step: NULL argument");
    if (n_segs <= 0 || n_segs > 24) return tfail(AHK_ERR_ARG, "%s", "ahk_train_*_step: 1 to 24 segments");
    DeviceGuard guard(t->device);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    Segments sg{};
    sg.count = n_segs;
    for (int i = 0; i < n_segs; ++i) sg.seg[i] = segs[i];
    double *norm2 = t->sums + 128, *penalty = t->sums + 129;
    TR_CUDA(cudaMemsetAsync(norm2, 0, 16, s));
    grad_norm_kernel<<<t->sms * 4, 256, 0, s>>>(params_dev, grads_dev, sg, l2, norm2, penalty);
    if (adam) step_kernel<true><<<t->sms * 4, 256, 0, s>>>(params_dev, grads_dev, m1_dev, m2_dev, sg, lr_t, momentum, clipnorm, l2, beta1, beta2, eps, norm2, batch_stats_dev, bn_momentum, loss_dev, penalty);
    else step_kernel<false><<<t->sms * 4, 256, 0, s>>>(params_dev, grads_dev, m1_dev, m2_dev, sg, lr_t, momentum, clipnorm, l2, beta1, beta2, eps, norm2, batch_stats_dev, bn_momentum, loss_dev, penalty);
    TR_CUDA(cudaGetLastError());
    t->launches += 2;
    return AHK_OK;
}

int ahk_train_sgd_step(ahk_trainer *t, float *params_dev, const float *grads_dev, float *velocity_dev, const ahk_param_segment *segs, int n_segs, float lr_t,
                       float momentum, float clipnorm, float l2, const float *batch_stats_dev, float bn_momentum, float *loss_dev, void *stream)
{
    return optimizer_step(false, t, params_dev, grads_dev, velocity_dev, nullptr, segs, n_segs, lr_t, momentum, clipnorm, l2, 0.0f, 0.0f, 0.0f, batch_stats_dev, bn_momentum,
                          loss_dev, stream);
}

int ahk_train_adam_step(ahk_trainer *t, float *params_dev, const float *grads_dev, float *m_dev, float *v_dev, const ahk_param_segment *segs, int n_segs, float lr_t,
                        float beta1, float beta2, float eps, float l2, const float *batch_stats_dev, float bn_momentum, float *loss_dev, void *stream)
{
    return optimizer_step(true, t, params_dev, grads_dev, m_dev, v_dev, segs, n_segs, lr_t, 0.0f, 0.0f, l2, beta1, beta2, eps, batch_stats_dev, bn_momentum, loss_dev, stream);
}

uint64_t ahk_trainer_launches(const ahk_trainer *t) { return t ? t->launches : 0; }

int ahk_trainer_copy_buffer(ahk_trainer *t, int which, float *dst_dev, size_t count, void *stream)
{
    if (!t || !dst_dev || which < 0 || which > 7) return tfail(AHK_ERR_ARG, "%s", "ahk_trainer_copy_buffer: bad argument");
    const float *src[8] = {t->dz5, t->dz4, t->da3, t->dz3, t->da2, t->dz2, t->da1, t->dz1};
    DeviceGuard guard(t->device);
    TR_CUDA(cudaMemcpyAsync(dst_dev, src[which], count * 4, cudaMemcpyDeviceToDevice, static_cast<cudaStream_t>(stream)));
    return AHK_OK;
}

} // extern "C"
