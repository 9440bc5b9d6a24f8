// ahk_policy.cu — K3: forward pass of the reference's 'rgb' CNN policy on the frame ring, on 5th-gen tensor cores.
//
// Network (ai/supervised/keras/models.py:20-65, Keras defaults: valid padding, bias, BatchNorm eps 1e-3 AFTER the ReLU):
//   (9,128,128) -> conv1 5xk8 s3 + ReLU + BN -> (5,41,41) -> conv2 32xk4 s2 + ReLU + BN -> (32,19,19)
//   -> conv3 64xk4 s2 + ReLU + BN -> (64,8,8) -> flatten (C,H,W) -> dense 512 ReLU -> dense 512 ReLU -> dense 10.
// Used by the reference at batch 1 per env step (ai/reinforcement/rl_agents.py:78-98, ai/supervised/keras/test.py:42-43);
// here at batch = number of envs, reading the uint8 ring the rasteriser just wrote (no float observation is built).
//
// Every layer is a GEMM  D[M x N] = act(A[M x K] * B[K x N] + bias)  issued as tcgen05.mma (kind::f16, bf16 operands,
// fp32 accumulators in TMEM), one 128-row tile per CTA:
//   * B (weights) is pre-packed on the host into the canonical no-swizzle K-major core-matrix layout, one contiguous
//     blob per 64-wide K chunk, and lands in shared memory with ONE cp.async.bulk per chunk (mbarrier complete_tx);
//   * A is gathered by the CTA's 128 threads (thread = tile row) straight into the same canonical layout — im2col on
//     the fly for the convolutions; for conv1 from a bf16 copy of the needed input rows staged in shared memory,
//     so each uint8 is converted once;
//   * two smem stages; tcgen05.commit -> mbarrier recycles a stage when the MMAs that read it have retired;
//   * epilogue: tcgen05.ld 32x32b (thread = TMEM lane = tile row), bias, ReLU, bf16, 16-byte stores.
// Folded on the host, exactly: the input normalisation (x-128)/128 into conv1, and each BatchNorm (an affine map applied
// after the ReLU) into the NEXT layer's weights and bias.
#include "../../include/ahk.h"
#include "ahk_kernels.cuh"
#include "ahk_tc.cuh"
#include "ahk_gemm.cuh"

#include <cuda.h>
#include <cudaTypedefs.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

namespace ahk {
namespace policy {

constexpr int kBM = 128;     // rows per tile = UMMA M
constexpr int kBK = 64;      // K per pipeline stage = 8 core matrices of 8 bf16
constexpr int kThreads = 128;

// ---- geometry of the network (models.py:20-65)
constexpr int kC1 = 41, kP1 = kC1 * kC1; // conv1 output side / pixels
constexpr int kC2 = 19, kP2 = kC2 * kC2;
constexpr int kC3 = 8, kP3 = kC3 * kC3;
constexpr int kAct1C = 8;                // conv1 channels 5 padded to 8 (one 16-byte K core per tap for conv2)

enum Layer { kConv1 = 0, kConv2, kConv3, kDense1, kDense2, kOut,
             kG1, kG2, kG3, kG4, kGDense1 }; // the 'gray-diff' network (models.py:67-116); its 'out' layer is kOut

// ---- geometry of the 'gray-diff' network: (1,128,128) -> 16 k8 s4 (31x31) -> 24 k4 s2 (14x14) -> 32 k4 s1 (11x11)
//      -> 48 k4 s1 (8x8) -> flatten 3072 -> dense 512 -> 10
constexpr int kGS1 = 31, kGP1 = kGS1 * kGS1, kGS2 = 14, kGP2 = kGS2 * kGS2, kGS3 = 11, kGP3 = kGS3 * kGS3, kGS4 = 8, kGP4 = kGS4 * kGS4;

template <int L> struct Traits;
template <> struct Traits<kConv2> { static constexpr int N = 32, K = 128, NSplit = 1; };  // K = 16 taps x 8 (5 real)
template <> struct Traits<kConv3> { static constexpr int N = 64, K = 512, NSplit = 1; };  // K = 16 taps x 32
template <> struct Traits<kDense1> { static constexpr int N = 256, K = 4096, NSplit = 2; }; // 512 outputs in two halves
template <> struct Traits<kDense2> { static constexpr int N = 256, K = 512, NSplit = 2; };
template <> struct Traits<kOut> { static constexpr int N = 16, K = 512, NSplit = 1; };     // 10 real outputs
template <> struct Traits<kG1> { static constexpr int N = 16, K = 64, NSplit = 1; };       // K = 8 rows x 8 pixels of the gray image
template <> struct Traits<kG2> { static constexpr int N = 32, K = 256, NSplit = 1; };      // 24 real outputs; K = 16 taps x 16
template <> struct Traits<kG3> { static constexpr int N = 32, K = 512, NSplit = 1; };      // K = 16 taps x 32 (24 real)
template <> struct Traits<kG4> { static constexpr int N = 48, K = 512, NSplit = 1; };      // K = 16 taps x 32
template <> struct Traits<kGDense1> { static constexpr int N = 256, K = 3072, NSplit = 2; }; // K = 64 pixels x 48

struct Args {
    const uint8_t *frames;      // conv1: the ring [3][n][3][128][128]
    const __nv_bfloat16 *in;    // other layers: input activations
    const uint4 *weights;       // packed B blobs: [NSplit][K/64] chunks of N*128 bytes
    const float *bias;          // [NSplit * N]
    __nv_bfloat16 *out;         // output activations (bf16)
    float *logits;              // kOut: float [n][10]
    uint8_t *action;            // kOut: argmax
    int n;                      // samples
    int m_total;                // valid rows of the flattened GEMM (conv1: per sample)
    int head;
    int depth;                  // ring slots (3, or a deep ring: ahk_stack)
    const uint8_t *warm;        // deep ring: warm-up frames of each env's latest reset, else nullptr (stack_frame, ahk_kernels.cuh)
    const uint32_t *meta;       // deep ring: state word 15 per env
    int pdl;                    // programmatic dependent launch of the NEXT layer's kernel: 0 off, 1 trigger after the prologue, 2 trigger at the end
    int no_swap;                // conv_tm_kernel A/B (AHK_CONV2_SWAP=0): every lane fetches its two 16-byte halves in the same order
};

using namespace ahk::tc; // inline-PTX wrappers: ahk_tc.cuh

// ------------------------------------------------------------------------------------------------ A gathers
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi)
{ // exact for the integers 0..255; general values: truncation is never used (see cvt below for activations)
    return __byte_perm(__float_as_uint(lo), __float_as_uint(hi), 0x7632);
}

// 4 uint8 -> 4 bf16 (exact): 0x4B0000xx is 8388608 + xx as a float
__device__ __forceinline__ uint2 u8x4_to_bf16x4(uint32_t w)
{
    const float f0 = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7440)) - 8388608.0f;
    const float f1 = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7441)) - 8388608.0f;
    const float f2 = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7442)) - 8388608.0f;
    const float f3 = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7443)) - 8388608.0f;
    return make_uint2(pack_bf16(f0, f1), pack_bf16(f2, f3));
}

// A row of the GEMM as 16-byte K cores, gathered straight from the previous layer's activations (im2col on the fly):
// r0 = element offset of the row's first input element, kc = K core index (8 elements each)
template <int L>
__device__ __forceinline__ uint4 gather(const Args &a, bool row_ok, int r0, int kc)
{
    if (!row_ok) return make_uint4(0, 0, 0, 0);
    if (L == kConv2) { // kc = tap ; r0 = element offset of act1[s][2Y*41 + 2X][0]
        const int ky = kc >> 2, kx = kc & 3;
        return __ldg(reinterpret_cast<const uint4 *>(a.in + r0 + (ky * kC1 + kx) * kAct1C));
    }
    if (L == kConv3) { // kc = tap * 4 + channel block ; r0 = element offset of act2[s][2Y*19 + 2X][0]
        const int tap = kc >> 2, cb = kc & 3, ky = tap >> 2, kx = tap & 3;
        return __ldg(reinterpret_cast<const uint4 *>(a.in + r0 + (ky * kC2 + kx) * 32 + cb * 8));
    }
    if (L == kG1) { // kc = ky ; r0 = element offset of gray[s][4Y][4X]: 8 pixels, 8-byte aligned only
        const uint2 lo = __ldg(reinterpret_cast<const uint2 *>(a.in + r0 + kc * kDim)), hi = __ldg(reinterpret_cast<const uint2 *>(a.in + r0 + kc * kDim + 4));
        return make_uint4(lo.x, lo.y, hi.x, hi.y);
    }
    if (L == kG2) { // kc = tap * 2 + half ; r0 = element offset of g1[s][2Y*31 + 2X][0]
        const int tap = kc >> 1, ky = tap >> 2, kx = tap & 3;
        return __ldg(reinterpret_cast<const uint4 *>(a.in + r0 + (ky * kGS1 + kx) * 16 + (kc & 1) * 8));
    }
    if (L == kG3 || L == kG4) { // kc = tap * 4 + channel block ; stride 1
        const int tap = kc >> 2, ky = tap >> 2, kx = tap & 3;
        return __ldg(reinterpret_cast<const uint4 *>(a.in + r0 + (ky * (L == kG3 ? kGS2 : kGS3) + kx) * 32 + (kc & 3) * 8));
    }
    // dense: r0 = element offset of in[row][0]
    return __ldg(reinterpret_cast<const uint4 *>(a.in + r0 + kc * 8));
}

// ------------------------------------------------------------------------------------------------ the GEMM tile kernel
// conv2, conv3 and the dense layers.  Persistent over 128-row tiles (barriers, TMEM and bias set up once per CTA).
// Gather mapping: one warp instruction fetches 8 rows x 4 consecutive K cores (64 contiguous bytes per row; conv2: four
// adjacent pixels of one kernel row) — 8 lines per load instead of the 32 of a row-per-thread mapping — and stores them
// to 8 distinct rows-within-core-matrix, i.e. across all banks, with the core matrices still 128-byte aligned
// (an odd K-core pitch that de-conflicted a row-per-8-lanes mapping made the MMA's own A reads straddle lines: slower).
constexpr int kALbo = 16 * 128;       // bytes between K cores of the A stage (= LBO of its descriptor)
constexpr int kAStage = 8 * kALbo;

template <int L>
__global__ void __launch_bounds__(kThreads) layer_kernel(Args a)
{
    using T = Traits<L>;
    constexpr int N = T::N, NCH = T::K / kBK;
    constexpr int kBBytes = N * kBK * 2;
    constexpr uint32_t kTmemCols = N <= 32 ? 32 : (N <= 64 ? 64 : (N <= 128 ? 128 : 256));

    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *s_a = smem;                      // [2][kAStage]
    uint8_t *s_b = smem + 2 * kAStage;        // [2][kBBytes]
    __shared__ __align__(8) uint64_t s_full[2], s_free[2], s_done;
    __shared__ uint32_t s_tmem;
    __shared__ float s_bias[N];

    const int tid = threadIdx.x, warp = tid >> 5;
    const int split = blockIdx.y;             // which N half (dense layers)

    if (tid == 0) {
        mbar_init(&s_full[0], 1); mbar_init(&s_full[1], 1);
        mbar_init(&s_free[0], 1); mbar_init(&s_free[1], 1);
        mbar_init(&s_done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < N) s_bias[tid] = a.bias[split * N + tid];
    if (N > kThreads) for (int i = tid + kThreads; i < N; i += kThreads) s_bias[i] = a.bias[split * N + i];
    if (warp == 0) tmem_alloc(&s_tmem, kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;
    pdl_wait(); // programmatic dependent launch (ahk_tc.cuh): the previous layer's output is read from here on

    const uint4 *wsrc = a.weights + static_cast<size_t>(split) * NCH * (kBBytes / 16);
    constexpr uint32_t kIdesc = instr_desc(N);
    const int n_tiles = (a.m_total + kBM - 1) / kBM;
    // gather: this thread's K core of every chunk, and its 8 rows (i = 0..7): row group 2 i + (warp >> 1), row tid & 7
    const int kc_lane = (warp & 1) * 4 + ((tid & 31) >> 3), g = (warp >> 1) * 8 + (tid & 7);
    uint32_t chg = 0, it = 0;                  // chunks / tiles processed so far by this CTA (barrier phases)

    // The gather runs one chunk ahead of its use (across tile boundaries too): the next chunk's loads are issued right
    // after this chunk's registers have been stored and are in flight during the barrier, the MMA issue and the epilogue.
    int r0[8];
    unsigned ok = 0u;
    auto rows_of = [&](int tile) { // per-row constants of the gather for `tile`
        ok = 0u;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int row = tile * kBM + i * 16 + g;
            r0[i] = 0;
            if (row < a.m_total) {
                ok |= 1u << i;
                if (L == kConv2) {
                    const int s = row / kP2, p = row - s * kP2, y = p / kC2, x = p - y * kC2;
                    r0[i] = (s * kP1 + 2 * y * kC1 + 2 * x) * kAct1C;
                } else if (L == kConv3) {
                    const int s = row / kP3, p = row - s * kP3, y = p / kC3, x = p - y * kC3;
                    r0[i] = (s * kP2 + 2 * y * kC2 + 2 * x) * 32;
                } else if (L == kG1) {
                    const int s = row / kGP1, p = row - s * kGP1, y = p / kGS1, x = p - y * kGS1;
                    r0[i] = s * (kDim * kDim) + 4 * y * kDim + 4 * x;
                } else if (L == kG2) {
                    const int s = row / kGP2, p = row - s * kGP2, y = p / kGS2, x = p - y * kGS2;
                    r0[i] = (s * kGP1 + 2 * y * kGS1 + 2 * x) * 16;
                } else if (L == kG3) {
                    const int s = row / kGP3, p = row - s * kGP3, y = p / kGS3, x = p - y * kGS3;
                    r0[i] = (s * kGP2 + y * kGS2 + x) * 32;
                } else if (L == kG4) {
                    const int s = row / kGP4, p = row - s * kGP4, y = p / kGS4, x = p - y * kGS4;
                    r0[i] = (s * kGP3 + y * kGS3 + x) * 32;
                } else {
                    r0[i] = row * T::K;
                }
            }
        }
    };
    uint4 v[8];
    int pf_tile = blockIdx.x, pf_ch = 0; // the chunk held in v / the next one to fetch
    auto prefetch = [&]() {
        if (pf_tile >= n_tiles) return;
        if (pf_ch == 0) rows_of(pf_tile);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = gather<L>(a, (ok >> i) & 1u, r0[i], pf_ch * 8 + kc_lane);
        if (++pf_ch == NCH) { pf_ch = 0; pf_tile += gridDim.x; }
    };
    prefetch();

#pragma unroll 1
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
#pragma unroll 1
        for (int ch = 0; ch < NCH; ++ch, ++chg) {
            const uint32_t st = chg & 1u, use = chg >> 1;
            if (chg >= 2u) mbar_wait(&s_free[st], (use - 1u) & 1u); // the MMAs that read this stage have retired
            if (tid == 0) {
                mbar_expect_tx(&s_full[st], kBBytes);
                bulk_load(s_b + st * kBBytes, wsrc + static_cast<size_t>(ch) * (kBBytes / 16), kBBytes, &s_full[st]);
            }
            // canonical layout [K core][row group][row % 8]: row i*16 + g sits at g*16 + i*256 inside its K core
            uint8_t *dst = s_a + st * kAStage + kc_lane * kALbo + g * 16;
#pragma unroll
            for (int i = 0; i < 8; ++i) *reinterpret_cast<uint4 *>(dst + i * 256) = v[i];
            prefetch();
            fence_async_smem();
            __syncthreads();
            if (warp == 0) { // the whole warp runs this with uniform operands and one elected lane issues (see conv1t_kernel)
                mbar_wait(&s_full[st], use & 1u);
                tc_fence_after();
                const uint64_t ad0 = smem_desc(smem_u32(s_a + st * kAStage), kALbo, 128), bd0 = smem_desc(smem_u32(s_b + st * kBBytes), N / 8 * 128, 128);
                if (elect_one()) {
#pragma unroll
                    for (int j = 0; j < kBK / 16; ++j)
                        umma_bf16(tmem, ad0 + static_cast<uint64_t>(j * 2 * kALbo / 16), bd0 + static_cast<uint64_t>(j * 2 * (N / 8 * 128) / 16), kIdesc, (ch | j) != 0);
                    umma_commit(&s_free[st]);
                    if (ch == NCH - 1) umma_commit(&s_done);
                }
                __syncwarp();
            }
        }
        mbar_wait(&s_done, it & 1u);
        tc_fence_after();

        // ---- epilogue: thread = TMEM lane = tile row
        const int row = tile * kBM + tid;
        const bool row_ok = row < a.m_total;
        const uint32_t lane_addr = tmem + (static_cast<uint32_t>(warp * 32) << 16);
        if (L == kOut) {
            float v[16];
            tmem_load16(lane_addr, v);
            if (row_ok) {
                int best = 0;
                float best_v = v[0] + s_bias[0];
#pragma unroll
                for (int j = 0; j < AHK_N_ACTIONS; ++j) {
                    const float q = v[j] + s_bias[j];
                    if (a.logits) a.logits[static_cast<size_t>(row) * AHK_N_ACTIONS + j] = q;
                    if (q > best_v) { best_v = q; best = j; }
                }
                if (a.action) a.action[row] = static_cast<uint8_t>(best);
            }
        } else {
            constexpr int kOutPitch = T::NSplit * N; // elements per output row
            __nv_bfloat16 *out = a.out + static_cast<size_t>(row) * kOutPitch + split * N;
#pragma unroll 1
            for (int c0 = 0; c0 < N; c0 += 16) {
                float v[16];
                tmem_load16(lane_addr + c0, v);
                uint32_t packed[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float lo = fmaxf(v[2 * j] + s_bias[c0 + 2 * j], 0.0f), hi = fmaxf(v[2 * j + 1] + s_bias[c0 + 2 * j + 1], 0.0f);
                    const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
                    packed[j] = *reinterpret_cast<const uint32_t *>(&h);
                }
                if (row_ok) {
                    *reinterpret_cast<uint4 *>(out + c0) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
                    *reinterpret_cast<uint4 *>(out + c0 + 8) = make_uint4(packed[4], packed[5], packed[6], packed[7]);
                }
            }
        }
        tc_fence_before();
        __syncthreads(); // every lane of the accumulator has been read: the next tile's first MMA may overwrite it
        tc_fence_after();
    }
    if (warp == 0) tmem_dealloc(tmem, kTmemCols);
}

template <int L> constexpr int smem_bytes() { return 2 * kAStage + 2 * Traits<L>::N * kBK * 2; }

// grid.x = min(tiles, resident CTAs on the device): the kernel loops over its tiles
template <int L>
cudaError_t launch(const Args &a, int tiles, int grid_y, int sms, cudaStream_t s, bool pdl = false)
{
    static int per_sm = 0;
    if (per_sm == 0) {
        cudaError_t e = cudaFuncSetAttribute(layer_kernel<L>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes<L>());
        if (e != cudaSuccess) return e;
        // resident CTAs per SM: shared memory (227 KB, + 1 KB reserved and ~1.3 KB static per CTA), tensor memory (512 columns)
        constexpr int tmem_cols = Traits<L>::N <= 32 ? 32 : (Traits<L>::N <= 64 ? 64 : (Traits<L>::N <= 128 ? 128 : 256));
        per_sm = (227 * 1024) / (smem_bytes<L>() + 3 * 1024);
        if (per_sm > 512 / tmem_cols) per_sm = 512 / tmem_cols;
        if (per_sm > 8) per_sm = 8;
        if (per_sm < 1) per_sm = 1;
    }
    const int resident = sms * per_sm / grid_y;
    return launch_kernel(layer_kernel<L>, dim3(tiles < resident ? tiles : (resident > 0 ? resident : 1), grid_y), dim3(kThreads), smem_bytes<L>(), s, pdl, a);
}

// ------------------------------------------------------------------------------------------------ conv1, Toeplitz form
// conv1 (9 -> 5 channels, 8x8, stride 3: models.py:22-24) is 47 % of the network's MACs and hostile to an im2col GEMM:
// N = 5, and every output pixel's 576-element A row would be built in shared memory by the CUDA cores (1.9 MB of
// shared-memory writes per sample; the first version of this file did that and spent 6.2 ms per 8192 samples in it).
// Here the A operand is the IMAGE ITSELF, bf16, never expanded:
//   * M row m = (input-row index i, 8-pixel block xb); the K = 16 window of one MMA is the 16 consecutive pixels
//     8 xb .. 8 xb + 15 of one input row of one plane — in the canonical no-swizzle K-major layout that is simply the
//     raw row (one core-matrix row = 8 pixels = 16 bytes, next M row = next 16 bytes) with LBO = 16 bytes: the second
//     K core matrix of row m IS the first of row m + 1.
//   * Input rows are stored by residue class (row mod 3): class j, index i holds input row 3 i + j, so that
//     (i, half row) -> address is linear (SBO = 128) and ONE A tile (plane, class j) serves the three kernel rows
//     ky = j, j + 3, j + 6: the row contributes to output rows y = i, i - 1, i - 2.
//   * Output pixels whose 8-wide window lies inside the 16-pixel window: x = 8q + 3 ph + jj with xb = 3q + ph, jj < 3,
//     at offset ph + 3 jj.  The stride-3 phase differs per row and one MMA applies one B to all rows, so B carries all
//     of it: N = 144 = 3 kernel-row groups x 3 phases x (3 outputs x 5 channels, padded to 16) — a banded (Toeplitz)
//     copy of three 8-tap kernel rows.  27 MMAs (9 planes x 3 classes) per 8 input-row indices = 6 finished output rows;
//     an MMA of this shape costs ~70 cycles whatever N <= 144 is (it is bound by streaming its 4 KB A tile from shared
//     memory; measured with N = 48 and 72 MMAs per tile: tensor pipe 74 % busy, 1.15 ms per 8192 samples).
//   * The epilogue keeps the 16 columns of its row's phase in each kernel-row group and adds the three groups across
//     rows (y = i - kv) through a small shared-memory exchange, then bias, ReLU, bf16, [pixel][8] layout for conv2.
// Persistent, warp-specialised CTA (one per SM): 8 producer warps convert uint8 ring rows to bf16 slabs (one frame of
// the 3-deep stack = 3 colour planes x 24 rows per stage, 3 stages, next slab's loads in flight during the conversion),
// 1 warp issues the MMAs, 4 warps run the epilogue out of a double-buffered TMEM accumulator.
namespace c1 {
constexpr int kProducers = 256, kCta = kProducers + 32 + 128;
constexpr int kStages = 3;
constexpr int kITile = 8, kYTile = 6, kTilesPerSample = 7; // 8 input-row indices -> 6 finished output rows; 7 x 6 >= 41
constexpr int kRows = 3 * kITile;                          // input rows per slab plane
constexpr int kRowBytes = kDim * 2, kClassBytes = kITile * kRowBytes, kPlaneBytes = 3 * kClassBytes, kSlabBytes = 3 * kPlaneBytes;
constexpr int kSlabPitch = kSlabBytes + 256;               // + zeroed pad: the last row's K window spills 16 bytes
constexpr int kN = 144;
constexpr int kWBlock = 16 * kN * 2;                       // one (plane, class) B operand: K = 16 x N = 144 bf16
constexpr int kWBytes = 27 * kWBlock;                      // 124 416 B, resident in shared memory
constexpr int kExBytes = 2 * kYTile * 16 * 16 * 4;         // exchange of the kv = 1, 2 partial sums: [2][6][16 xb][16] float, 
constexpr int kSmem = kWBytes + kStages * kSlabPitch + kExBytes; // 192 768 B
constexpr int kVecsPerSlab = 3 * kRows * 8;                // 16-pixel vectors per slab
constexpr int kTmemCols = 512;                             // 2 accumulators x 256 columns (144 used)
} // namespace c1

// 16 bytes, read once; volatile so that it stays where it is written, ahead of the barrier waits (an ld.global.nc from
// __ldg may be sunk past them to its first use, which defeats the prefetch)
__device__ __forceinline__ uint4 load_stream(const uint8_t *p)
{
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

__global__ void __launch_bounds__(c1::kCta, 1) conv1_kernel(Args a)
{
    using namespace c1;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *s_w = smem;
    uint8_t *s_slab = smem + kWBytes;
    float *s_ex = reinterpret_cast<float *>(smem + kWBytes + kStages * kSlabPitch);
    __shared__ __align__(8) uint64_t s_wfull, s_full[kStages], s_empty[kStages], s_tfull[2], s_tempty[2];
    __shared__ uint32_t s_tmem;
    __shared__ float s_bias[8];

    const int tid = threadIdx.x, warp = tid >> 5;
    const int items = a.n * kTilesPerSample;
    if (tid == 0) {
        mbar_init(&s_wfull, 1);
        for (int i = 0; i < kStages; ++i) { mbar_init(&s_full[i], kProducers); mbar_init(&s_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&s_tfull[i], 1); mbar_init(&s_tempty[i], 128); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < 8) s_bias[tid] = tid < 5 ? a.bias[tid] : 0.0f;
    for (int i = tid; i < kStages * 16; i += kCta) // the pad behind each slab
        *reinterpret_cast<uint4 *>(s_slab + (i >> 4) * kSlabPitch + kSlabBytes + (i & 15) * 16) = make_uint4(0, 0, 0, 0);
    if (warp == kProducers / 32) tmem_alloc(&s_tmem, kTmemCols);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;

    if (tid < kProducers) {
        // ---------------------------------------------------------------- producers: uint8 ring rows -> bf16 slabs
        // per-thread constants: up to three 16-pixel vectors of every slab
        int src_off[3], dst_off[3], lrow[3];
#pragma unroll
        for (int t = 0; t < 3; ++t) {
            const int v = tid + t * kProducers;
            const int colour = v / (kRows * 8), rem = v - colour * (kRows * 8), lr = rem >> 3, xv = rem & 7;
            src_off[t] = (colour * kDim + lr) * kDim + xv * 16;
            dst_off[t] = colour * kPlaneBytes + (lr % 3) * kClassBytes + (lr / 3) * kRowBytes + xv * 32;
            lrow[t] = v < kVecsPerSlab ? lr : kDim; // kDim: never valid
        }
        const int swap = (tid >> 2) & 1; // xv >= 4
        const int my_items = blockIdx.x < items ? (items - 1 - blockIdx.x) / gridDim.x + 1 : 0;
        // all three slabs (stack ages) of tile k: 9 vectors per thread
        auto load = [&](int k, uint4 (&px)[3][3]) {
            const int item = blockIdx.x + k * gridDim.x;
            const int sample = item / kTilesPerSample, r_first = 3 * (item - sample * kTilesPerSample) * kYTile;
#pragma unroll
            for (int age = 0; age < 3; ++age) {
                const uint8_t *src = stack_frame(a.frames, a.warm, a.meta, a.depth, static_cast<size_t>(a.n), a.head, static_cast<size_t>(sample), age) + r_first * kDim;
#pragma unroll
                for (int t = 0; t < 3; ++t) {
                    px[age][t] = make_uint4(0, 0, 0, 0);
                    if (r_first + lrow[t] < kDim) px[age][t] = load_stream(src + src_off[t]);
                }
            }
        };
        uint32_t f = 0; // slabs filled so far
        // Tile k + 1's loads (27 KB per CTA) are issued before tile k is converted and stay in flight meanwhile: with one
        // slab in flight the kernel ran at load latency (9 KB per microsecond per SM: 1.15 ms per 8192 samples).  The two
        // register sets swap roles by unrolling, so nothing waits on a load it does not need yet.
        auto convert = [&](int k, uint4 (&cur)[3][3], uint4 (&nxt)[3][3]) {
            // Every vector of `cur` is touched before the next tile's loads are issued: the hardware tracks loads with a
            // handful of scoreboard counters, and a new load sharing a counter with a vector still to be converted would
            // make that conversion wait for the NEW load (seen as a full DRAM latency per tile in the stall samples).
            uint32_t touched = 0u;
#pragma unroll
            for (int age = 0; age < 3; ++age)
#pragma unroll
                for (int t = 0; t < 3; ++t) touched ^= cur[age][t].x ^ cur[age][t].y ^ cur[age][t].z ^ cur[age][t].w;
            if (touched == 0x9E3779B9u && a.n < 0) s_bias[7] = 1.0f; // never true; keeps the touches alive
            if (k + 1 < my_items) load(k + 1, nxt);
#pragma unroll
            for (int age = 0; age < 3; ++age, ++f) {
                const uint32_t st = f % kStages;
                if (f >= kStages) mbar_wait(&s_empty[st], ((f / kStages) - 1u) & 1u);
                uint8_t *slab = s_slab + st * kSlabPitch;
#pragma unroll
                for (int t = 0; t < 3; ++t) {
                    if (lrow[t] < kDim) {
                        const uint4 c = cur[age][t];
                        const uint2 q0 = u8x4_to_bf16x4(c.x), q1 = u8x4_to_bf16x4(c.y), q2 = u8x4_to_bf16x4(c.z), q3 = u8x4_to_bf16x4(c.w);
                        uint4 *dst = reinterpret_cast<uint4 *>(slab + dst_off[t]);
                        // lanes 4..7 of each row store their halves in the other order: one instruction then covers all
                        // eight 16-byte bank groups instead of four
                        const uint4 lo = make_uint4(q0.x, q0.y, q1.x, q1.y), hi = make_uint4(q2.x, q2.y, q3.x, q3.y);
                        dst[swap] = swap ? hi : lo;
                        dst[swap ^ 1] = swap ? lo : hi;
                    }
                }
                fence_async_smem();
                mbar_arrive(&s_full[st]);
            }
        };
        uint4 pa[3][3], pb[3][3];
        if (my_items) load(0, pa);
#pragma unroll 1
        for (int k = 0; k < my_items; k += 2) {
            convert(k, pa, pb);
            if (k + 1 < my_items) convert(k + 1, pb, pa);
        }
    } else if (warp == kProducers / 32) {
        // ---------------------------------------------------------------- MMA issue (one thread)
        if ((tid & 31) == 0) {
            mbar_expect_tx(&s_wfull, kWBytes);
            for (int i = 0; i < 9; ++i) bulk_load(s_w + i * (kWBytes / 9), reinterpret_cast<const uint8_t *>(a.weights) + i * (kWBytes / 9), kWBytes / 9, &s_wfull);
            mbar_wait(&s_wfull, 0);
            constexpr uint32_t kIdesc = instr_desc(kN);
            const uint32_t w_base = smem_u32(s_w);
            uint32_t fill = 0, tile = 0;
            for (int item = blockIdx.x; item < items; item += gridDim.x, ++tile) {
                const uint32_t buf = tile & 1u;
                if (tile >= 2) mbar_wait(&s_tempty[buf], ((tile >> 1) - 1) & 1);
                tc_fence_after();
                for (int age = 0; age < 3; ++age, ++fill) {
                    const uint32_t st = fill % kStages;
                    mbar_wait(&s_full[st], (fill / kStages) & 1);
                    tc_fence_after();
                    const uint32_t slab = smem_u32(s_slab + st * kSlabPitch);
#pragma unroll
                    for (int colour = 0; colour < 3; ++colour) {
#pragma unroll
                        for (int cls = 0; cls < 3; ++cls) {
                            // M row (i, xb) = input row 3 i + cls, pixels 8 xb ..: K step = 16 bytes, 8-row groups 128 bytes apart
                            const uint64_t ad = smem_desc(slab + colour * kPlaneBytes + cls * kClassBytes, 16, 128);
                            const uint64_t bd = smem_desc(w_base + ((age * 3 + colour) * 3 + cls) * kWBlock, (kN / 8) * 128, 128);
                            umma_bf16(tmem + buf * 256, ad, bd, kIdesc, (age | colour | cls) != 0);
                        }
                    }
                    umma_commit(&s_empty[st]);
                }
                umma_commit(&s_tfull[buf]);
            }
        }
    } else {
        // ---------------------------------------------------------------- epilogue: 4 warps, thread = TMEM lane = M row
        // (one warp per scheduler, so its instruction count is its time: the phase select is branch-free — as divergent
        // branches it made this role the kernel's bottleneck — and the TMEM loads are waited for three at a time)
        const int q4 = warp & 3;                       // the TMEM lane quarter this warp may read
        const int m = q4 * 32 + (tid & 31);
        const int il = m >> 4, xb = m & 15, ph = xb % 3, xq = xb / 3;
        const int is_ph0 = ph == 0, is_ph2 = ph == 2;
        float *ex = s_ex;
        uint32_t tile = 0;
        for (int item = blockIdx.x; item < items; item += gridDim.x, ++tile) {
            const int sample = item / kTilesPerSample, y0 = (item - sample * kTilesPerSample) * kYTile;
            const uint32_t buf = tile & 1u;
            mbar_wait(&s_tfull[buf], (tile >> 1) & 1);
            tc_fence_after();
            const uint32_t taddr = tmem + buf * 256 + (static_cast<uint32_t>(q4 * 32) << 16);
            float part[3][16]; // [kernel-row group kv][3 outputs x 5 channels] of this row's phase
#pragma unroll
            for (int kv = 0; kv < 3; ++kv) {
                uint32_t v0[16], v1[16], v2[16];
                tmem_load16_async(taddr + kv * 48, v0);
                tmem_load16_async(taddr + kv * 48 + 16, v1);
                tmem_load16_async(taddr + kv * 48 + 32, v2);
                tmem_wait_ld();
#pragma unroll
                for (int e = 0; e < 16; ++e) part[kv][e] = __uint_as_float(select(v2[e], select(v0[e], v1[e], is_ph0), is_ph2));
            }
            tc_fence_before();
            mbar_arrive(&s_tempty[buf]); // the accumulator is in registers: the MMA warp may overwrite it
            // kernel-row group kv of input-row index i belongs to output row i - kv
#pragma unroll
            for (int kv = 1; kv < 3; ++kv) {
                const int yl = il - kv;
                if (yl >= 0 && yl < kYTile) {
                    float4 *dst = reinterpret_cast<float4 *>(ex + (((kv - 1) * kYTile + yl) * 16 + xb) * 16);
#pragma unroll
                    for (int e = 0; e < 4; ++e) dst[e] = make_float4(part[kv][4 * e], part[kv][4 * e + 1], part[kv][4 * e + 2], part[kv][4 * e + 3]);
                }
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");
            const int y = y0 + il;
            if (il < kYTile && y < kC1) {
                const float4 *p1 = reinterpret_cast<const float4 *>(ex + ((0 * kYTile + il) * 16 + xb) * 16);
                const float4 *p2 = reinterpret_cast<const float4 *>(ex + ((1 * kYTile + il) * 16 + xb) * 16);
                float sum[16];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float4 u = p1[e], w = p2[e];
                    sum[4 * e] = part[0][4 * e] + u.x + w.x;
                    sum[4 * e + 1] = part[0][4 * e + 1] + u.y + w.y;
                    sum[4 * e + 2] = part[0][4 * e + 2] + u.z + w.z;
                    sum[4 * e + 3] = part[0][4 * e + 3] + u.w + w.w;
                }
                const int x0 = 8 * xq + 3 * ph;
                uint4 *out = reinterpret_cast<uint4 *>(a.out + (static_cast<size_t>(sample) * kP1 + y * kC1 + x0) * kAct1C);
#pragma unroll
                for (int j = 0; j < 3; ++j) {
                    if ((j < 2 || ph < 2) && x0 + j < kC1) {
                        float c[5];
#pragma unroll
                        for (int ch = 0; ch < 5; ++ch) c[ch] = fmaxf(sum[j * 5 + ch] + s_bias[ch], 0.0f);
                        const __nv_bfloat162 p01 = __floats2bfloat162_rn(c[0], c[1]), p23 = __floats2bfloat162_rn(c[2], c[3]), p4 = __floats2bfloat162_rn(c[4], 0.0f);
                        out[j] = make_uint4(*reinterpret_cast<const uint32_t *>(&p01), *reinterpret_cast<const uint32_t *>(&p23), *reinterpret_cast<const uint32_t *>(&p4), 0u);
                    }
                }
            }
            asm volatile("bar.sync 1, 128;" ::: "memory"); // the exchange buffer is free for the next tile
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kProducers / 32) tmem_dealloc(tmem, kTmemCols);
}

cudaError_t launch_conv1(const Args &a, int grid, cudaStream_t s)
{
    static bool prepared = false;
    if (!prepared) {
        cudaError_t e = cudaFuncSetAttribute(conv1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, c1::kSmem);
        if (e != cudaSuccess) return e;
        prepared = true;
    }
    conv1_kernel<<<grid, c1::kCta, c1::kSmem, s>>>(a);
    return cudaGetLastError();
}


// ------------------------------------------------------------------------------------------------ conv1, phase-separated
// The Toeplitz form above spends two thirds of its columns on stride phases a row does not have (N = 144 = 3 kernel-row
// groups x 3 PHASES x 16; a row (i, xb) only uses the 48 columns of phase xb mod 3), and its K window starts 16 bytes
// into a 128-byte core matrix, so every second core matrix straddles two shared-memory lines.  Here the M rows of one
// MMA all have the SAME phase, so N = 48 (3 kernel-row groups x (3 outputs x 5 channels, padded to 16)):
//   * shared layout per colour plane and row class c = row mod 3: four arrays A0..A3 of [q = 0..4][il = 0..22] x 16 bytes
//     (one 8-pixel block of one input row, fp16), A_a holding blocks 3q + a — A3 repeats blocks 3, 6, 9, 12 of A0 next
//     to block 15 (+25 % of the conversion stores).  For phase ph the K = 16 window of row (q, il) is block 3q + ph in
//     A_ph followed by block 3q + ph + 1 in A_{ph+1}: K core 1 lies one array pitch (LBO = 2048) behind K core 0, both
//     128-byte aligned, and the 128 M rows are simply 128 consecutive 16-byte rows of the array (SBO = 128): r = 23 q + il.
//     Outputs x = 8q + 3 ph + jj (jj < 3) sit at offset ph + 3 jj of the window; x = 40 = 8*4 + 6 + 2 is phase 2's third
//     output of q = 4, so five blocks per phase suffice (5 x 23 = 115 of 128 rows).
//   * a work item is HALF a sample: input-row indices i0 .. i0 + 22 (i0 = 0 or 20) give output rows 0..19 or 20..40 — one
//     M tile per phase, three accumulators of 48 columns per item, double-buffered in TMEM (2 x 256 columns), nine
//     24 KB stages (one colour plane of one stack age each) of 9 MMAs per item: 162 MMAs of N = 48 per sample instead
//     of 189 of N = 144.
//   * the A operand is fp16, not bf16: 0..255 is exact in both, and uint8 -> fp16 is one PRMT + one HSUB2 per two pixels
//     (0x6400 | v is the half 1024 + v) against five instructions through float32; the weights are fp16 too (kind::f16
//     takes either; |w| <= 0.09 stays in fp16's normal range with 11 significant bits against bf16's 8) and the 1/128
//     of the input normalisation moves from the weights to the epilogue.
// Roles as before: 8 producer warps (global uint8 -> registers -> fp16 -> shared, three tasks of 1 KB in flight per
// warp), 1 MMA warp, 4 epilogue warps (kernel-row groups summed across rows through a 14 KB exchange).
namespace c1p {
constexpr int kConverters = 256, kMmaWarp = kConverters / 32, kLoadWarp = kMmaWarp + 1, kCta = kConverters + 64 + 128;
constexpr int kStages = 2;                                      // converted A stages (MMA of plane f overlaps conversion of f + 1)
constexpr int kRaw = 4;                                         // raw uint8 stages in flight: 4 x 8.6 KB per SM keeps HBM busy
constexpr int kRowsQ = 23, kQ = 5, kRowsUsed = kRowsQ * kQ;     // 115 M rows per phase
constexpr int kAP = 128 * 16;                                   // array pitch = LBO: 128 rows of 16 bytes
constexpr int kClassBytes = 4 * kAP, kStageBytes = 3 * kClassBytes; // 24 576 B: one colour plane of one half sample
constexpr int kRawRows = 3 * kRowsQ;                             // 69 image rows of one colour plane: ONE tiled TMA copy per stage
constexpr int kRawBytes = 9 * 1024;                             // 8 832 B of pixels, 128-byte swizzle: stages are 1024-byte aligned
constexpr int kN = 48;
constexpr int kWBlock = 16 * kN * 2;                            // one (plane, class, phase) B operand
constexpr int kWBlocks = 9 * 3 * 3, kWBytes = kWBlocks * kWBlock; // 124 416 B, resident
constexpr int kExBytes = 2 * kRowsQ * kQ * 16 * 4;              // exchange of the kv = 1, 2 partial sums
constexpr int kSmem = 1024 + kRaw * kRawBytes + kWBytes + kStages * kStageBytes + kExBytes; // 226 176 B (1 KB of alignment slack)
constexpr int kTmemCols = 512, kBufCols = 256;
} // namespace c1p

// instruction descriptor: D f32, A/B fp16, both K-major, M = 128
__host__ __device__ constexpr uint32_t instr_desc_f16(int n) { return 1u << 4 | static_cast<uint32_t>(n >> 3) << 17 | (128u >> 4) << 24; }

// 4 uint8 -> 4 fp16 (exact): 0x6400 | v is 1024 + v as a half
__device__ __forceinline__ uint2 u8x4_to_f16x4(uint32_t w)
{
    const uint32_t lo = __byte_perm(w, 0x64646464u, 0x4140), hi = __byte_perm(w, 0x64646464u, 0x4342);
    const __half2 k = __halves2half2(__ushort_as_half(0x6400), __ushort_as_half(0x6400));
    const __half2 a = __hsub2(*reinterpret_cast<const __half2 *>(&lo), k), b = __hsub2(*reinterpret_cast<const __half2 *>(&hi), k);
    return make_uint2(*reinterpret_cast<const uint32_t *>(&a), *reinterpret_cast<const uint32_t *>(&b));
}

__global__ void __launch_bounds__(c1p::kCta, 1) conv1p_kernel(Args a, const __grid_constant__ CUtensorMap tm_ring, const __grid_constant__ CUtensorMap tm_warm)
{
    using namespace c1p;
    extern __shared__ __align__(128) uint8_t smem_unaligned[];
    uint8_t *s_raw = smem_unaligned + ((1024u - (smem_u32(smem_unaligned) & 1023u)) & 1023u); // swizzle atoms are 1024-byte aligned
    uint8_t *s_w = s_raw + kRaw * kRawBytes;
    uint8_t *s_stage = s_w + kWBytes;
    float *s_ex = reinterpret_cast<float *>(s_stage + kStages * kStageBytes);
    __shared__ __align__(8) uint64_t s_wfull, s_rfull[kRaw], s_rempty[kRaw], s_full[kStages], s_empty[kStages], s_tfull[2], s_tempty[2];
    __shared__ uint32_t s_tmem;
    __shared__ float s_bias[8];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int items = a.n * 2;
    const int my_items = static_cast<int>(blockIdx.x) < items ? (items - 1 - static_cast<int>(blockIdx.x)) / static_cast<int>(gridDim.x) + 1 : 0;
    if (tid == 0) {
        mbar_init(&s_wfull, 1);
        for (int i = 0; i < kRaw; ++i) { mbar_init(&s_rfull[i], 1); mbar_init(&s_rempty[i], kConverters / 32); }
        for (int i = 0; i < kStages; ++i) { mbar_init(&s_full[i], kConverters / 32); mbar_init(&s_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&s_tfull[i], 1); mbar_init(&s_tempty[i], 128); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < 8) s_bias[tid] = tid < 5 ? a.bias[tid] : 0.0f;
    for (int i = tid; i < kStages * kStageBytes / 16; i += kCta) // rows 115..127 of every array are never written: keep them finite
        reinterpret_cast<uint4 *>(s_stage)[i] = make_uint4(0, 0, 0, 0);
    if (warp == kMmaWarp) tmem_alloc(&s_tmem, kTmemCols);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;

    if (warp == kLoadWarp) {
        // ---------------------------------------------------------------- loader: one tiled TMA copy per stage
        // Stage f = (item, plane): the 69 image rows 3 i0 .. 3 i0 + 68 of one colour plane (128 x 69 bytes, 128-byte swizzle:
        // the 16-byte chunk k of row r lands at chunk k ^ (r & 7), so that lanes = every third row read conflict-free).
        // The ring and the warm-up side buffer are two tensor maps over [frame][colour][row] x 128 bytes; the frame indices
        // of the NEXT item (a dependent load of the episode word when the ring is deep) are fetched while this item's
        // nine planes are issued.  Row 128 of the last band (half 1) is the next plane's first row or out of bounds
        // (zero fill): the converters never use it.
        if (lane == 0) {
            auto frame_rows = [&](int t, int (&p)[3]) { // row index of (frame, colour 0, row 0); bit 31: warm-up buffer
                const int sample = (static_cast<int>(blockIdx.x) + t * static_cast<int>(gridDim.x)) >> 1;
                const int steps = a.warm != nullptr ? static_cast<int>(a.meta[sample] & 0xFFFFu) : 0x7FFFFFFF;
#pragma unroll
                for (int age = 0; age < 3; ++age)
                    p[age] = age >= steps ? static_cast<int>((static_cast<uint32_t>(2 - (age - steps)) * a.n + sample) * (3u * kDim) | 0x80000000u)
                                          : ((a.head + a.depth - age) % a.depth * a.n + sample) * (3 * kDim);
            };
            int cur[3] = {0, 0, 0}, nxt[3] = {0, 0, 0};
            if (my_items > 0) frame_rows(0, cur);
            uint32_t fill = 0;
            for (int t = 0; t < my_items; ++t) {
                if (t + 1 < my_items) frame_rows(t + 1, nxt);
                const int item = static_cast<int>(blockIdx.x) + t * static_cast<int>(gridDim.x);
                const int row0 = (item & 1) ? 60 : 0;
#pragma unroll
                for (int plane = 0; plane < 9; ++plane, ++fill) {
                    const uint32_t r = fill % kRaw;
                    if (fill >= kRaw) mbar_wait(&s_rempty[r], ((fill / kRaw) - 1u) & 1u);
                    mbar_expect_tx(&s_rfull[r], kRawRows * kDim);
                    const int f = cur[plane / 3];
                    const int row = (f & 0x7FFFFFFF) + (plane % 3) * kDim + row0;
                    const CUtensorMap *tm = f < 0 ? &tm_warm : &tm_ring;
                    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                                 :: "r"(smem_u32(s_raw + r * kRawBytes)), "l"(reinterpret_cast<uint64_t>(tm)), "r"(0), "r"(row), "r"(smem_u32(&s_rfull[r])) : "memory");
                }
#pragma unroll
                for (int age = 0; age < 3; ++age) cur[age] = nxt[age];
            }
        }
    } else if (warp < kMmaWarp) {
        // ---------------------------------------------------------------- converters: uint8 rows -> the fp16 phase arrays
        // warp w owns the 16-pixel chunk w of every row (8-pixel blocks 2 w and 2 w + 1); lane = input-row index il, whose
        // three rows 3 il + c are the row classes c
        const bool active = lane < kRowsQ;
        uint32_t src[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) src[c] = (3 * lane + c) * kDim + ((warp ^ ((3 * lane + c) & 7)) << 4);
        uint32_t fill = 0;
        for (int t = 0; t < my_items; ++t) {
            const int item = static_cast<int>(blockIdx.x) + t * static_cast<int>(gridDim.x);
            const bool short_row = (item & 1) && lane == kRowsQ - 1; // class 2 of il = 22 is image row 128: zeros
#pragma unroll 1
            for (int plane = 0; plane < 9; ++plane, ++fill) {
                const uint32_t r = fill % kRaw, st = fill % kStages;
                mbar_wait(&s_rfull[r], (fill / kRaw) & 1u);
                uint4 px[3];
#pragma unroll
                for (int c = 0; c < 3; ++c) px[c] = active ? *reinterpret_cast<const uint4 *>(s_raw + r * kRawBytes + src[c]) : make_uint4(0, 0, 0, 0);
                if (short_row) px[2] = make_uint4(0, 0, 0, 0);
                if (fill >= kStages) mbar_wait(&s_empty[st], ((fill / kStages) - 1u) & 1u);
                if (active) {
                    uint8_t *base = s_stage + st * kStageBytes + lane * 16;
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        const uint32_t w[4] = {px[c].x, px[c].y, px[c].z, px[c].w};
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            const uint2 lo = u8x4_to_f16x4(w[2 * h]), hi = u8x4_to_f16x4(w[2 * h + 1]);
                            const uint4 v = make_uint4(lo.x, lo.y, hi.x, hi.y);
                            const int b = 2 * warp + h, arr = b % 3, q = b / 3;
                            if (b <= 14) *reinterpret_cast<uint4 *>(base + c * kClassBytes + arr * kAP + q * (kRowsQ * 16)) = v;
                            if (arr == 0 && b >= 3) *reinterpret_cast<uint4 *>(base + c * kClassBytes + 3 * kAP + (q - 1) * (kRowsQ * 16)) = v; // A3: blocks 3, 6, 9, 12, 15
                        }
                    }
                }
                fence_async_smem();
                __syncwarp();
                // the raw stage is released only here, after every pixel register has been consumed: an arrive issued right behind
                // the loads does not wait for their data, and the loader's next copy overwrote rows still being read (seen as
                // run-to-run differences at 8192 samples, tools/conv1_ab.py)
                if (lane == 0) { mbar_arrive(&s_full[st]); mbar_arrive(&s_rempty[r]); }
            }
        }
    } else if (warp == kMmaWarp) {
        // ---------------------------------------------------------------- MMA issue (one thread)
        if (lane == 0) {
            mbar_expect_tx(&s_wfull, kWBytes);
            for (int i = 0; i < 9; ++i) bulk_load(s_w + i * (kWBytes / 9), reinterpret_cast<const uint8_t *>(a.weights) + i * (kWBytes / 9), kWBytes / 9, &s_wfull);
            mbar_wait(&s_wfull, 0);
            constexpr uint32_t kIdesc = instr_desc_f16(kN);
            const uint32_t w_base = smem_u32(s_w);
            uint32_t fill = 0;
            for (int t = 0; t < my_items; ++t) {
                const uint32_t buf = t & 1u;
                if (t >= 2) mbar_wait(&s_tempty[buf], ((t >> 1) - 1) & 1);
                tc_fence_after();
                for (int plane = 0; plane < 9; ++plane, ++fill) {
                    const uint32_t st = fill % kStages;
                    mbar_wait(&s_full[st], (fill / kStages) & 1);
                    tc_fence_after();
                    const uint32_t stage = smem_u32(s_stage + st * kStageBytes);
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
#pragma unroll
                        for (int ph = 0; ph < 3; ++ph) {
                            const uint64_t ad = smem_desc(stage + c * kClassBytes + ph * kAP, kAP, 128);
                            const uint64_t bd = smem_desc(w_base + ((plane * 3 + c) * 3 + ph) * kWBlock, (kN / 8) * 128, 128);
                            umma_bf16(tmem + buf * kBufCols + ph * kN, ad, bd, kIdesc, (plane | c) != 0);
                        }
                    }
                    umma_commit(&s_empty[st]);
                }
                umma_commit(&s_tfull[buf]);
            }
        }
    } else {
        // ---------------------------------------------------------------- epilogue: 4 warps, thread = TMEM lane = M row 23 q + il
        const int q4 = warp & 3;
        const int m = q4 * 32 + lane;
        const int q = m / kRowsQ, il = m - q * kRowsQ;
        const bool row_ok = m < kRowsUsed;
        for (int t = 0; t < my_items; ++t) {
            const int item = static_cast<int>(blockIdx.x) + t * static_cast<int>(gridDim.x);
            const int sample = item >> 1, y0 = (item & 1) ? 20 : 0, ny = (item & 1) ? 21 : 20;
            const uint32_t buf = t & 1u;
            mbar_wait(&s_tfull[buf], (t >> 1) & 1);
            tc_fence_after();
#pragma unroll 1
            for (int ph = 0; ph < 3; ++ph) {
                const uint32_t taddr = tmem + buf * kBufCols + ph * kN + (static_cast<uint32_t>(q4 * 32) << 16);
                uint32_t v0[16], v1[16], v2[16]; // kernel-row groups kv = 0, 1, 2: [3 outputs x 5 channels]
                tmem_load16_async(taddr, v0);
                tmem_load16_async(taddr + 16, v1);
                tmem_load16_async(taddr + 32, v2);
                tmem_wait_ld();
                if (ph == 2) { // the item's accumulators are all in registers: the MMA warp may overwrite them
                    tc_fence_before();
                    mbar_arrive(&s_tempty[buf]);
                }
                // kernel-row group kv of input-row index il belongs to output row il - kv
                if (row_ok && il >= 1 && il - 1 < ny) {
                    float4 *dst = reinterpret_cast<float4 *>(s_ex + ((0 * kRowsQ + (il - 1)) * kQ + q) * 16);
#pragma unroll
                    for (int e = 0; e < 4; ++e) dst[e] = make_float4(__uint_as_float(v1[4 * e]), __uint_as_float(v1[4 * e + 1]), __uint_as_float(v1[4 * e + 2]), __uint_as_float(v1[4 * e + 3]));
                }
                if (row_ok && il >= 2 && il - 2 < ny) {
                    float4 *dst = reinterpret_cast<float4 *>(s_ex + ((1 * kRowsQ + (il - 2)) * kQ + q) * 16);
#pragma unroll
                    for (int e = 0; e < 4; ++e) dst[e] = make_float4(__uint_as_float(v2[4 * e]), __uint_as_float(v2[4 * e + 1]), __uint_as_float(v2[4 * e + 2]), __uint_as_float(v2[4 * e + 3]));
                }
                asm volatile("bar.sync 1, 128;" ::: "memory");
                if (row_ok && il < ny) {
                    const float4 *p1 = reinterpret_cast<const float4 *>(s_ex + ((0 * kRowsQ + il) * kQ + q) * 16);
                    const float4 *p2 = reinterpret_cast<const float4 *>(s_ex + ((1 * kRowsQ + il) * kQ + q) * 16);
                    float sum[16];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float4 u = p1[e], w = p2[e];
                        sum[4 * e] = __uint_as_float(v0[4 * e]) + u.x + w.x;
                        sum[4 * e + 1] = __uint_as_float(v0[4 * e + 1]) + u.y + w.y;
                        sum[4 * e + 2] = __uint_as_float(v0[4 * e + 2]) + u.z + w.z;
                        sum[4 * e + 3] = __uint_as_float(v0[4 * e + 3]) + u.w + w.w;
                    }
                    const int y = y0 + il, x0 = 8 * q + 3 * ph;
                    uint4 *out = reinterpret_cast<uint4 *>(a.out + (static_cast<size_t>(sample) * kP1 + y * kC1 + x0) * kAct1C);
#pragma unroll
                    for (int j = 0; j < 3; ++j) {
                        if (j < 2 || ph < 2 || q == kQ - 1) { // phase 2's third output is the next block's first, except x = 40
                            float c[5];
#pragma unroll
                            for (int ch = 0; ch < 5; ++ch) c[ch] = fmaxf(sum[j * 5 + ch] * (1.0f / 128.0f) + s_bias[ch], 0.0f); // (x - 128) / 128: processor.py:39-42
                            const __nv_bfloat162 p01 = __floats2bfloat162_rn(c[0], c[1]), p23 = __floats2bfloat162_rn(c[2], c[3]), p4 = __floats2bfloat162_rn(c[4], 0.0f);
                            out[j] = make_uint4(*reinterpret_cast<const uint32_t *>(&p01), *reinterpret_cast<const uint32_t *>(&p23), *reinterpret_cast<const uint32_t *>(&p4), 0u);
                        }
                    }
                }
                asm volatile("bar.sync 1, 128;" ::: "memory"); // the exchange buffer is free for the next phase
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) tmem_dealloc(tmem, kTmemCols);
}

// [frame][colour][row] x 128 bytes as a 2-D uint8 tensor, box = 69 rows, 128-byte swizzle
static cudaError_t frame_tensor_map(CUtensorMap *tm, const uint8_t *base, size_t frames)
{
    static PFN_cuTensorMapEncodeTiled_v12000 encode = nullptr;
    if (encode == nullptr) {
        cudaDriverEntryPointQueryResult q;
        void *fn = nullptr;
        cudaError_t e = cudaGetDriverEntryPointByVersion("cuTensorMapEncodeTiled", &fn, 12000, cudaEnableDefault, &q);
        if (e != cudaSuccess) return e;
        if (q != cudaDriverEntryPointSuccess || fn == nullptr) return cudaErrorNotSupported;
        encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
    }
    const cuuint64_t dims[2] = {static_cast<cuuint64_t>(kDim), static_cast<cuuint64_t>(frames) * 3 * kDim};
    const cuuint64_t strides[1] = {static_cast<cuuint64_t>(kDim)};
    const cuuint32_t box[2] = {static_cast<cuuint32_t>(kDim), static_cast<cuuint32_t>(c1p::kRawRows)};
    const cuuint32_t estr[2] = {1, 1};
    if (dims[1] >= (1ull << 31)) return cudaErrorInvalidValue;
    const CUresult r = encode(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<uint8_t *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}

// the tensor maps of conv1p_kernel / conv1t_kernel are a pure function of (pointer, frame count): keep the last pair per calling thread
struct FrameMaps { const uint8_t *frames = nullptr, *warm = nullptr; int n = 0, depth = 0; CUtensorMap ring, side; };
static cudaError_t frame_maps(const Args &a, const FrameMaps **out)
{
    static thread_local FrameMaps c;
    if (c.frames != a.frames || c.warm != a.warm || c.n != a.n || c.depth != a.depth) {
        cudaError_t e;
        if ((e = frame_tensor_map(&c.ring, a.frames, static_cast<size_t>(a.depth) * a.n)) != cudaSuccess) return e;
        if (a.warm != nullptr) { if ((e = frame_tensor_map(&c.side, a.warm, static_cast<size_t>(3) * a.n)) != cudaSuccess) return e; }
        else c.side = c.ring;
        c.frames = a.frames; c.warm = a.warm; c.n = a.n; c.depth = a.depth;
    }
    *out = &c;
    return cudaSuccess;
}

cudaError_t launch_conv1p(const Args &a, int grid, cudaStream_t s)
{
    cudaError_t e = cudaFuncSetAttribute(conv1p_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, c1p::kSmem); // per device
    if (e != cudaSuccess) return e;
    const FrameMaps *m = nullptr;
    if ((e = frame_maps(a, &m)) != cudaSuccess) return e;
    conv1p_kernel<<<grid, c1p::kCta, c1p::kSmem, s>>>(a, m->ring, m->side);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ conv1, A operand in TMEM
// ncu of conv1p_kernel (profiles/conv1p_r2c_ncu.txt): tensor pipe 18 %, "tc" pipe 50 %, converters busy 65-80 % of the
// time, a third of that in fence.proxy.async + mbarrier.  Two limits meet there: an MMA whose A operand comes from shared
// memory reads it at 64 B/clk — 4 KB = 64 cycles per M128 x K16 tile whatever N is, against a floor of 128 N / 256 = 24
// cycles at N = 48 — and every converted byte goes register -> shared (generic proxy) -> fence -> tensor core.
// Here the converters write the A tiles straight into TENSOR MEMORY (tcgen05.st, 256 B/clk, no proxy fence) and the MMAs
// take A from TMEM (tcgen05.mma [d], [a], b-desc: only the 1.5 KB B block is read from shared memory):
//   * same work items, M rows (r = 23 q + il), phases, weights (pack_conv1_phase) and epilogue as conv1p_kernel;
//   * TMEM lane r of tile (class c, phase ph) holds the 16 fp16 pixels 24 q + 8 ph .. + 15 of image row 3 il + c in 8
//     columns (two K elements per 32-bit column).  A thread owns one lane: per class it reads its row's 32 pixels
//     24 q .. 24 q + 31 from the swizzled raw stage (four 8-byte loads), converts each pixel ONCE (16 registers) and stores
//     the three overlapping phase windows — registers 4 ph .. 4 ph + 7 — with three tcgen05.st.32x32b.x8;
//   * 512 TMEM columns = 2 accumulator buffers x 3 phases x 48 + 3 A stages x 9 tiles x 8;
//   * 8 converter warps = 2 groups x 4 lane quarters, the groups taking alternate planes; no converted stage in shared
//     memory, so 8 raw stages (72 KB per SM) of TMA loads are in flight.
namespace c1t {
// kGroups converter groups of 4 warps (one per TMEM lane quarter), then the MMA warp, the loader warp, kEpiSets x 4 epilogue warps;
// the group count is a template parameter of the kernel (A/B switch AHK_CONV1_GROUPS=2|3|4)
constexpr int kEpiSets = 2;
constexpr int cta_threads(int groups) { return groups * 128 + 64 + kEpiSets * 128; }
constexpr int kAStages = 4, kRaw = 8;
constexpr int kRowsQ = c1p::kRowsQ, kQ = c1p::kQ, kRowsUsed = c1p::kRowsUsed;
constexpr int kRawRows = c1p::kRawRows, kRawBytes = c1p::kRawBytes;
constexpr int kN = 48, kWBlock = c1p::kWBlock, kWBytes = c1p::kWBytes, kExBytes = c1p::kExBytes;
constexpr int kSmem = 1024 + kRaw * kRawBytes + kWBytes + kEpiSets * kExBytes; // 228 608 B
constexpr int kTmemCols = 512, kBufCols = 3 * kN, kACol = 2 * kBufCols, kAClassCols = 16, kAStageCols = 3 * kAClassCols; // 288 + 4 x 48 = 480
// TMEM lane m <-> M row (q, il): lanes are laid out [il / 8][q][il % 8], so that a warp's 32 lanes are 4 pixel blocks x 8
// consecutive rows: their 8-byte loads from the swizzled stage hit 32 different (chunk, half) slots and the exchange rows
// of the epilogue are contiguous.  120 lanes, 115 of them rows.
__device__ __forceinline__ bool lane_row(int m, int &q, int &il)
{
    const int g = m / 40, rem = m - g * 40;
    q = rem >> 3;
    il = g * 8 + (rem & 7);
    return m < 120 && il < kRowsQ;
}
} // namespace c1t

// D[tmem] (+)= A[tmem] * B[smem]: A is lanes 0..127 x 8 columns (16 fp16), issued by ONE thread
__device__ __forceinline__ void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        :: "r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_f16_ts_bf16(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
{
    umma_f16_ts(tmem_d, tmem_a, bdesc, idesc, accumulate); // the operand types are in the instruction descriptor
}
// 8 consecutive 32-bit columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_store8(uint32_t taddr, uint32_t r0, uint32_t r1, uint32_t r2, uint32_t r3, uint32_t r4, uint32_t r5, uint32_t r6, uint32_t r7)
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                 :: "r"(taddr), "r"(r0), "r"(r1), "r"(r2), "r"(r3), "r"(r4), "r"(r5), "r"(r6), "r"(r7) : "memory");
}
__device__ __forceinline__ void tmem_store16(uint32_t taddr, const uint32_t (&r)[16])
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
                 :: "r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
                    "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}
// 4 uint8 -> 4 fp16 denormals (v * 2^-24, exact): zero-extended bytes
__device__ __forceinline__ uint2 u8x4_to_f16x4_denormal(uint32_t w) { return make_uint2(__byte_perm(w, 0u, 0x4140), __byte_perm(w, 0u, 0x4342)); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// kDenormal: pixels as fp16 DENORMALS (bit pattern 0x00vv = v * 2^-24): one PRMT per two pixels, the 2^24 goes to the epilogue
template <bool kDenormal, int kGroups>
__global__ void __launch_bounds__(c1t::cta_threads(kGroups), 1) conv1t_kernel(Args a, const __grid_constant__ CUtensorMap tm_ring, const __grid_constant__ CUtensorMap tm_warm)
{
    using namespace c1t;
    constexpr int kMmaWarp = kGroups * 4, kLoadWarp = kMmaWarp + 1;
    extern __shared__ __align__(128) uint8_t smem_unaligned[];
    uint8_t *s_raw = smem_unaligned + ((1024u - (smem_u32(smem_unaligned) & 1023u)) & 1023u); // swizzle atoms are 1024-byte aligned
    uint8_t *s_w = s_raw + kRaw * kRawBytes;
    float *s_ex = reinterpret_cast<float *>(s_w + kWBytes);
    __shared__ __align__(8) uint64_t s_wfull, s_rfull[kRaw], s_rempty[kRaw], s_afull[kAStages], s_aempty[kAStages], s_tfull[2], s_tempty[2];
    __shared__ uint32_t s_tmem;
    __shared__ float s_bias[8];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int items = a.n * 2;
    const int my_items = static_cast<int>(blockIdx.x) < items ? (items - 1 - static_cast<int>(blockIdx.x)) / static_cast<int>(gridDim.x) + 1 : 0;
    if (tid == 0) {
        mbar_init(&s_wfull, 1);
        for (int i = 0; i < kRaw; ++i) { mbar_init(&s_rfull[i], 1); mbar_init(&s_rempty[i], 4); }
        for (int i = 0; i < kAStages; ++i) { mbar_init(&s_afull[i], 4); mbar_init(&s_aempty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&s_tfull[i], 1); mbar_init(&s_tempty[i], 128); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < 8) s_bias[tid] = tid < 5 ? a.bias[tid] : 0.0f;
    if (warp == kMmaWarp) tmem_alloc(&s_tmem, kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;
    if (a.pdl == 1) pdl_trigger(); // conv2 of the same forward pass may move in as this grid's CTAs retire (ahk_tc.cuh)

    if (warp == kLoadWarp) {
        // ---------------------------------------------------------------- loader: one tiled TMA copy per plane (as conv1p_kernel)
        if (lane == 0) {
            auto frame_rows = [&](int t, int (&p)[3]) { // row index of (frame, colour 0, row 0); bit 31: warm-up buffer
                const int sample = (static_cast<int>(blockIdx.x) + t * static_cast<int>(gridDim.x)) >> 1;
                const int steps = a.warm != nullptr ? static_cast<int>(a.meta[sample] & 0xFFFFu) : 0x7FFFFFFF;
#pragma unroll
                for (int age = 0; age < 3; ++age)
                    p[age] = age >= steps ? static_cast<int>((static_cast<uint32_t>(2 - (age - steps)) * a.n + sample) * (3u * kDim) | 0x80000000u)
                                          : ((a.head + a.depth - age) % a.depth * a.n + sample) * (3 * kDim);
            };
            int cur[3] = {0, 0, 0}, nxt[3] = {0, 0, 0};
            if (my_items > 0) frame_rows(0, cur);
            uint32_t fill = 0;
            for (int t = 0; t < my_items; ++t) {
                if (t + 1 < my_items) frame_rows(t + 1, nxt);
                const int item = static_cast<int>(blockIdx.x) + t * static_cast<int>(gridDim.x);
                const int row0 = (item & 1) ? 60 : 0;
#pragma unroll
                for (int plane = 0; plane < 9; ++plane, ++fill) {
                    const uint32_t r = fill % kRaw;
                    if (fill >= kRaw) mbar_wait_sleep(&s_rempty[r], ((fill / kRaw) - 1u) & 1u);
                    mbar_expect_tx(&s_rfull[r], kRawRows * kDim);
                    const int f = cur[plane / 3];
                    const int row = (f & 0x7FFFFFFF) + (plane % 3) * kDim + row0;
                    const CUtensorMap *tm = f < 0 ? &tm_warm : &tm_ring;
                    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                                 :: "r"(smem_u32(s_raw + r * kRawBytes)), "l"(reinterpret_cast<uint64_t>(tm)), "r"(0), "r"(row), "r"(smem_u32(&s_rfull[r])) : "memory");
                }
#pragma unroll
                for (int age = 0; age < 3; ++age) cur[age] = nxt[age];
            }
        }
    } else if (warp < kMmaWarp) {
        // ---------------------------------------------------------------- converters: raw uint8 rows -> fp16 A tiles in TMEM
        // warp = (group, lane quarter); thread = TMEM lane r = 23 q + il; group g takes the planes with fill % kGroups == g
        const int quarter = warp & 3, group = warp >> 2;
        int q, il;
        const bool active = lane_row(quarter * 32 + lane, q, il);
        uint32_t off[3][4]; // the four 8-pixel pieces 24 q + 8 j of image row 3 il + c in the swizzled stage
#pragma unroll
        for (int c = 0; c < 3; ++c)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int row = 3 * il + c, o = 24 * q + 8 * j;
                off[c][j] = active ? row * kDim + ((((o >> 4) ^ (row & 7)) << 4) | (o & 8)) : 0u;
            }
        const uint32_t lane_base = tmem + (static_cast<uint32_t>(quarter * 32) << 16) + kACol;
        // (lanes r >= 115 and, in the second half of a sample, image row 128 hold whatever the copy brought: finite fp16
        // either way, and they only reach accumulator rows / kernel rows (ky = 8: zero weights) that are never used)
        const uint32_t fills = static_cast<uint32_t>(my_items) * 9u;
#pragma unroll 1
        for (uint32_t fill = group; fill < fills; fill += kGroups) {
            const uint32_t rs = fill % kRaw, st = fill % kAStages;
            mbar_wait_sleep(&s_rfull[rs], (fill / kRaw) & 1u);
            uint2 px[3][4];
            const uint8_t *raw = s_raw + rs * kRawBytes;
#ifndef AHK_C1T_NO_CONVERT // timing experiment (tools/build_variant.sh): the pipeline without the converters' loads, conversions and stores
#pragma unroll
            for (int c = 0; c < 3; ++c)
#pragma unroll
                for (int j = 0; j < 4; ++j) px[c][j] = *reinterpret_cast<const uint2 *>(raw + off[c][j]);
#endif
            if (fill >= kAStages) mbar_wait_sleep(&s_aempty[st], ((fill / kAStages) - 1u) & 1u);
            tc_fence_after();
            const uint32_t stage = lane_base + st * kAStageCols;
#ifndef AHK_C1T_NO_CONVERT
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                // the class's 32 pixels, converted once, as ONE 16-column strip: phase ph reads its 16-pixel window from columns 4 ph .. 4 ph + 7
                uint32_t h[16];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const uint2 lo = kDenormal ? u8x4_to_f16x4_denormal(px[c][j].x) : u8x4_to_f16x4(px[c][j].x);
                    const uint2 hi = kDenormal ? u8x4_to_f16x4_denormal(px[c][j].y) : u8x4_to_f16x4(px[c][j].y);
                    h[4 * j] = lo.x; h[4 * j + 1] = lo.y; h[4 * j + 2] = hi.x; h[4 * j + 3] = hi.y;
                }
                tmem_store16(stage + c * kAClassCols, h);
            }
#else
            (void)stage; (void)raw; (void)px;
#endif
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) { mbar_arrive(&s_afull[st]); mbar_arrive(&s_rempty[rs]); } // every pixel register has been consumed: the raw stage is free
        }
    } else if (warp == kMmaWarp) {
        // ---------------------------------------------------------------- MMA issue
        // The WHOLE warp runs the loop and one elected lane issues: with every operand warp-uniform the compiler feeds the
        // MMA from uniform registers; under `if (lane == 0)` it wrapped each MMA in an election loop of ~13 instructions
        // (~100 cycles per MMA for the single issuing thread — the limit of conv1p_kernel and of the first version of this one).
        if (lane == 0) {
            mbar_expect_tx(&s_wfull, kWBytes);
            for (int i = 0; i < 9; ++i) bulk_load(s_w + i * (kWBytes / 9), reinterpret_cast<const uint8_t *>(a.weights) + i * (kWBytes / 9), kWBytes / 9, &s_wfull);
        }
        __syncwarp();
        mbar_wait_sleep(&s_wfull, 0);
        constexpr uint32_t kIdesc = instr_desc_f16(kN);
        const uint64_t w_desc = smem_desc(smem_u32(s_w), (kN / 8) * 128, 128);
        uint32_t fill = 0;
        for (int t = 0; t < my_items; ++t) {
            const uint32_t buf = t & 1u;
            if (t >= 2) mbar_wait_sleep(&s_tempty[buf], ((t >> 1) - 1) & 1);
            tc_fence_after();
            const uint32_t d_base = tmem + buf * kBufCols;
#pragma unroll 1
            for (int plane = 0; plane < 9; ++plane, ++fill) {
                const uint32_t st = fill % kAStages;
                mbar_wait_sleep(&s_afull[st], (fill / kAStages) & 1);
                tc_fence_after();
                const uint32_t a_stage = tmem + kACol + st * kAStageCols;
                const uint64_t bd0 = w_desc + static_cast<uint64_t>(plane * (9 * kWBlock / 16));
                if (elect_one()) {
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
#pragma unroll
                        for (int ph = 0; ph < 3; ++ph)
#ifndef AHK_C1T_NO_MMA // timing experiment (tools/build_variant.sh): the pipeline without its MMAs
                            umma_f16_ts(d_base + ph * kN, a_stage + c * kAClassCols + ph * 4, bd0 + static_cast<uint64_t>((c * 3 + ph) * (kWBlock / 16)), kIdesc, (plane | c) != 0);
#else
                            ;
#endif
                    }
                    umma_commit(&s_aempty[st]);
                    if (plane == 8) umma_commit(&s_tfull[buf]);
                }
                __syncwarp();
            }
        }
    } else {
        // ---------------------------------------------------------------- epilogue: 4 warps, thread = TMEM lane = M row 23 q + il
        // two sets of 4 warps, set s taking the items of accumulator buffer s (one warp per scheduler runs at the latency of
        // its own instruction chain: with a single set the epilogue was busy 70 % of the time)
        const int q4 = warp & 3, set = (warp - (kLoadWarp + 1)) >> 2;
        int q, il;
        const bool row_ok = lane_row(q4 * 32 + lane, q, il);
        s_ex += set * (kExBytes / 4);
        for (int t = set; t < my_items; t += kEpiSets) {
            const int item = static_cast<int>(blockIdx.x) + t * static_cast<int>(gridDim.x);
            const int sample = item >> 1, y0 = (item & 1) ? 20 : 0, ny = (item & 1) ? 21 : 20;
            const uint32_t buf = t & 1u;
            mbar_wait_sleep(&s_tfull[buf], (t >> 1) & 1);
            tc_fence_after();
#ifdef AHK_C1T_NO_EPI // timing experiment (tools/build_variant.sh): the pipeline without the epilogue's TMEM loads, exchange and stores
            tc_fence_before();
            mbar_arrive(&s_tempty[buf]);
            (void)sample; (void)y0; (void)ny; (void)row_ok; (void)q; (void)il;
            continue;
#endif
#pragma unroll 1
            for (int ph = 0; ph < 3; ++ph) {
                const uint32_t taddr = tmem + buf * kBufCols + ph * kN + (static_cast<uint32_t>(q4 * 32) << 16);
                uint32_t v0[16], v1[16], v2[16]; // kernel-row groups kv = 0, 1, 2: [3 outputs x 5 channels]
                tmem_load16_async(taddr, v0);
                tmem_load16_async(taddr + 16, v1);
                tmem_load16_async(taddr + 32, v2);
                tmem_wait_ld();
                if (ph == 2) { // the item's accumulators are all in registers: the MMA warp may overwrite them
                    tc_fence_before();
                    mbar_arrive(&s_tempty[buf]);
                }
                // kernel-row group kv of input-row index il belongs to output row il - kv
                // (exchange layout [kv][e][q][il] x 16 bytes: consecutive lanes = consecutive il = consecutive 16 bytes; as
                // [il][q][16 floats] rows were 320 bytes apart and every 128-bit access took 4x its wavefronts — the shared-memory
                // pipe was 77 % busy, two thirds of it here)
                float4 *ex4 = reinterpret_cast<float4 *>(s_ex);
                if (row_ok && il >= 1 && il - 1 < ny) {
#pragma unroll
                    for (int e = 0; e < 4; ++e) ex4[((0 * 4 + e) * kQ + q) * kRowsQ + il - 1] = make_float4(__uint_as_float(v1[4 * e]), __uint_as_float(v1[4 * e + 1]), __uint_as_float(v1[4 * e + 2]), __uint_as_float(v1[4 * e + 3]));
                }
                if (row_ok && il >= 2 && il - 2 < ny) {
#pragma unroll
                    for (int e = 0; e < 4; ++e) ex4[((1 * 4 + e) * kQ + q) * kRowsQ + il - 2] = make_float4(__uint_as_float(v2[4 * e]), __uint_as_float(v2[4 * e + 1]), __uint_as_float(v2[4 * e + 2]), __uint_as_float(v2[4 * e + 3]));
                }
                asm volatile("bar.sync %0, 128;" :: "r"(set + 1) : "memory");
                if (row_ok && il < ny) {
                    float sum[16];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float4 u = ex4[((0 * 4 + e) * kQ + q) * kRowsQ + il], w = ex4[((1 * 4 + e) * kQ + q) * kRowsQ + il];
                        sum[4 * e] = __uint_as_float(v0[4 * e]) + u.x + w.x;
                        sum[4 * e + 1] = __uint_as_float(v0[4 * e + 1]) + u.y + w.y;
                        sum[4 * e + 2] = __uint_as_float(v0[4 * e + 2]) + u.z + w.z;
                        sum[4 * e + 3] = __uint_as_float(v0[4 * e + 3]) + u.w + w.w;
                    }
                    const int y = y0 + il, x0 = 8 * q + 3 * ph;
                    uint4 *out = reinterpret_cast<uint4 *>(a.out + (static_cast<size_t>(sample) * kP1 + y * kC1 + x0) * kAct1C);
#pragma unroll
                    for (int j = 0; j < 3; ++j) {
                        if (j < 2 || ph < 2 || q == kQ - 1) { // phase 2's third output is the next block's first, except x = 40
                            float c[5];
#pragma unroll
                            for (int ch = 0; ch < 5; ++ch) c[ch] = fmaxf(sum[j * 5 + ch] * (kDenormal ? 131072.0f : 1.0f / 128.0f) + s_bias[ch], 0.0f); // (x - 128) / 128: processor.py:39-42 (2^17 = 2^24 / 128)
                            const __nv_bfloat162 p01 = __floats2bfloat162_rn(c[0], c[1]), p23 = __floats2bfloat162_rn(c[2], c[3]), p4 = __floats2bfloat162_rn(c[4], 0.0f);
                            out[j] = make_uint4(*reinterpret_cast<const uint32_t *>(&p01), *reinterpret_cast<const uint32_t *>(&p23), *reinterpret_cast<const uint32_t *>(&p4), 0u);
                        }
                    }
                }
                asm volatile("bar.sync %0, 128;" :: "r"(set + 1) : "memory"); // the exchange buffer is free for the next phase
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (a.pdl == 2) pdl_trigger();
    if (warp == kMmaWarp) tmem_dealloc(tmem, kTmemCols);
}

template <bool kDenormal, int kGroups>
cudaError_t launch_conv1t_as(const Args &a, int grid, cudaStream_t s)
{
    cudaError_t e = cudaFuncSetAttribute(conv1t_kernel<kDenormal, kGroups>, cudaFuncAttributeMaxDynamicSharedMemorySize, c1t::kSmem); // per device
    if (e != cudaSuccess) return e;
    const FrameMaps *m = nullptr;
    if ((e = frame_maps(a, &m)) != cudaSuccess) return e;
    conv1t_kernel<kDenormal, kGroups><<<grid, c1t::cta_threads(kGroups), c1t::kSmem, s>>>(a, m->ring, m->side);
    return cudaGetLastError();
}

cudaError_t launch_conv1t(const Args &a, int grid, bool denormal, cudaStream_t s, int groups = 4)
{
    if (!denormal) return launch_conv1t_as<false, 2>(a, grid, s);
    if (groups == 3) return launch_conv1t_as<true, 3>(a, grid, s);
    if (groups == 2) return launch_conv1t_as<true, 2>(a, grid, s);
    return launch_conv1t_as<true, 4>(a, grid, s);
}

// ------------------------------------------------------------------------------------------------ conv2 / conv3, pipelined
// The same GEMM tiles as layer_kernel, as a persistent warp-specialised pipeline (the layer_kernel form of these two
// layers sat at 5-6 % tensor-pipe activity: every chunk was gathered, stored, barriered and multiplied in sequence):
//   * the layer's whole B operand (conv2 8 KB, conv3 64 KB) is resident in shared memory, loaded once per CTA;
//   * 4 producer warps gather A chunks with cp.async (16 bytes per row and K core, straight into the canonical layout,
//     zero-fill past the last row), kDepth chunks in flight per thread and no register staging; a chunk is published
//     (fence.proxy.async + mbarrier arrive) once this thread's copies of it have landed;
//   * 1 warp issues the MMAs and recycles stages with tcgen05.commit; 4 warps run the epilogue out of a double-buffered
//     TMEM accumulator while the next tile is being multiplied.
namespace ws {
constexpr int kProducers = 128, kCta = kProducers + 32 + 128;
// A stages / chunks in flight per producer thread besides the one being issued.  conv2 (8 KB of weights, 2 chunks per
// tile) runs three CTAs per SM with a short pipeline each, conv3 (64 KB of weights) one CTA with a deep one.
template <int L> struct Cfg { static constexpr int kStages = 8, kDepth = 6, kPerSm = 1; };
template <> struct Cfg<kConv2> { static constexpr int kStages = 4, kDepth = 2, kPerSm = 3; };
template <> struct Cfg<kG2> { static constexpr int kStages = 4, kDepth = 2, kPerSm = 2; }; // 16 KB of weights, 4 chunks per tile
// K-core pitch of the A stage: 16 bytes more than the 16 row groups, so that the 8 lanes that fetch one row's 8 K cores
// (128 contiguous global bytes: one L1 wavefront instead of two half-used ones) land in 8 different bank groups
constexpr int kLbo = 16 * 128 + 16, kStage = 8 * kLbo;
template <int L> constexpr int smem_bytes() { return Cfg<L>::kStages * kStage + Traits<L>::K * Traits<L>::N * 2; }
} // namespace ws

__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src, uint32_t bytes)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" :: "r"(dst), "l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory"); }

template <int L>
__global__ void __launch_bounds__(ws::kCta, ws::Cfg<L>::kPerSm) conv_ws_kernel(Args a)
{
    using namespace ws;
    constexpr int kStages = Cfg<L>::kStages, kDepth = Cfg<L>::kDepth;
    using T = Traits<L>;
    constexpr int N = T::N, NCH = T::K / kBK;
    constexpr int kBBytes = N * kBK * 2;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *s_a = smem;                       // [kStages][kAStage]
    uint8_t *s_b = smem + kStages * kStage;   // [NCH][kBBytes], resident
    __shared__ __align__(8) uint64_t s_wfull, s_full[kStages], s_empty[kStages], s_tfull[2], s_tempty[2];
    __shared__ uint32_t s_tmem;
    __shared__ float s_bias[N];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n_tiles = (a.m_total + kBM - 1) / kBM;
    const int my_tiles = static_cast<int>(blockIdx.x) < n_tiles ? (n_tiles - 1 - blockIdx.x) / gridDim.x + 1 : 0;
    if (tid == 0) {
        mbar_init(&s_wfull, 1);
        for (int i = 0; i < kStages; ++i) { mbar_init(&s_full[i], kProducers); mbar_init(&s_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&s_tfull[i], 1); mbar_init(&s_tempty[i], 128); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < N) s_bias[tid] = a.bias[tid];
    constexpr uint32_t kBufCols = N <= 32 ? 32 : (N <= 64 ? 64 : 128); // one accumulator; the allocation (two of them) must be a power of two
    if (warp == kProducers / 32) tmem_alloc(&s_tmem, 2 * kBufCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;

    if (tid < kProducers) {
        // ---------------------------------------------------------------- producers
        const int kc_lane = lane & 7, g = warp * 4 + (lane >> 3); // rows i * 16 + g, i = 0..7
        const uint32_t dst0 = smem_u32(s_a) + kc_lane * kLbo + g * 16;
        const int chunks = my_tiles * NCH;
        int r0[8];
        unsigned ok = 0u;
        auto publish = [&](int c) {
            fence_async_smem();
            mbar_arrive(&s_full[c % kStages]);
        };
#pragma unroll 1
        for (int c = 0; c < chunks; ++c) {
            const int ch = c % NCH;
            if (ch == 0) { // per-row constants of the new tile
                const int tile = blockIdx.x + (c / NCH) * gridDim.x;
                ok = 0u;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int row = tile * kBM + i * 16 + g;
                    r0[i] = 0;
                    if (row < a.m_total) {
                        ok |= 1u << i;
                        if (L == kConv2) {
                            const int s = row / kP2, p = row - s * kP2, y = p / kC2, x = p - y * kC2;
                            r0[i] = (s * kP1 + 2 * y * kC1 + 2 * x) * kAct1C;
                        } else if (L == kConv3) {
                            const int s = row / kP3, p = row - s * kP3, y = p / kC3, x = p - y * kC3;
                            r0[i] = (s * kP2 + 2 * y * kC2 + 2 * x) * 32;
                        } else if (L == kG2) { // 16 channels, stride 2
                            const int s = row / kGP2, p = row - s * kGP2, y = p / kGS2, x = p - y * kGS2;
                            r0[i] = (s * kGP1 + 2 * y * kGS1 + 2 * x) * 16;
                        } else if (L == kG3) { // 32 channels (24 real), stride 1
                            const int s = row / kGP3, p = row - s * kGP3, y = p / kGS3, x = p - y * kGS3;
                            r0[i] = (s * kGP2 + y * kGS2 + x) * 32;
                        } else { // kG4
                            const int s = row / kGP4, p = row - s * kGP4, y = p / kGS4, x = p - y * kGS4;
                            r0[i] = (s * kGP3 + y * kGS3 + x) * 32;
                        }
                    }
                }
            }
            const int st = c % kStages;
            if (c >= kStages) mbar_wait(&s_empty[st], ((c / kStages) - 1) & 1);
            const int kc = ch * 8 + kc_lane;
            int off;
            // the same K order as gather<L>: a chunk of 8 K cores is 128 contiguous bytes of the input (conv2, kG2: the four
            // pixels of one kernel row; 32-channel inputs: two adjacent pixels)
            if (L == kConv2) off = ((kc >> 2) * kC1 + (kc & 3)) * kAct1C;
            else if (L == kG2) { const int tap = kc >> 1; off = ((tap >> 2) * kGS1 + (tap & 3)) * 16 + (kc & 1) * 8; }
            else {
                constexpr int kSide = L == kConv3 ? kC2 : (L == kG3 ? kGS2 : kGS3);
                const int tap = kc >> 2;
                off = ((tap >> 2) * kSide + (tap & 3)) * 32 + (kc & 3) * 8;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const bool v = (ok >> i) & 1u;
                cp_async16(dst0 + st * kStage + i * 256, a.in + (v ? r0[i] + off : 0), v ? 16u : 0u);
            }
            cp_async_commit();
            if (c >= kDepth) { cp_async_wait<kDepth>(); publish(c - kDepth); }
        }
        cp_async_wait<0>();
        for (int c = chunks > kDepth ? chunks - kDepth : 0; c < chunks; ++c) publish(c);
    } else if (warp == kProducers / 32) {
        // ---------------------------------------------------------------- MMA issue (one thread)
        // (the whole warp runs the loop with uniform operands, one elected lane issues: see conv1t_kernel)
        if (lane == 0) {
            mbar_expect_tx(&s_wfull, NCH * kBBytes);
            for (int i = 0; i < NCH; ++i) bulk_load(s_b + i * kBBytes, reinterpret_cast<const uint8_t *>(a.weights) + i * kBBytes, kBBytes, &s_wfull);
        }
        __syncwarp();
        mbar_wait(&s_wfull, 0);
        constexpr uint32_t kIdesc = instr_desc(N);
        const uint64_t b_desc = smem_desc(smem_u32(s_b), N / 8 * 128, 128);
        int c = 0;
        for (int t = 0; t < my_tiles; ++t) {
            const uint32_t buf = t & 1;
            if (t >= 2) mbar_wait(&s_tempty[buf], ((t >> 1) - 1) & 1);
            tc_fence_after();
#pragma unroll 1
            for (int ch = 0; ch < NCH; ++ch, ++c) {
                const int st = c % kStages;
                mbar_wait(&s_full[st], (c / kStages) & 1);
                tc_fence_after();
                const uint64_t ad0 = smem_desc(smem_u32(s_a + st * kStage), kLbo, 128), bd0 = b_desc + static_cast<uint64_t>(ch * (kBBytes / 16));
                if (elect_one()) {
#pragma unroll
                    for (int j = 0; j < kBK / 16; ++j)
                        umma_bf16(tmem + buf * kBufCols, ad0 + static_cast<uint64_t>(j * 2 * kLbo / 16), bd0 + static_cast<uint64_t>(j * 2 * (N / 8 * 128) / 16), kIdesc, (ch | j) != 0);
                    umma_commit(&s_empty[st]);
                    if (ch == NCH - 1) umma_commit(&s_tfull[buf]);
                }
                __syncwarp();
            }
        }
    } else {
        // ---------------------------------------------------------------- epilogue: thread = TMEM lane = tile row
        const int q4 = warp & 3;
        for (int t = 0; t < my_tiles; ++t) {
            const int tile = blockIdx.x + t * gridDim.x;
            const uint32_t buf = t & 1;
            const int row = tile * kBM + q4 * 32 + lane;
            mbar_wait(&s_tfull[buf], (t >> 1) & 1);
            tc_fence_after();
            const uint32_t taddr = tmem + buf * kBufCols + (static_cast<uint32_t>(q4 * 32) << 16);
            uint32_t v[N];
#pragma unroll
            for (int c0 = 0; c0 < N; c0 += 16) tmem_load16_async(taddr + c0, *reinterpret_cast<uint32_t (*)[16]>(&v[c0]));
            tmem_wait_ld();
            tc_fence_before();
            mbar_arrive(&s_tempty[buf]);
            if (row < a.m_total) {
                uint4 *out = reinterpret_cast<uint4 *>(a.out + static_cast<size_t>(row) * N);
#pragma unroll
                for (int c0 = 0; c0 < N; c0 += 8) {
                    uint32_t pk[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float lo = fmaxf(__uint_as_float(v[c0 + 2 * j]) + s_bias[c0 + 2 * j], 0.0f);
                        const float hi = fmaxf(__uint_as_float(v[c0 + 2 * j + 1]) + s_bias[c0 + 2 * j + 1], 0.0f);
                        const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
                        pk[j] = *reinterpret_cast<const uint32_t *>(&h);
                    }
                    out[c0 / 8] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kProducers / 32) tmem_dealloc(tmem, 2 * kBufCols);
}

template <int L>
cudaError_t launch_ws(const Args &a, int tiles, int sms, cudaStream_t s)
{
    static int per_sm = 0;
    if (per_sm == 0) {
        cudaError_t e = cudaFuncSetAttribute(conv_ws_kernel<L>, cudaFuncAttributeMaxDynamicSharedMemorySize, ws::smem_bytes<L>());
        if (e != cudaSuccess) return e;
        per_sm = ws::Cfg<L>::kPerSm;
    }
    const int resident = sms * per_sm;
    conv_ws_kernel<L><<<tiles < resident ? tiles : resident, ws::kCta, ws::smem_bytes<L>(), s>>>(a);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ conv2, parity arrays
// conv2 (5 -> 32 channels, 4x4, stride 2: models.py:30-32) on act1 [sample][41 x 41][8 channels] (16 bytes per pixel).
// Both earlier forms moved every act1 pixel to shared memory four times — conv_ws_kernel's cp.async gather (0.155 ms per
// 8192 samples, bound by shared-memory wavefronts) and the tensor-map im2col (0.21-0.25 ms, bound by L2 -> SM traffic:
// 1.2 GB for 220 MB of activations).  Here a sample is copied ONCE (one 26.9 KB bulk copy) and rearranged by four warps
// into four PARITY ARRAYS P[y & 1][x & 1][y >> 1][x >> 1] (24 slots per row of 21): with the stride of 2 folded into the
// parity, output pixel m = y2 * 24 + x2 needs, for tap (ky, kx), slot m + (ky >> 1) * 24 + (kx >> 1) of array
// P[ky & 1][kx & 1] — LINEAR in m.  So the A operand of every MMA is a plain window of an array: rows 16 bytes apart
// (no-swizzle K-major core matrices), the two K cores of an MMA = taps kx and kx + 2 = neighbouring slots (LBO = 16: the
// second core of row m is the first of row m + 1).  8 MMAs (4 ky x 2 kx parities) of K = 16 per 128-row tile, 4 tiles per
// sample (456 rows, 361 of them output pixels), weights (8 KB) resident, nothing is gathered.
namespace c2p {
constexpr int kCta = 320;                                   // warp 0 loader, 1-4 rearrange, 5 MMA, 6-9 epilogue
constexpr int kRawBytes = kP1 * kAct1C * 2, kRawPitch = 27008; // 26 896 B per sample
constexpr int kSlots = 24, kRowsP = 21, kArrBytes = kRowsP * kSlots * 16, kParPitch = 33536; // 4 arrays + 1 152 B the last tile reads past them
constexpr int kWBytes = 128 * 32 * 2;
constexpr int kSmem = 128 + 2 * kRawPitch + 2 * kParPitch + kWBytes; // 129 408 B
constexpr int kTmemCols = 256;                              // 2 buffers x 4 tiles x 32 columns
} // namespace c2p

__global__ void __launch_bounds__(c2p::kCta, 1) conv2p_kernel(Args a)
{
    using namespace c2p;
    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint8_t *s_base = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
    uint8_t *s_rawbuf = s_base, *s_par = s_base + 2 * kRawPitch, *s_w = s_par + 2 * kParPitch;
    __shared__ __align__(8) uint64_t s_wfull, s_rfull[2], s_rempty[2], s_pfull[2], s_pempty[2], s_tfull[2], s_tempty[2];
    __shared__ uint32_t s_tmem;
    __shared__ float s_bias[32];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int my_items = static_cast<int>(blockIdx.x) < a.n ? (a.n - 1 - static_cast<int>(blockIdx.x)) / static_cast<int>(gridDim.x) + 1 : 0;
    if (tid == 0) {
        mbar_init(&s_wfull, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(&s_rfull[i], 1); mbar_init(&s_rempty[i], 128); mbar_init(&s_pfull[i], 128); mbar_init(&s_pempty[i], 1);
            mbar_init(&s_tfull[i], 1); mbar_init(&s_tempty[i], 128);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < 32) s_bias[tid] = a.bias[tid];
    for (int i = tid; i < 2 * kParPitch / 16; i += kCta) reinterpret_cast<uint4 *>(s_par)[i] = make_uint4(0, 0, 0, 0); // unused slots stay finite
    if (warp == 5) tmem_alloc(&s_tmem, kTmemCols);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;

    if (warp == 0) {
        // ---------------------------------------------------------------- loader: one bulk copy per sample
        if (lane == 0) {
            mbar_expect_tx(&s_wfull, kWBytes);
            bulk_load(s_w, a.weights, kWBytes, &s_wfull);
            for (int t = 0; t < my_items; ++t) {
                const uint32_t rb = t & 1u;
                if (t >= 2) mbar_wait_sleep(&s_rempty[rb], ((t >> 1) - 1) & 1);
                const int sample = static_cast<int>(blockIdx.x) + t * static_cast<int>(gridDim.x);
                mbar_expect_tx(&s_rfull[rb], kRawBytes);
                bulk_load(s_rawbuf + rb * kRawPitch, a.in + static_cast<size_t>(sample) * kP1 * kAct1C, kRawBytes, &s_rfull[rb]);
            }
        }
    } else if (warp < 5) {
        // ---------------------------------------------------------------- rearrange: pixel (y, x) -> P[y & 1][x & 1][y >> 1][x >> 1]
        const int rt = tid - 32;
        for (int t = 0; t < my_items; ++t) {
            const uint32_t b = t & 1u;
            mbar_wait_sleep(&s_rfull[b], (t >> 1) & 1);
            if (t >= 2) mbar_wait_sleep(&s_pempty[b], ((t >> 1) - 1) & 1);
            const uint4 *src = reinterpret_cast<const uint4 *>(s_rawbuf + b * kRawPitch);
            uint8_t *dst = s_par + b * kParPitch;
#pragma unroll 2
            for (int i = rt; i < kP1; i += 128) {
                const int y = i / kC1, x = i - y * kC1;
                *reinterpret_cast<uint4 *>(dst + ((y & 1) * 2 + (x & 1)) * kArrBytes + ((y >> 1) * kSlots + (x >> 1)) * 16) = src[i];
            }
            fence_async_smem();
            mbar_arrive(&s_pfull[b]);
            mbar_arrive(&s_rempty[b]); // the stores above consumed every loaded pixel: the raw buffer is free
        }
    } else if (warp == 5) {
        // ---------------------------------------------------------------- MMA issue (whole warp, one elected lane)
        mbar_wait_sleep(&s_wfull, 0);
        constexpr uint32_t kIdesc = instr_desc(32);
        const uint64_t w_desc = smem_desc(smem_u32(s_w), 4 * 128, 128);
        for (int t = 0; t < my_items; ++t) {
            const uint32_t b = t & 1u;
            if (t >= 2) mbar_wait_sleep(&s_tempty[b], ((t >> 1) - 1) & 1);
            mbar_wait_sleep(&s_pfull[b], (t >> 1) & 1);
            tc_fence_after();
            const uint64_t a_desc = smem_desc(smem_u32(s_par + b * kParPitch), 16, 128);
            if (elect_one()) {
#pragma unroll
                for (int tile = 0; tile < 4; ++tile) {
#pragma unroll
                    for (int q = 0; q < 8; ++q) { // q = ky * 2 + kx parity
                        const int ky = q >> 1, kxp = q & 1;
                        const uint32_t a_off = ((ky & 1) * 2 + kxp) * kArrBytes + ((ky >> 1) * kSlots + tile * 128) * 16;
                        // B block q of the packed weights: blob q / 4 (64 K each), K cores (q % 4) * 2 and + 1
                        const uint32_t b_off = (q >> 2) * (64 * 32 * 2) + (q & 3) * 2 * (4 * 128);
                        umma_bf16(tmem + b * 128 + tile * 32, a_desc + static_cast<uint64_t>(a_off >> 4), w_desc + static_cast<uint64_t>(b_off >> 4), kIdesc, q != 0);
                    }
                }
                umma_commit(&s_pempty[b]);
                umma_commit(&s_tfull[b]);
            }
            __syncwarp();
        }
    } else {
        // ---------------------------------------------------------------- epilogue: thread = TMEM lane = row of each of the 4 tiles
        const int q4 = warp & 3, r = q4 * 32 + lane;
        for (int t = 0; t < my_items; ++t) {
            const uint32_t b = t & 1u;
            const int sample = static_cast<int>(blockIdx.x) + t * static_cast<int>(gridDim.x);
            mbar_wait_sleep(&s_tfull[b], (t >> 1) & 1);
            tc_fence_after();
#pragma unroll 1
            for (int tile = 0; tile < 4; ++tile) {
                const uint32_t taddr = tmem + b * 128 + tile * 32 + (static_cast<uint32_t>(q4 * 32) << 16);
                uint32_t v[32];
                tmem_load16_async(taddr, *reinterpret_cast<uint32_t (*)[16]>(&v[0]));
                tmem_load16_async(taddr + 16, *reinterpret_cast<uint32_t (*)[16]>(&v[16]));
                tmem_wait_ld();
                if (tile == 3) { // the sample's accumulators are in registers
                    tc_fence_before();
                    mbar_arrive(&s_tempty[b]);
                }
                const int m = tile * 128 + r, y2 = m / kSlots, x2 = m - y2 * kSlots;
                if (y2 < kC2 && x2 < kC2) {
                    uint4 *out = reinterpret_cast<uint4 *>(a.out + (static_cast<size_t>(sample) * kP2 + y2 * kC2 + x2) * 32);
#pragma unroll
                    for (int c0 = 0; c0 < 32; c0 += 8) {
                        uint32_t pk[4];
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const float lo = fmaxf(__uint_as_float(v[c0 + 2 * j]) + s_bias[c0 + 2 * j], 0.0f);
                            const float hi = fmaxf(__uint_as_float(v[c0 + 2 * j + 1]) + s_bias[c0 + 2 * j + 1], 0.0f);
                            const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
                            pk[j] = *reinterpret_cast<const uint32_t *>(&h);
                        }
                        out[c0 / 8] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 5) tmem_dealloc(tmem, kTmemCols);
}

cudaError_t launch_conv2p(const Args &a, int sms, cudaStream_t s)
{
    cudaError_t e = cudaFuncSetAttribute(conv2p_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, c2p::kSmem);
    if (e != cudaSuccess) return e;
    conv2p_kernel<<<a.n < sms ? a.n : sms, c2p::kCta, c2p::kSmem, s>>>(a);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ conv2 / conv3, A in TMEM
// The im2col of a small convolution built where the tensor core reads it: an item (conv2: one sample of act1, 26.9 KB;
// conv3: two samples of act2, 46 KB) is bulk-copied to shared memory ONCE, and converter threads — thread = TMEM lane =
// output pixel — copy each MMA's 32-byte A row (two adjacent taps of 8 channels for conv2, half a tap's 32 channels for
// conv3) from the raw sample straight into tensor memory (two 16-byte loads, one tcgen05.st.x8).  The MMAs take A from TMEM
// (floor 128 N / 256 cycles instead of the 64 cycles an A tile costs from shared memory), B = the layer's packed weights,
// resident.  Activations cross L2 -> SM once instead of four times.  Shared-memory bank conflicts of the converter loads decided
// both layers (a 16-byte load runs in phases of eight lanes, one phase wants eight different 16-byte bank groups):
//   conv2: neighbouring lanes are 32 bytes apart -> four bank groups, twice; lanes 4-7 of each phase fetch their two halves in the
//          other order.  0.122 -> 0.100 ms per 8192 samples (parity arrays 0.116, cp.async gather 0.155, tensor-map im2col 0.21-0.25).
//   conv3: pixels are 64 bytes and neighbouring lanes 128 bytes apart -> ONE bank group per phase with a plain copy (0.165 ms, 80 % of
//          the shared-memory pipe, profiles/conv_tm_r2v_ncu.txt).  The item is now fetched as three tiled TMA copies of 128-byte rows
//          with the 128-byte swizzle (chunk c of row r at c ^ (r & 7)): the eight lanes of a phase land in eight bank groups.
//          0.064 ms against 0.084 for the TMA-fed GEMM (same box), so both layers run here by default; AHK_CONVTM=2 / 0 for A/B.
namespace ctm {
template <int L> struct Cfg;
template <> struct Cfg<kConv2> {
    static constexpr int N = 32, kMmas = 8, kGroup = 8, kTiles = 3, kAStages = 4;     // MMAs per tile, MMAs per A stage, tiles per item
    static constexpr int kSamples = 1, kRowsOut = kP2, kItemBytes = kP1 * kAct1C * 2, kRawPitch = 27008, kWBytes = 128 * 32 * 2;
};
template <> struct Cfg<kConv3> {
    static constexpr int N = 64, kMmas = 32, kGroup = 16, kTiles = 1, kAStages = 3;
    // the item arrives as three tiled TMA copies of 128 rows of 128 bytes (two pixels) with the 128-byte swizzle: 361 rows used
    static constexpr int kSamples = 2, kRowsOut = 2 * kP3, kItemBytes = 2 * kP2 * 32 * 2, kRawPitch = 3 * 128 * 128, kWBytes = 512 * 64 * 2;
};
constexpr int kCta = 256 + 64 + 128; // 8 converter warps (2 groups x 4 lane quarters), loader, MMA, 4 epilogue warps
// raw item buffers: with two, the copy of item t + 2 could only start once item t was converted whole — one item time to cover the
// latency of a 27 KB copy (ncu: DRAM at 49 % with no unit busy, profiles/conv_tm_r2s5b_ncu.txt).  conv2: 4 buffers, 0.0965 -> 0.0927 ms per
// 8192 samples; conv3 (48 KB items): 3 buffers were SLOWER than 2 (0.0644 vs 0.0605 ms, same box) and it keeps 2
template <int L> constexpr int default_raw() { return L == kConv2 ? 4 : 2; }
template <int L> constexpr int smem_bytes(int raw) { return 1024 + raw * Cfg<L>::kRawPitch + Cfg<L>::kWBytes; }
} // namespace ctm

template <int L, int kRaw>
__global__ void __launch_bounds__(ctm::kCta, 1) conv_tm_kernel(Args a, const __grid_constant__ CUtensorMap tm_in)
{
    using C = ctm::Cfg<L>;
    constexpr int N = C::N, kStagesPerItem = C::kTiles * (C::kMmas / C::kGroup), kACols = C::kGroup * 8, kDCols = C::kTiles * N;
    constexpr int kABase = 2 * kDCols;
    static_assert(kABase + C::kAStages * kACols <= 512, "TMEM");
    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint8_t *s_base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u); // swizzle atoms are 1024-byte aligned
    uint8_t *s_in = s_base, *s_w = s_base + kRaw * C::kRawPitch;
    __shared__ __align__(8) uint64_t s_wfull, s_rfull[kRaw], s_rempty[kRaw], s_afull[C::kAStages], s_aempty[C::kAStages], s_tfull[2], s_tempty[2];
    __shared__ uint32_t s_tmem;
    __shared__ float s_bias[N];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int items = (a.n + C::kSamples - 1) / C::kSamples;
    const int my_items = static_cast<int>(blockIdx.x) < items ? (items - 1 - static_cast<int>(blockIdx.x)) / static_cast<int>(gridDim.x) + 1 : 0;
    if (tid == 0) {
        mbar_init(&s_wfull, 1);
        for (int i = 0; i < kRaw; ++i) { mbar_init(&s_rfull[i], 1); mbar_init(&s_rempty[i], 8); }
        for (int i = 0; i < 2; ++i) { mbar_init(&s_tfull[i], 1); mbar_init(&s_tempty[i], 128); }
        for (int i = 0; i < C::kAStages; ++i) { mbar_init(&s_afull[i], 4); mbar_init(&s_aempty[i], 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < N) s_bias[tid] = a.bias[tid];
    if (warp == 8) tmem_alloc(&s_tmem, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;
    // programmatic dependent launch (ahk_tc.cuh): the weights are fetched under the previous layer's tail, its output after the wait
    if (a.pdl == 1) pdl_trigger();
    if (warp == 9 && lane == 0) {
        mbar_expect_tx(&s_wfull, C::kWBytes);
        for (int i = 0; i < C::kWBytes; i += 8192) bulk_load(s_w + i, reinterpret_cast<const uint8_t *>(a.weights) + i, 8192, &s_wfull);
    }
    pdl_wait();

    if (warp == 9) {
        // ---------------------------------------------------------------- loader: one bulk copy per item
        if (lane == 0) {
            for (int t = 0; t < my_items; ++t) {
                const uint32_t rb = static_cast<uint32_t>(t) % kRaw;
                if (t >= kRaw) mbar_wait_sleep(&s_rempty[rb], ((t / kRaw) - 1) & 1);
                const int item = static_cast<int>(blockIdx.x) + t * static_cast<int>(gridDim.x);
                const int samples = (item + 1) * C::kSamples <= a.n ? C::kSamples : a.n - item * C::kSamples;
                const uint32_t bytes = static_cast<uint32_t>(samples) * (C::kItemBytes / C::kSamples);
                if (L == kConv3) { // act2 as rows of 128 bytes: rows [361 item, + 384), swizzled; rows past the end of act2 are zero-filled
                    mbar_expect_tx(&s_rfull[rb], C::kRawPitch);
                    for (int j = 0; j < 3; ++j) ahk::tg::tma_2d(s_in + rb * C::kRawPitch + j * (128 * 128), &tm_in, 0, item * kP2 + j * 128, &s_rfull[rb]);
                    continue;
                }
                mbar_expect_tx(&s_rfull[rb], bytes);
                bulk_load(s_in + rb * C::kRawPitch, reinterpret_cast<const uint8_t *>(a.in) + static_cast<size_t>(item) * C::kItemBytes, bytes, &s_rfull[rb]);
            }
        }
    } else if (warp < 8) {
        // ---------------------------------------------------------------- converters: raw sample -> A rows in TMEM
        const int quarter = warp & 3, group = warp >> 2, r = quarter * 32 + lane;
        const uint32_t lane_base = tmem + (static_cast<uint32_t>(quarter * 32) << 16) + kABase;
        const bool swap_halves = ((lane >> 2) & 1) && !a.no_swap;
        uint32_t f = 0; // A stages so far (both groups count all of them)
        for (int t = 0; t < my_items; ++t) {
            const uint32_t rb = static_cast<uint32_t>(t) % kRaw;
            mbar_wait_sleep(&s_rfull[rb], (t / kRaw) & 1);
            const uint8_t *raw = s_in + rb * C::kRawPitch;
#pragma unroll 1
            for (int sidx = 0; sidx < kStagesPerItem; ++sidx, ++f) {
                if (static_cast<int>(f & 1u) != group) continue;
                const int tile = sidx / (C::kMmas / C::kGroup), q0 = (sidx % (C::kMmas / C::kGroup)) * C::kGroup;
                // the row's base offset in the raw item
                uint32_t off;
                if (L == kConv2) {
                    int p = tile * 128 + r;
                    p = p < kP2 ? p : 0;
                    const int y = p / kC2, x = p - y * kC2;
                    off = ((2 * y) * kC1 + 2 * x) * (kAct1C * 2);
                } else {
                    const int smp = r >> 6, p = r & 63, y = p >> 3, x = p & 7;
                    off = smp * (kP2 * 64) + ((2 * y) * kC2 + 2 * x) * 64;
                }
                const uint32_t st = f % C::kAStages;
                if (f >= static_cast<uint32_t>(C::kAStages)) mbar_wait_sleep(&s_aempty[st], ((f / C::kAStages) - 1u) & 1u);
                tc_fence_after();
                const uint32_t stage = lane_base + st * kACols;
#pragma unroll
                for (int i = 0; i < C::kGroup; ++i) {
                    const int q = q0 + i;
                    uint32_t tapoff;
                    if (L == kConv2) tapoff = ((q >> 1) * kC1 + (q & 1) * 2) * (kAct1C * 2);           // taps (ky, 2h), (ky, 2h + 1): 32 contiguous bytes
                    else { const int tap = q >> 1; tapoff = ((tap >> 2) * kC2 + (tap & 3)) * 64 + (q & 1) * 32; } // half of a tap's 32 channels
                    // conv2: neighbouring lanes are 32 bytes apart, so the eight lanes of a 16-byte load phase would cover only four
                    // of the eight 16-byte bank groups, twice; lanes 4-7 of each phase fetch their two halves in the other order
                    const uint32_t first = (L == kConv2 && swap_halves) ? 16u : 0u;
                    uint32_t at0 = off + tapoff + first, at1 = off + tapoff + (first ^ 16u);
                    if (L == kConv3) { // 128-byte swizzle of the TMA copy: 16-byte chunk c of row r sits at chunk c ^ (r & 7).  Neighbouring lanes are
                        // neighbouring rows (output pixels two input pixels = 128 bytes apart), so a load phase covers all eight bank groups
                        at0 ^= ((at0 >> 7) & 7u) << 4;
                        at1 ^= ((at1 >> 7) & 7u) << 4;
                    }
                    const uint4 va = *reinterpret_cast<const uint4 *>(raw + at0), vb = *reinterpret_cast<const uint4 *>(raw + at1);
                    const uint4 v0 = first ? vb : va, v1 = first ? va : vb;
                    tmem_store8(stage + i * 8, v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w);
                }
                tmem_wait_st();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&s_afull[st]);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&s_rempty[rb]); // this warp's loads of the item have all been consumed by their stores
        }
    } else if (warp == 8) {
        // ---------------------------------------------------------------- MMA issue (whole warp, one elected lane)
        mbar_wait_sleep(&s_wfull, 0);
        constexpr uint32_t kIdesc = instr_desc(N);
        const uint64_t w_desc = smem_desc(smem_u32(s_w), (N / 8) * 128, 128);
        uint32_t f = 0;
        for (int t = 0; t < my_items; ++t) {
            const uint32_t b = t & 1u;
            if (t >= 2) mbar_wait_sleep(&s_tempty[b], ((t >> 1) - 1) & 1);
            tc_fence_after();
#pragma unroll 1
            for (int sidx = 0; sidx < kStagesPerItem; ++sidx, ++f) {
                const int tile = sidx / (C::kMmas / C::kGroup), q0 = (sidx % (C::kMmas / C::kGroup)) * C::kGroup;
                const uint32_t st = f % C::kAStages;
                mbar_wait_sleep(&s_afull[st], (f / C::kAStages) & 1u);
                tc_fence_after();
                const uint32_t a_stage = tmem + kABase + st * kACols, d = tmem + b * kDCols + tile * N;
                // B of MMA q: blob q / 4 (64 K each), K cores (q % 4) * 2 and + 1
                const uint64_t bd0 = w_desc + static_cast<uint64_t>(((q0 >> 2) * (64 * N * 2)) >> 4);
                if (elect_one()) {
#pragma unroll
                    for (int i = 0; i < C::kGroup; ++i)
                        umma_f16_ts_bf16(d, a_stage + i * 8, bd0 + static_cast<uint64_t>((((i >> 2) * (64 * N * 2)) + (i & 3) * 2 * ((N / 8) * 128)) >> 4), kIdesc, (q0 + i) != 0);
                    umma_commit(&s_aempty[st]);
                    if (sidx == kStagesPerItem - 1) umma_commit(&s_tfull[b]);
                }
                __syncwarp();
            }
        }
    } else {
        // ---------------------------------------------------------------- epilogue: thread = TMEM lane = output pixel of each tile
        const int q4 = warp & 3, r = q4 * 32 + lane;
        for (int t = 0; t < my_items; ++t) {
            const uint32_t b = t & 1u;
            const int item = static_cast<int>(blockIdx.x) + t * static_cast<int>(gridDim.x);
            mbar_wait_sleep(&s_tfull[b], (t >> 1) & 1);
            tc_fence_after();
#pragma unroll 1
            for (int tile = 0; tile < C::kTiles; ++tile) {
                const uint32_t taddr = tmem + b * kDCols + tile * N + (static_cast<uint32_t>(q4 * 32) << 16);
                uint32_t v[N];
#pragma unroll
                for (int c0 = 0; c0 < N; c0 += 16) tmem_load16_async(taddr + c0, *reinterpret_cast<uint32_t (*)[16]>(&v[c0]));
                tmem_wait_ld();
                if (tile == C::kTiles - 1) { // the item's accumulators are in registers
                    tc_fence_before();
                    mbar_arrive(&s_tempty[b]);
                }
                const int p = tile * 128 + r;
                const long long row = static_cast<long long>(item) * C::kRowsOut + p;
                if (p < C::kRowsOut && row < static_cast<long long>(a.n) * (C::kRowsOut / C::kSamples)) {
                    uint4 *out = reinterpret_cast<uint4 *>(a.out + row * N);
#pragma unroll
                    for (int c0 = 0; c0 < N; c0 += 8) {
                        uint32_t pk[4];
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const float lo = fmaxf(__uint_as_float(v[c0 + 2 * j]) + s_bias[c0 + 2 * j], 0.0f);
                            const float hi = fmaxf(__uint_as_float(v[c0 + 2 * j + 1]) + s_bias[c0 + 2 * j + 1], 0.0f);
                            const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
                            pk[j] = *reinterpret_cast<const uint32_t *>(&h);
                        }
                        out[c0 / 8] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (a.pdl == 2) pdl_trigger();
    if (warp == 8) tmem_dealloc(tmem, 512);
}

template <int L, int kRaw>
cudaError_t launch_tm_as(const Args &a, const CUtensorMap &tm_in, int sms, cudaStream_t s, bool pdl)
{
    constexpr int bytes = ctm::smem_bytes<L>(kRaw);
    static_assert(bytes <= 227 * 1024, "raw stages do not fit");
    cudaError_t e = cudaFuncSetAttribute(conv_tm_kernel<L, kRaw>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return e;
    const int items = (a.n + ctm::Cfg<L>::kSamples - 1) / ctm::Cfg<L>::kSamples;
    return launch_kernel(conv_tm_kernel<L, kRaw>, dim3(items < sms ? items : sms), dim3(ctm::kCta), bytes, s, pdl, a, tm_in);
}

// two_raw: the double-buffered form (A/B switch AHK_CONVTM_RAW=2)
template <int L>
cudaError_t launch_tm(const Args &a, const CUtensorMap &tm_in, int sms, cudaStream_t s, bool pdl = false, bool two_raw = false)
{
    return two_raw ? launch_tm_as<L, 2>(a, tm_in, sms, s, pdl) : launch_tm_as<L, ctm::default_raw<L>()>(a, tm_in, sms, s, pdl);
}

// ------------------------------------------------------------------------------------------------ conv2 -> conv3 fused
// One kernel for both layers of an INFERENCE forward: act2 (23 KB per sample) never leaves the SM — conv2's epilogue writes it
// into shared memory in the 128-byte-swizzled row layout conv_tm_kernel<conv3> reads, instead of 143 MB of stores + 189 MB of
// loads per 8192 samples through HBM, and one launch less.  Same roles, converters, MMAs and arithmetic as the two
// conv_tm_kernel instantiations (results are bit-identical); a work item is a PAIR of samples (one conv3 tile).
//   shared memory: act2, a ring of 3 SAMPLE buffers (69 KB: conv2's epilogue writes the next pair's first sample while conv3 of this pair
//                  is still being converted; with one pair buffer the converters waited 15 % of the time for act2, profiles/conv23_r2s5n_ncu.txt)
//                  | conv3 weights 64 KB | conv2 weights 8 KB | ring of 3 act1 samples (79 KB)
//   tensor memory: conv2 accumulators, 6 tile slots x 32 columns (handed over slot by slot) | conv3 accumulators 2 x 64 |
//                  3 A stages x 64 columns (8 MMAs each)
//   every role walks the same software-pipelined stage sequence: conv2(0); then per step conv2(t + 1), conv3(t) — conv3 of a
//   pair is converted while the next pair's conv2 accumulators wait in their slots, and conv2's epilogue of pair t + 1 starts as
//   soon as the converters have read pair t's act2.
namespace c23 {
// 8 converter warps (2 groups x 4 lane quarters), MMA warp, loader warp, kEpi sets of 4 epilogue warps.  Two sets (default): with one the
// epilogue warps were busy 77 % of the kernel and the converters waited 16 % of it for act2 (ncu role breakdown of call s5t)
constexpr int cta_threads(int epi) { return 256 + 64 + epi * 128; }
constexpr int kRaw = 3, kRawPitch = 27008, kAct2Ring = 3, kAct2Sample = 23552, kAct2Bytes = kAct2Ring * kAct2Sample; // 361 pixels x 64 B = 23 104, rounded up to a multiple of 1024
constexpr int kW2Bytes = 128 * 32 * 2, kW3Bytes = 512 * 64 * 2;
constexpr int kSmem = 1024 + kAct2Bytes + kW3Bytes + kW2Bytes + kRaw * kRawPitch;
constexpr int kD2 = 0, kD3 = 192, kA = 320, kACols = 64, kAStages = 3;
static_assert(kSmem <= 227 * 1024 && kA + kAStages * kACols == 512, "budgets");
} // namespace c23

template <int kEpi>
__global__ void __launch_bounds__(c23::cta_threads(kEpi), 1) conv23_kernel(Args a, const uint4 *__restrict__ w3, const float *__restrict__ b3, __nv_bfloat16 *__restrict__ out3)
{
    using namespace c23;
    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint8_t *s_act2 = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u); // swizzle atoms are 1024-byte aligned
    uint8_t *s_w3 = s_act2 + kAct2Bytes, *s_w2 = s_w3 + kW3Bytes, *s_in = s_w2 + kW2Bytes;
    __shared__ __align__(8) uint64_t s_wfull, s_rfull[kRaw], s_rempty[kRaw], s_afull[kAStages], s_aempty[kAStages], s_t2full[6], s_t2empty[6], s_t3full[2],
        s_t3empty[2], s_a2full[kAct2Ring], s_a2empty[kAct2Ring];
    __shared__ uint32_t s_tmem;
    __shared__ float s_bias2[32], s_bias3[64];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int items = (a.n + 1) / 2;
    const int my_items = static_cast<int>(blockIdx.x) < items ? (items - 1 - static_cast<int>(blockIdx.x)) / static_cast<int>(gridDim.x) + 1 : 0;
    if (tid == 0) {
        mbar_init(&s_wfull, 1);
        for (int i = 0; i < kAct2Ring; ++i) { mbar_init(&s_a2full[i], 128 * kEpi); mbar_init(&s_a2empty[i], 8); }
        for (int i = 0; i < kRaw; ++i) { mbar_init(&s_rfull[i], 1); mbar_init(&s_rempty[i], 8); }
        for (int i = 0; i < kAStages; ++i) { mbar_init(&s_afull[i], 4); mbar_init(&s_aempty[i], 1); }
        for (int i = 0; i < 6; ++i) { mbar_init(&s_t2full[i], 1); mbar_init(&s_t2empty[i], 128); }
        for (int i = 0; i < 2; ++i) { mbar_init(&s_t3full[i], 1); mbar_init(&s_t3empty[i], 128); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < 32) s_bias2[tid] = a.bias[tid];
    if (tid >= 64 && tid < 128) s_bias3[tid - 64] = b3[tid - 64];
    if (warp == 8) tmem_alloc(&s_tmem, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;
    if (a.pdl == 1) pdl_trigger();
    if (warp == 9 && lane == 0) { // both weight sets: not the previous kernel's output, fetched before the wait
        mbar_expect_tx(&s_wfull, kW2Bytes + kW3Bytes);
        bulk_load(s_w2, a.weights, kW2Bytes, &s_wfull);
        for (int i = 0; i < kW3Bytes; i += 8192) bulk_load(s_w3 + i, reinterpret_cast<const uint8_t *>(w3) + i, 8192, &s_wfull);
    }
    pdl_wait();

    if (warp == 9) {
        // ---------------------------------------------------------------- loader: one bulk copy per SAMPLE of act1
        if (lane == 0) {
            for (int u = 0; u < 2 * my_items; ++u) {
                const uint32_t rb = static_cast<uint32_t>(u) % kRaw;
                if (u >= kRaw) mbar_wait_sleep(&s_rempty[rb], ((u / kRaw) - 1) & 1);
                const int item = static_cast<int>(blockIdx.x) + (u >> 1) * static_cast<int>(gridDim.x);
                int sample = item * 2 + (u & 1);
                sample = sample < a.n ? sample : a.n - 1; // the missing half of an odd batch's last pair: its rows are never stored
                mbar_expect_tx(&s_rfull[rb], kP1 * kAct1C * 2);
                bulk_load(s_in + rb * kRawPitch, reinterpret_cast<const uint8_t *>(a.in) + static_cast<size_t>(sample) * (kP1 * kAct1C * 2), kP1 * kAct1C * 2, &s_rfull[rb]);
            }
        }
    } else if (warp < 8) {
        // ---------------------------------------------------------------- converters: A rows of both layers -> TMEM
        const int quarter = warp & 3, r = quarter * 32 + lane;
        const uint32_t group = static_cast<uint32_t>(warp >> 2);
        const uint32_t lane_base = tmem + (static_cast<uint32_t>(quarter * 32) << 16) + kA;
        const bool swap_halves = ((lane >> 2) & 1) && !a.no_swap;
        uint32_t f = 0; // A stages so far (both groups count all of them)
        for (int s = 0; s <= my_items; ++s) {
            if (s < my_items) {
#pragma unroll 1
                for (int j = 0; j < 2; ++j) { // conv2 of pair s: three tiles per sample
                    const uint32_t u = 2u * static_cast<uint32_t>(s) + j, rb = u % kRaw;
                    mbar_wait_sleep(&s_rfull[rb], (u / kRaw) & 1u);
                    const uint8_t *raw = s_in + rb * kRawPitch;
#pragma unroll 1
                    for (int k = 0; k < 3; ++k, ++f) {
                        if ((f & 1u) != group) continue;
                        int p = k * 128 + r;
                        p = p < kP2 ? p : 0;
                        const int y = p / kC2, x = p - y * kC2;
                        const uint32_t off = ((2 * y) * kC1 + 2 * x) * (kAct1C * 2);
                        const uint32_t st = f % kAStages;
                        if (f >= static_cast<uint32_t>(kAStages)) mbar_wait_sleep(&s_aempty[st], ((f / kAStages) - 1u) & 1u);
                        tc_fence_after();
                        const uint32_t stage = lane_base + st * kACols;
#pragma unroll
                        for (int i = 0; i < 8; ++i) { // taps (ky, 2h), (ky, 2h + 1): 32 contiguous bytes; lanes 4-7 of a load phase swap their halves
                            const uint32_t tapoff = ((i >> 1) * kC1 + (i & 1) * 2) * (kAct1C * 2), first = swap_halves ? 16u : 0u;
                            const uint4 va = *reinterpret_cast<const uint4 *>(raw + off + tapoff + first), vb = *reinterpret_cast<const uint4 *>(raw + off + tapoff + (first ^ 16u));
                            const uint4 v0 = first ? vb : va, v1 = first ? va : vb;
                            tmem_store8(stage + i * 8, v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w);
                        }
                        tmem_wait_st();
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&s_afull[st]);
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&s_rempty[rb]); // this warp's loads of the sample have been consumed by their stores
                }
            }
            if (s >= 1) { // conv3 of pair s - 1, from act2 in shared memory
                const int t = s - 1;
                const uint32_t u0 = 2u * static_cast<uint32_t>(t), u1 = u0 + 1u; // the pair's samples: act2 ring slots u % 3
                mbar_wait_sleep(&s_a2full[u0 % kAct2Ring], (u0 / kAct2Ring) & 1u);
                mbar_wait_sleep(&s_a2full[u1 % kAct2Ring], (u1 / kAct2Ring) & 1u);
                const int smp = r >> 6, pp = r & 63, y = pp >> 3, x = pp & 7;
                const uint32_t off = ((smp ? u1 : u0) % kAct2Ring) * kAct2Sample + ((2 * y) * kC2 + 2 * x) * 64;
#pragma unroll 1
                for (int g = 0; g < 4; ++g, ++f) {
                    if ((f & 1u) != group) continue;
                    const uint32_t st = f % kAStages;
                    if (f >= static_cast<uint32_t>(kAStages)) mbar_wait_sleep(&s_aempty[st], ((f / kAStages) - 1u) & 1u);
                    tc_fence_after();
                    const uint32_t stage = lane_base + st * kACols;
#pragma unroll
                    for (int i = 0; i < 8; ++i) { // half of a tap's 32 channels; addresses follow the 128-byte swizzle conv2's epilogue wrote
                        const int q = g * 8 + i, tap = q >> 1;
                        uint32_t at0 = off + ((tap >> 2) * kC2 + (tap & 3)) * 64 + (q & 1) * 32, at1 = at0 + 16;
                        at0 ^= ((at0 >> 7) & 7u) << 4;
                        at1 ^= ((at1 >> 7) & 7u) << 4;
                        const uint4 v0 = *reinterpret_cast<const uint4 *>(s_act2 + at0), v1 = *reinterpret_cast<const uint4 *>(s_act2 + at1);
                        tmem_store8(stage + i * 8, v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w);
                    }
                    tmem_wait_st();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&s_afull[st]);
                }
                __syncwarp();
                if (lane == 0) { mbar_arrive(&s_a2empty[u0 % kAct2Ring]); mbar_arrive(&s_a2empty[u1 % kAct2Ring]); } // act2 of the pair has been read by this warp
            }
        }
    } else if (warp == 8) {
        // ---------------------------------------------------------------- MMA issue (whole warp, one elected lane)
        mbar_wait_sleep(&s_wfull, 0);
        constexpr uint32_t kIdesc2 = instr_desc(32), kIdesc3 = instr_desc(64);
        const uint64_t w2_desc = smem_desc(smem_u32(s_w2), (32 / 8) * 128, 128), w3_desc = smem_desc(smem_u32(s_w3), (64 / 8) * 128, 128);
        uint32_t f = 0;
        for (int s = 0; s <= my_items; ++s) {
            if (s < my_items) {
#pragma unroll 1
                for (int slot = 0; slot < 6; ++slot, ++f) {
                    if (s >= 1) mbar_wait_sleep(&s_t2empty[slot], (s - 1) & 1);
                    const uint32_t st = f % kAStages;
                    mbar_wait_sleep(&s_afull[st], (f / kAStages) & 1u);
                    tc_fence_after();
                    const uint32_t a_stage = tmem + kA + st * kACols, d = tmem + kD2 + slot * 32;
                    if (elect_one()) {
#pragma unroll
                        for (int i = 0; i < 8; ++i)
                            umma_f16_ts_bf16(d, a_stage + i * 8, w2_desc + static_cast<uint64_t>((((i >> 2) * (64 * 32 * 2)) + (i & 3) * 2 * ((32 / 8) * 128)) >> 4), kIdesc2, i != 0);
                        umma_commit(&s_aempty[st]);
                        umma_commit(&s_t2full[slot]);
                    }
                    __syncwarp();
                }
            }
            if (s >= 1) {
                const int t = s - 1;
                const uint32_t b = t & 1u;
                if (t >= 2) mbar_wait_sleep(&s_t3empty[b], ((t >> 1) - 1) & 1);
                tc_fence_after();
#pragma unroll 1
                for (int g = 0; g < 4; ++g, ++f) {
                    const uint32_t st = f % kAStages;
                    mbar_wait_sleep(&s_afull[st], (f / kAStages) & 1u);
                    tc_fence_after();
                    const uint32_t a_stage = tmem + kA + st * kACols, d = tmem + kD3 + b * 64;
                    const uint64_t bd0 = w3_desc + static_cast<uint64_t>(((2 * g) * (64 * 64 * 2)) >> 4); // MMA q = 8 g + i: blob q / 4
                    if (elect_one()) {
#pragma unroll
                        for (int i = 0; i < 8; ++i)
                            umma_f16_ts_bf16(d, a_stage + i * 8, bd0 + static_cast<uint64_t>((((i >> 2) * (64 * 64 * 2)) + (i & 3) * 2 * ((64 / 8) * 128)) >> 4), kIdesc3, (g | i) != 0);
                        umma_commit(&s_aempty[st]);
                        if (g == 3) umma_commit(&s_t3full[b]);
                    }
                    __syncwarp();
                }
            }
        }
    } else {
        // ---------------------------------------------------------------- epilogues: thread = TMEM lane = output pixel of a tile
        // set e of kEpi takes the conv2 tile slots with slot % kEpi == e and the conv3 accumulators of the pairs with t % kEpi == e
        const int q4 = warp & 3, r = q4 * 32 + lane, set = (warp - 10) >> 2;
        const uint32_t lane_addr = tmem + (static_cast<uint32_t>(q4 * 32) << 16);
        auto epilogue2 = [&](int t) { // conv2 accumulators of pair t -> bias, ReLU, bf16 -> act2 in shared memory (swizzled rows of two pixels)
#pragma unroll 1
            for (int slot = 0; slot < 6; ++slot) {
                if (slot % kEpi != set) continue;
                const uint32_t u = 2u * static_cast<uint32_t>(t) + (slot >= 3 ? 1u : 0u), ring = u % kAct2Ring; // sample and its act2 ring slot
                const int s0 = slot >= 3 ? 3 : 0; // this set's first and last tile of the sample
                const int first_mine = s0 + ((set - s0 % kEpi) % kEpi + kEpi) % kEpi;
                int last_mine = first_mine;
                while (last_mine + kEpi < s0 + 3) last_mine += kEpi;
                if (slot == first_mine && u >= static_cast<uint32_t>(kAct2Ring)) mbar_wait_sleep(&s_a2empty[ring], ((u / kAct2Ring) - 1u) & 1u);
                mbar_wait_sleep(&s_t2full[slot], t & 1);
                tc_fence_after();
                uint32_t v[32];
                tmem_load16_async(lane_addr + kD2 + slot * 32, *reinterpret_cast<uint32_t (*)[16]>(&v[0]));
                tmem_load16_async(lane_addr + kD2 + slot * 32 + 16, *reinterpret_cast<uint32_t (*)[16]>(&v[16]));
                tmem_wait_ld();
                tc_fence_before();
                mbar_arrive(&s_t2empty[slot]);
                const int j = slot / 3, pix = (slot - 3 * j) * 128 + r;
                if (pix < kP2) {
                    const uint32_t lin = ring * kAct2Sample + static_cast<uint32_t>(pix) * 64u;
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        uint32_t pk[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float lo = fmaxf(__uint_as_float(v[c * 8 + 2 * e]) + s_bias2[c * 8 + 2 * e], 0.0f);
                            const float hi = fmaxf(__uint_as_float(v[c * 8 + 2 * e + 1]) + s_bias2[c * 8 + 2 * e + 1], 0.0f);
                            const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
                            pk[e] = *reinterpret_cast<const uint32_t *>(&h);
                        }
                        uint32_t at = lin + c * 16u;
                        at ^= ((at >> 7) & 7u) << 4;
                        *reinterpret_cast<uint4 *>(s_act2 + at) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                    }
                }
                if (slot == last_mine) mbar_arrive(&s_a2full[ring]); // release: this thread's rows of the sample are in shared memory
            }
        };
        auto epilogue3 = [&](int t) { // conv3 accumulators of pair t -> act3 in global memory
            const uint32_t b = t & 1u;
            const int item = static_cast<int>(blockIdx.x) + t * static_cast<int>(gridDim.x);
            mbar_wait_sleep(&s_t3full[b], (t >> 1) & 1);
            tc_fence_after();
            uint32_t v[64];
#pragma unroll
            for (int c0 = 0; c0 < 64; c0 += 16) tmem_load16_async(lane_addr + kD3 + b * 64 + c0, *reinterpret_cast<uint32_t (*)[16]>(&v[c0]));
            tmem_wait_ld();
            tc_fence_before();
            mbar_arrive(&s_t3empty[b]);
            const long long row = static_cast<long long>(item) * 128 + r;
            if (row < static_cast<long long>(a.n) * kP3) {
                uint4 *out = reinterpret_cast<uint4 *>(out3 + row * 64);
#pragma unroll
                for (int c0 = 0; c0 < 64; c0 += 8) {
                    uint32_t pk[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float lo = fmaxf(__uint_as_float(v[c0 + 2 * e]) + s_bias3[c0 + 2 * e], 0.0f);
                        const float hi = fmaxf(__uint_as_float(v[c0 + 2 * e + 1]) + s_bias3[c0 + 2 * e + 1], 0.0f);
                        const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
                        pk[e] = *reinterpret_cast<const uint32_t *>(&h);
                    }
                    out[c0 / 8] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                }
            }
        };
        if (my_items > 0) epilogue2(0);
        for (int t = 0; t < my_items; ++t) {
            if (t + 1 < my_items) epilogue2(t + 1);
            if (t % kEpi == set) epilogue3(t);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (a.pdl == 2) pdl_trigger();
    if (warp == 8) tmem_dealloc(tmem, 512);
}

// a: conv2's arguments (in = act1, weights, bias); w3 / b3 / out3: conv3's packed weights, bias and act3
template <int kEpi>
cudaError_t launch_conv23_as(const Args &a, const uint4 *w3, const float *b3, __nv_bfloat16 *out3, int sms, cudaStream_t s, bool pdl)
{
    cudaError_t e = cudaFuncSetAttribute(conv23_kernel<kEpi>, cudaFuncAttributeMaxDynamicSharedMemorySize, c23::kSmem);
    if (e != cudaSuccess) return e;
    const int items = (a.n + 1) / 2;
    return launch_kernel(conv23_kernel<kEpi>, dim3(items < sms ? items : sms), dim3(c23::cta_threads(kEpi)), c23::kSmem, s, pdl, a, w3, b3, out3);
}

cudaError_t launch_conv23(const Args &a, const uint4 *w3, const float *b3, __nv_bfloat16 *out3, int sms, cudaStream_t s, bool pdl, int epi = 2)
{
    return epi == 1 ? launch_conv23_as<1>(a, w3, b3, out3, sms, s, pdl) : launch_conv23_as<2>(a, w3, b3, out3, sms, s, pdl);
}

// ------------------------------------------------------------------------------------------------ gray-diff input
// DataProcessor's 'gray-diff' observation (processor.py:53-64) straight from the ring, as the network's bf16 input:
// per colour the uint8 wrap-around double difference oldest - middle - newest, 0.2989 R + 0.5870 G + 0.1140 B,
// (x - 128) / 128.  One thread per 16 pixels.
__global__ void __launch_bounds__(256) gray_diff_kernel(const uint8_t *__restrict__ frames, __nv_bfloat16 *__restrict__ out, int n, int head,
                                                        int depth, const uint8_t *__restrict__ warm, const uint32_t *__restrict__ meta)
{
    const size_t total = static_cast<size_t>(n) * 1024;
    for (size_t idx = blockIdx.x * 256ull + threadIdx.x; idx < total; idx += gridDim.x * 256ull) {
        const size_t env = idx >> 10, v = idx & 1023;
        uint4 q[3][3]; // [age][colour]
#pragma unroll
        for (int age = 0; age < 3; ++age) {
            const uint4 *f = reinterpret_cast<const uint4 *>(stack_frame(frames, warm, meta, depth, static_cast<size_t>(n), head, env, age));
#pragma unroll
            for (int c = 0; c < 3; ++c) q[age][c] = __ldg(f + c * 1024 + v);
        }
        uint32_t packed[8];
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            float g[4];
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                float ch[3];
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const uint32_t o = (&q[2][c].x)[w] >> (8 * b), m = (&q[1][c].x)[w] >> (8 * b), nw = (&q[0][c].x)[w] >> (8 * b);
                    ch[c] = static_cast<float>((o - m - nw) & 0xFFu);
                }
                g[b] = (0.2989f * ch[0] + 0.5870f * ch[1] + 0.1140f * ch[2] - 128.0f) * (1.0f / 128.0f);
            }
            const __nv_bfloat162 lo = __floats2bfloat162_rn(g[0], g[1]), hi = __floats2bfloat162_rn(g[2], g[3]);
            packed[2 * w] = *reinterpret_cast<const uint32_t *>(&lo);
            packed[2 * w + 1] = *reinterpret_cast<const uint32_t *>(&hi);
        }
        uint4 *dst = reinterpret_cast<uint4 *>(out + (env * 1024 + v) * 16);
        dst[0] = make_uint4(packed[0], packed[1], packed[2], packed[3]);
        dst[1] = make_uint4(packed[4], packed[5], packed[6], packed[7]);
    }
}

// ------------------------------------------------------------------------------------------------ device: fold + pack
// ahk_policy_set_weights_dev: the same folding and packing as the host path below (load_rgb_weights), from float32
// Keras-layout weights that already live in device memory — after an optimizer step nothing has to cross PCIe.
struct DevWeights { const float *f[24]; }; // ahk_policy_weights with device pointers, same order

__device__ __forceinline__ float bn_scale(const float *gamma, const float *var, int c) { return __fdiv_rn(gamma[c], __fsqrt_rn(__fadd_rn(var[c], 1e-3f))); }
__device__ __forceinline__ float bn_shift(const float *gamma, const float *beta, const float *mean, const float *var, int c)
{
    return __fsub_rn(beta[c], __fmul_rn(mean[c], bn_scale(gamma, var, c)));
}
__device__ __forceinline__ uint16_t dev_bf16(float v) { return __bfloat16_as_ushort(__float2bfloat16_rn(v)); }

// layer 1 conv2, 2 conv3, 3 dense1, 4 dense2, 5 out: folded value of GEMM element (k, col), 0 in the padding
template <int L> __device__ __forceinline__ float folded(const DevWeights &w, int k, int col)
{
    if (L == 1) { // k = (ky*4+kx)*8 + c, BN1 folded
        const int tap = k >> 3, c = k & 7;
        return c < 5 ? __fmul_rn(w.f[6][(tap * 5 + c) * 32 + col], bn_scale(w.f[2], w.f[5], c)) : 0.0f;
    }
    if (L == 6) { // conv2 in conv2p_kernel's K order: k = ((ky * 2 + kx parity) * 2 + j) * 8 + c  <->  tap (ky, kx = parity + 2 j)
        const int c = k & 7, j = (k >> 3) & 1, kxp = (k >> 4) & 1, ky = k >> 5, tap = ky * 4 + kxp + 2 * j;
        return c < 5 ? __fmul_rn(w.f[6][(tap * 5 + c) * 32 + col], bn_scale(w.f[2], w.f[5], c)) : 0.0f;
    }
    if (L == 2) return __fmul_rn(w.f[12][k * 64 + col], bn_scale(w.f[8], w.f[11], k & 31)); // k = tap*32 + c, BN2 folded
    if (L == 3) { // ours k = pixel*64 + c; Keras flatten index c*64 + pixel; BN3 folded
        const int px = k >> 6, c = k & 63;
        return __fmul_rn(w.f[18][static_cast<size_t>(c * 64 + px) * 512 + col], bn_scale(w.f[14], w.f[17], c));
    }
    if (L == 4) return w.f[20][k * 512 + col];
    return col < 10 ? w.f[22][k * 10 + col] : 0.0f;
}

template <int L> __global__ void __launch_bounds__(256) pack_b_kernel(DevWeights w, uint16_t *__restrict__ out, int k_total, int n_pad, int n_split)
{
    const int total = k_total * n_pad * n_split, groups = n_pad / 8;
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
        const int n = idx % n_pad, k = (idx / n_pad) % k_total, sp = idx / (n_pad * k_total); // consecutive threads: consecutive columns
        const int ch = k / kBK, kc = (k % kBK) / 8, ke = k % 8;
        const size_t at = ((static_cast<size_t>(sp) * (k_total / kBK) + ch) * 8 + kc) * groups * 64 + static_cast<size_t>(n / 8) * 64 + (n % 8) * 8 + ke;
        out[at] = dev_bf16(folded<L>(w, k, sp * n_pad + n));
    }
}

// the same folded weights as a plain [N][K] bf16 matrix (the B operand of the TMA-fed GEMM, ahk_gemm.cuh)
template <int L> __global__ void __launch_bounds__(256) pack_nk_kernel(DevWeights w, uint16_t *__restrict__ out, int k_total, int n_total)
{
    const int total = k_total * n_total;
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
        const int n = idx % n_total, k = idx / n_total; // consecutive threads: consecutive columns of the Keras matrix
        out[static_cast<size_t>(n) * k_total + k] = dev_bf16(folded<L>(w, k, n));
    }
}

// biases: layer 0 conv1 b - sum(W) (normalisation folded), 1..3 b + sum(W * BN shift of the input channel), 4, 5 copies
// dense1's kernel summed over the 64 pixels of each input channel: w3sum[c][o] = sum_px W[c * 64 + px][o] (what a constant
// BatchNorm shift of channel c contributes to output o) — 64 coalesced rows per thread instead of a 4096-long loop in fold_bias_kernel
__global__ void __launch_bounds__(512) dense1_px_sums_kernel(const float *__restrict__ w, float *__restrict__ w3sum)
{
    const int c = blockIdx.x, o = threadIdx.x;
    double acc = 0.0;
    for (int px = 0; px < 64; ++px) acc += static_cast<double>(w[static_cast<size_t>(c * 64 + px) * 512 + o]);
    w3sum[c * 512 + o] = static_cast<float>(acc);
}

__global__ void __launch_bounds__(128) fold_bias_kernel(DevWeights w, float *b0, float *b1, float *b2, float *b3, float *b4, float *b5, const float *w3sum, int only = -1)
{
    const int layer = blockIdx.y, o = blockIdx.x * blockDim.x + threadIdx.x;
    if (only >= 0 && layer != only) return;
    if (layer == 0) {
        if (o >= 16) return;
        double sum = 0.0;
        if (o < 5) for (int i = 0; i < 576; ++i) sum += static_cast<double>(w.f[0][i * 5 + o]);
        b0[o] = o < 5 ? __fsub_rn(w.f[1][o], static_cast<float>(sum)) : 0.0f;
    } else if (layer == 1) {
        if (o >= 32) return;
        double add = 0.0;
        for (int i = 0; i < 80; ++i) add += static_cast<double>(w.f[6][i * 32 + o]) * static_cast<double>(bn_shift(w.f[2], w.f[3], w.f[4], w.f[5], i % 5));
        b1[o] = __fadd_rn(w.f[7][o], static_cast<float>(add));
    } else if (layer == 2) {
        if (o >= 64) return;
        double add = 0.0;
        for (int i = 0; i < 512; ++i) add += static_cast<double>(w.f[12][i * 64 + o]) * static_cast<double>(bn_shift(w.f[8], w.f[9], w.f[10], w.f[11], i & 31));
        b2[o] = __fadd_rn(w.f[13][o], static_cast<float>(add));
    } else if (layer == 3) {
        if (o >= 512) return;
        double add = 0.0;
        for (int c = 0; c < 64; ++c) add += static_cast<double>(w3sum[c * 512 + o]) * static_cast<double>(bn_shift(w.f[14], w.f[15], w.f[16], w.f[17], c));
        b3[o] = __fadd_rn(w.f[19][o], static_cast<float>(add));
    } else if (layer == 4) {
        if (o < 512) b4[o] = w.f[21][o];
    } else {
        if (o < 16) b5[o] = o < 10 ? w.f[23][o] : 0.0f;
    }
}

// conv1 in Toeplitz form (pack_conv1_toeplitz below): one thread per (plane, class, kernel-row group, phase, output, channel, kx)
__global__ void __launch_bounds__(256) pack_conv1_kernel(DevWeights w, uint16_t *__restrict__ out)
{
    constexpr int N = 144;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 9 * 3 * 3 * 3 * 3 * 5 * 8) return;
    int r = idx;
    const int kx = r % 8; r /= 8;
    const int ch = r % 5; r /= 5;
    const int j = r % 3; r /= 3;
    const int ph = r % 3; r /= 3;
    const int kv = r % 3; r /= 3;
    const int cls = r % 3; r /= 3;
    const int plane = r;
    const int ky = cls + 3 * kv;
    const int k = kx + ph + 3 * j, n = kv * 48 + ph * 16 + j * 5 + ch;
    if (ky >= 8 || k >= 16) return;
    const size_t at = static_cast<size_t>(plane * 3 + cls) * (16 * N) + (k / 8) * (8 * N) + (n / 8) * 64 + (n % 8) * 8 + (k % 8);
    out[at] = dev_bf16(__fdiv_rn(w.f[0][((ky * 8 + kx) * 9 + plane) * 5 + ch], 128.0f));
}

// conv1, phase-separated form (conv1p_kernel): 81 blocks (plane, class, phase) of [K = 16][N = 48] fp16, unscaled weights
__global__ void __launch_bounds__(256) pack_conv1p_kernel(DevWeights w, uint16_t *__restrict__ out)
{
    constexpr int N = 48;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 9 * 3 * 3 * 3 * 3 * 5 * 8) return;
    int r = idx;
    const int kx = r % 8; r /= 8;
    const int ch = r % 5; r /= 5;
    const int j = r % 3; r /= 3;
    const int ph = r % 3; r /= 3;
    const int kv = r % 3; r /= 3;
    const int cls = r % 3; r /= 3;
    const int plane = r;
    const int ky = cls + 3 * kv;
    const int k = kx + ph + 3 * j, n = kv * 16 + j * 5 + ch;
    if (ky >= 8) return;
    const size_t at = static_cast<size_t>((plane * 3 + cls) * 3 + ph) * (16 * N) + (k / 8) * (8 * N) + (n / 8) * 64 + (n % 8) * 8 + (k % 8);
    out[at] = __half_as_ushort(__float2half_rn(w.f[0][((ky * 8 + kx) * 9 + plane) * 5 + ch]));
}

// ------------------------------------------------------------------------------------------------ batch statistics
// Training-mode BatchNormalization (model.fit, rl_agents.py:189,276) normalises with the batch's own per-channel mean and
// (biased) variance of the layer's post-ReLU output.  act: bf16 [rows][kCPad]; sums: double [2][kCPad] (zeroed).
template <int kCPad>
__global__ void __launch_bounds__(256) bn_sums_kernel(const __nv_bfloat16 *__restrict__ act, long long rows, double *__restrict__ sums)
{
    constexpr int kGroups = kCPad / 8; // 16-byte vectors per row
    __shared__ float s_part[256][17];
    const int tid = threadIdx.x, grp = tid % kGroups;
    float sum[8] = {0, 0, 0, 0, 0, 0, 0, 0}, sq[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const long long row_stride = static_cast<long long>(gridDim.x) * (256 / kGroups);
    for (long long row = static_cast<long long>(blockIdx.x) * (256 / kGroups) + tid / kGroups; row < rows; row += row_stride) {
        const uint4 v = __ldg(reinterpret_cast<const uint4 *>(act + row * kCPad + grp * 8));
        const uint32_t wds[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float lo = __uint_as_float(wds[j] << 16), hi = __uint_as_float(wds[j] & 0xFFFF0000u);
            sum[2 * j] += lo; sq[2 * j] += lo * lo;
            sum[2 * j + 1] += hi; sq[2 * j + 1] += hi * hi;
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) { s_part[tid][j] = sum[j]; s_part[tid][8 + j] = sq[j]; }
    __syncthreads();
    if (tid < kCPad * 2) { // one thread per (channel, sum | sum of squares)
        const int c = tid % kCPad, which = tid / kCPad, g = c / 8, j = c % 8;
        double acc = 0.0;
        for (int t = g; t < 256; t += kGroups) acc += static_cast<double>(s_part[t][which * 8 + j]);
        atomicAdd(sums + which * kCPad + c, acc);
    }
}
// mean / biased variance of the batch -> stats[0 .. c) and stats[kStatStride .. ), and the sums cleared for the next use
__global__ void bn_finalize_kernel(double *__restrict__ sums, int c_pad, long long count, float *__restrict__ mean, float *__restrict__ var)
{
    const int c = threadIdx.x;
    if (c >= c_pad) return;
    const double m = sums[c] / static_cast<double>(count), v = sums[c_pad + c] / static_cast<double>(count) - m * m;
    mean[c] = static_cast<float>(m);
    var[c] = static_cast<float>(v > 0.0 ? v : 0.0);
    sums[c] = 0.0;
    sums[c_pad + c] = 0.0;
}

// ------------------------------------------------------------------------------------------------ host: fold + pack
uint16_t to_bf16(float f)
{ // round to nearest even
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7F800000u) == 0x7F800000u) return static_cast<uint16_t>(u >> 16);
    u += 0x7FFFu + ((u >> 16) & 1u);
    return static_cast<uint16_t>(u >> 16);
}

// W: [K][n_real] row-major (fp32, already folded) -> bf16 blobs [n_split][K/64][K core 0..7][n_pad/8 groups][8 n][8 k]
std::vector<uint16_t> pack_b(const std::vector<float> &w, int k_total, int n_real, int n_pad, int n_split)
{
    std::vector<uint16_t> out(static_cast<size_t>(k_total) * n_pad * n_split, 0);
    const int groups = n_pad / 8;
    for (int sp = 0; sp < n_split; ++sp)
        for (int k = 0; k < k_total; ++k)
            for (int n = 0; n < n_pad; ++n) {
                const int col = sp * n_pad + n;
                const float v = col < n_real ? w[static_cast<size_t>(k) * n_real + col] : 0.0f;
                const int ch = k / kBK, kc = (k % kBK) / 8, ke = k % 8;
                const size_t at = ((static_cast<size_t>(sp) * (k_total / kBK) + ch) * 8 + kc) * groups * 64 +
                                  static_cast<size_t>(n / 8) * 64 + (n % 8) * 8 + ke;
                out[at] = to_bf16(v);
            }
    return out;
}


// W: [K][n_real] row-major (fp32, already folded) -> plain [n_real][K] bf16 (the B operand of the TMA-fed GEMM)
std::vector<uint16_t> pack_nk(const std::vector<float> &w, int k_total, int n_real)
{
    std::vector<uint16_t> out(static_cast<size_t>(k_total) * n_real);
    for (int k = 0; k < k_total; ++k)
        for (int n = 0; n < n_real; ++n) out[static_cast<size_t>(n) * k_total + k] = to_bf16(w[static_cast<size_t>(k) * n_real + n]);
    return out;
}

// conv1, Toeplitz form: W [576][5] folded (k = plane*64 + ky*8 + kx) -> 27 blocks (plane, class) of the K-major canonical
// B operand [K = 16][N = 144]: element (n, k) at (k/8) * 2304 + (n/8) * 128 + (n%8) * 16 + (k%8) * 2 bytes;
// n = kv*48 + phase*16 + j*5 + channel holds W[plane][ky = class + 3 kv][kx = k - (phase + 3j)][channel]
// (0 outside the 8 taps and for ky = 8).
std::vector<uint16_t> pack_conv1_toeplitz(const std::vector<float> &wf)
{
    constexpr int N = 144;
    std::vector<uint16_t> out(static_cast<size_t>(27) * 16 * N, 0);
    for (int plane = 0; plane < 9; ++plane)
        for (int cls = 0; cls < 3; ++cls)
            for (int kv = 0; kv < 3; ++kv) {
                const int ky = cls + 3 * kv;
                if (ky >= 8) continue;
                for (int ph = 0; ph < 3; ++ph)
                    for (int j = 0; j < 3; ++j)
                        for (int ch = 0; ch < 5; ++ch)
                            for (int kx = 0; kx < 8; ++kx) {
                                const int k = kx + ph + 3 * j, n = kv * 48 + ph * 16 + j * 5 + ch;
                                if (k >= 16) continue; // phase 2, third output: belongs to the next block's first
                                const size_t at = static_cast<size_t>(plane * 3 + cls) * (16 * N) + (k / 8) * (8 * N) + (n / 8) * 64 + (n % 8) * 8 + (k % 8);
                                out[at] = to_bf16(wf[(static_cast<size_t>(plane * 8 + ky) * 8 + kx) * 5 + ch]);
                            }
            }
    return out;
}

// conv1, phase-separated form: W [576][5] UNSCALED (k = plane*64 + ky*8 + kx) -> 81 blocks (plane, class, phase) of the
// K-major canonical B operand [K = 16][N = 48] in fp16: n = kv*16 + j*5 + channel holds
// W[plane][ky = class + 3 kv][kx = k - (phase + 3 j)][channel] (0 outside the 8 taps and for ky = 8)
std::vector<uint16_t> pack_conv1_phase(const std::vector<float> &wf)
{
    constexpr int N = 48;
    std::vector<uint16_t> out(static_cast<size_t>(81) * 16 * N, 0);
    for (int plane = 0; plane < 9; ++plane)
        for (int cls = 0; cls < 3; ++cls)
            for (int ph = 0; ph < 3; ++ph)
                for (int kv = 0; kv < 3; ++kv) {
                    const int ky = cls + 3 * kv;
                    if (ky >= 8) continue;
                    for (int j = 0; j < 3; ++j)
                        for (int ch = 0; ch < 5; ++ch)
                            for (int kx = 0; kx < 8; ++kx) {
                                const int k = kx + ph + 3 * j, n = kv * 16 + j * 5 + ch;
                                const size_t at = static_cast<size_t>((plane * 3 + cls) * 3 + ph) * (16 * N) + (k / 8) * (8 * N) + (n / 8) * 64 + (n % 8) * 8 + (k % 8);
                                out[at] = __half_as_ushort(__float2half_rn(wf[(static_cast<size_t>(plane * 8 + ky) * 8 + kx) * 5 + ch]));
                            }
                }
    return out;
}

} // namespace policy
} // namespace ahk

// ------------------------------------------------------------------------------------------------ C ABI
using namespace ahk::policy;

struct ahk_policy {
    int device = 0, max_batch = 0;
    uint16_t *w[6] = {};
    float *b[6] = {};
    uint16_t *w1t = nullptr; // conv1 in Toeplitz form (conv1_kernel, A/B switch AHK_CONV1=n144)
    uint16_t *w2p = nullptr; // conv2 in conv2p_kernel's K order (pack_b blobs)
    bool conv2_parity = true; // A/B switch AHK_CONV2=ws: conv_ws_kernel instead of conv2p_kernel
    // conv_tm_kernel (A operand built in TMEM from a once-copied item): the default for conv2 (0.100 ms per 8192 samples) and conv3
    // (0.064 ms; the TMA-fed GEMM 0.084).  A/B switch AHK_CONVTM=0 (neither) | 2 (conv2 only) | 23 (both, default)
    bool conv_tm = true, conv_tm3 = true;
    // the layer kernels of one inference forward chained by programmatic dependent launch: 2 (default) = a kernel releases the next
    // layer's launch when its own work is done, so that launch latency, barrier set-up, tensor-memory allocation and the weight
    // fetch run under its teardown (forward of 8192 samples 0.5156 -> 0.5040 ms, same box); 1 = released right after the prologue
    // (0.529 ms: the early CTAs of the next layer take SM resources from the running one); 0 = ordinary launches.  AHK_POLICY_PDL
    int pdl = 2;
    bool conv2_no_swap = false;
    int conv1_groups = 4; // converter groups of conv1t_kernel (forward of 8192 samples 0.4543 / 0.4507 / 0.4492 ms with 2 / 3 / 4, same box); A/B switch AHK_CONV1_GROUPS=2|3|4
    // conv2 -> conv3 in one kernel, act2 in shared memory (conv23_kernel): bit-identical to the two kernels, forward of 8192 samples
    // 0.4542 -> 0.4370 ms, rollout step 0.586 -> 0.556 ms (same box), 0.33 GB less HBM traffic; inference forwards only.  A/B switch AHK_FUSE23=0
    int fuse23_epi = 2; // epilogue warp sets of conv23_kernel; A/B switch AHK_FUSE23_EPI=1|2
    bool fuse23 = true, fuse23_profiled = false; // AHK_FUSE23=2: also in the profiled forward (its 'conv2' time is then the fused kernel's, 'conv3' ~ 0)
    bool convtm_two_raw = false; // A/B switch AHK_CONVTM_RAW=2: conv_tm_kernel<conv2> with two raw item buffers instead of 4
    uint16_t *w1p = nullptr; // conv1 in phase-separated form (conv1p_kernel, the default), fp16, unscaled
    bool conv1_n144 = false;
    bool conv1_normal = false; // A/B switch AHK_CONV1=normal: conv1t_kernel with the pixels as NORMAL fp16 numbers (1024 + v - 1024)
    bool conv1_smem = false; // A/B switch AHK_CONV1=smem: conv1p_kernel (A operand converted into shared memory) instead of conv1t_kernel (A in TMEM)
    cudaEvent_t ev[7] = {};  // ahk_policy_forward_profiled: around each of the six layer kernels
    // the 'gray-diff' network (ahk_policy_gray_create): weights/biases of its six layers, input image and activations
    bool gray = false;
    uint16_t *gw[6] = {};
    float *gb[6] = {};
    __nv_bfloat16 *gimg = nullptr, *ga1 = nullptr, *ga2 = nullptr, *ga3 = nullptr, *ga4 = nullptr, *gh1 = nullptr;
    int sms = 148;
    bool legacy_conv = false; // A/B switch AHK_CONV23=tile: conv2/conv3 through layer_kernel instead of conv_ws_kernel
    // conv2, conv3, dense1, dense2 on the TMA-fed GEMM (ahk_gemm.cuh; the default; AHK_GEMM=gather selects the cp.async /
    // register-gather kernels): weights as plain [N][K] bf16 matrices, tensor maps over them and over the activations
    bool tma_gemm = true, tma_conv2 = false;
    uint16_t *wn[5] = {};    // [1] conv2 32 x 128, [2] conv3 64 x 512, [3] dense1 512 x 4096, [4] dense2 512 x 512
    CUtensorMap tm_w[5], tm_a[5];
    CUtensorMap tm_rows2; // act2 as a matrix of 128-byte rows (two 32-channel pixels each): conv_tm_kernel<conv3>'s item copies
    __nv_bfloat16 *act1 = nullptr, *act2 = nullptr, *act3 = nullptr, *h1 = nullptr, *h2 = nullptr;
    float *w3sum = nullptr;    // [64][512]: dense1_px_sums_kernel
    double *bn_sums = nullptr; // ahk_policy_forward_train: [2][64] running sums of the layer being normalised
    uint64_t launches = 0;
};

namespace {

int pfail(int code, const char *fmt, const char *detail)
{
    char text[512];
    snprintf(text, sizeof text, fmt, detail);
    ahk::set_error_text(text); // the thread-local buffer behind ahk_last_error() lives in ahk_api.cu
    return code;
}

#define POL_CUDA(expr)                                                                       \
    do {                                                                                     \
        cudaError_t err__ = (expr);                                                          \
        if (err__ != cudaSuccess) return pfail(AHK_ERR_CUDA, #expr ": %s", cudaGetErrorString(err__)); \
    } while (0)

int upload(const std::vector<uint16_t> &w, const std::vector<float> &b, uint16_t **wd, float **bd)
{
    if (!*wd) POL_CUDA(cudaMalloc(wd, w.size() * 2)); // ahk_policy_set_weights overwrites the same buffers (same sizes)
    POL_CUDA(cudaMemcpy(*wd, w.data(), w.size() * 2, cudaMemcpyHostToDevice));
    if (!*bd) POL_CUDA(cudaMalloc(bd, b.size() * 4));
    POL_CUDA(cudaMemcpy(*bd, b.data(), b.size() * 4, cudaMemcpyHostToDevice));
    return AHK_OK;
}

int upload_nk(const std::vector<uint16_t> &w, uint16_t **wd)
{
    if (!*wd) POL_CUDA(cudaMalloc(wd, w.size() * 2));
    POL_CUDA(cudaMemcpy(*wd, w.data(), w.size() * 2, cudaMemcpyHostToDevice));
    return AHK_OK;
}

} // namespace

extern "C" {

// Folds and packs the 'rgb' network's Keras weights and uploads them (allocating on first use).
static int load_rgb_weights(ahk_policy *p, const ahk_policy_weights *w)
{
    // BatchNorm after the ReLU is the affine map y = s * r + t per channel (eps = 1e-3, Keras default)
    auto affine = [](const float *gamma, const float *beta, const float *mean, const float *var, int c, std::vector<float> &s, std::vector<float> &t) {
        s.resize(c); t.resize(c);
        for (int i = 0; i < c; ++i) {
            s[i] = gamma[i] / std::sqrt(var[i] + 1e-3f);
            t[i] = beta[i] - mean[i] * s[i];
        }
    };
    std::vector<float> s1, t1, s2, t2, s3, t3;
    affine(w->bn1_gamma, w->bn1_beta, w->bn1_mean, w->bn1_var, 5, s1, t1);
    affine(w->bn2_gamma, w->bn2_beta, w->bn2_mean, w->bn2_var, 32, s2, t2);
    affine(w->bn3_gamma, w->bn3_beta, w->bn3_mean, w->bn3_var, 64, s3, t3);

    int rc = AHK_OK;

    { // conv1: Keras kernel [ky][kx][c][o]; GEMM k = c*64 + ky*8 + kx; (x-128)/128 folded: W/128, b - sum(W)
        std::vector<float> wf(576 * 5), wraw(576 * 5), bf(16, 0.0f);
        for (int o = 0; o < 5; ++o) {
            double sum = 0.0;
            for (int ky = 0; ky < 8; ++ky) for (int kx = 0; kx < 8; ++kx) for (int c = 0; c < 9; ++c) {
                const float v = w->conv1_w[((ky * 8 + kx) * 9 + c) * 5 + o];
                wf[(c * 64 + ky * 8 + kx) * 5 + o] = v / 128.0f;
                wraw[(c * 64 + ky * 8 + kx) * 5 + o] = v;
                sum += v;
            }
            bf[o] = w->conv1_b[o] - static_cast<float>(sum);
        }
        rc = upload(pack_conv1_toeplitz(wf), bf, &p->w1t, &p->b[0]);
        if (rc == AHK_OK) { // the phase-separated form: fp16, unscaled (the 1/128 is applied to the accumulator)
            const std::vector<uint16_t> wp = pack_conv1_phase(wraw);
            if (!p->w1p) POL_CUDA(cudaMalloc(&p->w1p, wp.size() * 2));
            POL_CUDA(cudaMemcpy(p->w1p, wp.data(), wp.size() * 2, cudaMemcpyHostToDevice));
        }
    }
    if (rc == AHK_OK) { // conv2: k = (ky*4+kx)*8 + c, BN1 folded
        std::vector<float> wf(128 * 32, 0.0f), bf(32);
        for (int o = 0; o < 32; ++o) {
            double add = 0.0;
            for (int ky = 0; ky < 4; ++ky) for (int kx = 0; kx < 4; ++kx) for (int c = 0; c < 5; ++c) {
                const float v = w->conv2_w[((ky * 4 + kx) * 5 + c) * 32 + o];
                wf[((ky * 4 + kx) * 8 + c) * 32 + o] = v * s1[c];
                add += static_cast<double>(v) * t1[c];
            }
            bf[o] = w->conv2_b[o] + static_cast<float>(add);
        }
        rc = upload(pack_b(wf, 128, 32, 32, 1), bf, &p->w[1], &p->b[1]);
        if (rc == AHK_OK) rc = upload_nk(pack_nk(wf, 128, 32), &p->wn[1]);
        if (rc == AHK_OK) { // the same matrix in conv2p_kernel's K order
            std::vector<float> wp(128 * 32);
            for (int k = 0; k < 128; ++k) {
                const int c = k & 7, j = (k >> 3) & 1, kxp = (k >> 4) & 1, ky = k >> 5, src = ((ky * 4 + kxp + 2 * j) * 8 + c);
                for (int o = 0; o < 32; ++o) wp[k * 32 + o] = wf[src * 32 + o];
            }
            rc = upload_nk(pack_b(wp, 128, 32, 32, 1), &p->w2p);
        }
    }
    if (rc == AHK_OK) { // conv3: k = (ky*4+kx)*32 + c, BN2 folded
        std::vector<float> wf(512 * 64), bf(64);
        for (int o = 0; o < 64; ++o) {
            double add = 0.0;
            for (int ky = 0; ky < 4; ++ky) for (int kx = 0; kx < 4; ++kx) for (int c = 0; c < 32; ++c) {
                const float v = w->conv3_w[((ky * 4 + kx) * 32 + c) * 64 + o];
                wf[((ky * 4 + kx) * 32 + c) * 64 + o] = v * s2[c];
                add += static_cast<double>(v) * t2[c];
            }
            bf[o] = w->conv3_b[o] + static_cast<float>(add);
        }
        rc = upload(pack_b(wf, 512, 64, 64, 1), bf, &p->w[2], &p->b[2]);
        if (rc == AHK_OK) rc = upload_nk(pack_nk(wf, 512, 64), &p->wn[2]);
    }
    if (rc == AHK_OK) { // dense1: Keras flatten index f = c*64 + pixel; ours k = pixel*64 + c; BN3 folded
        std::vector<float> wf(static_cast<size_t>(4096) * 512), bf(512);
        std::vector<double> add(512, 0.0);
        for (int c = 0; c < 64; ++c) for (int px = 0; px < 64; ++px) {
            const float *src = w->dense1_w + static_cast<size_t>(c * 64 + px) * 512;
            float *dst = wf.data() + static_cast<size_t>(px * 64 + c) * 512;
            for (int o = 0; o < 512; ++o) { dst[o] = src[o] * s3[c]; add[o] += static_cast<double>(src[o]) * t3[c]; }
        }
        for (int o = 0; o < 512; ++o) bf[o] = w->dense1_b[o] + static_cast<float>(add[o]);
        rc = upload(pack_b(wf, 4096, 512, 256, 2), bf, &p->w[3], &p->b[3]);
        if (rc == AHK_OK) rc = upload_nk(pack_nk(wf, 4096, 512), &p->wn[3]);
    }
    if (rc == AHK_OK) {
        std::vector<float> wf(w->dense2_w, w->dense2_w + 512 * 512), bf(w->dense2_b, w->dense2_b + 512);
        rc = upload(pack_b(wf, 512, 512, 256, 2), bf, &p->w[4], &p->b[4]);
        if (rc == AHK_OK) rc = upload_nk(pack_nk(wf, 512, 512), &p->wn[4]);
    }
    if (rc == AHK_OK) {
        std::vector<float> wf(w->out_w, w->out_w + 512 * 10), bf(16, 0.0f);
        for (int o = 0; o < 10; ++o) bf[o] = w->out_b[o];
        rc = upload(pack_b(wf, 512, 10, 16, 1), bf, &p->w[5], &p->b[5]);
    }
    return rc;
}

int ahk_policy_create(int device, int max_batch, const ahk_policy_weights *w, ahk_policy **out)
{
    if (!w || !out || max_batch <= 0 || max_batch > 131072) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_create: bad argument (max_batch must be in [1, 131072])");
    const float *const *fields = reinterpret_cast<const float *const *>(w);
    for (int i = 0; i < 24; ++i)
        if (!fields[i]) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_create: every weight pointer must be non-NULL");
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return pfail(AHK_ERR_NO_DEVICE, "%s", "no CUDA device: the policy has no CPU fallback");
    POL_CUDA(cudaSetDevice(device));
    ahk_policy *p = new ahk_policy();
    p->device = device;
    p->max_batch = max_batch;
    cudaDeviceGetAttribute(&p->sms, cudaDevAttrMultiProcessorCount, device);
    if (const char *sw = getenv("AHK_CONV23")) p->legacy_conv = strcmp(sw, "tile") == 0;
    if (const char *sw = getenv("AHK_CONV2")) p->conv2_parity = strcmp(sw, "ws") != 0;
    if (const char *sw = getenv("AHK_CONVTM")) { p->conv_tm = strcmp(sw, "0") != 0; p->conv_tm3 = strcmp(sw, "23") == 0; } // "2": conv2 only, "0": neither
    if (const char *sw = getenv("AHK_POLICY_PDL")) p->pdl = atoi(sw);
    if (const char *sw = getenv("AHK_CONV2_SWAP")) p->conv2_no_swap = strcmp(sw, "0") == 0;
    if (const char *sw = getenv("AHK_CONV1_GROUPS")) p->conv1_groups = atoi(sw);
    if (const char *sw = getenv("AHK_CONVTM_RAW")) p->convtm_two_raw = strcmp(sw, "2") == 0;
    if (const char *sw = getenv("AHK_FUSE23_EPI")) p->fuse23_epi = atoi(sw);
    if (const char *sw = getenv("AHK_FUSE23")) { p->fuse23 = strcmp(sw, "0") != 0; p->fuse23_profiled = strcmp(sw, "2") == 0; }
    if (const char *sw = getenv("AHK_CONV1")) { p->conv1_n144 = strcmp(sw, "n144") == 0; p->conv1_smem = strcmp(sw, "smem") == 0; p->conv1_normal = strcmp(sw, "normal") == 0; }
    int rc = load_rgb_weights(p, w);
    const size_t nb = static_cast<size_t>(max_batch);
    cudaError_t err = cudaSuccess;
    for (int i = 0; i < 7 && err == cudaSuccess; ++i) err = cudaEventCreate(&p->ev[i]);
    if (rc == AHK_OK) {
        err = cudaMalloc(&p->act1, nb * kP1 * kAct1C * 2);
        if (err == cudaSuccess) err = cudaMalloc(&p->act2, nb * kP2 * 32 * 2 + 128); // + 128: tm_rows2 rounds an odd pixel count up to whole 128-byte rows
        if (err == cudaSuccess) err = cudaMalloc(&p->act3, nb * 4096 * 2);
        if (err == cudaSuccess) err = cudaMalloc(&p->h1, nb * 512 * 2);
        if (err == cudaSuccess) err = cudaMalloc(&p->h2, nb * 512 * 2);
    }
    if (const char *sw = getenv("AHK_GEMM")) { p->tma_gemm = strcmp(sw, "gather") != 0; p->tma_conv2 = strcmp(sw, "tma2") == 0; }
    if (rc == AHK_OK && err == cudaSuccess) { // tensor maps of the TMA-fed GEMM layers (ahk_gemm.cuh): weights [N][K], activations as im2col views
        namespace tg = ahk::tg;
        err = tg::matrix_map(&p->tm_w[1], p->wn[1], 32, 128, 32, 32, 64);
        if (err == cudaSuccess) err = tg::matrix_map(&p->tm_w[2], p->wn[2], 64, 512, 64, 64);
        if (err == cudaSuccess) err = tg::matrix_map(&p->tm_w[3], p->wn[3], 512, 4096, 64, 256);
        if (err == cudaSuccess) err = tg::matrix_map(&p->tm_w[4], p->wn[4], 512, 512, 64, 256);
        if (err == cudaSuccess) { // conv2 reads act1 [sample][41 x 41][8]: element (tap offset + 4 pixels x 8 channels), x2, y2 (stride 2 pixels / rows), sample
            const unsigned long long dims[4] = {static_cast<unsigned long long>(kP1) * kAct1C, kC2, kC2, nb}, strides[3] = {2 * kAct1C * 2, 2 * kC1 * kAct1C * 2, static_cast<unsigned long long>(kP1) * kAct1C * 2};
            const unsigned box[4] = {32, kC2, 6, 1};
            err = tg::encode_bf16_map(&p->tm_a[1], p->act1, 4, dims, strides, box, 64);
        }
        if (err == cudaSuccess) { // conv3 reads act2 [sample][19 x 19][32]: element (tap offset + 2 pixels x 32 channels), x3, y3, sample
            const unsigned long long dims[4] = {static_cast<unsigned long long>(kP2) * 32, kC3, kC3, nb}, strides[3] = {2 * 32 * 2, 2 * kC2 * 32 * 2, static_cast<unsigned long long>(kP2) * 32 * 2};
            const unsigned box[4] = {64, kC3, kC3, 2};
            err = tg::encode_bf16_map(&p->tm_a[2], p->act2, 4, dims, strides, box, 128);
        }
        if (err == cudaSuccess) err = tg::matrix_map(&p->tm_rows2, p->act2, (nb * kP2 + 1) / 2, 64, 64, 128);
        if (err == cudaSuccess) err = tg::matrix_map(&p->tm_a[3], p->act3, nb, 4096, 64, 128);
        if (err == cudaSuccess) err = tg::matrix_map(&p->tm_a[4], p->h1, nb, 512, 64, 128);
    }
    if (rc != AHK_OK || err != cudaSuccess) {
        ahk_policy_destroy(p);
        return rc != AHK_OK ? rc : pfail(AHK_ERR_CUDA, "ahk_policy_create: %s", cudaGetErrorString(err));
    }
    *out = p;
    return AHK_OK;
}

int ahk_policy_set_weights(ahk_policy *p, const ahk_policy_weights *w)
{
    if (!p || !w) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_set_weights: NULL argument");
    if (p->gray) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_set_weights: handle holds the 'gray-diff' network");
    const float *const *fields = reinterpret_cast<const float *const *>(w);
    for (int i = 0; i < 24; ++i)
        if (!fields[i]) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_set_weights: every weight pointer must be non-NULL");
    POL_CUDA(cudaSetDevice(p->device));
    POL_CUDA(cudaDeviceSynchronize()); // forward passes in flight still read the old weights
    return load_rgb_weights(p, w);
}

int ahk_policy_set_weights_dev(ahk_policy *p, const ahk_policy_weights *w_dev, void *stream)
{
    if (!p || !w_dev) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_set_weights_dev: NULL argument");
    if (p->gray) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_set_weights_dev: handle holds the 'gray-diff' network");
    DevWeights w;
    const float *const *fields = reinterpret_cast<const float *const *>(w_dev);
    for (int i = 0; i < 24; ++i) {
        if (!fields[i]) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_set_weights_dev: every weight pointer must be non-NULL");
        w.f[i] = fields[i];
    }
    POL_CUDA(cudaSetDevice(p->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    POL_CUDA(cudaMemsetAsync(p->w1t, 0, static_cast<size_t>(27) * 16 * 144 * 2, s)); // the Toeplitz bands leave zeros between them
    pack_conv1_kernel<<<(9 * 3 * 3 * 3 * 3 * 5 * 8 + 255) / 256, 256, 0, s>>>(w, p->w1t);
    POL_CUDA(cudaMemsetAsync(p->w1p, 0, static_cast<size_t>(c1p::kWBytes), s));
    pack_conv1p_kernel<<<(9 * 3 * 3 * 3 * 3 * 5 * 8 + 255) / 256, 256, 0, s>>>(w, p->w1p);
    pack_b_kernel<1><<<16, 256, 0, s>>>(w, p->w[1], 128, 32, 1);
    pack_b_kernel<6><<<16, 256, 0, s>>>(w, p->w2p, 128, 32, 1);
    pack_b_kernel<2><<<128, 256, 0, s>>>(w, p->w[2], 512, 64, 1);
    pack_b_kernel<3><<<p->sms * 8, 256, 0, s>>>(w, p->w[3], 4096, 256, 2);
    pack_b_kernel<4><<<p->sms * 4, 256, 0, s>>>(w, p->w[4], 512, 256, 2);
    pack_b_kernel<5><<<32, 256, 0, s>>>(w, p->w[5], 512, 16, 1);
    pack_nk_kernel<1><<<16, 256, 0, s>>>(w, p->wn[1], 128, 32);
    pack_nk_kernel<2><<<128, 256, 0, s>>>(w, p->wn[2], 512, 64);
    pack_nk_kernel<3><<<p->sms * 8, 256, 0, s>>>(w, p->wn[3], 4096, 512);
    pack_nk_kernel<4><<<p->sms * 4, 256, 0, s>>>(w, p->wn[4], 512, 512);
    if (!p->w3sum) POL_CUDA(cudaMalloc(&p->w3sum, 64 * 512 * sizeof(float)));
    dense1_px_sums_kernel<<<64, 512, 0, s>>>(w.f[18], p->w3sum);
    fold_bias_kernel<<<dim3(4, 6), 128, 0, s>>>(w, p->b[0], p->b[1], p->b[2], p->b[3], p->b[4], p->b[5], p->w3sum);
    POL_CUDA(cudaGetLastError());
    p->launches += 13;
    return AHK_OK;
}

int ahk_policy_gray_create(int device, int max_batch, const ahk_policy_gray_weights *w, ahk_policy **out)
{
    if (!w || !out || max_batch <= 0 || max_batch > 131072) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_gray_create: bad argument (max_batch must be in [1, 131072])");
    const float *const *fields = reinterpret_cast<const float *const *>(w);
    for (int i = 0; i < 28; ++i)
        if (!fields[i]) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_gray_create: every weight pointer must be non-NULL");
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return pfail(AHK_ERR_NO_DEVICE, "%s", "no CUDA device: the policy has no CPU fallback");
    POL_CUDA(cudaSetDevice(device));
    ahk_policy *p = new ahk_policy();
    p->device = device;
    p->max_batch = max_batch;
    p->gray = true;
    if (const char *sw = getenv("AHK_CONV23")) p->legacy_conv = strcmp(sw, "tile") == 0; // A/B: conv2-conv4 through layer_kernel
    cudaDeviceGetAttribute(&p->sms, cudaDevAttrMultiProcessorCount, device);

    // BatchNorm after the ReLU: y = s * r + t per channel, folded into the next layer
    auto affine = [](const float *gamma, const float *beta, const float *mean, const float *var, int c, std::vector<float> &s, std::vector<float> &t) {
        s.resize(c); t.resize(c);
        for (int i = 0; i < c; ++i) { s[i] = gamma[i] / std::sqrt(var[i] + 1e-3f); t[i] = beta[i] - mean[i] * s[i]; }
    };
    // a 4x4 convolution as GEMM weights [16 taps x cin_pad][cout]: k = tap * cin_pad + c, previous BatchNorm folded in
    auto conv4x4 = [](const float *kw, const float *kb, int cin, int cin_pad, int cout, int cout_pad, const std::vector<float> &s,
                      const std::vector<float> &t, std::vector<float> &wf, std::vector<float> &bf) {
        wf.assign(static_cast<size_t>(16) * cin_pad * cout, 0.0f);
        bf.assign(cout_pad, 0.0f);
        for (int o = 0; o < cout; ++o) {
            double add = 0.0;
            for (int tap = 0; tap < 16; ++tap) for (int c = 0; c < cin; ++c) {
                const float v = kw[(tap * cin + c) * cout + o];
                wf[(static_cast<size_t>(tap) * cin_pad + c) * cout + o] = v * s[c];
                add += static_cast<double>(v) * t[c];
            }
            bf[o] = kb[o] + static_cast<float>(add);
        }
    };
    std::vector<float> s1, t1, s2, t2, s3, t3, s4, t4, wf, bf;
    affine(w->bn1_gamma, w->bn1_beta, w->bn1_mean, w->bn1_var, 16, s1, t1);
    affine(w->bn2_gamma, w->bn2_beta, w->bn2_mean, w->bn2_var, 24, s2, t2);
    affine(w->bn3_gamma, w->bn3_beta, w->bn3_mean, w->bn3_var, 32, s3, t3);
    affine(w->bn4_gamma, w->bn4_beta, w->bn4_mean, w->bn4_var, 48, s4, t4);
    int rc = AHK_OK;
    { // conv1: Keras kernel [ky][kx][1][o] is already [k = ky*8 + kx][o]; the input arrives normalised
        wf.assign(w->conv1_w, w->conv1_w + 64 * 16);
        bf.assign(w->conv1_b, w->conv1_b + 16);
        rc = upload(pack_b(wf, 64, 16, 16, 1), bf, &p->gw[0], &p->gb[0]);
    }
    if (rc == AHK_OK) { conv4x4(w->conv2_w, w->conv2_b, 16, 16, 24, 32, s1, t1, wf, bf); rc = upload(pack_b(wf, 256, 24, 32, 1), bf, &p->gw[1], &p->gb[1]); }
    if (rc == AHK_OK) { conv4x4(w->conv3_w, w->conv3_b, 24, 32, 32, 32, s2, t2, wf, bf); rc = upload(pack_b(wf, 512, 32, 32, 1), bf, &p->gw[2], &p->gb[2]); }
    if (rc == AHK_OK) { conv4x4(w->conv4_w, w->conv4_b, 32, 32, 48, 48, s3, t3, wf, bf); rc = upload(pack_b(wf, 512, 48, 48, 1), bf, &p->gw[3], &p->gb[3]); }
    if (rc == AHK_OK) { // dense1: Keras flatten index f = c*64 + pixel; ours k = pixel*48 + c; BN4 folded
        wf.assign(static_cast<size_t>(3072) * 512, 0.0f);
        bf.assign(512, 0.0f);
        std::vector<double> add(512, 0.0);
        for (int c = 0; c < 48; ++c) for (int px = 0; px < 64; ++px) {
            const float *src = w->dense1_w + static_cast<size_t>(c * 64 + px) * 512;
            float *dst = wf.data() + static_cast<size_t>(px * 48 + c) * 512;
            for (int o = 0; o < 512; ++o) { dst[o] = src[o] * s4[c]; add[o] += static_cast<double>(src[o]) * t4[c]; }
        }
        for (int o = 0; o < 512; ++o) bf[o] = w->dense1_b[o] + static_cast<float>(add[o]);
        rc = upload(pack_b(wf, 3072, 512, 256, 2), bf, &p->gw[4], &p->gb[4]);
    }
    if (rc == AHK_OK) {
        wf.assign(w->out_w, w->out_w + 512 * 10);
        bf.assign(16, 0.0f);
        for (int o = 0; o < 10; ++o) bf[o] = w->out_b[o];
        rc = upload(pack_b(wf, 512, 10, 16, 1), bf, &p->gw[5], &p->gb[5]);
    }
    const size_t nb = static_cast<size_t>(max_batch);
    cudaError_t err = cudaSuccess;
    for (int i = 0; i < 7 && err == cudaSuccess; ++i) err = cudaEventCreate(&p->ev[i]);
    if (rc == AHK_OK) {
        if (err == cudaSuccess) err = cudaMalloc(&p->gimg, nb * ahk::kDim * ahk::kDim * 2 + 64); // + slack: the last row's 16-byte window
        if (err == cudaSuccess) err = cudaMalloc(&p->ga1, nb * kGP1 * 16 * 2);
        if (err == cudaSuccess) err = cudaMalloc(&p->ga2, nb * kGP2 * 32 * 2);
        if (err == cudaSuccess) err = cudaMalloc(&p->ga3, nb * kGP3 * 32 * 2);
        if (err == cudaSuccess) err = cudaMalloc(&p->ga4, nb * kGP4 * 48 * 2);
        if (err == cudaSuccess) err = cudaMalloc(&p->gh1, nb * 512 * 2);
    }
    if (rc != AHK_OK || err != cudaSuccess) {
        ahk_policy_destroy(p);
        return rc != AHK_OK ? rc : pfail(AHK_ERR_CUDA, "ahk_policy_gray_create: %s", cudaGetErrorString(err));
    }
    *out = p;
    return AHK_OK;
}

void ahk_policy_destroy(ahk_policy *p)
{
    if (!p) return;
    cudaSetDevice(p->device);
    for (int i = 0; i < 6; ++i) { if (p->w[i]) cudaFree(p->w[i]); if (p->b[i]) cudaFree(p->b[i]); }
    if (p->w1t) cudaFree(p->w1t);
    for (int i = 0; i < 5; ++i) if (p->wn[i]) cudaFree(p->wn[i]);
    if (p->bn_sums) cudaFree(p->bn_sums);
    if (p->w3sum) cudaFree(p->w3sum);
    if (p->w1p) cudaFree(p->w1p);
    if (p->w2p) cudaFree(p->w2p);
    for (int i = 0; i < 7; ++i) if (p->ev[i]) cudaEventDestroy(p->ev[i]);
    for (int i = 0; i < 6; ++i) { if (p->gw[i]) cudaFree(p->gw[i]); if (p->gb[i]) cudaFree(p->gb[i]); }
    cudaFree(p->gimg); cudaFree(p->ga1); cudaFree(p->ga2); cudaFree(p->ga3); cudaFree(p->ga4); cudaFree(p->gh1);
    cudaFree(p->act1); cudaFree(p->act2); cudaFree(p->act3); cudaFree(p->h1); cudaFree(p->h2);
    delete p;
}

// One layer of the 'rgb' network on n samples (a: frames / ring geometry filled in): 0 conv1 ... 5 out
// pdl: launch with programmatic stream serialization behind the previous layer's kernel (inference forward only: in
// model.fit's forward the kernel before a layer folds that layer's weights, which these kernels fetch before their wait)
// fuse23: conv2 and conv3 as ONE kernel (conv23_kernel; inference only: model.fit's forward needs act2 in HBM for its batch statistics)
static int rgb_layer(ahk_policy *p, Args a, int layer, float *logits, uint8_t *action, cudaStream_t s, bool pdl = false, bool fuse23 = false)
{
    namespace tg = ahk::tg;
    const int n = a.n;
    auto tiles = [](long long rows) { return static_cast<int>((rows + kBM - 1) / kBM); };
    tg::Args g{};
    g.k_splits = 1; g.m_total = n; g.scale = 1.0f; g.pdl = a.pdl;
    fuse23 = fuse23 && p->conv_tm && p->conv_tm3 && p->conv2_parity && !p->legacy_conv && p->tma_gemm && !p->tma_conv2;
    switch (layer) {
    case 0:
        a.bias = p->b[0]; a.out = p->act1; a.m_total = kP1;
        if (p->conv1_n144) {
            a.weights = reinterpret_cast<const uint4 *>(p->w1t);
            const int items = n * c1::kTilesPerSample;
            POL_CUDA(launch_conv1(a, items < p->sms ? items : p->sms, s));
        } else {
            a.weights = reinterpret_cast<const uint4 *>(p->w1p);
            const int items = n * 2;
            POL_CUDA(p->conv1_smem ? launch_conv1p(a, items < p->sms ? items : p->sms, s) : launch_conv1t(a, items < p->sms ? items : p->sms, !p->conv1_normal, s, p->conv1_groups));
        }
        return AHK_OK;
    case 1:
        // conv2 stays on the cp.async pipeline: as a tensor-map im2col its overlapping 64-byte windows read every act1 pixel
        // four times over L2 -> SM (1.2 GB per 8192 samples: 0.21-0.25 ms against 0.155; AHK_GEMM=tma2 runs it for the record)
        a.in = p->act1; a.weights = reinterpret_cast<const uint4 *>(p->w[1]); a.bias = p->b[1]; a.out = p->act2; a.m_total = n * kP2;
        g.tiles_m = n * 4; g.blocks_n = 1; g.k_chunks = 4; g.n_valid = 32; g.bias = p->b[1]; g.out_bf16 = p->act2; g.out_pitch = 32;
        if (fuse23) POL_CUDA(launch_conv23(a, reinterpret_cast<const uint4 *>(p->w[2]), p->b[2], p->act3, p->sms, s, pdl, p->fuse23_epi));
        else if (p->conv_tm && p->conv2_parity && !p->legacy_conv && !(p->tma_gemm && p->tma_conv2)) POL_CUDA(launch_tm<kConv2>(a, p->tm_rows2, p->sms, s, pdl, p->convtm_two_raw)); // (conv2 takes its sample by a plain bulk copy: the map is unused)
        else if (p->conv2_parity && !p->legacy_conv && !(p->tma_gemm && p->tma_conv2)) {
            a.weights = reinterpret_cast<const uint4 *>(p->w2p);
            POL_CUDA(launch_conv2p(a, p->sms, s));
        } else if (p->tma_gemm && p->tma_conv2) POL_CUDA((tg::launch<32, 64, tg::kConv2, false, tg::kBiasReluBf16, 8, 4>(g, p->tm_a[1], p->tm_w[1], p->sms, s)));
        else if (p->legacy_conv) POL_CUDA(launch<kConv2>(a, tiles(static_cast<long long>(n) * kP2), 1, p->sms, s));
        else POL_CUDA(launch_ws<kConv2>(a, tiles(static_cast<long long>(n) * kP2), p->sms, s));
        return AHK_OK;
    case 2:
        a.in = p->act2; a.weights = reinterpret_cast<const uint4 *>(p->w[2]); a.bias = p->b[2]; a.out = p->act3; a.m_total = n * kP3;
        g.tiles_m = (n + 1) / 2; g.blocks_n = 1; g.k_chunks = 8; g.n_valid = 64; g.bias = p->b[2]; g.out_bf16 = p->act3; g.out_pitch = 64;
        if (fuse23) return AHK_OK; // done by layer 1's kernel
        if (p->conv_tm3 && p->tma_gemm && !p->legacy_conv) POL_CUDA(launch_tm<kConv3>(a, p->tm_rows2, p->sms, s, pdl, p->convtm_two_raw));
        else if (p->tma_gemm) POL_CUDA((tg::launch<64, 128, tg::kConv3, false, tg::kBiasReluBf16, 6, 8>(g, p->tm_a[2], p->tm_w[2], p->sms, s, pdl)));
        else if (p->legacy_conv) POL_CUDA(launch<kConv3>(a, tiles(static_cast<long long>(n) * kP3), 1, p->sms, s));
        else POL_CUDA(launch_ws<kConv3>(a, tiles(static_cast<long long>(n) * kP3), p->sms, s));
        return AHK_OK;
    case 3:
        a.in = p->act3; a.weights = reinterpret_cast<const uint4 *>(p->w[3]); a.bias = p->b[3]; a.out = p->h1; a.m_total = n;
        g.tiles_m = tiles(n); g.blocks_n = 2; g.k_chunks = 64; g.n_valid = 256; g.bias = p->b[3]; g.out_bf16 = p->h1; g.out_pitch = 512;
        if (p->tma_gemm) POL_CUDA((tg::launch<256, 128, tg::kLinear, false, tg::kBiasReluBf16, 4>(g, p->tm_a[3], p->tm_w[3], p->sms, s, pdl)));
        else POL_CUDA(launch<kDense1>(a, tiles(n), 2, p->sms, s));
        return AHK_OK;
    case 4:
        a.in = p->h1; a.weights = reinterpret_cast<const uint4 *>(p->w[4]); a.bias = p->b[4]; a.out = p->h2; a.m_total = n;
        g.tiles_m = tiles(n); g.blocks_n = 2; g.k_chunks = 8; g.n_valid = 256; g.bias = p->b[4]; g.out_bf16 = p->h2; g.out_pitch = 512;
        if (p->tma_gemm) POL_CUDA((tg::launch<256, 128, tg::kLinear, false, tg::kBiasReluBf16, 4>(g, p->tm_a[4], p->tm_w[4], p->sms, s, pdl)));
        else POL_CUDA(launch<kDense2>(a, tiles(n), 2, p->sms, s));
        return AHK_OK;
    default:
        a.in = p->h2; a.weights = reinterpret_cast<const uint4 *>(p->w[5]); a.bias = p->b[5]; a.out = nullptr; a.logits = logits; a.action = action; a.m_total = n;
        POL_CUDA(launch<kOut>(a, tiles(n), 1, p->sms, s, pdl));
        return AHK_OK;
    }
}

static int policy_forward(ahk_policy *p, const ahk_stack &stack, float *logits, uint8_t *action, void *stream, bool timed)
{
    const uint8_t *frames = stack.frames;
    const int n = stack.n, head = stack.head;
    if (!p || !frames) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_forward: policy/frames is NULL");
    if (n <= 0 || n > p->max_batch) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_forward: n must be in [1, max_batch]");
    if (stack.depth < 3 || head < 0 || head >= stack.depth) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_forward: head must be in [0, depth), depth >= 3");
    if (stack.warm && !stack.meta) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_forward: a deep ring needs the meta words");
    int prev_device = -1;
    POL_CUDA(cudaGetDevice(&prev_device));
    if (prev_device != p->device) POL_CUDA(cudaSetDevice(p->device));
    struct Restore { int prev, cur; ~Restore() { if (prev != cur) cudaSetDevice(prev); } } restore{prev_device, p->device};
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    auto tiles = [](long long rows) { return static_cast<int>((rows + kBM - 1) / kBM); };
    Args a{};
    a.n = n;
    a.head = head;
    a.depth = stack.depth;
    a.warm = stack.warm;
    a.meta = stack.meta;
    if (p->gray) { // the 'gray-diff' network: input image + six GEMM-tile layers
        const long long ln = n;
        if (timed) POL_CUDA(cudaEventRecord(p->ev[0], s));
        const int blocks = static_cast<int>((ln * 1024 + 255) / 256);
        gray_diff_kernel<<<blocks < p->sms * 16 ? blocks : p->sms * 16, 256, 0, s>>>(frames, p->gimg, n, head, stack.depth, stack.warm, stack.meta);
        POL_CUDA(cudaGetLastError());
        a.in = p->gimg; a.weights = reinterpret_cast<const uint4 *>(p->gw[0]); a.bias = p->gb[0]; a.out = p->ga1; a.m_total = n * kGP1;
        POL_CUDA(launch<kG1>(a, tiles(ln * kGP1), 1, p->sms, s));
        if (timed) POL_CUDA(cudaEventRecord(p->ev[1], s));
        a.in = p->ga1; a.weights = reinterpret_cast<const uint4 *>(p->gw[1]); a.bias = p->gb[1]; a.out = p->ga2; a.m_total = n * kGP2;
        POL_CUDA(p->legacy_conv ? launch<kG2>(a, tiles(ln * kGP2), 1, p->sms, s) : launch_ws<kG2>(a, tiles(ln * kGP2), p->sms, s));
        if (timed) POL_CUDA(cudaEventRecord(p->ev[2], s));
        a.in = p->ga2; a.weights = reinterpret_cast<const uint4 *>(p->gw[2]); a.bias = p->gb[2]; a.out = p->ga3; a.m_total = n * kGP3;
        POL_CUDA(p->legacy_conv ? launch<kG3>(a, tiles(ln * kGP3), 1, p->sms, s) : launch_ws<kG3>(a, tiles(ln * kGP3), p->sms, s));
        if (timed) POL_CUDA(cudaEventRecord(p->ev[3], s));
        a.in = p->ga3; a.weights = reinterpret_cast<const uint4 *>(p->gw[3]); a.bias = p->gb[3]; a.out = p->ga4; a.m_total = n * kGP4;
        POL_CUDA(p->legacy_conv ? launch<kG4>(a, tiles(ln * kGP4), 1, p->sms, s) : launch_ws<kG4>(a, tiles(ln * kGP4), p->sms, s));
        if (timed) POL_CUDA(cudaEventRecord(p->ev[4], s));
        a.in = p->ga4; a.weights = reinterpret_cast<const uint4 *>(p->gw[4]); a.bias = p->gb[4]; a.out = p->gh1; a.m_total = n;
        POL_CUDA(launch<kGDense1>(a, tiles(n), 2, p->sms, s));
        if (timed) POL_CUDA(cudaEventRecord(p->ev[5], s));
        a.in = p->gh1; a.weights = reinterpret_cast<const uint4 *>(p->gw[5]); a.bias = p->gb[5]; a.out = nullptr; a.logits = logits; a.action = action; a.m_total = n;
        POL_CUDA(launch<kOut>(a, tiles(n), 1, p->sms, s));
        if (timed) POL_CUDA(cudaEventRecord(p->ev[6], s));
        p->launches += 7;
        return AHK_OK;
    }
    if (timed) POL_CUDA(cudaEventRecord(p->ev[0], s));
    a.frames = frames;
    for (int layer = 0; layer < 6; ++layer) {
        a.pdl = timed ? 0 : p->pdl;
        a.no_swap = p->conv2_no_swap;
        if (int rc = rgb_layer(p, a, layer, logits, action, s, a.pdl != 0 && layer > 0, p->fuse23 && (!timed || p->fuse23_profiled))) return rc; // profiled: one kernel per layer unless AHK_FUSE23=2
        if (timed) POL_CUDA(cudaEventRecord(p->ev[layer + 1], s));
    }
    p->launches += (p->fuse23 && (!timed || p->fuse23_profiled) && p->conv_tm && p->conv_tm3 && p->conv2_parity && !p->legacy_conv && p->tma_gemm && !p->tma_conv2) ? 5 : 6;
    return AHK_OK;
}

// model.fit's forward pass: see include/ahk.h
int ahk_policy_forward_train(ahk_policy *p, const ahk_policy_weights *w_dev, const uint8_t *frames_dev, int n, int head, float *logits_dev,
                             float *batch_stats_dev, void *stream)
{
    if (!p || !w_dev || !frames_dev || !logits_dev || !batch_stats_dev) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_forward_train: NULL argument");
    if (p->gray) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_forward_train: handle holds the 'gray-diff' network");
    if (n <= 0 || n > p->max_batch || head < 0 || head > 2) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_forward_train: n must be in [1, max_batch], head in [0, 3)");
    if (p->conv1_n144 || p->legacy_conv || !p->tma_gemm) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_forward_train: runs on the default kernels only (unset AHK_CONV1 / AHK_CONV23 / AHK_GEMM)");
    DevWeights w;
    const float *const *fields = reinterpret_cast<const float *const *>(w_dev);
    for (int i = 0; i < 24; ++i) {
        if (!fields[i]) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_forward_train: every weight pointer must be non-NULL");
        w.f[i] = fields[i];
    }
    int prev_device = -1;
    POL_CUDA(cudaGetDevice(&prev_device));
    if (prev_device != p->device) POL_CUDA(cudaSetDevice(p->device));
    struct Restore { int prev, cur; ~Restore() { if (prev != cur) cudaSetDevice(prev); } } restore{prev_device, p->device};
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (!p->w3sum) POL_CUDA(cudaMalloc(&p->w3sum, 64 * 512 * sizeof(float)));
    if (!p->bn_sums) {
        POL_CUDA(cudaMalloc(&p->bn_sums, 2 * 64 * sizeof(double)));
        POL_CUDA(cudaMemsetAsync(p->bn_sums, 0, 2 * 64 * sizeof(double), s));
    }
    // the BatchNorm moving statistics of the weight set are replaced by the batch statistics, layer by layer
    float *mean = batch_stats_dev, *var = batch_stats_dev + AHK_BN_STATS;
    w.f[4] = mean; w.f[5] = var; w.f[10] = mean + 8; w.f[11] = var + 8; w.f[16] = mean + 40; w.f[17] = var + 40;
    Args a{};
    a.n = n; a.head = head; a.depth = 3; a.frames = frames_dev;
    const int stat_grid = p->sms * 4;
    // conv1 (no BatchNorm before it)
    POL_CUDA(cudaMemsetAsync(p->w1p, 0, static_cast<size_t>(c1p::kWBytes), s));
    pack_conv1p_kernel<<<(9 * 3 * 3 * 3 * 3 * 5 * 8 + 255) / 256, 256, 0, s>>>(w, p->w1p);
    fold_bias_kernel<<<dim3(4, 6), 128, 0, s>>>(w, p->b[0], p->b[1], p->b[2], p->b[3], p->b[4], p->b[5], p->w3sum, 0);
    if (int rc = rgb_layer(p, a, 0, nullptr, nullptr, s)) return rc;
    bn_sums_kernel<8><<<stat_grid, 256, 0, s>>>(p->act1, static_cast<long long>(n) * kP1, p->bn_sums);
    bn_finalize_kernel<<<1, 64, 0, s>>>(p->bn_sums, 8, static_cast<long long>(n) * kP1, mean, var);
    // conv2 with BatchNorm 1 (batch statistics) folded in
    pack_b_kernel<1><<<16, 256, 0, s>>>(w, p->w[1], 128, 32, 1);
    pack_b_kernel<6><<<16, 256, 0, s>>>(w, p->w2p, 128, 32, 1);
    fold_bias_kernel<<<dim3(4, 6), 128, 0, s>>>(w, p->b[0], p->b[1], p->b[2], p->b[3], p->b[4], p->b[5], p->w3sum, 1);
    if (int rc = rgb_layer(p, a, 1, nullptr, nullptr, s)) return rc;
    bn_sums_kernel<32><<<stat_grid, 256, 0, s>>>(p->act2, static_cast<long long>(n) * kP2, p->bn_sums);
    bn_finalize_kernel<<<1, 64, 0, s>>>(p->bn_sums, 32, static_cast<long long>(n) * kP2, mean + 8, var + 8);
    // conv3
    pack_nk_kernel<2><<<128, 256, 0, s>>>(w, p->wn[2], 512, 64);
    pack_b_kernel<2><<<128, 256, 0, s>>>(w, p->w[2], 512, 64, 1);
    fold_bias_kernel<<<dim3(4, 6), 128, 0, s>>>(w, p->b[0], p->b[1], p->b[2], p->b[3], p->b[4], p->b[5], p->w3sum, 2);
    if (int rc = rgb_layer(p, a, 2, nullptr, nullptr, s)) return rc;
    bn_sums_kernel<64><<<stat_grid, 256, 0, s>>>(p->act3, static_cast<long long>(n) * kP3, p->bn_sums);
    bn_finalize_kernel<<<1, 64, 0, s>>>(p->bn_sums, 64, static_cast<long long>(n) * kP3, mean + 40, var + 40);
    // dense layers
    pack_nk_kernel<3><<<p->sms * 8, 256, 0, s>>>(w, p->wn[3], 4096, 512);
    pack_nk_kernel<4><<<p->sms * 4, 256, 0, s>>>(w, p->wn[4], 512, 512);
    pack_b_kernel<5><<<32, 256, 0, s>>>(w, p->w[5], 512, 16, 1);
    dense1_px_sums_kernel<<<64, 512, 0, s>>>(w.f[18], p->w3sum);
    for (int layer = 3; layer < 6; ++layer) fold_bias_kernel<<<dim3(4, 6), 128, 0, s>>>(w, p->b[0], p->b[1], p->b[2], p->b[3], p->b[4], p->b[5], p->w3sum, layer);
    for (int layer = 3; layer < 6; ++layer)
        if (int rc = rgb_layer(p, a, layer, logits_dev, nullptr, s)) return rc;
    POL_CUDA(cudaGetLastError());
    p->launches += 25;
    return AHK_OK;
}

__global__ void __launch_bounds__(256) bf16_to_f32_kernel(const __nv_bfloat16 *__restrict__ src, float *__restrict__ dst, size_t count)
{
    for (size_t i = blockIdx.x * 256ull + threadIdx.x; i < count; i += gridDim.x * 256ull) dst[i] = __bfloat162float(src[i]);
}

int ahk_policy_copy_activation(ahk_policy *p, int which, float *dst_dev, size_t count, void *stream)
{
    if (!p || !dst_dev || which < 0 || which > 4 || p->gray) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_copy_activation: bad argument");
    const __nv_bfloat16 *src[5] = {p->act1, p->act2, p->act3, p->h1, p->h2};
    const size_t per[5] = {static_cast<size_t>(kP1) * kAct1C, static_cast<size_t>(kP2) * 32, 4096, 512, 512};
    if (count > per[which] * static_cast<size_t>(p->max_batch)) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_copy_activation: count exceeds the buffer");
    POL_CUDA(cudaSetDevice(p->device));
    bf16_to_f32_kernel<<<p->sms * 4, 256, 0, static_cast<cudaStream_t>(stream)>>>(src[which], dst_dev, count);
    POL_CUDA(cudaGetLastError());
    return AHK_OK;
}

int ahk_policy_get_activations(ahk_policy *p, ahk_policy_acts *out)
{
    if (!p || !out) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_get_activations: NULL argument");
    if (p->gray) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_get_activations: handle holds the 'gray-diff' network");
    out->act1 = p->act1; out->act2 = p->act2; out->act3 = p->act3; out->h1 = p->h1; out->h2 = p->h2;
    out->max_batch = p->max_batch;
    return AHK_OK;
}

static ahk_stack plain_ring(const uint8_t *frames, int n, int head)
{
    ahk_stack st{};
    st.frames = frames;
    st.n = n;
    st.depth = 3;
    st.head = head;
    return st;
}

int ahk_policy_forward(ahk_policy *p, const uint8_t *frames, int n, int head, float *logits, uint8_t *action, void *stream)
{
    return policy_forward(p, plain_ring(frames, n, head), logits, action, stream, false);
}

int ahk_policy_forward_stack(ahk_policy *p, const ahk_stack *stack, float *logits, uint8_t *action, void *stream)
{
    if (!stack) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_forward_stack: stack is NULL");
    return policy_forward(p, *stack, logits, action, stream, false);
}

static int forward_profiled(ahk_policy *p, const ahk_stack &stack, float *logits, uint8_t *action, void *stream, float *layer_ms)
{
    if (!layer_ms) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_forward_profiled: layer_ms is NULL");
    if (int rc = policy_forward(p, stack, logits, action, stream, true)) return rc;
    POL_CUDA(cudaEventSynchronize(p->ev[6]));
    for (int i = 0; i < 6; ++i) POL_CUDA(cudaEventElapsedTime(layer_ms + i, p->ev[i], p->ev[i + 1]));
    return AHK_OK;
}

int ahk_policy_forward_profiled(ahk_policy *p, const uint8_t *frames, int n, int head, float *logits, uint8_t *action,
                                void *stream, float *layer_ms)
{
    return forward_profiled(p, plain_ring(frames, n, head), logits, action, stream, layer_ms);
}

int ahk_policy_forward_stack_profiled(ahk_policy *p, const ahk_stack *stack, float *logits, uint8_t *action, void *stream, float *layer_ms)
{
    if (!stack) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_forward_stack_profiled: stack is NULL");
    return forward_profiled(p, *stack, logits, action, stream, layer_ms);
}

} // extern "C"

// ------------------------------------------------------------------------------------------------ ahk_policy_forward_static
// The forward pass for callers that may not allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc):
// the same layer kernels as ahk_policy_forward, with the activations between them — the forward's scratch memory — in a
// file-scope array of the library's device image instead of the handle's cudaMalloc'ed buffers.  No counterpart in the reference.
//   * AHK_POLICY_STATIC_MAX_BATCH = 64 samples: act1 [64][41 x 41][8] + act2 [64][19 x 19][32] + act3 [64][4096] + h1, h2 [64][512]
//     bf16 = 3 855 616 bytes, held by every GPU that loads the library until the process exits.
//   * g_static_scratch is a HOST-SIDE STAND-IN in host code, not a device address: its address on the current device is taken
//     with cudaGetSymbolAddress once per device (static_setup, with that device current) and that pointer goes to the kernels.
//   * Every call on a device shares the array and finds the previous call's contents in it.  Nothing in it has to start at
//     zero: every activation row of samples [0, n) is written by the layer before the layer that reads it (ahk_policy_create
//     does not clear its buffers either); rows of samples >= n are only ever multiplied into accumulator rows that no
//     epilogue stores.  Two calls that overlap on the GPU (different streams, or different handles) RACE on the array.
namespace {

constexpr int kStaticBatch = AHK_POLICY_STATIC_MAX_BATCH;
constexpr size_t kStaticAct1 = 0;
constexpr size_t kStaticAct2 = kStaticAct1 + static_cast<size_t>(kStaticBatch) * kP1 * kAct1C * 2;
constexpr size_t kStaticAct3 = kStaticAct2 + static_cast<size_t>(kStaticBatch) * kP2 * 32 * 2 + 256; // + 256: as act2's + 128 in ahk_policy_create, aligned
constexpr size_t kStaticH1 = kStaticAct3 + static_cast<size_t>(kStaticBatch) * 4096 * 2;
constexpr size_t kStaticH2 = kStaticH1 + static_cast<size_t>(kStaticBatch) * 512 * 2;
constexpr size_t kStaticBytes = kStaticH2 + static_cast<size_t>(kStaticBatch) * 512 * 2;
static_assert(kStaticAct2 % 256 == 0 && kStaticAct3 % 256 == 0 && kStaticH1 % 256 == 0 && kStaticH2 % 256 == 0, "256-byte aligned regions");

static __device__ __align__(256) uint8_t g_static_scratch[kStaticBytes];

struct StaticSetup { // per device: the array's DEVICE address and the tensor maps of the TMA-fed layers over it
    bool ready = false;
    uint8_t *base = nullptr;
    CUtensorMap tm_a[5], tm_rows2;
};
constexpr int kStaticDevices = 64;
StaticSetup g_static_setup[kStaticDevices];
std::mutex g_static_mutex;

// once per device, with that device current
int static_setup(int device, const StaticSetup **out)
{
    namespace tg = ahk::tg;
    if (device < 0 || device >= kStaticDevices) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_forward_static: device ordinal beyond 63");
    std::lock_guard<std::mutex> lock(g_static_mutex);
    StaticSetup &st = g_static_setup[device];
    if (!st.ready) {
        void *ptr = nullptr;
        POL_CUDA(cudaGetSymbolAddress(&ptr, g_static_scratch));
        uint8_t *base = static_cast<uint8_t *>(ptr);
        const size_t nb = kStaticBatch;
        cudaError_t err = cudaSuccess;
        { // the same views as ahk_policy_create builds over its own buffers
            const unsigned long long dims[4] = {static_cast<unsigned long long>(kP1) * kAct1C, kC2, kC2, nb}, strides[3] = {2 * kAct1C * 2, 2 * kC1 * kAct1C * 2, static_cast<unsigned long long>(kP1) * kAct1C * 2};
            const unsigned box[4] = {32, kC2, 6, 1};
            err = tg::encode_bf16_map(&st.tm_a[1], base + kStaticAct1, 4, dims, strides, box, 64);
        }
        if (err == cudaSuccess) {
            const unsigned long long dims[4] = {static_cast<unsigned long long>(kP2) * 32, kC3, kC3, nb}, strides[3] = {2 * 32 * 2, 2 * kC2 * 32 * 2, static_cast<unsigned long long>(kP2) * 32 * 2};
            const unsigned box[4] = {64, kC3, kC3, 2};
            err = tg::encode_bf16_map(&st.tm_a[2], base + kStaticAct2, 4, dims, strides, box, 128);
        }
        if (err == cudaSuccess) err = tg::matrix_map(&st.tm_rows2, base + kStaticAct2, (nb * kP2 + 1) / 2, 64, 64, 128);
        if (err == cudaSuccess) err = tg::matrix_map(&st.tm_a[3], base + kStaticAct3, nb, 4096, 64, 128);
        if (err == cudaSuccess) err = tg::matrix_map(&st.tm_a[4], base + kStaticH1, nb, 512, 64, 128);
        if (err != cudaSuccess) return pfail(AHK_ERR_CUDA, "ahk_policy_forward_static: tensor maps over the static scratch: %s", cudaGetErrorString(err));
        st.base = base;
        st.ready = true;
    }
    *out = &st;
    return AHK_OK;
}

} // namespace

extern "C" int ahk_policy_forward_static(ahk_policy *p, const uint8_t *frames, int n, int head, float *logits, uint8_t *action, void *stream)
{
    if (!p || !frames || !logits || !action) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_forward_static: NULL argument");
    if (p->gray) return pfail(AHK_ERR_ARG, "%s", "ahk_policy_forward_static: handle holds the 'gray-diff' network (the static scratch is laid out for 'rgb')");
    if (n < 1 || n > kStaticBatch) {
        char text[160];
        snprintf(text, sizeof text, "n = %d is outside [1, AHK_POLICY_STATIC_MAX_BATCH = %d]: the static scratch holds %d samples; use ahk_policy_forward", n, kStaticBatch, kStaticBatch);
        return pfail(AHK_ERR_ARG, "ahk_policy_forward_static: %s", text);
    }
    int prev_device = -1;
    POL_CUDA(cudaGetDevice(&prev_device));
    if (prev_device != p->device) POL_CUDA(cudaSetDevice(p->device));
    struct Restore { int prev, cur; ~Restore() { if (prev != cur) cudaSetDevice(prev); } } restore{prev_device, p->device};
    const StaticSetup *st = nullptr;
    if (int rc = static_setup(p->device, &st)) return rc;
    // a host-side copy of the handle whose scratch pointers and activation maps are the static array's: weights, weight maps
    // and kernel selection stay the handle's, so the launches are those of ahk_policy_forward
    ahk_policy view = *p;
    view.max_batch = kStaticBatch;
    view.act1 = reinterpret_cast<__nv_bfloat16 *>(st->base + kStaticAct1);
    view.act2 = reinterpret_cast<__nv_bfloat16 *>(st->base + kStaticAct2);
    view.act3 = reinterpret_cast<__nv_bfloat16 *>(st->base + kStaticAct3);
    view.h1 = reinterpret_cast<__nv_bfloat16 *>(st->base + kStaticH1);
    view.h2 = reinterpret_cast<__nv_bfloat16 *>(st->base + kStaticH2);
    for (int i = 1; i <= 4; ++i) view.tm_a[i] = st->tm_a[i];
    view.tm_rows2 = st->tm_rows2;
    const int rc = policy_forward(&view, plain_ring(frames, n, head), logits, action, stream, false);
    p->launches = view.launches;
    return rc;
}
