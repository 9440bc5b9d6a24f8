// ahk_gemm.cuh — a TMA-fed tcgen05 GEMM:  D[M x N] = A[M x K] * B[N x K]^T, bf16 operands, fp32 accumulators in TMEM.
//
// Used by the CNN forward (conv2, conv3, dense1, dense2 of models.py:20-65 — ahk_policy.cu) and by the hand-written
// backward pass of the RL fit (rl_agents.py:189,276 — ahk_train.cu).  No CUDA core touches an operand:
//   * A and B tiles are fetched by the TMA engine (cp.async.bulk.tensor) straight into the swizzled layout the MMA reads.
//     For the convolutions the tensor map IS the im2col: its dimensions alias the activation tensor with overlapping
//     strides (output x / output y / sample), the kernel tap is the start coordinate, so one box = one K chunk of a
//     128-row tile, gathered by the copy engine.  (The cp.async gather kernels these replace were bound by shared-memory
//     wavefronts and load latency: conv2 0.155 ms at 28 % of DRAM throughput, profiles/policy_kernels_r2l_ncu.txt.)
//   * either operand may be K-major (rows of K: activations x weights) or MN-major (rows of M or N: the weight-gradient
//     GEMMs, whose reduction runs over the batch), selected per operand in the instruction descriptor;
//   * warp 0 = producer (one elected lane issues the copies), warp 1 = MMA issue (whole warp, warp-uniform operands, one
//     elected lane), warps 2-5 = epilogue out of a double-buffered accumulator; persistent over tiles.
#pragma once
#include "ahk_tc.cuh"

#include <cudaTypedefs.h>
#include <cuda_runtime.h>

namespace ahk {
namespace tg {

using namespace ahk::tc;

constexpr int kCta = 192;

// how tile index -> A box coordinates
enum AMode {
    kLinear = 0,   // 2-D map [K][M]: (kc * chunk, tile * 128)
    kConv2 = 1,    // 4-D map over act1 [sample][41 x 41][8]: element, x2 (stride 2 px), y2 (stride 2 rows), sample; box [32][19][6][1]
    kConv3 = 2,    // 4-D map over act2 [sample][19 x 19][32]: element, x3, y3, sample; box [64][8][8][2]
    kLinearMn = 3, // A is MN-major: 2-D map [M][K]: two boxes [64][chunk K rows] per chunk at (tile * 128 + 64 h, kc * chunk)
    // conv3 weight gradient: A = im2col(act2)^T, MN-major, through the conv3 map with ONE sample per box [64][8][8][1]: M tile = kernel
    // row ky (4 taps x 32 channels), the two boxes are its tap pairs, K chunk kc = sample kc (its 64 output pixels)
    kConv3Wgrad = 4,
    // conv3 data gradient, one parity class (Args::py, px) per launch: A = dz3 [sample][8][8][64] through a 4-D map, box [64][10][10][1]
    // at (0, -jx, -jy, sample): row r = yq * 10 + xq of the tile is input pixel (2 yq + py, 2 xq + px), chunk j = (jy, jx) its tap
    // (py + 2 jy, px + 2 jx); out-of-range output pixels are the copy engine's zero fill.  B rows = tap * 32 of the Keras kernel.
    kConv3Dgrad = 5,
};
// what the epilogue does with the accumulator
enum Epi {
    kBiasReluBf16 = 0, // out_bf16[row][n0 + c] = bf16(relu(acc + bias[n0 + c]))
    kBiasBf16 = 1,     // the same without the ReLU
    kMaskBf16 = 2,     // out_bf16[row][n0 + c] = bf16(acc) where mask_bf16[row][n0 + c] > 0, else 0    (ReLU backward on the fly)
    kAtomicF32 = 3,    // atomicAdd(out_f32[row][n0 + c], acc * scale)                                   (split-K weight gradients)
    kStoreF32 = 4,     // out_f32[row][n0 + c] = acc * scale
    kMaskBoth = 5,     // kMaskBf16 and the same values as float32 into out_f32
};

struct Args {
    int tiles_m, blocks_n, k_chunks; // work items = tiles_m * blocks_n * k_splits; K chunks per work item
    int k_splits;                    // split-K: chunk range [ks * k_chunks, (ks + 1) * k_chunks)
    int m_total;                     // valid rows (linear modes) or samples (conv modes)
    int n_valid;                     // valid columns of the last N block (others are full)
    const float *bias;
    __nv_bfloat16 *out_bf16;
    const __nv_bfloat16 *mask_bf16;
    float *out_f32;
    long long out_pitch;             // elements per output row
    float scale;
    int py, px;                      // kConv3Dgrad: parity class of the input pixels
    int pdl;                         // programmatic dependent launch of the next kernel: 0 off, 1 trigger after the prologue, 2 at the end
};

// shared-memory matrix descriptor of a swizzled tile (cute::UMMA::SmemDescriptor, version 1): K-major: rows of `span` bytes
// (128 -> SWIZZLE_128B, 64 -> SWIZZLE_64B), 8-row groups 8 * span apart; MN-major (span 128 only): 64 MN elements per row,
// K rows 128 bytes apart, 8-row K groups at SBO = 1024, the next 64 MN elements at LBO
__device__ __forceinline__ uint64_t sw_desc(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t layout)
{
    return static_cast<uint64_t>((addr & 0x3FFFFu) >> 4) | static_cast<uint64_t>(lbo >> 4) << 16 | static_cast<uint64_t>(sbo >> 4) << 32 |
           1ull << 46 | static_cast<uint64_t>(layout) << 61;
}
// instruction descriptor: D f32, A/B bf16, M = 128, N, majors (0 = K, 1 = MN)
__host__ __device__ constexpr uint32_t idesc_bf16(int n, int a_mn, int b_mn)
{
    return 1u << 4 | 1u << 7 | 1u << 10 | static_cast<uint32_t>(a_mn) << 15 | static_cast<uint32_t>(b_mn) << 16 | static_cast<uint32_t>(n >> 3) << 17 | (128u >> 4) << 24;
}

__device__ __forceinline__ void tma_2d(void *dst, const CUtensorMap *tm, int c0, int c1, uint64_t *bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 :: "r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(tm)), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_4d(void *dst, const CUtensorMap *tm, int c0, int c1, int c2, int c3, uint64_t *bar)
{
    asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
                 :: "r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(tm)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(smem_u32(bar)) : "memory");
}

// kBRes > 0: the whole B operand (kBRes K chunks) is loaded once per CTA and stays in shared memory (conv layers: blocks_n = 1)
template <int N, int kSpan, int kAMode, bool kBMn, int kBRes = 0> struct Cfg {
    static constexpr bool kAMn = kAMode == kLinearMn || kAMode == kConv3Wgrad;
    static constexpr int kChunkK = kAMn || kBMn ? 64 : kSpan / 2;                                  // K elements per chunk
    static constexpr int kARows = kAMode == kConv2 ? 114 : (kAMode == kConv3Dgrad ? 100 : 128);
    static constexpr int kABytes = kAMn ? 2 * 64 * 128 : 128 * kSpan;                            // smem footprint of an A stage (conv2: 114 rows arrive)
    static constexpr int kATx = kAMn ? 2 * 64 * 128 : kARows * kSpan;                            // bytes the copies deliver
    static constexpr int kBBytes = kBMn ? (N / 64) * 64 * 128 : N * kSpan;
    static constexpr int kStageBytes = (kABytes + (kBRes ? 0 : kBBytes) + 1023) / 1024 * 1024;
    static constexpr int kResBytes = kBRes * kBBytes;
    static constexpr int kTmemCols = 2 * N <= 32 ? 32 : (2 * N <= 64 ? 64 : (2 * N <= 128 ? 128 : (2 * N <= 256 ? 256 : 512)));
    static constexpr int kBufCols = kTmemCols / 2;
};

template <int N, int kSpan, int kAMode, bool kBMn, int kEpi, int kStages, int kBRes>
__global__ void __launch_bounds__(kCta, 1) gemm_tma_kernel(Args g, const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b)
{
    using C = Cfg<N, kSpan, kAMode, kBMn, kBRes>;
    static_assert(kSpan == 128 || (kSpan == 64 && !C::kAMn && !kBMn), "MN-major tiles use the 128-byte swizzle");
    static_assert(N % 16 == 0 && N >= 16 && N <= 256 && (!kBMn || N % 64 == 0), "N");
    extern __shared__ __align__(128) uint8_t smem_unaligned[];
    uint8_t *s_res = smem_unaligned + ((1024u - (smem_u32(smem_unaligned) & 1023u)) & 1023u); // swizzle atoms are 1024-byte aligned
    uint8_t *s_stage = s_res + C::kResBytes;
    __shared__ __align__(8) uint64_t s_full[kStages], s_empty[kStages], s_tfull[2], s_tempty[2], s_bfull;
    __shared__ uint32_t s_tmem;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int items = g.tiles_m * g.blocks_n * g.k_splits;
    const int my_items = static_cast<int>(blockIdx.x) < items ? (items - 1 - static_cast<int>(blockIdx.x)) / static_cast<int>(gridDim.x) + 1 : 0;
    if (tid == 0) {
        for (int i = 0; i < kStages; ++i) { mbar_init(&s_full[i], 1); mbar_init(&s_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&s_tfull[i], 1); mbar_init(&s_tempty[i], 128); }
        mbar_init(&s_bfull, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) tmem_alloc(&s_tmem, C::kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;
    constexpr uint32_t kLayout = kSpan == 128 ? 2u : 4u; // SWIZZLE_128B / SWIZZLE_64B
    // programmatic dependent launch (ahk_tc.cuh): the resident weights are not the previous kernel's output and are fetched
    // before the wait; everything else — A tiles, streamed B, the epilogue's stores — comes after it
    if (g.pdl == 1) pdl_trigger();
    if (kBRes && my_items > 0 && tid == 0) {
        mbar_expect_tx(&s_bfull, C::kResBytes);
        for (int kc = 0; kc < kBRes; ++kc) tma_2d(s_res + kc * C::kBBytes, &tm_b, kc * C::kChunkK, 0, &s_bfull);
    }
    pdl_wait();

    // work item -> (M tile, N block, K split): N blocks of one M tile are neighbours, so CTAs running side by side share A in L2
    auto decode = [&](int it, int &tm_i, int &nb, int &ks) {
        const int item = static_cast<int>(blockIdx.x) + it * static_cast<int>(gridDim.x);
        nb = item % g.blocks_n;
        const int rest = item / g.blocks_n;
        ks = rest % g.k_splits;
        tm_i = rest / g.k_splits;
    };

    if (warp == 0) {
        // ---------------------------------------------------------------- producer
        if (lane == 0) {
            uint32_t fill = 0;
            for (int it = 0; it < my_items; ++it) {
                int tm_i, nb, ks;
                decode(it, tm_i, nb, ks);
                for (int kc = ks * g.k_chunks; kc < (ks + 1) * g.k_chunks; ++kc, ++fill) {
                    const uint32_t st = fill % kStages;
                    if (fill >= static_cast<uint32_t>(kStages)) mbar_wait_sleep(&s_empty[st], ((fill / kStages) - 1u) & 1u);
                    uint8_t *sa = s_stage + st * C::kStageBytes, *sb = sa + C::kABytes;
                    mbar_expect_tx(&s_full[st], C::kATx + (kBRes ? 0 : C::kBBytes));
                    if (kAMode == kLinear) tma_2d(sa, &tm_a, kc * C::kChunkK, tm_i * 128, &s_full[st]);
                    else if (kAMode == kLinearMn) {
                        tma_2d(sa, &tm_a, tm_i * 128, kc * C::kChunkK, &s_full[st]);
                        tma_2d(sa + 64 * 128, &tm_a, tm_i * 128 + 64, kc * C::kChunkK, &s_full[st]);
                    } else if (kAMode == kConv3Wgrad) {
                        tma_4d(sa, &tm_a, (tm_i * 19) * 32, 0, 0, kc, &s_full[st]);
                        tma_4d(sa + 64 * 128, &tm_a, (tm_i * 19 + 2) * 32, 0, 0, kc, &s_full[st]);
                    } else if (kAMode == kConv3Dgrad) tma_4d(sa, &tm_a, 0, -(kc & 1), -(kc >> 1), tm_i, &s_full[st]);                   // kc = jy * 2 + jx
                    else if (kAMode == kConv2) tma_4d(sa, &tm_a, kc * (41 * 8), 0, (tm_i & 3) * 6, tm_i >> 2, &s_full[st]);             // kc = ky
                    else tma_4d(sa, &tm_a, ((kc >> 1) * 19 + (kc & 1) * 2) * 32, 0, 0, tm_i * 2, &s_full[st]);                           // kc = ky * 2 + half
                    if (kBRes) continue;
                    if (kAMode == kConv3Dgrad) tma_2d(sb, &tm_b, 0, ((g.py + 2 * (kc >> 1)) * 4 + g.px + 2 * (kc & 1)) * 32, &s_full[st]);
                    else if (kBMn) {
#pragma unroll
                        for (int h = 0; h < N / 64; ++h) tma_2d(sb + h * 64 * 128, &tm_b, nb * N + h * 64, kc * C::kChunkK, &s_full[st]);
                    } else tma_2d(sb, &tm_b, kc * C::kChunkK, nb * N, &s_full[st]);
                }
            }
        }
    } else if (warp == 1) {
        // ---------------------------------------------------------------- MMA issue (whole warp, one elected lane)
        constexpr uint32_t kIdesc = idesc_bf16(N, C::kAMn, kBMn);
        uint32_t fill = 0;
        if (kBRes && my_items > 0) mbar_wait_sleep(&s_bfull, 0);
        for (int it = 0; it < my_items; ++it) {
            const uint32_t buf = it & 1u;
            if (it >= 2) mbar_wait_sleep(&s_tempty[buf], ((it >> 1) - 1) & 1);
            tc_fence_after();
#pragma unroll 1
            for (int kc = 0; kc < g.k_chunks; ++kc, ++fill) {
                const uint32_t st = fill % kStages;
                mbar_wait_sleep(&s_full[st], (fill / kStages) & 1u);
                tc_fence_after();
                const uint32_t sa = smem_u32(s_stage + st * C::kStageBytes), sb = kBRes ? smem_u32(s_res + kc * C::kBBytes) : sa + C::kABytes;
                // K-major: 8-row groups 8 * span apart, a K = 16 step is 32 bytes inside the swizzle span.
                // MN-major: K rows of 128 bytes (64 MN elements), 8-row K groups 1024 apart (SBO), next 64 MN elements at LBO
                // = one box; a K = 16 step is two K groups = 2048 bytes.
                const uint64_t ad = C::kAMn ? sw_desc(sa, 64 * 128, 1024, 2u) : sw_desc(sa, 16, 8 * kSpan, kLayout);
                const uint64_t bd = kBMn ? sw_desc(sb, 64 * 128, 1024, 2u) : sw_desc(sb, 16, 8 * kSpan, kLayout);
                constexpr uint32_t kAStep = C::kAMn ? 2048 / 16 : 32 / 16, kBStep = kBMn ? 2048 / 16 : 32 / 16;
                if (elect_one()) {
#pragma unroll
                    for (int j = 0; j < C::kChunkK / 16; ++j)
                        umma_bf16(tmem + buf * C::kBufCols, ad + static_cast<uint64_t>(j * kAStep), bd + static_cast<uint64_t>(j * kBStep), kIdesc, (kc | j) != 0);
                    umma_commit(&s_empty[st]);
                    if (kc == g.k_chunks - 1) umma_commit(&s_tfull[buf]);
                }
                __syncwarp();
            }
        }
    } else {
        // ---------------------------------------------------------------- epilogue: thread = TMEM lane = tile row
        const int q4 = warp & 3, r = q4 * 32 + lane;
        for (int it = 0; it < my_items; ++it) {
            int tm_i, nb, ks;
            decode(it, tm_i, nb, ks);
            const uint32_t buf = it & 1u;
            long long row;
            bool row_ok;
            if (kAMode == kConv2) { // tile = (sample, 6 output rows): r = y * 19 + x
                const int y = r / 19, x = r - y * 19, y2 = (tm_i & 3) * 6 + y;
                row_ok = r < 114 && y2 < 19;
                row = static_cast<long long>(tm_i >> 2) * 361 + y2 * 19 + x;
            } else if (kAMode == kConv3Dgrad) { // tile = sample: r = yq * 10 + xq -> input pixel (2 yq + py, 2 xq + px) of the 19 x 19 plane
                const int yq = r / 10, xq = r - yq * 10, y = 2 * yq + g.py, x = 2 * xq + g.px;
                row_ok = r < 100 && y < 19 && x < 19;
                row = static_cast<long long>(tm_i) * 361 + y * 19 + x;
            } else {
                row = static_cast<long long>(tm_i) * 128 + r;
                row_ok = row < (kAMode == kConv3 ? static_cast<long long>(g.m_total) * 64 : static_cast<long long>(g.m_total));
            }
            mbar_wait_sleep(&s_tfull[buf], (it >> 1) & 1);
            tc_fence_after();
            const uint32_t taddr = tmem + buf * C::kBufCols + (static_cast<uint32_t>(q4 * 32) << 16);
            const int n0 = nb * N;
            const int n_here = nb == g.blocks_n - 1 ? g.n_valid : N;
#pragma unroll 1
            for (int c0 = 0; c0 < N; c0 += 32) {
                uint32_t v[32];
                tmem_load16_async(taddr + c0, *reinterpret_cast<uint32_t (*)[16]>(&v[0]));
                if (c0 + 16 < N) tmem_load16_async(taddr + c0 + 16, *reinterpret_cast<uint32_t (*)[16]>(&v[16]));
                tmem_wait_ld();
                if (c0 + 32 >= N) { // the accumulator is in registers: the MMA warp may overwrite it
                    tc_fence_before();
                    mbar_arrive(&s_tempty[buf]);
                }
                if (!row_ok) continue;
                constexpr int kCols = N < 32 ? N : 32;
                if (kEpi == kBiasReluBf16 || kEpi == kBiasBf16 || kEpi == kMaskBf16 || kEpi == kMaskBoth) {
                    constexpr bool kMasked = kEpi == kMaskBf16 || kEpi == kMaskBoth;
                    __nv_bfloat16 *out = g.out_bf16 + row * g.out_pitch + n0 + c0;
                    const __nv_bfloat16 *mask = kMasked ? g.mask_bf16 + row * g.out_pitch + n0 + c0 : nullptr;
#pragma unroll
                    for (int c = 0; c < kCols; c += 8) {
                        if (c0 + c >= n_here) break;
                        uint32_t pk[4];
                        uint4 mk = make_uint4(0, 0, 0, 0);
                        if (kMasked) mk = *reinterpret_cast<const uint4 *>(mask + c);
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            float lo = __uint_as_float(v[c + 2 * j]), hi = __uint_as_float(v[c + 2 * j + 1]);
                            if (kMasked) {
                                const uint32_t m2 = (&mk.x)[j]; // two bf16: positive iff sign clear and not zero
                                lo = (m2 & 0x7FFFu) != 0u && (m2 & 0x8000u) == 0u ? lo : 0.0f;
                                hi = (m2 & 0x7FFF0000u) != 0u && (m2 & 0x80000000u) == 0u ? hi : 0.0f;
                                if (kEpi == kMaskBoth) *reinterpret_cast<float2 *>(g.out_f32 + row * g.out_pitch + n0 + c0 + c + 2 * j) = make_float2(lo, hi);
                            } else {
                                lo += g.bias[n0 + c0 + c + 2 * j];
                                hi += g.bias[n0 + c0 + c + 2 * j + 1];
                                if (kEpi == kBiasReluBf16) { lo = fmaxf(lo, 0.0f); hi = fmaxf(hi, 0.0f); }
                            }
                            const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
                            pk[j] = *reinterpret_cast<const uint32_t *>(&h);
                        }
                        *reinterpret_cast<uint4 *>(out + c) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                    }
                } else {
                    float *out = g.out_f32 + row * g.out_pitch + n0 + c0;
#pragma unroll
                    for (int c = 0; c < kCols; ++c) {
                        if (c0 + c >= n_here) break;
                        const float val = __uint_as_float(v[c]) * g.scale;
                        if (kEpi == kAtomicF32) atomicAdd(out + c, val);
                        else out[c] = val;
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (g.pdl == 2) pdl_trigger();
    if (warp == 1) tmem_dealloc(tmem, C::kTmemCols);
}

template <int N, int kSpan, int kAMode, bool kBMn, int kStages, int kBRes> constexpr int smem_bytes()
{
    return 1024 + Cfg<N, kSpan, kAMode, kBMn, kBRes>::kResBytes + kStages * Cfg<N, kSpan, kAMode, kBMn, kBRes>::kStageBytes;
}

template <int N, int kSpan, int kAMode, bool kBMn, int kEpi, int kStages, int kBRes = 0>
cudaError_t launch(const Args &g, const CUtensorMap &tm_a, const CUtensorMap &tm_b, int sms, cudaStream_t s, bool pdl = false)
{
    constexpr int bytes = smem_bytes<N, kSpan, kAMode, kBMn, kStages, kBRes>();
    static_assert(bytes <= 227 * 1024, "stages do not fit");
    cudaError_t e = cudaFuncSetAttribute(gemm_tma_kernel<N, kSpan, kAMode, kBMn, kEpi, kStages, kBRes>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return e;
    const int items = g.tiles_m * g.blocks_n * g.k_splits;
    if (items <= 0) return cudaSuccess;
    return launch_kernel(gemm_tma_kernel<N, kSpan, kAMode, kBMn, kEpi, kStages, kBRes>, dim3(items < sms ? items : sms), dim3(kCta), bytes, s, pdl, g, tm_a, tm_b);
}

// ---- tensor maps (host).  bf16 elements; dims[0] is the contiguous one; strides in BYTES for dims 1.. (multiples of 16)
inline cudaError_t encode_bf16_map(CUtensorMap *tm, const void *base, int rank, const unsigned long long *dims, const unsigned long long *strides_bytes,
                                   const unsigned *box, int span_bytes)
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
    cuuint64_t d[5], st[4];
    cuuint32_t b[5], es[5];
    for (int i = 0; i < rank; ++i) { d[i] = dims[i]; b[i] = box[i]; es[i] = 1; }
    for (int i = 0; i + 1 < rank; ++i) st[i] = strides_bytes[i];
    const CUresult r = encode(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void *>(base), d, st, b, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              span_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}
// row-major [rows][cols] bf16 matrix, box = [box_cols][box_rows]
inline cudaError_t matrix_map(CUtensorMap *tm, const void *base, unsigned long long rows, unsigned long long cols, unsigned box_cols, unsigned box_rows, int span_bytes = 128)
{
    const unsigned long long dims[2] = {cols, rows}, strides[1] = {cols * 2};
    const unsigned box[2] = {box_cols, box_rows};
    return encode_bf16_map(tm, base, 2, dims, strides, box, span_bytes);
}

} // namespace tg
} // namespace ahk
