// ahk_step.cu — the hot path as ONE kernel per step: physics + programmed AI + reward/labels (K1, K4) in the prologue,
// then the rasteriser (K2) — every new uint8 CHW 3x128x128 frame written to the HBM ring exactly once.
//
// Replaces AirHockeyEnv.step / AirHockeyEnv.reset (gym-air-hockey/gym_air_hockey/envs/air_hockey_env.py:26-155), the
// missing simulator's step and Pygame blits behind them, and DataProcessor.process_observation's resize / transpose /
// stack (gym-air-hockey/gym_air_hockey/processor.py:31-49).  NEAREST resize of a native 450x700 drawing == evaluating
// the drawing at the 128x128 sample points (SPEC.md §8), so nothing is drawn at native resolution.
//
// Layout of the work (B200: 148 SMs, one persistent CTA of 32 warps per SM):
//   * CTA b owns a contiguous block of environments.  One thread per env runs the game tick in registers on the SoA
//     state (coalesced) and leaves 16-byte draw records in shared memory — the step's only block-wide barrier.
//   * The table image at the sample points (48 KB) is fetched once per CTA with one bulk async copy (TMA engine,
//     mbarrier completion) while the physics runs.
//   * Rasterising is barrier-free and warp-independent.  A work unit is (frame, 16 output rows); a warp takes units
//     round-robin.  Per 4-row band each lane owns one 16-pixel vector position: it loads the table's three channel
//     vectors from shared memory, paints the sprite layers that reach its row as column intervals (integer-sqrt
//     half-widths and the exact inverse of the sample map come from small shared-memory tables; an interval becomes
//     a byte mask, a layer three LOP3s per word), and streams the result out with 128-bit st.global.cs.  Bands no
//     sprite reaches — most of them — are a straight shared->global copy.
// A pure write stream on this GPU peaks at 6.0-6.5 TB/s (tools/wprobe.cu, profiles/wprobe_r1b.json); the design goal
// is to keep the SM side far enough below its issue budget that the stores run at that rate.
#include "ahk_env.cuh"

namespace ahk {

namespace {

__device__ __forceinline__ constexpr uint32_t rgb(uint32_t r, uint32_t g, uint32_t b) { return r | g << 8 | b << 16; }

// SPEC.md §7 colours
constexpr uint32_t kLineColour = rgb(235, 235, 225);
constexpr uint32_t kPuckRim = rgb(150, 20, 30), kPuckBody = rgb(215, 40, 50), kPuckShine = rgb(240, 120, 130);
constexpr uint32_t kBotRim = rgb(20, 20, 24), kBotBody = rgb(40, 40, 48), kBotKnob = rgb(90, 90, 100), kBotShine = rgb(220, 220, 230);
constexpr uint32_t kTopRim = rgb(35, 20, 20), kTopBody = rgb(70, 40, 40), kTopKnob = rgb(110, 70, 70), kTopShine = rgb(225, 215, 215);
constexpr uint32_t kArmEdge = rgb(160, 110, 95), kArmInner = rgb(200, 150, 130);

// native sample column / row of output pixel j / i: Pillow NEAREST, int((j + 0.5) * 450 / 128) (SPEC.md §8)
__device__ __forceinline__ int sample_y(int i) { return ((2 * i + 1) * 175) >> 6; }

// Shared-memory lookup tables, built once per CTA:
//   col_lo[x]  = first output column whose sample column is >= x            (128: none)
//   col_hi1[x] = number of output columns whose sample column is <= x       (last such column + 1)
//   isqrt[v]   = floor(sqrt(v)), 0 <= v <= 576 (the largest squared radius)
struct Luts {
    uint8_t col_lo[kTableW + 2];
    uint8_t col_hi1[kTableW + 2];
    uint8_t isqrt[580];
};

constexpr int kRecordBytes = 3 * kStepThreads * 16;
constexpr int kStepSmem = kFrameBytes + kRecordBytes + static_cast<int>(sizeof(Luts));

__device__ __forceinline__ void build_luts(Luts &l)
{
    for (int x = threadIdx.x; x < kTableW; x += blockDim.x) {
        // sample_x(j) >= x  <=>  j >= ceil(128 x / 225) >> 1
        l.col_lo[x] = static_cast<uint8_t>(min(128, ((128 * x + 224) / 225) >> 1));
        l.col_hi1[x] = static_cast<uint8_t>(min(128, ((128 * (x + 1) + 224) / 225) >> 1));
    }
    for (int v = threadIdx.x; v <= 576; v += blockDim.x) {
        int h = static_cast<int>(sqrtf(static_cast<float>(v)));
        if (h * h > v) --h;
        if ((h + 1) * (h + 1) <= v) ++h;
        l.isqrt[v] = static_cast<uint8_t>(h);
    }
}

// 4 pixel bits -> 4 byte lanes of 0xFF
__device__ __forceinline__ uint32_t byte_lanes(uint32_t nibble) { return ((nibble * 0x00204081u) & 0x01010101u) * 0xFFu; }

__device__ __forceinline__ constexpr uint32_t splat(uint32_t colour, int channel) { return ((colour >> (8 * channel)) & 0xFFu) * 0x01010101u; }

// paint native columns x0..x1 of this lane's row with `colour`, inside the 16-pixel vector starting at column `base`
template <uint32_t kColour>
__device__ __forceinline__ void fill(uint32_t (&o)[3][4], const Luts &l, int x0, int x1, int base)
{
    x0 = max(x0, 0);
    x1 = min(x1, kTableW - 1);
    if (x0 > x1) return;
    const int lo = l.col_lo[x0], hi = static_cast<int>(l.col_hi1[x1]) - 1;
    const int a = max(lo - base, 0), b = min(hi - base, 15);
    if (a > b) return;
    const uint32_t m16 = (0xFFFFu >> (15 - b)) & (0xFFFFu << a);
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        const uint32_t m = byte_lanes((m16 >> (4 * g)) & 15u);
#pragma unroll
        for (int c = 0; c < 3; ++c) o[c][g] = (o[c][g] & ~m) | (splat(kColour, c) & m);
    }
}

// a disc of squared radius kR2 centred (cx, cy): on the row at vertical offset dy it covers cx - hw .. cx + hw
template <int kR2, uint32_t kColour>
__device__ __forceinline__ void disc(uint32_t (&o)[3][4], const Luts &l, int cx, int dy, int base)
{
    const int v = kR2 - dy * dy;
    if (v < 0) return;
    const int hw = l.isqrt[v];
    fill<kColour>(o, l, cx - hw, cx + hw, base);
}

// SPEC.md §7 items 4, 5 in painter's order: rim (d2 <= 576), body (<= 400), knob (<= 100), shine (3 px disc at (-3,-3))
template <uint32_t kRim, uint32_t kBody, uint32_t kKnob, uint32_t kShine>
__device__ __forceinline__ void mallet(uint32_t (&o)[3][4], const Luts &l, int cx, int cy, int y, int base)
{
    const int dy = y - cy;
    if (dy < -24 || dy > 24) return;
    disc<576, kRim>(o, l, cx, dy, base);
    disc<400, kBody>(o, l, cx, dy, base);
    disc<100, kKnob>(o, l, cx, dy, base);
    disc<9, kShine>(o, l, cx - 3, dy + 3, base);
}

__device__ __forceinline__ void mbarrier_wait(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n\t.reg .pred done;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 done, [%0], %1;\n\t"
        "@!done bra WAIT_%=;\n\t}"
        :: "r"(bar), "r"(parity) : "memory");
}

// One work unit: output rows 16q .. 16q+15 of one frame, by one warp.
__device__ __forceinline__ void draw_unit(const uint4 *__restrict__ s_table, const Luts &l, const uint4 rec, int q,
                                          uint8_t *__restrict__ frame, uint32_t env_id, uint2 key, int lane)
{
    uint4 *out = reinterpret_cast<uint4 *>(frame);
    if (rec.w & kDrawNoise) { // SPEC.md §7 last paragraph: a `random` episode's frame is Philox noise
#pragma unroll 1
        for (int t = 0; t < 12; ++t) {
            const int v = (t >> 2) * 1024 + q * 128 + (t & 3) * 32 + lane;
            uint4 r = philox(static_cast<uint32_t>(v), rec.x, env_id, 2u, key);
            r.x = __vsubus4(r.x, 0x01010101u); // per byte b -> (b * 255) >> 8 == max(b - 1, 0)
            r.y = __vsubus4(r.y, 0x01010101u);
            r.z = __vsubus4(r.z, 0x01010101u);
            r.w = __vsubus4(r.w, 0x01010101u);
            __stcs(out + v, r);
        }
        return;
    }

    const int ipx = static_cast<int>(rec.x & 0xFFFFu), ipy = static_cast<int>(rec.x >> 16);
    const int ibx = static_cast<int>(rec.y & 0xFFFFu), iby = static_cast<int>(rec.y >> 16);
    const int itx = static_cast<int>(rec.z & 0xFFFFu), ity = static_cast<int>(rec.z >> 16);
    const bool use_arm = rec.w & kUseArm, use_puck = rec.w & kUsePuck, use_top = rec.w & kUseTop;
    const int k = lane & 7, base = 16 * k;

#pragma unroll 1
    for (int band = 0; band < 4; ++band) {
        const int i0 = 16 * q + 4 * band;
        const int i = i0 + (lane >> 3), pos = i * 8 + k;
        const uint4 t0 = s_table[pos], t1 = s_table[1024 + pos], t2 = s_table[2048 + pos];

        // which sprites reach the band's rows at all (warp-uniform)
        const int yb0 = sample_y(i0), yb1 = sample_y(i0 + 3);
        const bool line_b = yb1 >= iby - 3 && yb0 <= iby + 2;
        const bool puck_b = use_puck && yb1 >= ipy - 18 && yb0 <= ipy + 18;
        const bool bot_b = yb1 >= iby - 24 && yb0 <= iby + 24;
        const bool top_b = use_top && yb1 >= ity - 24 && yb0 <= ity + 24;
        const bool arm_b = use_arm && yb0 <= ity + 10;
        if (!(line_b || puck_b || bot_b || top_b || arm_b)) { // the table shows through: straight copy
            __stcs(out + pos, t0);
            __stcs(out + 1024 + pos, t1);
            __stcs(out + 2048 + pos, t2);
            continue;
        }

        uint32_t o[3][4] = {{t0.x, t0.y, t0.z, t0.w}, {t1.x, t1.y, t1.z, t1.w}, {t2.x, t2.y, t2.z, t2.w}};
        const int y = sample_y(i);
        if (line_b && y >= iby - 3 && y <= iby + 2) { // item 2: the rail line spans the full width
#pragma unroll
            for (int c = 0; c < 3; ++c)
#pragma unroll
                for (int g = 0; g < 4; ++g) o[c][g] = splat(kLineColour, c);
        }
        if (puck_b) { // item 3: rim (d2 <= 324), body (<= 196), shine (4 px disc at (-5,-5))
            const int dy = y - ipy;
            if (dy >= -18 && dy <= 18) {
                disc<324, kPuckRim>(o, l, ipx, dy, base);
                disc<196, kPuckBody>(o, l, ipx, dy, base);
                disc<16, kPuckShine>(o, l, ipx - 5, dy + 5, base);
            }
        }
        if (bot_b) mallet<kBotRim, kBotBody, kBotKnob, kBotShine>(o, l, ibx, iby, y, base);
        if (top_b) mallet<kTopRim, kTopBody, kTopKnob, kTopShine>(o, l, itx, ity, y, base);
        if (arm_b && y <= ity + 10) { // item 6
            const int t = ity + 10 - y;
            const int xc = itx + ((t * 5) >> 5), hw = 26 + ((t * 3) >> 4);
            fill<kArmEdge>(o, l, xc - hw, xc + hw, base);
            fill<kArmInner>(o, l, xc - hw + 4, xc + hw - 4, base);
        }
        __stcs(out + pos, make_uint4(o[0][0], o[0][1], o[0][2], o[0][3]));
        __stcs(out + 1024 + pos, make_uint4(o[1][0], o[1][1], o[1][2], o[1][3]));
        __stcs(out + 2048 + pos, make_uint4(o[2][0], o[2][1], o[2][2], o[2][3]));
    }
}

} // namespace

// kMode 0: AirHockeyEnv.reset + draw; 1: AirHockeyEnv.step + draw; 2: draw only, records from the global draw list
// (the split path, after env_kernel).
template <int kMode>
__global__ void __launch_bounds__(kStepThreads, 1) step_kernel(StepArgs a)
{
    extern __shared__ __align__(128) uint8_t smem[];
    const uint4 *s_table = reinterpret_cast<const uint4 *>(smem);                 // [3][128][8] vectors
    uint4 *s_recs = reinterpret_cast<uint4 *>(smem + kFrameBytes);                // valid draw records of this round
    Luts &luts = *reinterpret_cast<Luts *>(smem + kFrameBytes + kRecordBytes);
    __shared__ __align__(8) uint64_t s_table_ready;
    __shared__ int s_count;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t bar = static_cast<uint32_t>(__cvta_generic_to_shared(&s_table_ready));
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(bar) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0) { // the table: one bulk async copy global -> shared (TMA engine), overlapping the physics below
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(kFrameBytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     :: "r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem))), "l"(a.ras.table), "r"(kFrameBytes), "r"(bar) : "memory");
    }
    build_luts(luts);
    bool table_ready = false;

    const int n = a.ras.n;
    const int lo = blockIdx.x * a.envs_per_cta, hi = min(n, lo + a.envs_per_cta);
    const size_t stride = static_cast<size_t>(n);

    for (int base = lo; base < hi; base += kStepThreads) {
        if (tid == 0) s_count = 0;
        __syncthreads(); // s_count cleared; previous round's records consumed; first round: tables built, barrier initialised

        const int i = base + tid;
        const bool live = i < hi;
        uint4 rec[3] = {make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0)};
        if (kMode < 2) {
            StatDelta d;
            if (live) env_update<kMode == 1>(a.env, i, rec, d);
            accumulate_stats<kMode == 1>(a.env.stats, d);
        } else if (live) {
            const uint4 *dl = reinterpret_cast<const uint4 *>(a.ras.drawlist);
#pragma unroll
            for (int b = 0; b < 3; ++b) rec[b] = __ldg(dl + b * stride + i);
        }
        // compact the valid records into shared memory: one atomic per warp per ring slot
#pragma unroll
        for (int b = 0; b < 3; ++b) {
            const bool valid = live && (rec[b].w & kDrawValid);
            const unsigned votes = __ballot_sync(0xFFFFFFFFu, valid);
            if (votes == 0u) continue;
            int at = 0;
            if (lane == 0) at = atomicAdd(&s_count, __popc(votes));
            at = __shfl_sync(0xFFFFFFFFu, at, 0) + __popc(votes & ((1u << lane) - 1u));
            if (valid) s_recs[at] = make_uint4(rec[b].x, rec[b].y, rec[b].z, (rec[b].w & 0xFFu) | static_cast<uint32_t>(tid) << 8 | static_cast<uint32_t>(b) << 20);
        }
        __syncthreads();
        const int units = s_count * 8;
        if (!table_ready) {
            mbarrier_wait(bar, 0u);
            table_ready = true;
        }
        for (int u = warp; u < units; u += kStepThreads / 32) {
            const uint4 rec_u = s_recs[u >> 3];
            const int env = base + static_cast<int>((rec_u.w >> 8) & 0x3FFu), back = static_cast<int>(rec_u.w >> 20); // back 0 = newest
            const int slot = (a.ras.head + 3 - back) % 3;
            uint8_t *frame = a.ras.frames + (static_cast<size_t>(slot) * stride + env) * kFrameBytes;
            draw_unit(s_table, luts, rec_u, u & 7, frame, a.ras.env_id_base + static_cast<uint32_t>(env), a.ras.key, lane);
        }
    }
    if (!table_ready) mbarrier_wait(bar, 0u); // never exit with the table copy still in flight
}

cudaError_t step_prepare()
{
    cudaError_t e = cudaFuncSetAttribute(step_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStepSmem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(step_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStepSmem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(step_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStepSmem);
    return e;
}

cudaError_t launch_step(const StepArgs &a, int mode, int grid, cudaStream_t stream)
{
    if (mode == 0) step_kernel<0><<<grid, kStepThreads, kStepSmem, stream>>>(a);
    else if (mode == 1) step_kernel<1><<<grid, kStepThreads, kStepSmem, stream>>>(a);
    else step_kernel<2><<<grid, kStepThreads, kStepSmem, stream>>>(a);
    return cudaGetLastError();
}

} // namespace ahk
