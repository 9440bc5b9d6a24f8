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
//     state (coalesced) and leaves a 16-byte draw record per frame to paint in shared memory — the step's only
//     block-wide barriers.  Running statistics are summed in shared memory and flushed once per CTA.
//   * The table image at the sample points (48 KB) is fetched once per CTA with one bulk async copy (TMA engine,
//     mbarrier completion) while the physics runs.
//   * Rasterising is barrier-free and warp-independent.  A work unit is (frame, 8 output rows) = 3 KB; warps take
//     units round-robin.  A unit no sprite reaches — two thirds of them — is a straight shared -> global copy (six
//     128-bit loads and st.global.cs per lane).  Otherwise the unit is composed in a per-warp 3 KB patch, sprite by
//     sprite in painter's order, with the lanes laid over the sprite's own bounding vectors: a row of a disc is a
//     column interval (integer-sqrt half-width and the exact inverse of the sample map come from small shared-memory
//     tables), an interval becomes a byte mask, a layer three LOP3s per word; then the patch is streamed out.
// A pure write stream on this GPU peaks at 6.0-6.5 TB/s (tools/wprobe.cu, profiles/wprobe_r1b.json).  Measured
// alternatives that lost: bulk-copying the table to each frame and overlaying the sprite vectors
// afterwards (partial rewrites cost more DRAM time than they save SM time), bulk copies for runs of untouched units,
// a global ticket pool for the tail, spreading the envs of a CTA one per warp, an even 148-way split of the envs.
// Consecutive step launches overlap by programmatic dependent launch.
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
static_assert(2 * 3 * kManyEnvs * 16 <= kRecordBytes, "ahk_step_many keeps two record buffers in the space of one");

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
__device__ __forceinline__ bool fill(uint32_t (&o)[3][4], const Luts &l, int x0, int x1, int base)
{
    x0 = max(x0, 0);
    x1 = min(x1, kTableW - 1);
    if (x0 > x1) return false;
    const int lo = l.col_lo[x0], hi = static_cast<int>(l.col_hi1[x1]) - 1;
    const int a = max(lo - base, 0), b = min(hi - base, 15);
    if (a > b) return false;
    const uint32_t m16 = (0xFFFFu >> (15 - b)) & (0xFFFFu << a);
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        const uint32_t m = byte_lanes((m16 >> (4 * g)) & 15u);
#pragma unroll
        for (int c = 0; c < 3; ++c) o[c][g] = (o[c][g] & ~m) | (splat(kColour, c) & m);
    }
    return true;
}

// a disc of squared radius kR2 centred (cx, cy): on the row at vertical offset dy it covers cx - hw .. cx + hw
template <int kR2, uint32_t kColour>
__device__ __forceinline__ bool disc(uint32_t (&o)[3][4], const Luts &l, int cx, int dy, int base)
{
    const int v = kR2 - dy * dy;
    if (v < 0) return false;
    const int hw = l.isqrt[v];
    return fill<kColour>(o, l, cx - hw, cx + hw, base);
}

// SPEC.md §7 items 4, 5 in painter's order: rim (d2 <= 576), body (<= 400), knob (<= 100), shine (3 px disc at (-3,-3))
template <uint32_t kRim, uint32_t kBody, uint32_t kKnob, uint32_t kShine>
__device__ __forceinline__ bool mallet(uint32_t (&o)[3][4], const Luts &l, int cx, int dy, int base)
{
    bool hit = disc<576, kRim>(o, l, cx, dy, base);
    hit |= disc<400, kBody>(o, l, cx, dy, base);
    hit |= disc<100, kKnob>(o, l, cx, dy, base);
    hit |= disc<9, kShine>(o, l, cx - 3, dy + 3, base);
    return hit;
}

__device__ __forceinline__ unsigned long long global_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
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

// ---- a work unit: kUnitRows output rows of one frame, by one warp, through a per-warp patch in shared memory
constexpr int kUnitRows = 8;
constexpr int kUnitVecs = kUnitRows * 8;         // vector positions per channel in a unit
constexpr int kUnitsPerFrame = kDim / kUnitRows; // 16
constexpr int kPatchBytes = 3 * kUnitVecs * 16;  // 3 KB per warp

// One sprite pass over the warp's patch, sprite-major: lane = (row lane >> 2 of the unit, vector k0 + (lane & 3)), so a
// sprite up to four vectors wide costs ONE pass per unit with every lane on a vector the sprite can reach.
// reach(y): does the sprite touch native row y at all; paint(o, y, base): apply its layers to the 16-pixel vector
// starting at column `base`, returning whether anything changed.  Passes run in painter's order (SPEC.md §7).
template <typename Reach, typename Paint>
__device__ __forceinline__ void sprite_pass(uint4 *patch, int i0, int k0, int lane, Reach reach, Paint paint)
{
    const int r = lane >> 2, k = k0 + (lane & 3);
    const int y = sample_y(i0 + r);
    if (k <= 7 && reach(y)) {
        const int at = r * 8 + k;
        const uint4 t0 = patch[at], t1 = patch[kUnitVecs + at], t2 = patch[2 * kUnitVecs + at];
        uint32_t o[3][4] = {{t0.x, t0.y, t0.z, t0.w}, {t1.x, t1.y, t1.z, t1.w}, {t2.x, t2.y, t2.z, t2.w}};
        if (paint(o, y, 16 * k)) {
            patch[at] = make_uint4(o[0][0], o[0][1], o[0][2], o[0][3]);
            patch[kUnitVecs + at] = make_uint4(o[1][0], o[1][1], o[1][2], o[1][3]);
            patch[2 * kUnitVecs + at] = make_uint4(o[2][0], o[2][1], o[2][2], o[2][3]);
        }
    }
    __syncwarp();
}

// first vector (0..7; 8 = none) a sprite whose leftmost native column is x0 can reach
__device__ __forceinline__ int first_vector(const Luts &l, int x0) { return l.col_lo[min(max(x0, 0), kTableW - 1)] >> 4; }

__device__ __forceinline__ void draw_unit(const uint4 *__restrict__ s_table, uint4 *__restrict__ patch, const Luts &l,
                                          const uint4 rec, int q, uint8_t *__restrict__ frame, uint32_t env_id, uint2 key, int lane)
{
    uint4 *out = reinterpret_cast<uint4 *>(frame);
    const int p0 = q * kUnitVecs + lane, p1 = p0 + 32; // this lane's two vector positions (per channel) of the unit
    if (rec.w & kDrawNoise) { // SPEC.md §7 last paragraph: a `random` episode's frame is Philox noise
#pragma unroll 1
        for (int t = 0; t < 6; ++t) {
            const int v = (t >> 1) * 1024 + ((t & 1) ? p1 : p0);
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

    // which sprites reach the unit's rows at all (warp-uniform)
    const int i0 = q * kUnitRows;
    const int yb0 = sample_y(i0), yb1 = sample_y(i0 + kUnitRows - 1);
    const bool line_u = yb1 >= iby - 3 && yb0 <= iby + 2;
    const bool puck_u = (rec.w & kUsePuck) && yb1 >= ipy - 18 && yb0 <= ipy + 18;
    const bool bot_u = yb1 >= iby - 24 && yb0 <= iby + 24;
    const bool top_u = (rec.w & kUseTop) && yb1 >= ity - 24 && yb0 <= ity + 24;
    const bool arm_u = (rec.w & kUseArm) && yb0 <= ity + 10;

    uint4 v[6];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        v[2 * c] = s_table[c * 1024 + p0];
        v[2 * c + 1] = s_table[c * 1024 + p1];
    }
    if (!(line_u || puck_u || bot_u || top_u || arm_u)) { // the table shows through: straight shared -> global copy
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            __stcs(out + c * 1024 + p0, v[2 * c]);
            __stcs(out + c * 1024 + p1, v[2 * c + 1]);
        }
        return;
    }

    __syncwarp(); // the previous unit's reads of the patch are done
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        patch[c * kUnitVecs + lane] = v[2 * c];
        patch[c * kUnitVecs + 32 + lane] = v[2 * c + 1];
    }
    __syncwarp();

    if (line_u) { // item 2: the rail line spans the full width; lane = (row, vector), two halves of the unit
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int at = h * 32 + lane, y = sample_y(i0 + (at >> 3));
            if (y >= iby - 3 && y <= iby + 2) {
#pragma unroll
                for (int c = 0; c < 3; ++c) { const uint32_t w = splat(kLineColour, c); patch[c * kUnitVecs + at] = make_uint4(w, w, w, w); }
            }
        }
        __syncwarp();
    }
    if (puck_u) // item 3: rim (d2 <= 324), body (<= 196), shine (4 px disc at (-5,-5))
        sprite_pass(patch, i0, first_vector(l, ipx - 18), lane,
                    [&](int y) { return y >= ipy - 18 && y <= ipy + 18; },
                    [&](uint32_t (&o)[3][4], int y, int base) {
                        const int dy = y - ipy;
                        bool hit = disc<324, kPuckRim>(o, l, ipx, dy, base);
                        hit |= disc<196, kPuckBody>(o, l, ipx, dy, base);
                        hit |= disc<16, kPuckShine>(o, l, ipx - 5, dy + 5, base);
                        return hit;
                    });
    if (bot_u) // item 4
        sprite_pass(patch, i0, first_vector(l, ibx - 24), lane,
                    [&](int y) { return y >= iby - 24 && y <= iby + 24; },
                    [&](uint32_t (&o)[3][4], int y, int base) { return mallet<kBotRim, kBotBody, kBotKnob, kBotShine>(o, l, ibx, y - iby, base); });
    if (top_u) // item 5
        sprite_pass(patch, i0, first_vector(l, itx - 24), lane,
                    [&](int y) { return y >= ity - 24 && y <= ity + 24; },
                    [&](uint32_t (&o)[3][4], int y, int base) { return mallet<kTopRim, kTopBody, kTopKnob, kTopShine>(o, l, itx, y - ity, base); });
    if (arm_u) { // item 6: a slanted band, widest at the top row of the unit; conservative column range, exact per lane
        const int t_max = ity + 10 - yb0;
        const int x_min = itx - 26 - ((t_max * 3) >> 4), x_max = itx + ((t_max * 5) >> 5) + 26 + ((t_max * 3) >> 4);
        const int k_min = first_vector(l, x_min);
        const int k_max = x_max < 0 ? -1 : (static_cast<int>(l.col_hi1[min(x_max, kTableW - 1)]) - 1) >> 4;
        for (int k0 = k_min; k0 <= k_max; k0 += 4)
            sprite_pass(patch, i0, k0, lane,
                        [&](int y) { return y <= ity + 10; },
                        [&](uint32_t (&o)[3][4], int y, int base) {
                            const int t = ity + 10 - y;
                            const int xc = itx + ((t * 5) >> 5), hw = 26 + ((t * 3) >> 4);
                            bool hit = fill<kArmEdge>(o, l, xc - hw, xc + hw, base);
                            hit |= fill<kArmInner>(o, l, xc - hw + 4, xc + hw - 4, base);
                            return hit;
                        });
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        v[2 * c] = patch[c * kUnitVecs + lane];
        v[2 * c + 1] = patch[c * kUnitVecs + 32 + lane];
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        __stcs(out + c * 1024 + p0, v[2 * c]);
        __stcs(out + c * 1024 + p1, v[2 * c + 1]);
    }
}

// Where the frame of a draw record goes.  back = 0: the newest frame, ring slot `head`.  back = 1, 2: the warm-up frames
// that a reset puts behind it in the stack (air_hockey_env.py:69-72) — in the 3-slot ring they rewrite the two older
// slots; in a deep ring they go to the env's side buffer (stack_frame() in ahk_kernels.cuh reads them back), so that the
// previous episode's last frames stay in the ring for replay.  An explicit reset (kReset: no step follows in the launch)
// files all three frames w1, w2, w3 (back 2, 1, 0) at side-buffer indices 0, 1, 2; a step that resets first drops w1 and
// files w2, w3 (back 2, 1) at indices 1, 2.
template <bool kReset>
__device__ __forceinline__ uint8_t *record_frame(const RasterArgs &r, int head, int env, int back)
{
    const size_t n = static_cast<size_t>(r.n);
    if (r.warm != nullptr && (kReset || back > 0))
        return r.warm + (static_cast<size_t>((kReset ? 2 : 3) - back) * n + static_cast<size_t>(env)) * kFrameBytes;
    return r.frames + (static_cast<size_t>((head + r.depth - back) % r.depth) * n + static_cast<size_t>(env)) * kFrameBytes;
}

constexpr int kLutOffset = kFrameBytes + kRecordBytes + (kStepThreads / 32) * kPatchBytes;
constexpr int kStepSmem = kLutOffset + static_cast<int>(sizeof(Luts));

} // namespace

// kMode 0: AirHockeyEnv.reset + draw; 1: AirHockeyEnv.step + draw; 2: draw only, records from the global draw list
// (the split path, after env_kernel).
// Adaptive partition, end of a step launch.  SMs do not drain their stores equally fast, and with overlapped launches
// the SM that finishes last also starts the next launch last: left alone, lateness feeds on itself.  So the shares are
// steered on what matters, the END time: the last warp of each CTA stamps %globaltimer; the last CTA of the grid moves
// every CTA's share of the envs against its lateness relative to the mean (an integral controller, gain 1/2, shares kept
// within +-50 % of even) and writes the boundaries a later launch will use.  Results do not depend on the partition:
// state lives in HBM, every env is ticked and drawn exactly once whoever owns it.
__device__ __forceinline__ void rebalance(const StepArgs &a, int slot, unsigned *warps_done, const unsigned long long *raster_start, int lane)
{
    unsigned k = 0;
    if (lane == 0) { __threadfence_block(); k = atomicAdd(warps_done, 1u); }
    k = __shfl_sync(0xFFFFFFFFu, k, 0);
    if (k != kStepThreads / 32 - 1) return; // this warp is the CTA's last
    const int grid = static_cast<int>(gridDim.x);
    unsigned finished = 0;
    unsigned long long end = 0;
    if (lane == 0) {
        end = global_ns();
        a.end_ns[slot] = end;
        __threadfence();
        finished = atomicAdd(a.done_count, 1u);
    }
    finished = __shfl_sync(0xFFFFFFFFu, finished, 0);
    if (finished != static_cast<unsigned>(grid) - 1u) return; // this CTA is the grid's last
    __threadfence();
    end = __shfl_sync(0xFFFFFFFFu, end, 0);
    const float span = fmaxf(static_cast<float>(end - *raster_start), 1000.0f); // ns this CTA spent rasterising: the scale
    constexpr int kPer = 8; // CTAs per lane: up to 256 CTAs
    float early[kPer], share[kPer], sum_early = 0.0f;
#pragma unroll
    for (int j = 0; j < kPer; ++j) {
        const int b = lane * kPer + j;
        early[j] = share[j] = 0.0f;
        if (b < grid) {
            const unsigned long long e = __ldcg(a.end_ns + b);
            early[j] = e < end ? static_cast<float>(end - e) : 0.0f; // ns before the last CTA
            share[j] = static_cast<float>(a.bounds[b + 1] - a.bounds[b]);
            sum_early += early[j];
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum_early += __shfl_xor_sync(0xFFFFFFFFu, sum_early, o);
    const float mean_early = sum_early / static_cast<float>(grid);
    const float even = static_cast<float>(a.ras.n) / static_cast<float>(grid);
    float mine = 0.0f;
#pragma unroll
    for (int j = 0; j < kPer; ++j) {
        const int b = lane * kPer + j;
        if (b < grid) {
            const float late = (mean_early - early[j]) / span; // > 0: later than the average CTA, as a fraction of the launch
            share[j] = fminf(fmaxf(share[j] - 0.5f * late * even, 0.5f * even), 1.5f * even);
            mine += share[j];
        }
    }
    float before = mine; // inclusive prefix over the lanes
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const float up = __shfl_up_sync(0xFFFFFFFFu, before, o);
        if (lane >= o) before += up;
    }
    const float all = __shfl_sync(0xFFFFFFFFu, before, 31);
    const float scale = static_cast<float>(a.ras.n) / all;
    float cum = before - mine;
    if (lane == 0) a.bounds_next[0] = 0;
#pragma unroll
    for (int j = 0; j < kPer; ++j) {
        const int b = lane * kPer + j;
        if (b < grid) {
            cum += share[j];
            a.bounds_next[b + 1] = b == grid - 1 ? a.ras.n : min(a.ras.n, __float2int_rn(cum * scale));
        }
    }
    if (lane == 0) *a.done_count = 0u;
}

// kTrace: the measurement build of the same kernel (ahk_set_trace) — per warp four %globaltimer stamps: 0 CTA start,
// 1 game tick done, 2 records + table staged, 3 last store issued.  The product kernel carries none of it.
template <int kMode, bool kTrace = false>
__global__ void __launch_bounds__(kStepThreads, 1) step_kernel(StepArgs a)
{
    extern __shared__ __align__(128) uint8_t smem[];
    const uint4 *s_table = reinterpret_cast<const uint4 *>(smem);                 // [3][128][8] vectors
    uint4 *s_recs = reinterpret_cast<uint4 *>(smem + kFrameBytes);                // valid draw records of this round
    Luts &luts = *reinterpret_cast<Luts *>(smem + kLutOffset);
    __shared__ __align__(8) uint64_t s_table_ready;
    __shared__ int s_count;
    __shared__ int s_next;
    __shared__ int s_count_old, s_next_old; // records that rewrite OLDER ring slots (after a reset), kept apart: see a.pdl == 3
    __shared__ unsigned s_stats[kStatWords];
    __shared__ unsigned s_warps_done;
    __shared__ unsigned long long s_raster_start;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned long long *trace = kTrace ? a.trace + (static_cast<size_t>(blockIdx.x) * (kStepThreads / 32) + warp) * 4 : nullptr;
    if (kTrace && lane == 0) trace[0] = global_ns();
    uint4 *s_patch = reinterpret_cast<uint4 *>(smem + kFrameBytes + kRecordBytes + warp * kPatchBytes); // this warp's scratch
    const uint32_t bar = static_cast<uint32_t>(__cvta_generic_to_shared(&s_table_ready));
    if (tid < kStatWords) s_stats[tid] = 0u;
    if (tid == 0) { s_warps_done = 0u; s_raster_start = 0ull; }
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
    bool table_ready = false, waited = false;
    // Caller-produced inputs (action labels, reset masks, overrides) are read by env_update below: what an earlier kernel
    // or copy on the stream wrote is only guaranteed visible after the wait.
    if (a.wait_first) { asm volatile("griddepcontrol.wait;" ::: "memory"); waited = true; }

    const int n = a.ras.n;
    const int part = static_cast<int>(blockIdx.x);
    int lo = part * a.envs_per_cta, hi = min(n, lo + a.envs_per_cta);
    if (a.bounds != nullptr) { lo = a.bounds[part]; hi = a.bounds[part + 1]; }
    const size_t stride = static_cast<size_t>(n);

    for (int base = lo; base < hi; base += kStepThreads) {
        if (tid == 0) { s_count = 0; s_next = 0; s_count_old = 0; s_next_old = 0; }
        __syncthreads(); // s_count cleared; previous round's records consumed; first round: tables built, barrier initialised

        const int i = base + tid;
        const bool live = i < hi;
        uint4 rec[3] = {make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0)};
        if (kMode < 2) {
            StatDelta d;
            if (live) env_update<kMode == 1>(a.env, i, rec, d);
            if (a.env.stats != nullptr) stat_to_shared<kMode == 1>(s_stats, d, live);
        } else if (live) {
            const uint4 *dl = reinterpret_cast<const uint4 *>(a.ras.drawlist);
#pragma unroll
            for (int b = 0; b < 3; ++b) rec[b] = __ldg(dl + b * stride + i);
        }
        if (a.pdl) __threadfence(); // this thread's state and outputs are out before the barrier below, hence before the trigger
        if (kTrace && lane == 0 && base == lo) trace[1] = global_ns();
        // compact the valid records into shared memory: one atomic per warp per ring slot.  Records of the newest frame
        // grow from the bottom of the list, records that rewrite the two older slots (warm-up frames after a reset) from
        // the top, so that the two kinds can be drawn in separate phases.
#pragma unroll
        for (int b = 0; b < 3; ++b) {
            const bool valid = live && (rec[b].w & kDrawValid);
            const unsigned votes = __ballot_sync(0xFFFFFFFFu, valid);
            if (votes == 0u) continue;
            const bool later = b != 0 || (rec[b].w & kDrawDefer); // phase 1: after the wait on the previous launch
            const unsigned votes_later = __ballot_sync(0xFFFFFFFFu, valid && later), votes_now = votes & ~votes_later;
            int at = 0, at_later = 0;
            if (lane == 0) {
                if (votes_now) at = atomicAdd(&s_count, __popc(votes_now));
                if (votes_later) at_later = atomicAdd(&s_count_old, __popc(votes_later));
            }
            at = __shfl_sync(0xFFFFFFFFu, at, 0) + __popc(votes_now & ((1u << lane) - 1u));
            at_later = __shfl_sync(0xFFFFFFFFu, at_later, 0) + __popc(votes_later & ((1u << lane) - 1u));
            if (later) at = 3 * kStepThreads - 1 - at_later;
            if (valid) s_recs[at] = make_uint4(rec[b].x, rec[b].y, rec[b].z, (rec[b].w & 0x1Fu) | static_cast<uint32_t>(tid) << 8 | static_cast<uint32_t>(b) << 20);
        }
        __syncthreads();
        // Programmatic dependent launch: once every CTA has ticked its last round the NEXT step kernel may be scheduled;
        // its CTAs move in as ours retire, and its prologue and game tick (which read only the state fenced above) run
        // while the rest of this grid is still draining its stores.
        if (a.pdl && base + kStepThreads >= hi) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
        // ahk_step_host_results: the host is waiting for reward / done / labels, not for the frames.  The last CTA to finish
        // its game ticks tells it so through mapped host memory; rasterising goes on underneath.
        // One thread of the LAST warp does it: the system-scope fence (a PCIe round trip; cumulative over the outputs the
        // other threads wrote before the barrier above) then delays that one warp's first unit claim, not the CTA's raster.
        if (a.env.h_tick_flag != nullptr && tid == kStepThreads - 32 && base + kStepThreads >= hi) {
            __threadfence_system();
            const unsigned slot = static_cast<unsigned>(a.env.tick_seq & 7ull);
            if (atomicAdd(a.env.tick_count + slot, 1u) == gridDim.x - 1u) {
                a.env.tick_count[slot] = 0u;
                __threadfence_system();
                *reinterpret_cast<volatile unsigned long long *>(a.env.h_tick_flag + slot) = a.env.tick_seq;
            }
        }
        const int units = s_count * kUnitsPerFrame, units_old = s_count_old * kUnitsPerFrame;
        if (!table_ready) {
            mbarrier_wait(bar, 0u);
            table_ready = true;
            // pdl 1: no frame is written before the previous launch has completed
            if (a.pdl == 1 && !waited) { asm volatile("griddepcontrol.wait;" ::: "memory"); waited = true; }
        }
        if (kTrace && lane == 0 && base == lo) trace[2] = global_ns();
        if (a.bounds_next != nullptr && tid == 0 && base == lo) s_raster_start = global_ns();
        // phase 0: the newest frame of every env, into ring slot `head` — a slot the previous launch does not touch.
        // phase 1: frames that replace the two older slots (reset envs only).  The previous launch may still be writing
        // those very frames, so with the wait deferred (pdl 3) it comes first, here.
#pragma unroll 1
        for (int phase = 0; phase < 2; ++phase) {
            const int total = phase == 0 ? units : units_old;
            if (phase == 1) {
                if (total == 0) break;
                if (!waited) { asm volatile("griddepcontrol.wait;" ::: "memory"); waited = true; }
            }
            int *next = phase == 0 ? &s_next : &s_next_old;
            for (int u = warp;; u += kStepThreads / 32) {
                if (a.dyn_units) { // claim the next unit: whichever warp is free takes it (composed units cost more than copies)
                    if (lane == 0) u = atomicAdd(next, 1);
                    u = __shfl_sync(0xFFFFFFFFu, u, 0);
                }
                if (u >= total) break;
                const int r = u / kUnitsPerFrame;
                const uint4 rec_u = s_recs[phase == 0 ? r : 3 * kStepThreads - 1 - r];
                const int env = base + static_cast<int>((rec_u.w >> 8) & 0x3FFu), back = static_cast<int>(rec_u.w >> 20); // back 0 = newest
                uint8_t *frame = record_frame<kMode == 0>(a.ras, a.ras.head, env, back);
                draw_unit(s_table, s_patch, luts, rec_u, u % kUnitsPerFrame, frame, a.ras.env_id_base + static_cast<uint32_t>(env), a.ras.key, lane);
            }
        }
        if (base + kStepThreads < hi) __syncthreads(); // every warp has read this round's record count before it is cleared
    }
    if (kTrace && lane == 0) trace[3] = global_ns();
    // pdl 3: the wait sits at the END.  It keeps completion in launch order (this grid cannot complete before the previous
    // one has), which is all that later launches and stream consumers rely on; DESIGN.md §3.5 has the hazard analysis.
    if (a.pdl == 3 && !waited) asm volatile("griddepcontrol.wait;" ::: "memory");
    if (kMode == 1 && a.bounds_next != nullptr) rebalance(a, part, &s_warps_done, &s_raster_start, lane);
    if (!table_ready) mbarrier_wait(bar, 0u); // never exit with the table copy still in flight
    if (kMode < 2 && a.env.stats != nullptr) {
        unsigned long long *st = reinterpret_cast<unsigned long long *>(a.env.stats);
        if (tid < kStatWords && s_stats[tid] != 0u)
            atomicAdd(st + tid, tid == 4 ? static_cast<unsigned long long>(static_cast<long long>(static_cast<int>(s_stats[tid]))) : static_cast<unsigned long long>(s_stats[tid]));
    }
}

// ------------------------------------------------------------------------------------------------ ahk_step_many
// T consecutive ACTION-FREE steps in ONE launch (generate_data.py:44-58: the programmed AI plays both sides, nothing outside
// the engine feeds a step).  Envs are CTA-private, so a CTA simply runs T steps of its own envs: the table and the lookup
// tables are staged once per launch instead of once per step, there is no launch boundary (no grid-wide drain, no cold
// instruction cache) between steps, and the game ticks run on the owner warps (one thread per env) UNDER the stores of
// earlier steps.  State goes through L2 every step as in the single-step kernel (same device code: env_update), so the
// results are those of T ahk_step calls, bit for bit.
//
// No block barrier between steps (round 2: one __syncthreads per step cost ~1 raster unit of idle time per warp per step,
// 7 % at 14 units per warp = 4096 envs).  Steps are decoupled through a ring of K record buffers and per-step counters in
// shared memory; a warp that runs out of units of step t goes straight on to step t + 1:
//   s_tick[t]  owner warps that have finished the game tick of step t  -> the records of step t are complete
//   s_count[t] records of step t;  s_next[t] unit claim counter;  s_left[t] warps that have left step t (all units drawn)
// * an owner warp entering step t first ticks up to step t + A (A = min(2, K - 1)); record slot t' % K is free once every
//   warp has left step t' - K;
// * the slowest and the fastest warp are at most K - 1 steps apart and K <= ring depth, so two steps in flight never write
//   the same ring slot with their NEWEST frames;
// * frames that a reset rewrites (back > 0: the two older ring slots, or the warm-up side buffer of a deep ring) wait until
//   every earlier step still in flight has been drawn, and the newest frame of an env that was reset w < K steps ago
//   waits for that step (its slot is one the reset rewrites when the ring has three slots) — release / acquire through the
//   s_left counters orders the stores to the same address.
constexpr int kManySteps = 512; // steps per launch (per-step counters in shared memory); longer windows take several launches

__device__ __forceinline__ void many_arrive(int *counter, int lane)
{
    __syncwarp();
    if (lane == 0) { __threadfence_block(); atomicAdd(counter, 1); }
}

__device__ __forceinline__ void many_wait(const int *counter, int want, int lane)
{
    if (lane == 0) {
        const volatile int *c = counter;
        while (*c < want) __nanosleep(32);
        __threadfence_block();
    }
    __syncwarp();
}

template <bool kBalance>
__global__ void __launch_bounds__(kStepThreads, 1) step_many_kernel(StepArgs a, int steps, StepLogArgs log, int ring_k)
{
    extern __shared__ __align__(128) uint8_t smem[];
    const uint4 *s_table = reinterpret_cast<const uint4 *>(smem);
    uint4 *s_recs = reinterpret_cast<uint4 *>(smem + kFrameBytes); // [K][3 * cnt32] records: step t uses buffer t % K
    Luts &luts = *reinterpret_cast<Luts *>(smem + kLutOffset);
    __shared__ __align__(8) uint64_t s_table_ready;
    __shared__ int s_count[kManySteps], s_next[kManySteps], s_tick[kManySteps], s_left[kManySteps];
    __shared__ unsigned s_stats[kStatWords];
    __shared__ unsigned s_warps_done;
    __shared__ unsigned long long s_raster_start;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint4 *s_patch = reinterpret_cast<uint4 *>(smem + kFrameBytes + kRecordBytes + warp * kPatchBytes);
    const uint32_t bar = static_cast<uint32_t>(__cvta_generic_to_shared(&s_table_ready));
    if (tid < kStatWords) s_stats[tid] = 0u;
    for (int i = tid; i < kManySteps; i += kStepThreads) { s_count[i] = 0; s_next[i] = 0; s_tick[i] = 0; s_left[i] = 0; }
    if (tid == 0) {
        s_warps_done = 0u;
        s_raster_start = 0ull;
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(bar) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(kFrameBytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     :: "r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem))), "l"(a.ras.table), "r"(kFrameBytes), "r"(bar) : "memory");
    }
    build_luts(luts);

    const int n = a.ras.n;
    const int part = static_cast<int>(blockIdx.x);
    int lo = part * a.envs_per_cta, hi = min(n, lo + a.envs_per_cta);
    if (a.bounds != nullptr) { lo = a.bounds[part]; hi = a.bounds[part + 1]; }
    const int cnt = hi - lo;                       // <= kManyEnvs (checked on the host)
    const int cnt32 = (cnt + 31) & ~31;
    const bool owner = tid < cnt32;                // whole warps: the statistics reduce with warp collectives
    const int owner_warps = cnt32 >> 5;
    const size_t stride = static_cast<size_t>(n);
    // K record buffers: as many as fit, at most the ring depth (see above), at most 8, at least 2
    int K = cnt32 > 0 ? (kRecordBytes / 16) / (3 * cnt32) : 8;
    K = min(min(K, 8), min(a.ras.depth, ring_k < 0 ? 2 : ring_k));
    K = max(K, 2);
    const int ahead = min(2, K - 1);
    const int slot_recs = 3 * cnt32;

    int ticked = 0, since_reset = 15; // owner warps: steps ticked so far; steps since this thread's env was last reset
    // the game tick of step t for this CTA's envs: records into buffer t % K, count in s_count[t]
    auto tick = [&](int t) {
        const bool live = tid < cnt;
        const int i = lo + tid;
        uint4 rec[3] = {make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0)};
        StatDelta d;
        if (live) {
            env_update<true>(a.env, i, rec, d);
            const size_t at = static_cast<size_t>(t) * stride + i;
            if (log.robot_label) log.robot_label[at] = static_cast<uint8_t>(d.robot_label);
            if (log.human_label) log.human_label[at] = static_cast<uint8_t>(d.human_label);
            if (log.done) log.done[at] = static_cast<uint8_t>(d.terminal);
            if (log.reward) log.reward[at] = 0.5f * static_cast<float>(d.reward2);
        }
        if (a.env.stats != nullptr) stat_to_shared<true>(s_stats, d, live);
        since_reset = ((rec[1].w | rec[2].w) & kDrawValid) ? 0 : min(since_reset + 1, 15);
        const uint32_t wait_for = (since_reset > 0 && since_reset < K) ? static_cast<uint32_t>(since_reset) : 0u;
        uint4 *recs = s_recs + (t % K) * slot_recs;
#pragma unroll
        for (int b = 0; b < 3; ++b) {
            const bool valid = live && (rec[b].w & kDrawValid);
            const unsigned votes = __ballot_sync(0xFFFFFFFFu, valid);
            if (votes == 0u) continue;
            int at = 0;
            if (lane == 0) at = atomicAdd(&s_count[t], __popc(votes));
            at = __shfl_sync(0xFFFFFFFFu, at, 0) + __popc(votes & ((1u << lane) - 1u));
            if (valid) recs[at] = make_uint4(rec[b].x, rec[b].y, rec[b].z, (rec[b].w & 0x1Fu) | static_cast<uint32_t>(tid) << 8 | static_cast<uint32_t>(b) << 20 |
                                                                        (b == 0 ? wait_for << 24 : 0u));
        }
        many_arrive(&s_tick[t], lane);
    };

    if (owner && steps > 0) { tick(0); ticked = 1; }
    mbarrier_wait(bar, 0u); // the table is staged
    __syncthreads();        // ... and the lookup tables are built
    if (kBalance && tid == 0) s_raster_start = global_ns();

#pragma unroll 1
    for (int t = 0; t < steps; ++t) {
        if (owner) {
            const int upto = min(steps, t + 1 + ahead);
            while (ticked < upto) {
                if (ticked >= K) many_wait(&s_left[ticked - K], kStepThreads / 32, lane); // the buffer's previous step is drawn
                tick(ticked);
                ++ticked;
            }
        }
        many_wait(&s_tick[t], owner_warps, lane);
        const int total = s_count[t] * kUnitsPerFrame;
        const uint4 *recs = s_recs + (t % K) * slot_recs;
        const int head = (a.ras.head + t) % a.ras.depth; // a.ras.head: the slot of the window's first step
        for (;;) {
            int u = 0;
            if (lane == 0) u = atomicAdd(&s_next[t], 1);
            u = __shfl_sync(0xFFFFFFFFu, u, 0);
            if (u >= total) break;
            const uint4 rec_u = recs[u / kUnitsPerFrame];
            const int env = lo + static_cast<int>((rec_u.w >> 8) & 0x3FFu), back = static_cast<int>((rec_u.w >> 20) & 3u);
            const int wait_for = static_cast<int>((rec_u.w >> 24) & 15u);
            if (back > 0) { // a reset's rewrite of older frames: after every earlier step still in flight
                for (int e = max(0, t - (K - 1)); e < t; ++e) many_wait(&s_left[e], kStepThreads / 32, lane);
            } else if (wait_for > 0 && t - wait_for >= 0) {
                many_wait(&s_left[t - wait_for], kStepThreads / 32, lane); // after the rewrites of the reset `wait_for` steps ago
            }
            uint8_t *frame = record_frame<false>(a.ras, head, env, back);
            draw_unit(s_table, s_patch, luts, rec_u, u % kUnitsPerFrame, frame, a.ras.env_id_base + static_cast<uint32_t>(env), a.ras.key, lane);
        }
        many_arrive(&s_left[t], lane);
        if (ring_k < 0) __syncthreads(); // A/B switch AHK_MANY_DEPTH=0: round 2's barrier per step
    }
    if (kBalance && a.bounds_next != nullptr) rebalance(a, part, &s_warps_done, &s_raster_start, lane);
    __syncthreads(); // every tick's statistics are in shared memory
    if (a.env.stats != nullptr) {
        unsigned long long *st = reinterpret_cast<unsigned long long *>(a.env.stats);
        if (tid < kStatWords && s_stats[tid] != 0u)
            atomicAdd(st + tid, tid == 4 ? static_cast<unsigned long long>(static_cast<long long>(static_cast<int>(s_stats[tid]))) : static_cast<unsigned long long>(s_stats[tid]));
    }
}

int many_launches(int steps) { return (steps + kManySteps - 1) / kManySteps; }

cudaError_t step_prepare()
{
    cudaError_t e = cudaFuncSetAttribute(step_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStepSmem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(step_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStepSmem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(step_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStepSmem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(step_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStepSmem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(step_many_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStepSmem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(step_many_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStepSmem);
    return e;
}

cudaError_t launch_step(const StepArgs &a, int mode, int grid, cudaStream_t stream)
{
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(static_cast<unsigned>(grid));
    cfg.blockDim = dim3(kStepThreads);
    cfg.dynamicSmemBytes = kStepSmem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = a.pdl ? 1 : 0;
    if (mode == 0) return cudaLaunchKernelEx(&cfg, step_kernel<0>, a);
    if (mode == 1 && a.trace) return cudaLaunchKernelEx(&cfg, step_kernel<1, true>, a);
    if (mode == 1) return cudaLaunchKernelEx(&cfg, step_kernel<1>, a);
    return cudaLaunchKernelEx(&cfg, step_kernel<2>, a);
}

// plain launches: no programmatic overlap with their neighbours (the launch boundary is paid once per kManySteps steps)
cudaError_t launch_step_many(const StepArgs &a0, int steps, const StepLogArgs &log0, int grid, cudaStream_t stream)
{
    static int ring_k = -1; // A/B switch AHK_MANY_DEPTH=2..8: record buffers in flight; 0: two buffers and a block barrier per step (round 2)
    if (ring_k < 0) { const char *sw = getenv("AHK_MANY_DEPTH"); ring_k = sw ? atoi(sw) : 8; if (ring_k < 2) ring_k = -1; }
    StepArgs a = a0;
    StepLogArgs log = log0;
    const size_t n = static_cast<size_t>(a.ras.n);
    for (int done = 0; done < steps; done += kManySteps) {
        const int part = steps - done < kManySteps ? steps - done : kManySteps;
        if (a.bounds_next != nullptr) step_many_kernel<true><<<grid, kStepThreads, kStepSmem, stream>>>(a, part, log, ring_k);
        else step_many_kernel<false><<<grid, kStepThreads, kStepSmem, stream>>>(a, part, log, ring_k);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
        a.ras.head = (a.ras.head + part) % a.ras.depth;
        a.bounds_next = nullptr; // the partition adapts on the window's first launch only
        if (log.robot_label) log.robot_label += part * n;
        if (log.human_label) log.human_label += part * n;
        if (log.done) log.done += part * n;
        if (log.reward) log.reward += part * n;
    }
    return cudaSuccess;
}

} // namespace ahk
