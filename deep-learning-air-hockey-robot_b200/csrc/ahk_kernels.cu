// ahk_kernels.cu — the environment hot path as sm_100a kernels.
//
//   env_kernel     K1+K4: per environment (one thread each, SoA state, coalesced): auto-reset, programmed AI,
//                  physics tick, reward / terminal / labels, running statistics, and a 16-byte draw record per frame
//                  to paint.  ≙ AirHockeyEnv.reset/step + AirHockey.reset/step (air_hockey_env.py:26-155).
//   raster_kernel  K2: persistent CTAs; the table's 128x128 samples staged once in shared memory (48 KB); per frame
//                  every thread composes 16-pixel vectors (table + analytic sprites, or Philox noise) and writes them
//                  with 128-bit stores, CHW uint8, straight into the ring slot — the frame goes to HBM exactly once.
//                  ≙ the Pygame blits + DataProcessor.process_observation's resize/transpose/stack
//                  (processor.py:31-49) fused: NEAREST resize == evaluating the scene at the 128x128 sample points.
//   observe_kernel materialises the reference-format observation (processor.py:52-64) from the ring on request.
#include "ahk_kernels.cuh"
#include "ahk_sim.cuh"

namespace ahk {

// ------------------------------------------------------------------------------------------------ K1 + K4
__device__ __forceinline__ void load_sim(const uint32_t *__restrict__ st, int n, int i, Sim &s, uint32_t &reset_count,
                                         uint32_t &noise_count, uint32_t &meta)
{
    auto f = [&](int w) { return __uint_as_float(st[static_cast<size_t>(w) * n + i]); };
    s.puck = Body{f(0), f(1), f(2), f(3)};
    s.bot = Body{f(4), f(5), f(6), f(7)};
    s.top = Body{f(8), f(9), f(10), f(11)};
    s.serve_count = st[static_cast<size_t>(12) * n + i];
    reset_count = st[static_cast<size_t>(13) * n + i];
    noise_count = st[static_cast<size_t>(14) * n + i];
    meta = st[static_cast<size_t>(15) * n + i];
}

__device__ __forceinline__ void store_sim(uint32_t *__restrict__ st, int n, int i, const Sim &s, uint32_t reset_count,
                                          uint32_t noise_count, uint32_t meta)
{
    auto f = [&](int w, float v) { st[static_cast<size_t>(w) * n + i] = __float_as_uint(v); };
    f(0, s.puck.x); f(1, s.puck.y); f(2, s.puck.vx); f(3, s.puck.vy);
    f(4, s.bot.x); f(5, s.bot.y); f(6, s.bot.vx); f(7, s.bot.vy);
    f(8, s.top.x); f(9, s.top.y); f(10, s.top.vx); f(11, s.top.vy);
    st[static_cast<size_t>(12) * n + i] = s.serve_count;
    st[static_cast<size_t>(13) * n + i] = reset_count;
    st[static_cast<size_t>(14) * n + i] = noise_count;
    st[static_cast<size_t>(15) * n + i] = meta;
}

// meta word: bits 0..15 episode_steps, 16..17 episode type, 18..20 use mask, 21 pending reset
__device__ __forceinline__ uint32_t make_meta(uint32_t steps, uint32_t episode, uint32_t use, uint32_t pending)
{
    return (steps > 0xFFFFu ? 0xFFFFu : steps) | episode << 16 | use << 18 | pending << 21;
}

template <bool kStep>
__global__ void __launch_bounds__(kEnvThreads) env_kernel(EnvArgs a)
{
    const int i = blockIdx.x * kEnvThreads + threadIdx.x;
    const bool live = i < a.n;
    const unsigned lane_mask = __activemask();

    uint64_t c_steps = 0, c_episodes = 0, c_top = 0, c_bottom = 0;
    int c_reward2 = 0;
    uint32_t robot_label = 0xFFu, human_label = 0xFFu;

    if (live) {
        Sim s;
        uint32_t reset_count, noise_count, meta;
        load_sim(a.state, a.n, i, s, reset_count, noise_count, meta);
        uint32_t steps = meta & 0xFFFFu, episode = (meta >> 16) & 3u, use = (meta >> 18) & 7u, pending = (meta >> 21) & 1u;
        const uint32_t env_id = a.env_id_base + static_cast<uint32_t>(i);
        const size_t n = static_cast<size_t>(a.n);

        // records for ring slots head (newest), head-1, head-2; invalid unless written below
        uint4 rec[3] = {make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0)};

        bool do_reset;
        if (kStep) do_reset = a.auto_reset && pending;
        else do_reset = a.reset_mask == nullptr || a.reset_mask[i] != 0;

        if (do_reset) { // ≙ AirHockeyEnv.reset, air_hockey_env.py:26-76 (SPEC.md §9)
            if (reset_count == 0u && s.serve_count == 0u) sim_reset(s, env_id, a.key); // construction serves once
            const uint4 r = philox(reset_count, 0u, env_id, 0u, a.key);
            reset_count += 1u;
            uint32_t forced = a.episode_override ? a.episode_override[i] : 255u;
            if (forced == 255u) {
                const uint32_t t = r.x % 20u;
                episode = t < 16u ? kEpNormal : (t < 19u ? kEpMissing : kEpRandom);
                use = kUseAll;
                if (episode == kEpMissing) { // air_hockey_env.py:39-63
                    const uint32_t combo = r.y % 6u;
                    use = (combo >= 4u ? kUseArm : 0u) | ((combo == 2u || combo == 3u) ? kUsePuck : 0u) | ((combo & 1u) ? kUseTop : 0u);
                }
            } else {
                episode = forced;
                use = a.use_override ? (a.use_override[i] & 7u) : kUseAll;
            }
            steps = 0u;
            pending = 0u;
            // warm-up frames, oldest first; with a step following, the oldest is pushed out of the 3-deep stack
            uint4 warm[3];
            if (episode != kEpRandom) {
                sim_reset(s, env_id, a.key);
#pragma unroll
                for (int k = 0; k < 3; ++k) { // stand, stand; use_object omitted => everything in use (air_hockey_env.py:69-72)
                    Tick t;
                    sim_tick(s, 4u, 4u, kUseAll, env_id, a.key, t);
                    warm[k] = draw_record(s, kUseAll);
                }
            } else {
#pragma unroll
                for (int k = 0; k < 3; ++k) warm[k] = noise_record(noise_count++);
            }
            if (kStep) { rec[1] = warm[2]; rec[2] = warm[1]; }
            else { rec[0] = warm[2]; rec[1] = warm[1]; rec[2] = warm[0]; }
            c_episodes = 1;
        }

        if (kStep) { // ≙ AirHockeyEnv.step, air_hockey_env.py:84-155
            const uint32_t robot = a.robot ? a.robot[i] : kActNone;
            const uint32_t human = a.human ? a.human[i] : kActNone;
            Tick t;
            sim_tick(s, robot, human, use, env_id, a.key, t);
            robot_label = static_cast<uint32_t>(t.rax * 3 + t.ray + 4); // processor.py:84-85
            human_label = static_cast<uint32_t>(t.hax * 3 + t.hay + 4);
            const int r2 = reward_x2(episode, t.info, robot_label);
            uint32_t terminal = (t.info >> 5) != 0u;
            if (episode == kEpRandom) { // air_hockey_env.py:147-151
                terminal = 0u;
                rec[0] = noise_record(noise_count++);
            } else {
                rec[0] = draw_record(s, use);
            }
            steps += 1u;
            if (a.auto_reset && (terminal || (a.max_episode_steps > 0 && steps >= static_cast<uint32_t>(a.max_episode_steps)))) pending = 1u;

            a.reward[i] = 0.5f * static_cast<float>(r2);
            a.done[i] = static_cast<uint8_t>(terminal);
            a.robot_label[i] = static_cast<uint8_t>(robot_label);
            a.human_label[i] = static_cast<uint8_t>(human_label);
            a.info[i] = static_cast<uint8_t>(t.info);
            c_steps = 1;
            c_reward2 = r2;
            c_top = (t.info >> 5) == kScoredTop;
            c_bottom = (t.info >> 5) == kScoredBottom;
        }

        store_sim(a.state, a.n, i, s, reset_count, noise_count, make_meta(steps, episode, use, pending));
        uint4 *dl = reinterpret_cast<uint4 *>(a.drawlist);
        dl[i] = rec[0];
        dl[n + i] = rec[1];
        dl[2 * n + i] = rec[2];
    }

    // running statistics (SURVEY.md §5 "Metrics"): warp-reduce, one atomic per warp per counter
    if (a.stats != nullptr) {
        const unsigned lane = threadIdx.x & 31u;
        const unsigned s_steps = __reduce_add_sync(lane_mask, static_cast<unsigned>(c_steps));
        const unsigned s_eps = __reduce_add_sync(lane_mask, static_cast<unsigned>(c_episodes));
        const unsigned s_top = __reduce_add_sync(lane_mask, static_cast<unsigned>(c_top));
        const unsigned s_bot = __reduce_add_sync(lane_mask, static_cast<unsigned>(c_bottom));
        const int s_rew = __reduce_add_sync(lane_mask, c_reward2);
        unsigned long long *st = reinterpret_cast<unsigned long long *>(a.stats);
        if (lane == 0) {
            if (s_steps) atomicAdd(st + 0, static_cast<unsigned long long>(s_steps));
            if (s_eps) atomicAdd(st + 1, static_cast<unsigned long long>(s_eps));
            if (s_top) atomicAdd(st + 2, static_cast<unsigned long long>(s_top));
            if (s_bot) atomicAdd(st + 3, static_cast<unsigned long long>(s_bot));
            if (s_rew) atomicAdd(st + 4, static_cast<unsigned long long>(static_cast<long long>(s_rew)));
        }
        if (kStep) {
#pragma unroll
            for (unsigned l = 0; l < 10; ++l) {
                const unsigned nr = __popc(__ballot_sync(lane_mask, robot_label == l));
                const unsigned nh = __popc(__ballot_sync(lane_mask, human_label == l));
                if (lane == 0 && nr) atomicAdd(st + 5 + l, static_cast<unsigned long long>(nr));
                if (lane == 0 && nh) atomicAdd(st + 15 + l, static_cast<unsigned long long>(nh));
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------ K2
constexpr uint32_t kNone = 0xFF000000u; // "table shows through"; colours are r | g<<8 | b<<16

__device__ __forceinline__ constexpr uint32_t rgb(uint32_t r, uint32_t g, uint32_t b) { return r | g << 8 | b << 16; }

struct Scene {
    int ipx, ipy, ibx, iby, itx, ity;
    uint32_t use;
};

// one disc-shaped mallet, SPEC.md §7 items 4 and 5
__device__ __forceinline__ uint32_t mallet_colour(int dx, int dy, uint32_t shine, uint32_t knob, uint32_t rim, uint32_t body)
{
    const int d2 = dx * dx + dy * dy;
    if (d2 > 576) return kNone;
    if ((dx + 3) * (dx + 3) + (dy + 3) * (dy + 3) <= 9) return shine;
    if (d2 <= 100) return knob;
    return d2 > 400 ? rim : body;
}

constexpr uint32_t kLayerLine = 1u, kLayerPuck = 2u, kLayerBot = 4u, kLayerTop = 8u, kLayerArm = 16u;

// what is constant along one output row (native row y)
struct Row {
    int y, arm_xc, arm_hw;
};

// colour of native pixel (x, row.y), top-most layer first (SPEC.md §7, painter's order reversed).  `layers` says which
// layers can reach this 16-pixel vector at all; the others are not even tested.
__device__ __forceinline__ uint32_t shade(const Scene &sc, const Row &row, uint32_t layers, int x)
{
    if (layers & kLayerArm) {
        const int a = abs(x - row.arm_xc);
        if (a <= row.arm_hw) return a > row.arm_hw - 4 ? rgb(160, 110, 95) : rgb(200, 150, 130);
    }
    if (layers & kLayerTop) {
        const uint32_t c = mallet_colour(x - sc.itx, row.y - sc.ity, rgb(225, 215, 215), rgb(110, 70, 70), rgb(35, 20, 20), rgb(70, 40, 40));
        if (c != kNone) return c;
    }
    if (layers & kLayerBot) {
        const uint32_t c = mallet_colour(x - sc.ibx, row.y - sc.iby, rgb(220, 220, 230), rgb(90, 90, 100), rgb(20, 20, 24), rgb(40, 40, 48));
        if (c != kNone) return c;
    }
    if (layers & kLayerPuck) {
        const int dx = x - sc.ipx, dy = row.y - sc.ipy, d2 = dx * dx + dy * dy;
        if (d2 <= 324) {
            if ((dx + 5) * (dx + 5) + (dy + 5) * (dy + 5) <= 16) return rgb(240, 120, 130);
            return d2 > 196 ? rgb(150, 20, 30) : rgb(215, 40, 50);
        }
    }
    return (layers & kLayerLine) ? rgb(235, 235, 225) : kNone;
}

// native sample column / row of output pixel j / i: Pillow NEAREST, (j + 0.5) * 450 / 128 truncated (SPEC.md §8)
__device__ __forceinline__ int sample_x(int j) { return ((2 * j + 1) * kTableW) >> 8; }
__device__ __forceinline__ int sample_y(int i) { return ((2 * i + 1) * kTableH) >> 8; }

template <int kChan, int kByte>
__device__ __forceinline__ uint32_t put_byte(uint32_t word, uint32_t colour)
{ // byte kByte of word := byte kChan of colour
    constexpr uint32_t sel = (0x3210u & ~(0xFu << (4 * kByte))) | (static_cast<uint32_t>(4 + kChan) << (4 * kByte));
    return __byte_perm(word, colour, sel);
}

__device__ __forceinline__ void store_vec(uint8_t *dst, const uint4 &v)
{ // streaming 128-bit store: the frame is not read again by this kernel
    __stcs(reinterpret_cast<uint4 *>(dst), v);
}

__global__ void __launch_bounds__(kRasterThreads, 4) raster_kernel(RasterArgs a)
{
    extern __shared__ uint4 s_table[]; // [3][128][8] vectors of 16 pixels: the table at the sample points
    for (int v = threadIdx.x; v < kFrameVecs; v += kRasterThreads) s_table[v] = a.table[v];
    __syncthreads();

    const int k = threadIdx.x & 7;   // which 16-pixel vector of the row
    const int r0 = threadIdx.x >> 3; // first row of this thread; rows r0, r0+32, r0+64, r0+96
    const int x_first = sample_x(16 * k), x_last = sample_x(16 * k + 15);
    const size_t n = static_cast<size_t>(a.n);
    const long long items = 3ll * a.n;

    for (long long item = blockIdx.x; item < items; item += gridDim.x) {
        const int back = static_cast<int>(item / a.n); // 0 newest
        const int env = static_cast<int>(item - static_cast<long long>(back) * a.n);
        const uint4 rec = __ldg(reinterpret_cast<const uint4 *>(a.drawlist) + item);
        if (!(rec.w & kDrawValid)) continue;
        const int slot = (a.head + 3 - back) % 3;
        uint8_t *frame = a.frames + (static_cast<size_t>(slot) * n + env) * kFrameBytes;

        if (rec.w & kDrawNoise) { // SPEC.md §7 last paragraph
            const uint32_t env_id = a.env_id_base + static_cast<uint32_t>(env);
#pragma unroll 1
            for (int v = threadIdx.x; v < kFrameVecs; v += kRasterThreads) {
                uint4 r = philox(static_cast<uint32_t>(v), rec.x, env_id, 2u, a.key);
                // per byte b -> (b * 255) >> 8 == max(b - 1, 0)
                r.x = __vsubus4(r.x, 0x01010101u);
                r.y = __vsubus4(r.y, 0x01010101u);
                r.z = __vsubus4(r.z, 0x01010101u);
                r.w = __vsubus4(r.w, 0x01010101u);
                store_vec(frame + static_cast<size_t>(v) * 16, r);
            }
            continue;
        }

        Scene sc;
        sc.ipx = static_cast<int>(rec.x & 0xFFFFu); sc.ipy = static_cast<int>(rec.x >> 16);
        sc.ibx = static_cast<int>(rec.y & 0xFFFFu); sc.iby = static_cast<int>(rec.y >> 16);
        sc.itx = static_cast<int>(rec.z & 0xFFFFu); sc.ity = static_cast<int>(rec.z >> 16);
        sc.use = rec.w & 7u;

#pragma unroll 1
        for (int pass = 0; pass < 4; ++pass) {
            const int i = r0 + 32 * pass;
            const int vec = i * 8 + k;
            uint4 out0 = s_table[vec], out1 = s_table[1024 + vec], out2 = s_table[2048 + vec];

            // which layers can touch this 16-pixel span at all?
            Row row;
            row.y = sample_y(i);
            row.arm_xc = 0;
            row.arm_hw = 0;
            uint32_t layers = (row.y >= sc.iby - 3 && row.y <= sc.iby + 2) ? kLayerLine : 0u; // the rail line spans the full width
            if ((sc.use & kUseArm) && row.y <= sc.ity + 10) {
                const int t = sc.ity + 10 - row.y;
                row.arm_xc = sc.itx + ((t * 5) >> 5);
                row.arm_hw = 26 + ((t * 3) >> 4);
                if (x_last >= row.arm_xc - row.arm_hw && x_first <= row.arm_xc + row.arm_hw) layers |= kLayerArm;
            }
            if (abs(row.y - sc.iby) <= 24 && x_last >= sc.ibx - 24 && x_first <= sc.ibx + 24) layers |= kLayerBot;
            if ((sc.use & kUseTop) && abs(row.y - sc.ity) <= 24 && x_last >= sc.itx - 24 && x_first <= sc.itx + 24) layers |= kLayerTop;
            if ((sc.use & kUsePuck) && abs(row.y - sc.ipy) <= 18 && x_last >= sc.ipx - 18 && x_first <= sc.ipx + 18) layers |= kLayerPuck;

            if (layers) {
                uint32_t *w0 = reinterpret_cast<uint32_t *>(&out0);
                uint32_t *w1 = reinterpret_cast<uint32_t *>(&out1);
                uint32_t *w2 = reinterpret_cast<uint32_t *>(&out2);
#pragma unroll
                for (int g = 0; g < 4; ++g) {
#define AHK_PIXEL(B)                                                                 \
    {                                                                                \
        const uint32_t c = shade(sc, row, layers, sample_x(16 * k + 4 * g + B));     \
        if (c != kNone) {                                                            \
            w0[g] = put_byte<0, B>(w0[g], c);                                        \
            w1[g] = put_byte<1, B>(w1[g], c);                                        \
            w2[g] = put_byte<2, B>(w2[g], c);                                        \
        }                                                                            \
    }
                    AHK_PIXEL(0) AHK_PIXEL(1) AHK_PIXEL(2) AHK_PIXEL(3)
#undef AHK_PIXEL
                }
            }
            store_vec(frame + static_cast<size_t>(vec) * 16, out0);
            store_vec(frame + static_cast<size_t>(1024 + vec) * 16, out1);
            store_vec(frame + static_cast<size_t>(2048 + vec) * 16, out2);
        }
    }
}

// ------------------------------------------------------------------------------------------------ observation
// One thread per 16 pixels of one output plane position. processor.py:52 ('rgb'), :53-64 ('gray-diff').
__global__ void __launch_bounds__(256) observe_kernel(ObserveArgs a)
{
    const size_t n = static_cast<size_t>(a.n);
    if (a.format == 2) { // gray-diff: float64 [n][1][128][128]
        const size_t total = n * 1024; // vectors per plane
        for (size_t idx = blockIdx.x * 256ull + threadIdx.x; idx < total; idx += gridDim.x * 256ull) {
            const size_t env = idx >> 10, v = idx & 1023;
            uint4 q[3][3]; // [age][channel]
#pragma unroll
            for (int age = 0; age < 3; ++age) {
                const int slot = (a.head + 3 - age) % 3;
                const uint4 *f = reinterpret_cast<const uint4 *>(a.frames + (static_cast<size_t>(slot) * n + env) * kFrameBytes);
#pragma unroll
                for (int c = 0; c < 3; ++c) q[age][c] = f[c * 1024 + v];
            }
            double *out = reinterpret_cast<double *>(a.out) + (env * 1024 + v) * 16;
#pragma unroll
            for (int p = 0; p < 16; ++p) {
                double ch[3];
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    auto byte = [&](const uint4 &u) { return (reinterpret_cast<const uint32_t *>(&u)[p >> 2] >> (8 * (p & 3))) & 255u; };
                    const uint32_t d = (byte(q[2][c]) - byte(q[1][c]) - byte(q[0][c])) & 255u; // uint8 wrap-around, processor.py:56
                    ch[c] = static_cast<double>(d);
                }
                const double gray = __dadd_rn(__dadd_rn(__dmul_rn(0.2989, ch[0]), __dmul_rn(0.5870, ch[1])), __dmul_rn(0.1140, ch[2]));
                out[p] = __ddiv_rn(__dsub_rn(gray, 128.0), 128.0);
            }
        }
        return;
    }
    // stacked 9-channel forms, newest first: uint8 or (x-128)/128 float32
    const size_t total = n * 3 * kFrameVecs;
    for (size_t idx = blockIdx.x * 256ull + threadIdx.x; idx < total; idx += gridDim.x * 256ull) {
        const size_t env = idx / (3 * kFrameVecs);
        const size_t rem = idx - env * (3 * kFrameVecs);
        const int age = static_cast<int>(rem / kFrameVecs);
        const size_t v = rem - static_cast<size_t>(age) * kFrameVecs;
        const int slot = (a.head + 3 - age) % 3;
        const uint4 q = *reinterpret_cast<const uint4 *>(a.frames + (static_cast<size_t>(slot) * n + env) * kFrameBytes + v * 16);
        if (a.format == 0) {
            reinterpret_cast<uint4 *>(a.out)[idx] = q;
        } else {
            float4 *out = reinterpret_cast<float4 *>(a.out) + idx * 4;
            const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                float4 f;
                f.x = __fdiv_rn(__fsub_rn(static_cast<float>(w[g] & 255u), 128.0f), 128.0f);
                f.y = __fdiv_rn(__fsub_rn(static_cast<float>((w[g] >> 8) & 255u), 128.0f), 128.0f);
                f.z = __fdiv_rn(__fsub_rn(static_cast<float>((w[g] >> 16) & 255u), 128.0f), 128.0f);
                f.w = __fdiv_rn(__fsub_rn(static_cast<float>(w[g] >> 24), 128.0f), 128.0f);
                out[g] = f;
            }
        }
    }
}

__global__ void set_pending_kernel(uint32_t *state, const uint8_t *pending, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t *m = state + static_cast<size_t>(15) * n + i;
    *m = (*m & ~(1u << 21)) | (pending[i] ? 1u << 21 : 0u);
}

// ------------------------------------------------------------------------------------------------ launchers
cudaError_t launch_env(const EnvArgs &a, bool step, cudaStream_t stream)
{
    const int blocks = (a.n + kEnvThreads - 1) / kEnvThreads;
    if (step) env_kernel<true><<<blocks, kEnvThreads, 0, stream>>>(a);
    else env_kernel<false><<<blocks, kEnvThreads, 0, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t raster_prepare()
{
    return cudaFuncSetAttribute(raster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFrameBytes);
}

cudaError_t launch_raster(const RasterArgs &a, int grid, cudaStream_t stream)
{
    raster_kernel<<<grid, kRasterThreads, kFrameBytes, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_observe(const ObserveArgs &a, int grid, cudaStream_t stream)
{
    observe_kernel<<<grid, 256, 0, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_set_pending(uint32_t *state, const uint8_t *pending, int n, cudaStream_t stream)
{
    set_pending_kernel<<<(n + 255) / 256, 256, 0, stream>>>(state, pending, n);
    return cudaGetLastError();
}

} // namespace ahk
