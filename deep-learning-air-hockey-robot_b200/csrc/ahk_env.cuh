// ahk_env.cuh — one environment's reset / step as a device function (registers only), shared by the fused step
// kernel (ahk_step.cu) and the stand-alone env kernel (ahk_kernels.cu).
//
// ≙ AirHockeyEnv.reset (air_hockey_env.py:26-76) and AirHockeyEnv.step (air_hockey_env.py:84-155) around the game
// rules of ahk_sim.cuh; SPEC.md §9.  State is structure-of-arrays in HBM: word w of env i at state[w * n + i], so a
// warp's 32 envs read and write 128 contiguous bytes per word.
#pragma once
#include "ahk_kernels.cuh"
#include "ahk_sim.cuh"

namespace ahk {

__device__ __forceinline__ void load_sim(const uint32_t *__restrict__ st, int n, int i, Sim &s, uint32_t &reset_count,
                                         uint32_t &noise_count, uint32_t &meta)
{
    // ld.global.cg: with overlapped launches and a moving partition another SM may have written these words while this
    // SM's L1 was live
    auto u = [&](int w) { return __ldcg(st + static_cast<size_t>(w) * n + i); };
    auto f = [&](int w) { return __uint_as_float(u(w)); };
    s.puck = Body{f(0), f(1), f(2), f(3)};
    s.bot = Body{f(4), f(5), f(6), f(7)};
    s.top = Body{f(8), f(9), f(10), f(11)};
    s.serve_count = u(12);
    reset_count = u(13);
    noise_count = u(14);
    meta = u(15);
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

// what one env adds to the running statistics this call
struct StatDelta {
    unsigned steps = 0, episodes = 0, goals_top = 0, goals_bottom = 0;
    int reward2 = 0;
    uint32_t robot_label = 0xFFu, human_label = 0xFFu, terminal = 0;
};

// Resets (if due) and/or steps env i; rec[0..2] are the draw records of ring slots head, head-1, head-2 (w == 0:
// nothing to paint there).  kStep: AirHockeyEnv.step, auto-reset first when armed; !kStep: AirHockeyEnv.reset.
// stamps (tools/prologue_probe.cu only; nullptr in the product, folded away): clock64 after the state load, the tick, the stores.
template <bool kStep>
__device__ __forceinline__ void env_update(const EnvArgs &a, int i, uint4 (&rec)[3], StatDelta &d, long long *stamps = nullptr)
{
    Sim s;
    uint32_t reset_count, noise_count, meta;
    // The caller's actions are fetched with the state, not where they are used: through ahk_step_host they sit in mapped host
    // memory, and a PCIe round trip issued now is over by the time the reset branch and the AI have run.
    uint32_t robot_in = kActNone, human_in = kActNone;
    if (kStep) {
        if (a.robot) robot_in = a.robot[i];
        if (a.human) human_in = a.human[i];
    }
    load_sim(a.state, a.n, i, s, reset_count, noise_count, meta);
    if (stamps) stamps[0] = static_cast<long long>(__float_as_uint(s.puck.x) + meta) * 0 + clock64();
    uint32_t steps = meta & 0xFFFFu, episode = (meta >> 16) & 3u, use = (meta >> 18) & 7u, pending = (meta >> 21) & 1u;
    const uint32_t env_id = a.env_id_base + static_cast<uint32_t>(i);
    const bool prev_step_reset = kStep && steps == 1u; // the previous step reset this env (or was its first): kDrawDefer
    rec[0] = rec[1] = rec[2] = make_uint4(0, 0, 0, 0);

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
#pragma unroll 1
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
        d.episodes = 1;
    }

    if (kStep) { // ≙ AirHockeyEnv.step, air_hockey_env.py:84-155
        const uint32_t robot = robot_in, human = human_in;
        Tick t;
        sim_tick(s, robot, human, use, env_id, a.key, t);
        if (stamps) stamps[1] = static_cast<long long>(t.info + __float_as_uint(s.puck.x)) * 0 + clock64();
        d.robot_label = static_cast<uint32_t>(t.rax * 3 + t.ray + 4); // processor.py:84-85
        d.human_label = static_cast<uint32_t>(t.hax * 3 + t.hay + 4);
        const int r2 = reward_x2(episode, t.info, d.robot_label);
        uint32_t terminal = (t.info >> 5) != 0u;
        if (episode == kEpRandom) { // air_hockey_env.py:147-151
            terminal = 0u;
            rec[0] = noise_record(noise_count++);
        } else {
            rec[0] = draw_record(s, use);
        }
        if (prev_step_reset) rec[0].w |= kDrawDefer; // also when this step resets again (max_episode_steps == 1)
        steps += 1u;
        if (a.auto_reset && (terminal || (a.max_episode_steps > 0 && steps >= static_cast<uint32_t>(a.max_episode_steps)))) pending = 1u;

        a.reward[i] = 0.5f * static_cast<float>(r2);
        a.done[i] = static_cast<uint8_t>(terminal);
        a.robot_label[i] = static_cast<uint8_t>(d.robot_label);
        a.human_label[i] = static_cast<uint8_t>(d.human_label);
        a.info[i] = static_cast<uint8_t>(t.info);
        if (a.h_reward) a.h_reward[i] = 0.5f * static_cast<float>(r2);
        if (a.h_done) a.h_done[i] = static_cast<uint8_t>(terminal);
        if (a.h_robot_label) a.h_robot_label[i] = static_cast<uint8_t>(d.robot_label);
        if (a.h_human_label) a.h_human_label[i] = static_cast<uint8_t>(d.human_label);
        d.steps = 1;
        d.terminal = terminal;
        d.reward2 = r2;
        d.goals_top = (t.info >> 5) == kScoredTop;
        d.goals_bottom = (t.info >> 5) == kScoredBottom;
    }

    store_sim(a.state, a.n, i, s, reset_count, noise_count, make_meta(steps, episode, use, pending));
    if (stamps) stamps[2] = clock64();
}

// running statistics (SURVEY.md §5 "Metrics"): warp-reduce, one atomic per warp per counter.  Call with the whole
// warp converged; lanes without an env pass a default StatDelta.
template <bool kStep>
__device__ __forceinline__ void accumulate_stats(uint64_t *stats, const StatDelta &d)
{
    if (stats == nullptr) return;
    constexpr unsigned full = 0xFFFFFFFFu;
    const unsigned lane = threadIdx.x & 31u;
    const unsigned s_steps = __reduce_add_sync(full, d.steps);
    const unsigned s_eps = __reduce_add_sync(full, d.episodes);
    if ((s_steps | s_eps) == 0u) return; // warp-uniform
    const unsigned s_top = __reduce_add_sync(full, d.goals_top);
    const unsigned s_bot = __reduce_add_sync(full, d.goals_bottom);
    const int s_rew = __reduce_add_sync(full, d.reward2);
    unsigned long long *st = reinterpret_cast<unsigned long long *>(stats);
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
            const unsigned nr = __popc(__ballot_sync(full, d.robot_label == l));
            const unsigned nh = __popc(__ballot_sync(full, d.human_label == l));
            if (lane == 0 && nr) atomicAdd(st + 5 + l, static_cast<unsigned long long>(nr));
            if (lane == 0 && nh) atomicAdd(st + 15 + l, static_cast<unsigned long long>(nh));
        }
    }
}

// The fused kernel's variant: the CTA collects its share in shared memory (s[kStatWords], same word order as
// ahk_buffers.stats) and adds it to the global sums once, at the end — a handful of global atomics per CTA per step
// whatever the env-to-lane mapping.  Call with the whole warp converged; `live`: this lane holds an env.
constexpr int kStatWords = 25;
template <bool kStep>
__device__ __forceinline__ void stat_to_shared(unsigned *s, const StatDelta &d, bool live)
{
    constexpr unsigned full = 0xFFFFFFFFu;
    const unsigned s_steps = __reduce_add_sync(full, d.steps);
    const unsigned s_eps = __reduce_add_sync(full, d.episodes);
    const unsigned s_top = __reduce_add_sync(full, d.goals_top);
    const unsigned s_bot = __reduce_add_sync(full, d.goals_bottom);
    const int s_rew = __reduce_add_sync(full, d.reward2);
    if ((threadIdx.x & 31u) == 0u) {
        if (s_steps) atomicAdd(s + 0, s_steps);
        if (s_eps) atomicAdd(s + 1, s_eps);
        if (s_top) atomicAdd(s + 2, s_top);
        if (s_bot) atomicAdd(s + 3, s_bot);
        if (s_rew) atomicAdd(s + 4, static_cast<unsigned>(s_rew)); // two's complement; sign-extended at the flush
    }
    if (kStep && live) {
        atomicAdd(s + 5 + d.robot_label, 1u);
        atomicAdd(s + 15 + d.human_label, 1u);
    }
}

} // namespace ahk
