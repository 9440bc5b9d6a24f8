// ahk_sim.cuh — device-side game rules (SPEC.md §4-§6, §9), written for registers: one environment per thread.
//
// Replaces, per environment, what the reference does in Python per call:
//   AirHockey.step / AirHockey.reset        (missing submodule; call sites air_hockey_env.py:67,71,89)
//   AirHockeyEnv.reset / AirHockeyEnv.step  (gym-air-hockey/gym_air_hockey/envs/air_hockey_env.py:26-155)
// Every float operation is an explicit round-to-nearest intrinsic so that ptxas can never contract a multiply and
// an add into an FMA: the CPU oracle and this code then agree bit for bit on the continuous state.
#pragma once
#include <cstdint>

namespace ahk {

// ---- SPEC.md §2, §3
constexpr float kRinkL = 25.0f, kRinkR = 425.0f, kRinkT = 25.0f, kRinkB = 675.0f;
constexpr float kMidY = 350.0f, kCornerR = 60.0f, kGoalL = 145.0f, kGoalR = 305.0f;
constexpr float kPuckR = 18.0f, kMalletR = 24.0f;
constexpr float kHomeX = 225.0f, kHomeBotY = 600.0f, kHomeTopY = 100.0f, kTargetR = 30.0f;
constexpr float kMalletAcc = 1.25f, kMalletFriction = 0.875f, kMalletMax = 9.0f;
constexpr float kPuckFriction = 0.99609375f, kPuckMax = 18.0f;
constexpr float kWallBounce = 1.875f, kWallE = 0.875f, kHitBounce = 1.9375f;
constexpr float kAiDeadband = 4.0f, kAiLookahead = 6.0f, kAiGuard = 40.0f, kAiSwing = 60.0f;

constexpr uint32_t kUseArm = 1u, kUsePuck = 2u, kUseTop = 4u, kUseAll = 7u;
constexpr uint32_t kEpNormal = 0u, kEpMissing = 1u, kEpRandom = 2u;
constexpr uint32_t kActNone = 255u;

// info byte (ahk.h): bits 0..4 flags, bits 5..6 scored
constexpr uint32_t kHit = 1u, kBorder = 2u, kPuckBottom = 4u, kCloser = 8u, kInTarget = 16u;
constexpr uint32_t kScoredTop = 1u, kScoredBottom = 2u;

// draw record flags (w3 low byte): bits 0..2 use mask, bit 3 noise frame, bit 4 record valid
constexpr uint32_t kDrawNoise = 8u, kDrawValid = 16u;
// the env was reset in the PREVIOUS step: that launch rewrote this very ring slot with a warm-up frame, so with overlapped
// launches (StepArgs::pdl == 3) this frame is drawn after griddepcontrol.wait, with the frames that rewrite older slots
constexpr uint32_t kDrawDefer = 32u;

__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float quo(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float root(float a) { return __fsqrt_rn(a); }
__device__ __forceinline__ float len2(float dx, float dy) { return add(mul(dx, dx), mul(dy, dy)); }

// ---- Philox4x32-10 (SPEC.md §4)
__device__ __forceinline__ uint4 philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint2 key)
{
    uint32_t k0 = key.x, k1 = key.y;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        c0 = hi1 ^ c1 ^ k0;
        c2 = hi0 ^ c3 ^ k1;
        c1 = lo1;
        c3 = lo0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
}

__device__ __forceinline__ float unit(uint32_t r) { return mul(__uint2float_rn(r >> 8), 5.9604644775390625e-08f); }

struct Body { float x, y, vx, vy; };

struct Sim {
    Body puck, bot, top;
    uint32_t serve_count;
};

struct Tick {
    int rax, ray, hax, hay; // the actions applied (AI fill-in included): GameInfo.robot_action / human_action
    uint32_t info;          // flags | scored << 5
};

// SPEC.md §5
__device__ __forceinline__ void sim_reset(Sim &s, uint32_t env_id, uint2 key)
{
    const uint4 r = philox(s.serve_count, 0u, env_id, 1u, key);
    s.serve_count += 1u;
    s.bot = Body{kHomeX, kHomeBotY, 0.0f, 0.0f};
    s.top = Body{kHomeX, kHomeTopY, 0.0f, 0.0f};
    s.puck.x = add(103.0f, mul(unit(r.x), 244.0f));
    s.puck.y = add(250.0f, mul(unit(r.y), 200.0f));
    s.puck.vx = mul(sub(unit(r.z), 0.5f), 8.0f);
    s.puck.vy = mul(sub(unit(r.w), 0.5f), 8.0f);
}

__device__ __forceinline__ void limit_speed(Body &b, float vmax)
{
    const float s2 = len2(b.vx, b.vy);
    if (s2 > mul(vmax, vmax)) {
        const float k = quo(vmax, root(s2));
        b.vx = mul(b.vx, k);
        b.vy = mul(b.vy, k);
    }
}

__device__ __forceinline__ int steer(float d) { return d > kAiDeadband ? 1 : (d < -kAiDeadband ? -1 : 0); }

// SPEC.md §6.1 — programmed opponent / labeler. bottom: own goal at larger y.
template <bool kBottom>
__device__ __forceinline__ void ai_move(const Body &puck, const Body &m, bool use_puck, int &ax, int &ay)
{
    ax = 0;
    ay = 0;
    if (!use_puck) return;
    const float home_y = kBottom ? kHomeBotY : kHomeTopY;
    float tx = kHomeX, ty = home_y;
    const bool own = kBottom ? (puck.y > kMidY) : (puck.y < kMidY);
    if (own) {
        const bool behind = kBottom ? (puck.y > m.y) : (puck.y < m.y);
        if (behind) {
            tx = add(puck.x, m.x < puck.x ? -kAiSwing : kAiSwing);
            ty = add(puck.y, kBottom ? kAiGuard : -kAiGuard);
        } else {
            tx = fminf(fmaxf(add(puck.x, mul(puck.vx, kAiLookahead)), kRinkL), kRinkR);
            ty = add(puck.y, mul(puck.vy, kAiLookahead));
        }
    }
    ax = steer(sub(tx, m.x));
    ay = steer(sub(ty, m.y));
}

// SPEC.md §6.3; returns true when the mallet touched its border
template <bool kBottom>
__device__ __forceinline__ bool drive_mallet(Body &m, int ax, int ay)
{
    constexpr float xmin = kRinkL + kMalletR, xmax = kRinkR - kMalletR;
    constexpr float ymin = kBottom ? kMidY + kMalletR : kRinkT + kMalletR;
    constexpr float ymax = kBottom ? kRinkB - kMalletR : kMidY - kMalletR;
    constexpr float cy = kBottom ? kRinkB - kCornerR : kRinkT + kCornerR;
    constexpr float lim = kCornerR - kMalletR;
    bool hit = false;
    m.vx = mul(add(m.vx, mul(static_cast<float>(ax), kMalletAcc)), kMalletFriction);
    m.vy = mul(add(m.vy, mul(static_cast<float>(ay), kMalletAcc)), kMalletFriction);
    limit_speed(m, kMalletMax);
    m.x = add(m.x, m.vx);
    m.y = add(m.y, m.vy);
    if (m.x < xmin) { m.x = xmin; m.vx = 0.0f; hit = true; }
    if (m.x > xmax) { m.x = xmax; m.vx = 0.0f; hit = true; }
    if (m.y < ymin) { m.y = ymin; m.vy = 0.0f; hit = true; }
    if (m.y > ymax) { m.y = ymax; m.vy = 0.0f; hit = true; }
    // the two rounded corners of this half; left first, then right (a mallet can only be in one of them)
    const bool past_y = kBottom ? (m.y > cy) : (m.y < cy);
#pragma unroll
    for (int right = 0; right < 2; ++right) {
        const float cx = right ? kRinkR - kCornerR : kRinkL + kCornerR;
        const bool past_x = right ? (m.x > cx) : (m.x < cx);
        if (past_x && past_y) {
            const float dx = sub(m.x, cx), dy = sub(m.y, cy);
            const float d2 = len2(dx, dy);
            if (d2 > lim * lim) {
                const float k = quo(lim, root(d2));
                m.x = add(cx, mul(dx, k));
                m.y = add(cy, mul(dy, k));
                m.vx = 0.0f;
                m.vy = 0.0f;
                hit = true;
            }
        }
    }
    return hit;
}

// SPEC.md §6.5
__device__ __forceinline__ bool strike(Body &p, const Body &m, float ny0)
{
    constexpr float R = kPuckR + kMalletR;
    const float dx = sub(p.x, m.x), dy = sub(p.y, m.y);
    const float d2 = len2(dx, dy);
    if (!(d2 < R * R)) return false;
    float nx = 0.0f, ny = ny0;
    if (d2 > 0.0f) {
        const float d = root(d2);
        nx = quo(dx, d);
        ny = quo(dy, d);
    }
    const float vn = add(mul(sub(p.vx, m.vx), nx), mul(sub(p.vy, m.vy), ny));
    if (vn < 0.0f) {
        const float j = mul(kHitBounce, vn);
        p.vx = sub(p.vx, mul(j, nx));
        p.vy = sub(p.vy, mul(j, ny));
    }
    p.x = add(m.x, mul(nx, R));
    p.y = add(m.y, mul(ny, R));
    return true;
}

// SPEC.md §6.6
__device__ __forceinline__ void rebound(Body &p)
{
    constexpr float top = kRinkT + kPuckR, bottom = kRinkB - kPuckR, left = kRinkL + kPuckR, right = kRinkR - kPuckR;
    constexpr float lim = kCornerR - kPuckR;
    const bool mouth = (p.x >= kGoalL + kPuckR) && (p.x <= kGoalR - kPuckR);
    if (!mouth && p.y < top) { p.y = top; if (p.vy < 0.0f) p.vy = mul(-p.vy, kWallE); }
    if (!mouth && p.y > bottom) { p.y = bottom; if (p.vy > 0.0f) p.vy = mul(-p.vy, kWallE); }
    if (p.x < left) { p.x = left; if (p.vx < 0.0f) p.vx = mul(-p.vx, kWallE); }
    if (p.x > right) { p.x = right; if (p.vx > 0.0f) p.vx = mul(-p.vx, kWallE); }
#pragma unroll
    for (int c = 0; c < 4; ++c) { // TL, TR, BL, BR
        const float cx = (c & 1) ? kRinkR - kCornerR : kRinkL + kCornerR;
        const float cy = (c & 2) ? kRinkB - kCornerR : kRinkT + kCornerR;
        const bool in_x = (c & 1) ? (p.x > cx) : (p.x < cx);
        const bool in_y = (c & 2) ? (p.y > cy) : (p.y < cy);
        if (in_x && in_y) {
            const float dx = sub(p.x, cx), dy = sub(p.y, cy);
            const float d2 = len2(dx, dy);
            if (d2 > lim * lim) {
                const float d = root(d2);
                const float nx = quo(dx, d), ny = quo(dy, d);
                const float vn = add(mul(p.vx, nx), mul(p.vy, ny));
                if (vn > 0.0f) {
                    const float j = mul(kWallBounce, vn);
                    p.vx = sub(p.vx, mul(j, nx));
                    p.vy = sub(p.vy, mul(j, ny));
                }
                p.x = add(cx, mul(nx, lim));
                p.y = add(cy, mul(ny, lim));
            }
        }
    }
}

__device__ __forceinline__ void label_to_vector(uint32_t label, int &ax, int &ay)
{ // processor.py:7-18: rows 0..8 are (l/3-1, l%3-1), row 9 is (0,0)
    const uint32_t l = label >= 9u ? 4u : label;
    ax = static_cast<int>(l / 3u) - 1;
    ay = static_cast<int>(l % 3u) - 1;
}

// SPEC.md §6
__device__ __forceinline__ void sim_tick(Sim &s, uint32_t robot, uint32_t human, uint32_t use, uint32_t env_id,
                                         uint2 key, Tick &t)
{
    const bool use_puck = (use & kUsePuck) != 0u, use_top = (use & kUseTop) != 0u;
    t.hax = 0;
    t.hay = 0;
    if (robot == kActNone) ai_move<true>(s.puck, s.bot, use_puck, t.rax, t.ray);
    else label_to_vector(robot, t.rax, t.ray);
    if (human != kActNone) label_to_vector(human, t.hax, t.hay);
    else if (use_top) ai_move<false>(s.puck, s.top, use_puck, t.hax, t.hay);

    const float before = len2(sub(s.bot.x, s.puck.x), sub(s.bot.y, s.puck.y));
    uint32_t info = drive_mallet<true>(s.bot, t.rax, t.ray) ? kBorder : 0u;
    if (use_top) drive_mallet<false>(s.top, t.hax, t.hay);

    if (use_puck) {
        s.puck.vx = mul(s.puck.vx, kPuckFriction);
        s.puck.vy = mul(s.puck.vy, kPuckFriction);
        limit_speed(s.puck, kPuckMax);
        s.puck.x = add(s.puck.x, s.puck.vx);
        s.puck.y = add(s.puck.y, s.puck.vy);
        if (strike(s.puck, s.bot, -1.0f)) info |= kHit;
        if (use_top) strike(s.puck, s.top, 1.0f);
        rebound(s.puck);
        limit_speed(s.puck, kPuckMax);
        if (s.puck.y < kRinkT) info |= kScoredBottom << 5;
        else if (s.puck.y > kRinkB) info |= kScoredTop << 5;
        if (len2(sub(s.bot.x, s.puck.x), sub(s.bot.y, s.puck.y)) < before) info |= kCloser;
        if (s.puck.y > kMidY) info |= kPuckBottom;
    }
    if (len2(sub(s.bot.x, kHomeX), sub(s.bot.y, kHomeBotY)) <= kTargetR * kTargetR) info |= kInTarget;
    if (info >> 5) sim_reset(s, env_id, key);
    t.info = info;
}

// air_hockey_env.py:95-145, as twice the reward (an exact small integer)
__device__ __forceinline__ int reward_x2(uint32_t episode, uint32_t info, uint32_t robot_label)
{
    if (episode == kEpNormal) {
        const uint32_t scored = info >> 5;
        if (scored == kScoredTop) return -2;
        if (scored == kScoredBottom) return 2;
        if (info & kHit) return 1;
        if (info & kBorder) return -1;
        if (info & kPuckBottom) return (info & kCloser) ? 1 : -1;
        if (info & kInTarget) return (robot_label == 4u || robot_label == 9u) ? 1 : -1;
        return -1;
    }
    if (episode == kEpMissing) return robot_label == 4u ? 1 : -1;
    return robot_label == 9u ? 1 : -1;
}

__device__ __forceinline__ int to_pixel(float v) { return __float2int_rz(add(v, 0.5f)); }

// what the rasteriser needs to draw one frame: three pixel positions + flags (SPEC.md §7)
__device__ __forceinline__ uint4 draw_record(const Sim &s, uint32_t use)
{
    auto pack = [](float x, float y) {
        return (static_cast<uint32_t>(to_pixel(x)) & 0xFFFFu) | (static_cast<uint32_t>(to_pixel(y)) << 16);
    };
    return make_uint4(pack(s.puck.x, s.puck.y), pack(s.bot.x, s.bot.y), pack(s.top.x, s.top.y), use | kDrawValid);
}

__device__ __forceinline__ uint4 noise_record(uint32_t noise_count)
{
    return make_uint4(noise_count, 0u, 0u, kDrawNoise | kDrawValid);
}

} // namespace ahk
