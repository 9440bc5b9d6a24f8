/*
 * oracle/ahk_oracle.c — CPU restatement of the air-hockey hot path.  TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this.
 * The product (deep-learning-air-hockey-robot_b200/) never links, imports or falls back to it.
 *
 * PARITY STATUS: "parity unpinned" for physics, programmed AI and pixels.  The simulator those live in
 * (python module `air_hockey`, submodule arakhmat/air-hockey @ b7c25ca61243fc69be5d6ad87d3142268772549b)
 * is missing from /root/reference (.SUBMODULES.json: status "missing"), and the reference ships no test
 * or golden vector for it.  What this file restates is therefore:
 *   - SPEC.md §4-§7            : this build's own game rules (scalar, native-resolution drawing), and
 *   - the reference's own code : the env wrapper   gym-air-hockey/gym_air_hockey/envs/air_hockey_env.py:26-155
 *                                the observation   gym-air-hockey/gym_air_hockey/processor.py:31-66,84-85
 *     which ARE pinned — tests/golden/ holds outputs of the unmodified reference wrapper + processor
 *     driven over this simulator (tests/golden/make_golden.py), and tests/test_oracle.py replays them.
 *
 * Build: oracle/Makefile (gcc -O2 -ffp-contract=off: every float op below is one IEEE binary32 rounding).
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_W 450
#define ORC_H 700
#define ORC_DIM 128
#define ORC_FRAME (3 * ORC_DIM * ORC_DIM)

/* SPEC.md §2, §3 */
static const float RINK_L = 25.0f, RINK_R = 425.0f, RINK_T = 25.0f, RINK_B = 675.0f;
static const float MID_Y = 350.0f, CORNER_R = 60.0f, GOAL_L = 145.0f, GOAL_R = 305.0f;
static const float PUCK_R = 18.0f, MALLET_R = 24.0f;
static const float HOME_X = 225.0f, HOME_BOT_Y = 600.0f, HOME_TOP_Y = 100.0f, TARGET_R = 30.0f;
static const float MALLET_ACC = 1.25f, MALLET_FRICTION = 0.875f, MALLET_MAX = 9.0f;
static const float PUCK_FRICTION = 0.99609375f, PUCK_MAX = 18.0f;
static const float WALL_BOUNCE = 1.875f, WALL_E = 0.875f, HIT_BOUNCE = 1.9375f;
static const float AI_DEADBAND = 4.0f, AI_LOOKAHEAD = 6.0f, AI_GUARD = 40.0f, AI_SWING = 60.0f;

enum { USE_ARM = 1, USE_PUCK = 2, USE_TOP = 4, USE_ALL = 7 };
enum { SCORED_NONE = 0, SCORED_TOP = 1, SCORED_BOTTOM = 2 };
enum { EP_NORMAL = 0, EP_MISSING = 1, EP_RANDOM = 2 };
enum { ACT_NONE = 255 };

/* ---------------------------------------------------------------- Philox4x32-10 (SPEC.md §4) */
void orc_philox(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4])
{
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
    for (int round = 0; round < 10; ++round) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

static float unit_float(uint32_t r) { return (float)(r >> 8) * 5.9604644775390625e-08f; }

/* ---------------------------------------------------------------- simulator (≙ missing `air_hockey.AirHockey`) */
typedef struct { float x, y, vx, vy; } orc_body;

typedef struct {
    orc_body puck, bot, top;
    uint32_t serve_count;
    uint32_t env_id;
    uint32_t key[2];
} orc_sim;

/* ≙ GameInfo, fields read at air_hockey_env.py:90-127 */
typedef struct {
    int32_t scored;
    int8_t robot_action[2];
    int8_t human_action[2];
    uint8_t puck_was_hit, hit_the_border, puck_is_at_the_bottom, distance_decreased, in_the_target;
} orc_info;

void orc_sim_init(orc_sim *s, uint64_t seed, uint32_t env_id)
{
    memset(s, 0, sizeof *s);
    s->env_id = env_id;
    s->key[0] = (uint32_t)seed;
    s->key[1] = (uint32_t)(seed >> 32);
}

/* SPEC.md §5; call site air_hockey_env.py:67 */
void orc_sim_reset(orc_sim *s)
{
    uint32_t ctr[4] = { s->serve_count, 0u, s->env_id, 1u }, r[4];
    orc_philox(ctr, s->key, r);
    s->serve_count += 1;
    s->bot.x = HOME_X; s->bot.y = HOME_BOT_Y; s->bot.vx = 0.0f; s->bot.vy = 0.0f;
    s->top.x = HOME_X; s->top.y = HOME_TOP_Y; s->top.vx = 0.0f; s->top.vy = 0.0f;
    s->puck.x = 103.0f + unit_float(r[0]) * 244.0f;
    s->puck.y = 250.0f + unit_float(r[1]) * 200.0f;
    s->puck.vx = (unit_float(r[2]) - 0.5f) * 8.0f;
    s->puck.vy = (unit_float(r[3]) - 0.5f) * 8.0f;
}

static void clamp_speed(orc_body *b, float vmax)
{
    float s2 = b->vx * b->vx + b->vy * b->vy;
    if (s2 > vmax * vmax) {
        float s = sqrtf(s2);
        float k = vmax / s;
        b->vx = b->vx * k;
        b->vy = b->vy * k;
    }
}

static int sign_with_deadband(float d)
{
    if (d > AI_DEADBAND) return 1;
    if (d < -AI_DEADBAND) return -1;
    return 0;
}

/* SPEC.md §6.1 */
static void ai_action(const orc_sim *s, const orc_body *m, float home_y, int dir, int use_puck, int8_t out[2])
{
    out[0] = 0; out[1] = 0;
    if (!use_puck) return;
    float tx = HOME_X, ty = home_y;
    int own = dir > 0 ? (s->puck.y > MID_Y) : (s->puck.y < MID_Y);
    if (own) {
        int behind = dir > 0 ? (s->puck.y > m->y) : (s->puck.y < m->y);
        if (behind) { /* go round the puck on the side the mallet is already on */
            tx = s->puck.x + (m->x < s->puck.x ? -AI_SWING : AI_SWING);
            ty = s->puck.y + (float)dir * AI_GUARD;
        } else {
            tx = s->puck.x + s->puck.vx * AI_LOOKAHEAD;
            tx = fminf(fmaxf(tx, RINK_L), RINK_R);
            ty = s->puck.y + s->puck.vy * AI_LOOKAHEAD;
        }
    }
    out[0] = (int8_t)sign_with_deadband(tx - m->x);
    out[1] = (int8_t)sign_with_deadband(ty - m->y);
}

/* SPEC.md §6.3 ; returns 1 if the mallet touched its border */
static int move_mallet(orc_body *m, const int8_t a[2], float ymin, float ymax, float corner_cy, int corner_below)
{
    const float xmin = RINK_L + MALLET_R, xmax = RINK_R - MALLET_R;
    int hit = 0;
    m->vx = (m->vx + (float)a[0] * MALLET_ACC) * MALLET_FRICTION;
    m->vy = (m->vy + (float)a[1] * MALLET_ACC) * MALLET_FRICTION;
    clamp_speed(m, MALLET_MAX);
    m->x = m->x + m->vx;
    m->y = m->y + m->vy;
    if (m->x < xmin) { m->x = xmin; m->vx = 0.0f; hit = 1; }
    if (m->x > xmax) { m->x = xmax; m->vx = 0.0f; hit = 1; }
    if (m->y < ymin) { m->y = ymin; m->vy = 0.0f; hit = 1; }
    if (m->y > ymax) { m->y = ymax; m->vy = 0.0f; hit = 1; }
    for (int side = 0; side < 2; ++side) {
        float cx = side == 0 ? RINK_L + CORNER_R : RINK_R - CORNER_R;
        int beyond_x = side == 0 ? (m->x < cx) : (m->x > cx);
        int beyond_y = corner_below ? (m->y > corner_cy) : (m->y < corner_cy);
        if (beyond_x && beyond_y) {
            float dx = m->x - cx, dy = m->y - corner_cy;
            float d2 = dx * dx + dy * dy;
            float lim = CORNER_R - MALLET_R;
            if (d2 > lim * lim) {
                float d = sqrtf(d2);
                float k = lim / d;
                m->x = cx + dx * k;
                m->y = corner_cy + dy * k;
                m->vx = 0.0f; m->vy = 0.0f;
                hit = 1;
            }
        }
    }
    return hit;
}

/* SPEC.md §6.5 */
static int hit_puck(orc_body *p, const orc_body *m, float ny_if_coincident)
{
    const float R = PUCK_R + MALLET_R;
    float dx = p->x - m->x, dy = p->y - m->y;
    float d2 = dx * dx + dy * dy;
    if (!(d2 < R * R)) return 0;
    float nx = 0.0f, ny = ny_if_coincident;
    if (d2 > 0.0f) {
        float d = sqrtf(d2);
        nx = dx / d;
        ny = dy / d;
    }
    float vn = (p->vx - m->vx) * nx + (p->vy - m->vy) * ny;
    if (vn < 0.0f) {
        float j = HIT_BOUNCE * vn;
        p->vx = p->vx - j * nx;
        p->vy = p->vy - j * ny;
    }
    p->x = m->x + nx * R;
    p->y = m->y + ny * R;
    return 1;
}

/* SPEC.md §6.6 */
static void puck_borders(orc_body *p)
{
    int mouth = (p->x >= GOAL_L + PUCK_R) && (p->x <= GOAL_R - PUCK_R);
    if (!mouth && p->y < RINK_T + PUCK_R) { p->y = RINK_T + PUCK_R; if (p->vy < 0.0f) p->vy = -p->vy * WALL_E; }
    if (!mouth && p->y > RINK_B - PUCK_R) { p->y = RINK_B - PUCK_R; if (p->vy > 0.0f) p->vy = -p->vy * WALL_E; }
    if (p->x < RINK_L + PUCK_R) { p->x = RINK_L + PUCK_R; if (p->vx < 0.0f) p->vx = -p->vx * WALL_E; }
    if (p->x > RINK_R - PUCK_R) { p->x = RINK_R - PUCK_R; if (p->vx > 0.0f) p->vx = -p->vx * WALL_E; }
    for (int c = 0; c < 4; ++c) { /* TL, TR, BL, BR */
        int right = c & 1, bottom = c >> 1;
        float cx = right ? RINK_R - CORNER_R : RINK_L + CORNER_R;
        float cy = bottom ? RINK_B - CORNER_R : RINK_T + CORNER_R;
        int in_x = right ? (p->x > cx) : (p->x < cx);
        int in_y = bottom ? (p->y > cy) : (p->y < cy);
        if (!(in_x && in_y)) continue;
        float dx = p->x - cx, dy = p->y - cy;
        float d2 = dx * dx + dy * dy;
        float lim = CORNER_R - PUCK_R;
        if (d2 > lim * lim) {
            float d = sqrtf(d2);
            float nx = dx / d, ny = dy / d;
            float vn = p->vx * nx + p->vy * ny;
            if (vn > 0.0f) {
                float j = WALL_BOUNCE * vn;
                p->vx = p->vx - j * nx;
                p->vy = p->vy - j * ny;
            }
            p->x = cx + nx * lim;
            p->y = cy + ny * lim;
        }
    }
}

static float dist2(const orc_body *a, const orc_body *b)
{
    float dx = a->x - b->x, dy = a->y - b->y;
    return dx * dx + dy * dy;
}

/* SPEC.md §6; call sites air_hockey_env.py:71 (use = USE_ALL) and :89.
 * robot/human: NULL = None (programmed AI plays that side). */
void orc_sim_step(orc_sim *s, const int8_t *robot, const int8_t *human, int use, orc_info *info)
{
    const int use_puck = (use & USE_PUCK) != 0, use_top = (use & USE_TOP) != 0;
    memset(info, 0, sizeof *info);

    if (robot) { info->robot_action[0] = robot[0]; info->robot_action[1] = robot[1]; }
    else ai_action(s, &s->bot, HOME_BOT_Y, +1, use_puck, info->robot_action);
    if (human) { info->human_action[0] = human[0]; info->human_action[1] = human[1]; }
    else if (use_top) ai_action(s, &s->top, HOME_TOP_Y, -1, use_puck, info->human_action);

    float d_before = dist2(&s->bot, &s->puck);

    info->hit_the_border = (uint8_t)move_mallet(&s->bot, info->robot_action, MID_Y + MALLET_R, RINK_B - MALLET_R,
                                                RINK_B - CORNER_R, 1);
    if (use_top)
        move_mallet(&s->top, info->human_action, RINK_T + MALLET_R, MID_Y - MALLET_R, RINK_T + CORNER_R, 0);

    if (use_puck) {
        s->puck.vx = s->puck.vx * PUCK_FRICTION;
        s->puck.vy = s->puck.vy * PUCK_FRICTION;
        clamp_speed(&s->puck, PUCK_MAX);
        s->puck.x = s->puck.x + s->puck.vx;
        s->puck.y = s->puck.y + s->puck.vy;

        info->puck_was_hit = (uint8_t)hit_puck(&s->puck, &s->bot, -1.0f);
        if (use_top) hit_puck(&s->puck, &s->top, 1.0f);

        puck_borders(&s->puck);
        clamp_speed(&s->puck, PUCK_MAX);

        if (s->puck.y < RINK_T) info->scored = SCORED_BOTTOM;
        else if (s->puck.y > RINK_B) info->scored = SCORED_TOP;

        info->distance_decreased = dist2(&s->bot, &s->puck) < d_before;
        info->puck_is_at_the_bottom = s->puck.y > MID_Y;
    }
    {
        float dx = s->bot.x - HOME_X, dy = s->bot.y - HOME_BOT_Y;
        info->in_the_target = (dx * dx + dy * dy) <= TARGET_R * TARGET_R;
    }
    if (info->scored != SCORED_NONE) orc_sim_reset(s);
}

/* ---------------------------------------------------------------- native frame (SPEC.md §7) */
static int pix(float v) { return (int)(v + 0.5f); }

static void put(uint8_t *frame, int x, int y, int r, int g, int b)
{
    if (x < 0 || x >= ORC_W || y < 0 || y >= ORC_H) return;
    uint8_t *p = frame + ((size_t)y * ORC_W + x) * 3;
    p[0] = (uint8_t)r; p[1] = (uint8_t)g; p[2] = (uint8_t)b;
}

static void draw_mallet(uint8_t *frame, int cx, int cy, const uint8_t pal[4][3])
{
    for (int dy = -24; dy <= 24; ++dy)
        for (int dx = -24; dx <= 24; ++dx) {
            int d2 = dx * dx + dy * dy;
            if (d2 > 576) continue;
            const uint8_t *c;
            if ((dx + 3) * (dx + 3) + (dy + 3) * (dy + 3) <= 9) c = pal[0];
            else if (d2 <= 100) c = pal[1];
            else if (d2 > 400) c = pal[2];
            else c = pal[3];
            put(frame, cx + dx, cy + dy, c[0], c[1], c[2]);
        }
}

/* frame: uint8[700][450][3]; table: same shape (the static table image, an input) */
void orc_draw(const orc_sim *s, int use, const uint8_t *table, uint8_t *frame)
{
    static const uint8_t PAL_BOT[4][3] = { {220,220,230}, {90,90,100}, {20,20,24}, {40,40,48} };
    static const uint8_t PAL_TOP[4][3] = { {225,215,215}, {110,70,70}, {35,20,20}, {70,40,40} };
    const int ibx = pix(s->bot.x), iby = pix(s->bot.y);
    const int itx = pix(s->top.x), ity = pix(s->top.y);
    const int ipx = pix(s->puck.x), ipy = pix(s->puck.y);

    memcpy(frame, table, (size_t)ORC_W * ORC_H * 3);
    for (int y = iby - 3; y <= iby + 2; ++y)
        for (int x = 0; x < ORC_W; ++x) put(frame, x, y, 235, 235, 225);
    if (use & USE_PUCK)
        for (int dy = -18; dy <= 18; ++dy)
            for (int dx = -18; dx <= 18; ++dx) {
                int d2 = dx * dx + dy * dy;
                if (d2 > 324) continue;
                if ((dx + 5) * (dx + 5) + (dy + 5) * (dy + 5) <= 16) put(frame, ipx + dx, ipy + dy, 240, 120, 130);
                else if (d2 > 196) put(frame, ipx + dx, ipy + dy, 150, 20, 30);
                else put(frame, ipx + dx, ipy + dy, 215, 40, 50);
            }
    draw_mallet(frame, ibx, iby, PAL_BOT);
    if (use & USE_TOP) draw_mallet(frame, itx, ity, PAL_TOP);
    if (use & USE_ARM)
        for (int y = 0; y <= ity + 10; ++y) {
            int t = ity + 10 - y;
            int xc = itx + ((t * 5) >> 5);
            int hw = 26 + ((t * 3) >> 4);
            for (int x = xc - hw; x <= xc + hw; ++x) {
                int a = x - xc; if (a < 0) a = -a;
                if (a > hw - 4) put(frame, x, y, 160, 110, 95);
                else put(frame, x, y, 200, 150, 130);
            }
        }
}

/* ---------------------------------------------------------------- observation (processor.py:31-66) */

/* Image.resize((128,128), NEAREST) then transpose(2,0,1)  (processor.py:34-36).
 * Restates Pillow's nearest-neighbour scaling: source index = (int)((dst + 0.5) * src_size / dst_size),
 * evaluated in double like Pillow's ImagingScaleAffine. */
void orc_resize_nearest_chw(const uint8_t *frame, int h, int w, uint8_t *out /* [3][128][128] */)
{
    const double sx = (double)w / ORC_DIM, sy = (double)h / ORC_DIM;
    for (int i = 0; i < ORC_DIM; ++i) {
        int yi = (int)((i + 0.5) * sy);
        for (int j = 0; j < ORC_DIM; ++j) {
            int xi = (int)((j + 0.5) * sx);
            const uint8_t *p = frame + ((size_t)yi * w + xi) * 3;
            for (int c = 0; c < 3; ++c) out[(c * ORC_DIM + i) * ORC_DIM + j] = p[c];
        }
    }
}

/* shift register, newest first (processor.py:47-49) */
void orc_push_frame(float *stack /* [9][128][128] */, const uint8_t *chw /* [3][128][128] */)
{
    memmove(stack + 2 * ORC_FRAME, stack + ORC_FRAME, ORC_FRAME * sizeof(float));
    memmove(stack + ORC_FRAME, stack, ORC_FRAME * sizeof(float));
    for (int k = 0; k < ORC_FRAME; ++k) stack[k] = (float)chw[k];
}

/* 'rgb' mode: (x - 128) / 128 in float32 (processor.py:39-42,52) */
void orc_observe_rgb(const float *stack, float *out /* [9][128][128] */)
{
    for (int k = 0; k < 3 * ORC_FRAME; ++k) out[k] = (stack[k] - 128.0f) / 128.0f;
}

/* 'gray-diff' mode: uint8 wrap-around oldest - middle - newest, luma, normalise; float64 (processor.py:53-64) */
void orc_observe_gray_diff(const float *stack, double *out /* [1][128][128] */)
{
    const int plane = ORC_DIM * ORC_DIM;
    for (int k = 0; k < plane; ++k) {
        uint8_t d[3];
        for (int c = 0; c < 3; ++c) {
            uint8_t newest = (uint8_t)stack[c * plane + k];
            uint8_t middle = (uint8_t)stack[ORC_FRAME + c * plane + k];
            uint8_t oldest = (uint8_t)stack[2 * ORC_FRAME + c * plane + k];
            d[c] = (uint8_t)(oldest - middle - newest);
        }
        double gray = 0.2989 * d[0] + 0.5870 * d[1] + 0.1140 * d[2];
        out[k] = (gray - 128.0) / 128.0;
    }
}

/* SPEC.md §7: noise frame of a `random` episode */
void orc_noise_frame(const uint32_t key[2], uint32_t env_id, uint32_t noise_count, uint8_t *chw)
{
    for (uint32_t v = 0; v < ORC_FRAME / 16; ++v) {
        uint32_t ctr[4] = { v, noise_count, env_id, 2u }, r[4];
        orc_philox(ctr, key, r);
        for (int k = 0; k < 16; ++k) {
            uint32_t byte = (r[k >> 2] >> (8 * (k & 3))) & 255u;
            chw[v * 16 + k] = (uint8_t)((byte * 255u) >> 8);
        }
    }
}

/* ---------------------------------------------------------------- env wrapper (air_hockey_env.py:26-155) */
static const int8_t ACTIONS[10][2] = { /* processor.py:7-18 */
    {-1,-1}, {-1,0}, {-1,1}, {0,-1}, {0,0}, {0,1}, {1,-1}, {1,0}, {1,1}, {0,0} };

int orc_action_to_label(const int8_t a[2]) { return a[0] * 3 + a[1] + 4; } /* processor.py:84-85 */

/* the reward table, air_hockey_env.py:95-145 */
float orc_reward(int episode, const orc_info *g, int robot_label)
{
    float reward = 0.0f;
    if (episode == EP_NORMAL) {
        if (g->scored == SCORED_TOP) reward = -1.0f;
        else if (g->scored == SCORED_BOTTOM) reward = 1.0f;
        else if (g->puck_was_hit) reward = 0.5f;
        else if (g->hit_the_border) reward = -0.5f;
        else if (g->puck_is_at_the_bottom) reward = g->distance_decreased ? 0.5f : -0.5f;
        else if (g->in_the_target) reward = (robot_label == 4 || robot_label == 9) ? 0.5f : -0.5f;
        else reward = -0.5f;
    } else if (episode == EP_MISSING) {
        reward = robot_label == 4 ? 0.5f : -0.5f;
    } else {
        reward = robot_label == 9 ? 0.5f : -0.5f;
    }
    if (reward < -1.0f) reward = -1.0f;
    if (reward > 1.0f) reward = 1.0f;
    return reward;
}

typedef struct {
    orc_sim sim;
    uint32_t reset_count, noise_count, episode_steps;
    int32_t episode, use, pending_reset;
    uint8_t ring[3][ORC_FRAME]; /* ring[0] newest */
    uint8_t *native;            /* scratch uint8[700][450][3] */
    const uint8_t *table;
    int skip_draw;              /* orc_run_trace: advance state and counters, leave the ring alone (long-horizon sweeps) */
} orc_env;

orc_env *orc_env_create(uint64_t seed, uint32_t env_id, const uint8_t *table)
{
    orc_env *e = (orc_env *)calloc(1, sizeof *e);
    if (!e) return NULL;
    orc_sim_init(&e->sim, seed, env_id);
    e->native = (uint8_t *)malloc((size_t)ORC_W * ORC_H * 3);
    e->table = table;
    e->use = USE_ALL;
    orc_sim_reset(&e->sim); /* construction serves once (SPEC.md §9), like AirHockey() before the wrapper's reset */
    return e;
}

void orc_env_destroy(orc_env *e) { if (e) { free(e->native); free(e); } }

static void env_push(orc_env *e, const uint8_t *chw)
{
    memcpy(e->ring[2], e->ring[1], ORC_FRAME);
    memcpy(e->ring[1], e->ring[0], ORC_FRAME);
    memcpy(e->ring[0], chw, ORC_FRAME);
}

static void env_render_push(orc_env *e, int use)
{
    uint8_t chw[ORC_FRAME];
    if (e->skip_draw) return;
    orc_draw(&e->sim, use, e->table, e->native);
    orc_resize_nearest_chw(e->native, ORC_H, ORC_W, chw);
    env_push(e, chw);
}

static void env_noise_push(orc_env *e)
{
    uint8_t chw[ORC_FRAME];
    e->noise_count += 1;
    if (e->skip_draw) return;
    orc_noise_frame(e->sim.key, e->sim.env_id, e->noise_count - 1, chw);
    env_push(e, chw);
}

/* SPEC.md §9 reset; episode < 0: sample (§4 purpose 0), else forced episode type and mask */
void orc_env_reset(orc_env *e, int episode, int use)
{
    uint32_t ctr[4] = { e->reset_count, 0u, e->sim.env_id, 0u }, r[4];
    orc_philox(ctr, e->sim.key, r);
    e->reset_count += 1;
    if (episode < 0) {
        uint32_t t = r[0] % 20u;
        episode = t < 16u ? EP_NORMAL : (t < 19u ? EP_MISSING : EP_RANDOM);
        use = USE_ALL;
        if (episode == EP_MISSING) {
            static const int COMBO[6] = { 0, USE_TOP, USE_PUCK, USE_PUCK | USE_TOP, USE_ARM, USE_ARM | USE_TOP };
            use = COMBO[r[1] % 6u]; /* air_hockey_env.py:39-63 */
        }
    }
    e->episode = episode;
    e->use = use;
    e->episode_steps = 0;
    e->pending_reset = 0;
    if (episode != EP_RANDOM) {
        orc_info info;
        orc_sim_reset(&e->sim);
        for (int k = 0; k < 3; ++k) { /* warm-up: stand, stand, use_object omitted (air_hockey_env.py:69-72) */
            orc_sim_step(&e->sim, ACTIONS[4], ACTIONS[4], USE_ALL, &info);
            env_render_push(e, USE_ALL);
        }
    } else {
        for (int k = 0; k < 3; ++k) env_noise_push(e);
    }
}

typedef struct {
    float reward;
    uint8_t terminal, robot_label, human_label, scored;
    uint8_t flags; /* bit0 puck_was_hit, 1 hit_the_border, 2 puck_is_at_the_bottom, 3 distance_decreased, 4 in_the_target */
} orc_step_out;

/* air_hockey_env.py:84-155; labels 0..9 or ACT_NONE */
void orc_env_step(orc_env *e, int robot_label, int human_label, orc_step_out *out)
{
    orc_info g;
    orc_sim_step(&e->sim, robot_label == ACT_NONE ? NULL : ACTIONS[robot_label],
                 human_label == ACT_NONE ? NULL : ACTIONS[human_label], e->use, &g);
    out->robot_label = (uint8_t)orc_action_to_label(g.robot_action);
    out->human_label = (uint8_t)orc_action_to_label(g.human_action);
    out->reward = orc_reward(e->episode, &g, out->robot_label);
    out->scored = (uint8_t)g.scored;
    out->terminal = g.scored != SCORED_NONE;
    out->flags = (uint8_t)(g.puck_was_hit | g.hit_the_border << 1 | g.puck_is_at_the_bottom << 2 |
                           g.distance_decreased << 3 | g.in_the_target << 4);
    if (e->episode == EP_RANDOM) {
        out->terminal = 0;
        env_noise_push(e);
    } else {
        env_render_push(e, e->use);
    }
    e->episode_steps += 1;
}

/* batched-engine semantics (SPEC.md §9 last paragraph): reset-if-pending, step, arm the next reset */
void orc_env_step_auto(orc_env *e, int robot_label, int human_label, int max_episode_steps, orc_step_out *out)
{
    if (e->pending_reset) orc_env_reset(e, -1, 0);
    orc_env_step(e, robot_label, human_label, out);
    if (out->terminal || (max_episode_steps > 0 && (int)e->episode_steps >= max_episode_steps)) e->pending_reset = 1;
}

void orc_env_set_pending(orc_env *e, int pending) { e->pending_reset = pending; }

/* accessors for ctypes */
const uint8_t *orc_env_ring(const orc_env *e) { return &e->ring[0][0]; }
void orc_env_get_state(const orc_env *e, float out[12], uint32_t meta[8])
{
    memcpy(out, &e->sim.puck, 12 * sizeof(float));
    meta[0] = e->sim.serve_count; meta[1] = e->reset_count; meta[2] = e->noise_count; meta[3] = e->episode_steps;
    meta[4] = (uint32_t)e->episode; meta[5] = (uint32_t)e->use; meta[6] = (uint32_t)e->pending_reset; meta[7] = e->sim.env_id;
}
void orc_env_set_state(orc_env *e, const float in[12], const uint32_t meta[8])
{
    memcpy(&e->sim.puck, in, 12 * sizeof(float));
    e->sim.serve_count = meta[0]; e->reset_count = meta[1]; e->noise_count = meta[2]; e->episode_steps = meta[3];
    e->episode = (int32_t)meta[4]; e->use = (int32_t)meta[5]; e->pending_reset = (int32_t)meta[6];
}
orc_sim *orc_env_sim(orc_env *e) { return &e->sim; }
const uint8_t *orc_env_native(const orc_env *e) { return e->native; }

/* the reference's float32 (9,128,128) observation of the current ring (processor.py:52) */
void orc_env_observe_rgb(const orc_env *e, float *out)
{
    for (int k = 0; k < 3 * ORC_FRAME; ++k) out[k] = ((float)(&e->ring[0][0])[k] - 128.0f) / 128.0f;
}

/* ---------------------------------------------------------------- timed CPU baseline (bench.py) */
/* Runs `steps` auto-reset steps (AI on both sides) for envs [0, n) and writes what the reference's
 * env.step() hands back per step: the float32 (9,128,128) observation, reward, terminal, two labels.
 * obs_out: float[n][9*128*128] (each env's latest).  Envs are split over `threads` pthreads.
 * Returns a checksum so the work cannot be optimised away. */
typedef struct {
    orc_env **envs; int lo, hi, steps, max_episode_steps;
    float *obs_out, *reward_out; uint8_t *done_out, *labels_out; uint64_t check;
} orc_job;

static void *run_job(void *arg)
{
    orc_job *j = (orc_job *)arg;
    uint64_t check = 0;
    for (int i = j->lo; i < j->hi; ++i) {
        orc_step_out o;
        for (int t = 0; t < j->steps; ++t) {
            orc_env_step_auto(j->envs[i], ACT_NONE, ACT_NONE, j->max_episode_steps, &o);
            if (j->obs_out) orc_env_observe_rgb(j->envs[i], j->obs_out + (size_t)i * 3 * ORC_FRAME);
            if (j->reward_out) j->reward_out[i] = o.reward;
            if (j->done_out) j->done_out[i] = o.terminal;
            if (j->labels_out) { j->labels_out[2 * i] = o.robot_label; j->labels_out[2 * i + 1] = o.human_label; }
            check += o.robot_label + 16u * o.human_label + 256u * o.terminal;
        }
    }
    j->check = check;
    return NULL;
}

/* A persistent pool of worker threads (bench.py: the reference arm and the cpu_baseline leg share it, so that neither
 * pays thread creation inside its timed region).  A job = one function over a slice of [0, n). */
typedef struct orc_pool {
    pthread_mutex_t mu;
    pthread_cond_t go, done;
    pthread_t *tids;
    int threads, generation, remaining, quit;
    void (*fn)(void *ctx, int lo, int hi, int worker);
    void *ctx;
    int n;
} orc_pool;

typedef struct { orc_pool *pool; int k; } orc_worker_arg;

static void *pool_worker(void *arg)
{
    orc_worker_arg *wa = (orc_worker_arg *)arg;
    orc_pool *p = wa->pool;
    const int k = wa->k;
    int seen = 0;
    free(wa);
    for (;;) {
        pthread_mutex_lock(&p->mu);
        while (p->generation == seen && !p->quit) pthread_cond_wait(&p->go, &p->mu);
        if (p->quit) { pthread_mutex_unlock(&p->mu); return NULL; }
        seen = p->generation;
        const int n = p->n;
        pthread_mutex_unlock(&p->mu);
        const int lo = (int)((long long)n * k / p->threads), hi = (int)((long long)n * (k + 1) / p->threads);
        if (hi > lo) p->fn(p->ctx, lo, hi, k);
        pthread_mutex_lock(&p->mu);
        if (--p->remaining == 0) pthread_cond_signal(&p->done);
        pthread_mutex_unlock(&p->mu);
    }
}

orc_pool *orc_pool_create(int threads)
{
    if (threads < 1) threads = 1;
    orc_pool *p = (orc_pool *)calloc(1, sizeof *p);
    if (!p) return NULL;
    pthread_mutex_init(&p->mu, NULL);
    pthread_cond_init(&p->go, NULL);
    pthread_cond_init(&p->done, NULL);
    p->threads = threads;
    p->tids = (pthread_t *)calloc((size_t)threads, sizeof *p->tids);
    for (int k = 0; k < threads; ++k) {
        orc_worker_arg *wa = (orc_worker_arg *)malloc(sizeof *wa);
        wa->pool = p; wa->k = k;
        pthread_create(&p->tids[k], NULL, pool_worker, wa);
    }
    return p;
}

void orc_pool_destroy(orc_pool *p)
{
    if (!p) return;
    pthread_mutex_lock(&p->mu);
    p->quit = 1;
    pthread_cond_broadcast(&p->go);
    pthread_mutex_unlock(&p->mu);
    for (int k = 0; k < p->threads; ++k) pthread_join(p->tids[k], NULL);
    pthread_mutex_destroy(&p->mu);
    pthread_cond_destroy(&p->go);
    pthread_cond_destroy(&p->done);
    free(p->tids);
    free(p);
}

int orc_pool_threads(const orc_pool *p) { return p ? p->threads : 0; }

static void pool_run(orc_pool *p, void (*fn)(void *, int, int, int), void *ctx, int n)
{
    pthread_mutex_lock(&p->mu);
    p->fn = fn; p->ctx = ctx; p->n = n;
    p->remaining = p->threads;
    p->generation += 1;
    pthread_cond_broadcast(&p->go);
    while (p->remaining > 0) pthread_cond_wait(&p->done, &p->mu);
    pthread_mutex_unlock(&p->mu);
}

typedef struct { orc_job job; uint64_t *checks; } orc_batch_ctx;

static void batch_slice(void *ctx, int lo, int hi, int worker)
{
    orc_batch_ctx *c = (orc_batch_ctx *)ctx;
    orc_job j = c->job;
    j.lo = lo; j.hi = hi;
    run_job(&j);
    c->checks[worker] = j.check;
}

/* orc_run_batch on a persistent pool: same work per env-step, no thread creation and no allocation inside the call */
uint64_t orc_pool_run_batch(orc_pool *p, orc_env **envs, int n, int steps, int max_episode_steps, float *obs_out,
                            float *reward_out, uint8_t *done_out, uint8_t *labels_out)
{
    uint64_t checks[1024] = { 0 }, check = 0;
    if (!p || p->threads > 1024) return 0;
    orc_batch_ctx c = { { envs, 0, 0, steps, max_episode_steps, obs_out, reward_out, done_out, labels_out, 0 }, checks };
    pool_run(p, batch_slice, &c, n);
    for (int k = 0; k < p->threads; ++k) check += checks[k];
    return check;
}

uint64_t orc_run_batch(orc_env **envs, int n, int steps, int max_episode_steps, int threads, float *obs_out,
                       float *reward_out, uint8_t *done_out, uint8_t *labels_out)
{
    if (threads < 1) threads = 1;
    if (threads > n) threads = n > 0 ? n : 1;
    orc_pool *p = orc_pool_create(threads);
    uint64_t check = orc_pool_run_batch(p, envs, n, steps, max_episode_steps, obs_out, reward_out, done_out, labels_out);
    orc_pool_destroy(p);
    return check;
}

/* ---------------------------------------------------------------- long-horizon trace (tools/stress_sweep.py, tests) */
/* Every env does `steps` steps (auto_reset: the batched engine's reset-if-pending protocol, else plain steps), AI on both
 * sides, and logs per step the engine's 16 state words — state_out[t][w][i], the SoA layout of ahk_buffers.state: float
 * bits of puck/bottom/top, serve/reset/noise counters, meta = episode_steps | episode << 16 | use << 18 | pending << 21 —
 * and out_out[t][i] = (reward * 2 + 2) | terminal << 3 | robot_label << 4 | human_label << 8 | flags << 12 | scored << 17.
 * robot / human: optional caller actions per step and env.
 * Frames are drawn only during the last `draw_last` steps (the ring at the end is then exact provided draw_last >= 3). */
typedef struct {
    orc_env **envs; int n, steps, max_episode_steps, auto_reset, draw_last;
    const uint8_t *robot, *human; /* [steps][n] action labels (ACT_NONE = programmed AI) or NULL = AI throughout */
    uint32_t *state_out, *out_out;
} orc_trace_ctx;

static void trace_slice(void *ctx, int lo, int hi, int worker)
{
    orc_trace_ctx *c = (orc_trace_ctx *)ctx;
    (void)worker;
    for (int i = lo; i < hi; ++i) {
        orc_env *e = c->envs[i];
        for (int t = 0; t < c->steps; ++t) {
            orc_step_out o;
            e->skip_draw = t < c->steps - c->draw_last;
            const int r = c->robot ? c->robot[(size_t)t * c->n + i] : ACT_NONE, h = c->human ? c->human[(size_t)t * c->n + i] : ACT_NONE;
            if (c->auto_reset) orc_env_step_auto(e, r, h, c->max_episode_steps, &o);
            else orc_env_step(e, r, h, &o);
            if (c->state_out) {
                uint32_t *w = c->state_out + (size_t)t * 16 * c->n + i;
                const uint32_t *f = (const uint32_t *)&e->sim.puck;
                for (int k = 0; k < 12; ++k) w[(size_t)k * c->n] = f[k];
                w[(size_t)12 * c->n] = e->sim.serve_count;
                w[(size_t)13 * c->n] = e->reset_count;
                w[(size_t)14 * c->n] = e->noise_count;
                w[(size_t)15 * c->n] = (e->episode_steps > 0xFFFFu ? 0xFFFFu : e->episode_steps) | (uint32_t)e->episode << 16 |
                                       (uint32_t)e->use << 18 | (uint32_t)e->pending_reset << 21;
            }
            if (c->out_out)
                c->out_out[(size_t)t * c->n + i] = (uint32_t)(int)(o.reward * 2.0f + 2.0f) | (uint32_t)o.terminal << 3 |
                                                   (uint32_t)o.robot_label << 4 | (uint32_t)o.human_label << 8 |
                                                   (uint32_t)o.flags << 12 | (uint32_t)o.scored << 17;
        }
        e->skip_draw = 0;
    }
}

void orc_run_trace(orc_pool *p, orc_env **envs, int n, int steps, int max_episode_steps, int auto_reset, int draw_last,
                   const uint8_t *robot, const uint8_t *human, uint32_t *state_out, uint32_t *out_out)
{
    orc_trace_ctx c = { envs, n, steps, max_episode_steps, auto_reset, draw_last, robot, human, state_out, out_out };
    pool_run(p, trace_slice, &c, n);
}

int orc_sizeof_sim(void) { return (int)sizeof(orc_sim); }

/* ---------------------------------------------------------------- ahk_cpu_*: the CPU twin of the engine's C ABI */
/* SURVEY.md §8(b), last row of "C symbols": "identical ahk_cpu_* entry points for the CPU baseline".  Same ahk_config /
 * ahk_buffers structs (include/ahk.h) and the same buffer layouts as the CUDA library, with HOST pointers: SoA state
 * [16][n], frame ring [3][n][3][128][128] with the newest frame in slot `head`, reward / done / labels / info [n], stats.
 * Each env is an orc_env of this file, stepped by a persistent thread pool.  Declared in include/ahk_cpu.h.  Test
 * infrastructure like the rest of oracle/: the product library never links or loads it. */
#include <stdio.h>
#include <unistd.h>
#include "../include/ahk_cpu.h"

struct ahk_cpu_env {
    ahk_config cfg;
    ahk_buffers buf;
    orc_env **envs;
    orc_pool *pool;
    uint8_t *table;
    int head, bound;
    uint64_t *part; /* per-worker statistics [threads][AHK_STATS_WORDS] */
    void *obs_out;  /* ahk_cpu_bind_observation: every step also writes each env's observation here ... */
    int obs_format; /* ... in this format, by the thread that stepped it */
};

static __thread char cpu_error[256];
const char *ahk_cpu_last_error(void) { return cpu_error; }
static int cpu_fail(int code, const char *text) { snprintf(cpu_error, sizeof cpu_error, "%s", text); return code; }

int ahk_cpu_create(const ahk_config *cfg, ahk_cpu_env **out)
{
    if (!cfg || !out) return cpu_fail(AHK_ERR_ARG, "cfg/out is NULL");
    if (cfg->n_envs < 1) return cpu_fail(AHK_ERR_ARG, "n_envs must be >= 1");
    if (cfg->ring_depth != 0 && cfg->ring_depth != 3) return cpu_fail(AHK_ERR_ARG, "the CPU twin keeps the 3-slot ring only");
    ahk_cpu_env *e = (ahk_cpu_env *)calloc(1, sizeof *e);
    if (!e) return cpu_fail(AHK_ERR_ARG, "out of memory");
    e->cfg = *cfg;
    int threads = (int)sysconf(_SC_NPROCESSORS_ONLN);
    const char *sw = getenv("AHK_CPU_THREADS");
    if (sw && atoi(sw) > 0) threads = atoi(sw);
    if (threads > cfg->n_envs) threads = cfg->n_envs;
    if (threads > 1024) threads = 1024;
    e->pool = orc_pool_create(threads);
    e->part = (uint64_t *)calloc((size_t)threads * AHK_STATS_WORDS, sizeof(uint64_t));
    e->head = 0; /* as ahk_create: a reset keeps the head, the first step writes slot 1 */
    *out = e;
    return AHK_OK;
}

void ahk_cpu_destroy(ahk_cpu_env *e)
{
    if (!e) return;
    if (e->envs) for (int i = 0; i < e->cfg.n_envs; ++i) orc_env_destroy(e->envs[i]);
    free(e->envs);
    orc_pool_destroy(e->pool);
    free(e->table);
    free(e->part);
    free(e);
}

int ahk_cpu_threads(const ahk_cpu_env *e) { return e ? orc_pool_threads(e->pool) : 0; }
int ahk_cpu_head(const ahk_cpu_env *e) { return e ? e->head : -1; }

int ahk_cpu_set_table_image(ahk_cpu_env *e, const uint8_t *rgb, int height, int width)
{
    if (!e || !rgb) return cpu_fail(AHK_ERR_ARG, "env/rgb is NULL");
    if (height != ORC_H || width != ORC_W) return cpu_fail(AHK_ERR_ARG, "the table image must be 700 x 450 x 3");
    if (e->envs) return cpu_fail(AHK_ERR_STATE, "the table image is already set");
    e->table = (uint8_t *)malloc((size_t)ORC_W * ORC_H * 3);
    memcpy(e->table, rgb, (size_t)ORC_W * ORC_H * 3);
    e->envs = (orc_env **)calloc((size_t)e->cfg.n_envs, sizeof *e->envs);
    for (int i = 0; i < e->cfg.n_envs; ++i) e->envs[i] = orc_env_create(e->cfg.seed, e->cfg.env_id_base + (uint32_t)i, e->table);
    return AHK_OK;
}

int ahk_cpu_bind_buffers(ahk_cpu_env *e, const ahk_buffers *b)
{
    if (!e || !b) return cpu_fail(AHK_ERR_ARG, "env/bufs is NULL");
    if (!b->state || !b->frames || !b->reward || !b->done || !b->robot_label || !b->human_label)
        return cpu_fail(AHK_ERR_ARG, "state, frames, reward, done and the two label buffers are required");
    e->buf = *b;
    e->bound = 1;
    return AHK_OK;
}

typedef struct {
    ahk_cpu_env *e;
    int mode;                 /* 0 reset, 1 step */
    const uint8_t *mask, *episode, *use, *robot, *human;
    const ahk_step_log *log;
    int steps;                /* mode 1: consecutive steps per env (ahk_cpu_step_many runs them env by env, like a CTA of the CUDA kernel) */
    int head0;                /* ring slot of the newest frame BEFORE the call */
} cpu_call;

static void cpu_observe_one(const orc_env *e, int i, int format, void *out);

static void cpu_publish(ahk_cpu_env *h, int i, int all_frames, const int slots[3])
{
    const orc_env *e = h->envs[i];
    const size_t n = (size_t)h->cfg.n_envs;
    uint32_t *w = h->buf.state + i;
    const uint32_t *f = (const uint32_t *)&e->sim.puck;
    for (int k = 0; k < 12; ++k) w[(size_t)k * n] = f[k];
    w[12 * n] = e->sim.serve_count;
    w[13 * n] = e->reset_count;
    w[14 * n] = e->noise_count;
    w[15 * n] = (e->episode_steps > 0xFFFFu ? 0xFFFFu : e->episode_steps) | (uint32_t)e->episode << 16 | (uint32_t)e->use << 18 |
                (uint32_t)e->pending_reset << 21;
    for (int age = 0; age < (all_frames ? 3 : 1); ++age)
        memcpy(h->buf.frames + ((size_t)slots[age] * n + (size_t)i) * ORC_FRAME, e->ring[age], ORC_FRAME);
}

static void cpu_slice(void *ctx, int lo, int hi, int worker)
{
    cpu_call *c = (cpu_call *)ctx;
    ahk_cpu_env *h = c->e;
    uint64_t *st = h->part + (size_t)worker * AHK_STATS_WORDS;
    const size_t n = (size_t)h->cfg.n_envs;
    for (int i = lo; i < hi; ++i) {
        orc_env *e = h->envs[i];
        int slots[3]; /* ring slots of the frames newest, middle, oldest */
        if (c->mode == 0) {
            if (c->mask && !c->mask[i]) continue;
            const int ep = c->episode && c->episode[i] != AHK_EP_SAMPLE ? c->episode[i] : -1;
            orc_env_reset(e, ep, ep < 0 ? 0 : (c->use ? c->use[i] : USE_ALL));
            for (int age = 0; age < 3; ++age) slots[age] = (c->head0 + 3 - age) % 3;
            cpu_publish(h, i, 1, slots);
            st[1] += 1; /* episodes: every reset counts, as in the CUDA engine */
            continue;
        }
        for (int t = 0; t < c->steps; ++t) { /* env-major: the env's state, ring and native frame stay in cache over its steps */
            orc_step_out o;
            const int was_pending = h->cfg.auto_reset && e->pending_reset;
            const int r = c->robot ? c->robot[i] : ACT_NONE, hu = c->human ? c->human[i] : ACT_NONE;
            if (h->cfg.auto_reset) orc_env_step_auto(e, r, hu, h->cfg.max_episode_steps, &o);
            else orc_env_step(e, r, hu, &o);
            for (int age = 0; age < 3; ++age) slots[age] = (c->head0 + 1 + t + 3 - age) % 3;
            cpu_publish(h, i, was_pending, slots);
            if (h->obs_out) cpu_observe_one(e, i, h->obs_format, h->obs_out); /* what the reference's env.step() returns, while the env is in cache */
            if (t == c->steps - 1) {
                h->buf.reward[i] = o.reward;
                h->buf.done[i] = o.terminal;
                h->buf.robot_label[i] = o.robot_label;
                h->buf.human_label[i] = o.human_label;
                if (h->buf.info) h->buf.info[i] = (uint8_t)(o.flags | o.scored << 5);
            }
            if (c->log) {
                const size_t at = (size_t)t * n + (size_t)i;
                if (c->log->robot_label) c->log->robot_label[at] = o.robot_label;
                if (c->log->human_label) c->log->human_label[at] = o.human_label;
                if (c->log->done) c->log->done[at] = o.terminal;
                if (c->log->reward) c->log->reward[at] = o.reward;
            }
            st[0] += 1;
            st[1] += (uint64_t)was_pending;
            st[2] += o.scored == SCORED_TOP;
            st[3] += o.scored == SCORED_BOTTOM;
            st[4] += (uint64_t)(int64_t)(int)(o.reward * 2.0f);
            st[5 + o.robot_label] += 1;
            st[15 + o.human_label] += 1;
        }
    }
}

static void cpu_run(ahk_cpu_env *e, cpu_call *c)
{
    pool_run(e->pool, cpu_slice, c, e->cfg.n_envs);
    if (e->buf.stats)
        for (int k = 0; k < orc_pool_threads(e->pool); ++k)
            for (int w = 0; w < AHK_STATS_WORDS; ++w) { e->buf.stats[w] += e->part[(size_t)k * AHK_STATS_WORDS + w]; e->part[(size_t)k * AHK_STATS_WORDS + w] = 0; }
}

static int cpu_ready(const ahk_cpu_env *e)
{
    if (!e) return cpu_fail(AHK_ERR_ARG, "env is NULL");
    if (!e->envs) return cpu_fail(AHK_ERR_STATE, "ahk_cpu_set_table_image has not been called");
    if (!e->bound) return cpu_fail(AHK_ERR_STATE, "ahk_cpu_bind_buffers has not been called");
    return AHK_OK;
}

int ahk_cpu_reset(ahk_cpu_env *e, const uint8_t *mask, const uint8_t *episode, const uint8_t *use, void *stream)
{
    (void)stream;
    int rc = cpu_ready(e);
    if (rc) return rc;
    cpu_call c = { e, 0, mask, episode, use, NULL, NULL, NULL, 0, e->head };
    cpu_run(e, &c);
    return AHK_OK;
}

int ahk_cpu_step(ahk_cpu_env *e, const uint8_t *robot, const uint8_t *human, void *stream)
{
    (void)stream;
    int rc = cpu_ready(e);
    if (rc) return rc;
    cpu_call c = { e, 1, NULL, NULL, NULL, robot, human, NULL, 1, e->head };
    cpu_run(e, &c);
    e->head = (e->head + 1) % 3;
    return AHK_OK;
}

int ahk_cpu_step_many(ahk_cpu_env *e, int steps, const ahk_step_log *log, void *stream)
{
    (void)stream;
    int rc = cpu_ready(e);
    if (rc) return rc;
    if (steps < 0) return cpu_fail(AHK_ERR_ARG, "steps must be >= 0");
    if (steps == 0) return AHK_OK;
    cpu_call c = { e, 1, NULL, NULL, NULL, NULL, NULL, log, steps, e->head };
    cpu_run(e, &c);
    e->head = (e->head + steps) % 3;
    return AHK_OK;
}

typedef struct { ahk_cpu_env *e; int format; void *out; } cpu_obs;

static void cpu_observe_one(const orc_env *e, int i, int format, void *out)
{
    if (format == AHK_OBS_U8) memcpy((uint8_t *)out + (size_t)i * 3 * ORC_FRAME, e->ring, 3 * ORC_FRAME);
    else if (format == AHK_OBS_RGB_F32) orc_env_observe_rgb(e, (float *)out + (size_t)i * 3 * ORC_FRAME);
    else {
        static __thread float stack[3 * ORC_FRAME]; /* the processor's float32 buffer (processor.py:23,45-49) */
        for (int k = 0; k < 3 * ORC_FRAME; ++k) stack[k] = (float)(&e->ring[0][0])[k];
        orc_observe_gray_diff(stack, (double *)out + (size_t)i * 128 * 128);
    }
}

static void cpu_obs_slice(void *ctx, int lo, int hi, int worker)
{
    cpu_obs *c = (cpu_obs *)ctx;
    (void)worker;
    for (int i = lo; i < hi; ++i) cpu_observe_one(c->e->envs[i], i, c->format, c->out);
}

/* ≙ ahk_observe: the stacked observation of every env in the reference's formats (processor.py:39-64), host memory */
int ahk_cpu_observe(ahk_cpu_env *e, int format, void *out, void *stream)
{
    (void)stream;
    int rc = cpu_ready(e);
    if (rc) return rc;
    if (!out || format < AHK_OBS_U8 || format > AHK_OBS_GRAY_DIFF_F64) return cpu_fail(AHK_ERR_ARG, "out is NULL or the format is unknown");
    cpu_obs c = { e, format, out };
    pool_run(e->pool, cpu_obs_slice, &c, e->cfg.n_envs);
    return AHK_OK;
}

/* The reference's env.step() RETURNS the processed observation (air_hockey_env.py:153-155, processor.py:39-64).  Bind a host
 * buffer and every ahk_cpu_step / ahk_cpu_step_many also writes each env's observation in `format` (as ahk_cpu_observe would),
 * by the thread that just stepped the env.  out_host NULL: unbind.  CPU twin only: on the GPU the frames stay uint8 in the ring
 * and ahk_observe materialises them on request. */
int ahk_cpu_bind_observation(ahk_cpu_env *e, int format, void *out_host)
{
    if (!e) return cpu_fail(AHK_ERR_ARG, "env is NULL");
    if (out_host && (format < AHK_OBS_U8 || format > AHK_OBS_GRAY_DIFF_F64)) return cpu_fail(AHK_ERR_ARG, "the format is unknown");
    e->obs_out = out_host;
    e->obs_format = format;
    return AHK_OK;
}
