// ahk_api.cu — the C ABI of include/ahk.h: handle management, argument checking, stream-ordered launches.
// There is no CPU implementation behind these entry points: without a CUDA device ahk_create fails.
#include "../../include/ahk.h"
#include "ahk_kernels.cuh"

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <unordered_map>
#include <vector>

namespace {

thread_local char g_error[512] = "";

int fail(int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof g_error, fmt, ap);
    va_end(ap);
    return code;
}

#define AHK_CUDA(expr)                                                                                      \
    do {                                                                                                    \
        cudaError_t err__ = (expr);                                                                         \
        if (err__ != cudaSuccess) return fail(AHK_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(err__));   \
    } while (0)

// Entry points run on the device of their handle and leave the caller's current device as they found it (two engines on
// different GPUs in one process, or a torch current device other than cfg.device, must not disturb each other).
struct DeviceGuard {
    int prev = -1;
    bool changed = false;
    cudaError_t status = cudaSuccess;
    explicit DeviceGuard(int device)
    {
        status = cudaGetDevice(&prev);
        if (status == cudaSuccess && prev != device) {
            status = cudaSetDevice(device);
            changed = status == cudaSuccess;
        }
    }
    ~DeviceGuard() { if (changed) cudaSetDevice(prev); }
    DeviceGuard(const DeviceGuard &) = delete;
    DeviceGuard &operator=(const DeviceGuard &) = delete;
};

#define AHK_ON_DEVICE(dev)                                                                                         \
    DeviceGuard guard__(dev);                                                                                      \
    if (guard__.status != cudaSuccess) return fail(AHK_ERR_CUDA, "cudaSetDevice(%d): %s", dev, cudaGetErrorString(guard__.status))

} // namespace

namespace ahk {
void set_error_text(const char *text) { snprintf(g_error, sizeof g_error, "%s", text); }
} // namespace ahk

struct ahk_env {
    ahk_config cfg{};
    ahk_buffers buf{};
    bool bound = false, table_set = false;
    uint4 *table_dev = nullptr;      // [3][128][8] vectors
    uint8_t *action_stage = nullptr; // [2][n] device staging for ahk_step_host
    cudaStream_t host_stream = nullptr;
    static constexpr int kHostDepth = 8;  // ahk_step_host_begin calls that may be in flight
    cudaEvent_t host_ev[kHostDepth] = {};
    uint64_t host_ticket = 0;
    // ahk_step_host_results: 8 mapped host words the step kernel signals through, their device alias, CTA counters
    unsigned long long *tick_flag_host = nullptr, *tick_flag_dev = nullptr;
    unsigned *tick_count_dev = nullptr;
    uint64_t tick_seq = 0;
    int host_early_ticket = -1; // >= 0: ahk_step_host_results returned before this launch (host_ev[ticket]) had completed
    cudaStream_t last_stream = nullptr; // stream of the previous fused launch
    bool zero_copy = true;                // A/B switch AHK_ZEROCOPY=0: always use the copy engine in ahk_step_host
    int head = 0;
    int depth = 3;              // ring slots (cfg.ring_depth; 0 = 3)
    int sms = 0;
    int raster_grid = 0;        // staged rasteriser (split path): one persistent CTA per SM
    int step_grid = 0, envs_per_cta = 0; // fused kernel: CTA b owns envs [b * envs_per_cta, (b + 1) * envs_per_cta)
    // A/B switches (measurements only): AHK_PATH=split runs env_kernel + a raster kernel instead of the fused kernel;
    // AHK_RASTER=staged picks the TMA-staged rasteriser for the split path (default: the band rasteriser, draw only)
    bool split = false, staged = false;
    uint64_t launches = 0;
    unsigned long long *trace = nullptr; // ahk_set_trace
    // adaptive partition (StepArgs): two boundary arrays (this launch / next launch), rates, finish counter
    int *bounds_dev[2] = {nullptr, nullptr};
    unsigned long long *end_dev = nullptr;
    unsigned *done_dev = nullptr;
    int bounds_cur = 0;
    unsigned adapt_tick = 0, adapt_every = 8;
    int switch_in = 0;
    bool balance = true; // A/B switch AHK_BALANCE=0
    bool many_kernel = true; // ahk_step_many through the persistent kernel; A/B switch AHK_MANY=loop: one launch per step
    int dyn_units = 1;   // warps claim raster units from a shared counter; A/B switch AHK_UNITS=static: round-robin
    int last_mode = -1;  // mode of the previous fused launch (step_args)
    int pdl = -1;        // programmatic dependent launch of consecutive step kernels; -1: by size (ahk_create); A/B switch AHK_PDL=0/1/3
    cudaEvent_t ev[3] = {nullptr, nullptr, nullptr}; // ahk_step_profiled: before env_kernel, between, after raster
};

namespace {

// Which engine launched the last step / reset kernel on each stream (any handle of this process).  The deferred
// griddepcontrol.wait (pdl 3) is only valid when the previous programmatic launch on the stream was THIS engine's own step
// ("a CTA of launch k runs => launch k-2 is complete" counts launches of one engine); two engines
// alternating on one stream see each other here and keep the wait in front of their first store (pdl 1).
std::mutex g_stream_mu;
std::unordered_map<cudaStream_t, const ahk_env *> g_stream_owner;

bool own_previous_launch(const ahk_env *e, cudaStream_t s)
{
    std::lock_guard<std::mutex> lock(g_stream_mu);
    auto it = g_stream_owner.find(s);
    const bool own = it != g_stream_owner.end() && it->second == e;
    g_stream_owner[s] = e;
    return own;
}

void forget_engine(const ahk_env *e)
{
    std::lock_guard<std::mutex> lock(g_stream_mu);
    for (auto it = g_stream_owner.begin(); it != g_stream_owner.end();)
        it = it->second == e ? g_stream_owner.erase(it) : std::next(it);
}

uint2 key_of(uint64_t seed) { return make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32)); }

int check_ready(const ahk_env *e)
{
    if (!e) return fail(AHK_ERR_ARG, "env is NULL");
    if (!e->bound) return fail(AHK_ERR_STATE, "ahk_bind_buffers has not been called");
    if (!e->table_set) return fail(AHK_ERR_STATE, "ahk_set_table_image has not been called");
    return AHK_OK;
}

ahk::EnvArgs env_args(const ahk_env *e)
{
    ahk::EnvArgs a{};
    a.state = e->buf.state;
    a.drawlist = e->buf.drawlist;
    a.reward = e->buf.reward;
    a.done = e->buf.done;
    a.robot_label = e->buf.robot_label;
    a.human_label = e->buf.human_label;
    a.info = e->buf.info;
    a.stats = e->buf.stats;
    a.key = key_of(e->cfg.seed);
    a.env_id_base = e->cfg.env_id_base;
    a.n = e->cfg.n_envs;
    a.auto_reset = e->cfg.auto_reset;
    a.max_episode_steps = e->cfg.max_episode_steps;
    return a;
}

ahk::RasterArgs raster_args(const ahk_env *e)
{
    ahk::RasterArgs r{};
    r.drawlist = e->buf.drawlist;
    r.table = e->table_dev;
    r.frames = e->buf.frames;
    r.warm = e->depth > 3 ? e->buf.warm : nullptr;
    r.depth = e->depth;
    r.key = key_of(e->cfg.seed);
    r.env_id_base = e->cfg.env_id_base;
    r.n = e->cfg.n_envs;
    r.head = e->head;
    return r;
}

ahk::StepArgs step_args(ahk_env *e, const ahk::EnvArgs &a, int mode, cudaStream_t s)
{
    ahk::StepArgs st{};
    st.env = a;
    st.ras = raster_args(e);
    st.envs_per_cta = e->envs_per_cta;
    // two trace slots, alternating per launch, so that overlapped consecutive launches can be laid side by side
    st.trace = e->trace ? e->trace + (e->launches & 1u) * static_cast<size_t>(e->step_grid) * (ahk::kStepThreads / 32) * 4 : nullptr;
    // Deferred wait (pdl 3) only between two consecutive STEP launches: a reset launch writes all three ring slots, so it
    // and the step after it keep the wait in front of their first store.
    const bool own_chain = own_previous_launch(e, s) && mode == 1 && e->last_mode == 1 && e->last_stream == s;
    st.pdl = (e->pdl == 3 && !own_chain) ? 1 : e->pdl;
    // Inputs another kernel (or copy) on the stream may have produced — action labels, reset masks, overrides, and the state
    // itself when the previous launch was not this engine's own step — are only guaranteed visible after
    // griddepcontrol.wait: the kernel then waits BEFORE its first load.  This costs nothing unless the predecessor is a
    // step kernel that triggered early.  Actions in mapped HOST memory (ahk_step_host*) are written by the host before the
    // launch and need no wait: only the engine's own action-free / host-action step-to-step chain runs its prologue under
    // the previous launch's tail.
    st.wait_first = st.pdl != 0 && (!own_chain || a.ext_inputs);
    e->last_mode = mode;
    e->last_stream = s;
    st.dyn_units = e->dyn_units;
    if (e->balance && !e->split) {
        // New boundaries are taken into use two launches after the one that computes them: with overlapped launches
        // the very next kernel may already be reading its boundaries while the last CTA of this one writes.
        if (e->switch_in > 0 && --e->switch_in == 0) e->bounds_cur ^= 1;
        st.bounds = e->bounds_dev[e->bounds_cur];
        // steps adapt, every 8th one (the last CTA's boundary computation is ~2 us on the launch's tail); a reset (three
        // frames per env, once) only uses the current boundaries
        if ((mode == 1 || mode == 3) && e->switch_in == 0 && (mode == 3 || (e->adapt_tick++ % e->adapt_every) == 0)) {
            st.bounds_next = e->bounds_dev[e->bounds_cur ^ 1];
            st.end_ns = e->end_dev;
            st.done_count = e->done_dev;
            e->switch_in = 2;
        }
    }
    return st;
}

// One pass of the hot path: mode 0 = reset (+ warm-up frames), 1 = step.  Default: ONE fused kernel.
int run(ahk_env *e, const ahk::EnvArgs &a, int mode, cudaStream_t s)
{
    // a launch that ahk_step_host_results left running on the engine's own stream: order this stream behind it
    if (e->host_early_ticket >= 0 && s != e->host_stream) AHK_CUDA(cudaStreamWaitEvent(s, e->host_ev[e->host_early_ticket], 0));
    ahk::StepArgs st = step_args(e, a, mode, s);
    if (!e->split) {
        AHK_CUDA(ahk::launch_step(st, mode, e->step_grid, s));
        e->launches += 1;
        return AHK_OK;
    }
    AHK_CUDA(ahk::launch_env(a, mode == 1, s));
    if (e->staged) AHK_CUDA(ahk::launch_raster_staged(st.ras, e->raster_grid, s));
    else AHK_CUDA(ahk::launch_step(st, 2, e->step_grid, s));
    e->launches += 2;
    return AHK_OK;
}

// Device-visible alias of a pinned (cudaHostAlloc / cudaHostRegister) host pointer, or nullptr for pageable memory.
template <typename T> T *mapped(T *host)
{
    if (!host) return nullptr;
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, host) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return at.type == cudaMemoryTypeHost ? static_cast<T *>(at.devicePointer) : nullptr;
}

} // namespace

extern "C" {

const char *ahk_last_error(void) { return g_error; }
const char *ahk_version(void) { return "ahk 0.1 (sm_100a)"; }

int ahk_create(const ahk_config *cfg, ahk_env **out)
{
    if (!cfg || !out) return fail(AHK_ERR_ARG, "cfg/out is NULL");
    if (cfg->n_envs <= 0) return fail(AHK_ERR_ARG, "n_envs must be positive, got %d", cfg->n_envs);
    if (cfg->max_episode_steps < 0 || cfg->max_episode_steps > 65535)
        return fail(AHK_ERR_ARG, "max_episode_steps must be in [0, 65535], got %d", cfg->max_episode_steps);
    if (cfg->ring_depth != 0 && (cfg->ring_depth < 3 || cfg->ring_depth > 65536))
        return fail(AHK_ERR_ARG, "ring_depth must be 0 (= 3) or in [3, 65536], got %d", cfg->ring_depth);
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
        return fail(AHK_ERR_NO_DEVICE, "no CUDA device: this engine has no CPU fallback");
    if (cfg->device < 0 || cfg->device >= count) return fail(AHK_ERR_ARG, "device %d out of range [0, %d)", cfg->device, count);
    AHK_ON_DEVICE(cfg->device);
    cudaDeviceProp prop{};
    AHK_CUDA(cudaGetDeviceProperties(&prop, cfg->device));
    if (prop.major < 10) return fail(AHK_ERR_NO_DEVICE, "device %d is sm_%d%d; this library is built for sm_100a only", cfg->device, prop.major, prop.minor);

    ahk_env *e = new (std::nothrow) ahk_env();
    if (!e) return fail(AHK_ERR_ARG, "out of host memory");
    e->cfg = *cfg;
    e->depth = cfg->ring_depth ? cfg->ring_depth : 3;
    if (const char *s = getenv("AHK_PATH")) e->split = strcmp(s, "split") == 0 && e->depth == 3; // the A/B path knows the 3-slot ring only
    if (const char *s = getenv("AHK_ADAPT_EVERY")) e->adapt_every = atoi(s) > 0 ? atoi(s) : 1;
    if (const char *s = getenv("AHK_PDL")) e->pdl = atoi(s);
    if (const char *s = getenv("AHK_UNITS")) e->dyn_units = strcmp(s, "static") != 0;
    if (const char *s = getenv("AHK_BALANCE")) e->balance = atoi(s) != 0;
    if (const char *s = getenv("AHK_RASTER")) e->staged = strcmp(s, "staged") == 0;
    if (const char *s = getenv("AHK_ZEROCOPY")) e->zero_copy = atoi(s) != 0;
    if (const char *s = getenv("AHK_MANY")) e->many_kernel = strcmp(s, "loop") != 0;
    e->sms = prop.multiProcessorCount;
    e->raster_grid = e->sms;
    e->envs_per_cta = (cfg->n_envs + e->sms - 1) / e->sms;
    e->step_grid = (cfg->n_envs + e->envs_per_cta - 1) / e->envs_per_cta;
    if (e->balance && !e->split && cfg->n_envs >= 4 * e->sms) e->step_grid = e->sms; // the adaptive partition uses every SM
    else e->balance = false;
    // Consecutive step kernels overlap (programmatic dependent launch): the next launch's CTAs move onto SMs as this
    // launch's CTAs retire and run their prologue and game tick under its tail.  pdl 1: griddepcontrol.wait before the
    // first frame store.  pdl 3 (default when the grid fills the device): the wait moves to the END of the kernel, so an SM
    // that retires early goes straight on with the next step's frames.  Why that is safe needs one CTA per SM and a grid
    // of at least one CTA per SM; smaller grids keep pdl 1.  A/B switch AHK_PDL=0/1/3.
    if (e->pdl < 0) e->pdl = 3;
    if (e->pdl == 3 && e->step_grid < e->sms) e->pdl = 1;
    cudaError_t err = cudaMalloc(&e->table_dev, ahk::kFrameBytes);
    if (err == cudaSuccess) err = cudaMalloc(&e->action_stage, 2 * static_cast<size_t>(cfg->n_envs));
    if (err == cudaSuccess && e->balance) {
        const int g = e->step_grid;
        std::vector<int> even(g + 1);
        for (int b = 0; b <= g; ++b) even[b] = static_cast<int>(static_cast<long long>(b) * cfg->n_envs / g);
        for (int k = 0; k < 2 && err == cudaSuccess; ++k) {
            err = cudaMalloc(&e->bounds_dev[k], (g + 1) * sizeof(int));
            if (err == cudaSuccess) err = cudaMemcpy(e->bounds_dev[k], even.data(), (g + 1) * sizeof(int), cudaMemcpyHostToDevice);
        }
        if (err == cudaSuccess) err = cudaMalloc(&e->end_dev, g * sizeof(unsigned long long));
        if (err == cudaSuccess) err = cudaMemset(e->end_dev, 0, g * sizeof(unsigned long long));
        if (err == cudaSuccess) err = cudaMalloc(&e->done_dev, sizeof(unsigned));
        if (err == cudaSuccess) err = cudaMemset(e->done_dev, 0, sizeof(unsigned));
    }
    if (err == cudaSuccess) err = cudaStreamCreateWithFlags(&e->host_stream, cudaStreamNonBlocking);
    if (err == cudaSuccess) err = cudaHostAlloc(&e->tick_flag_host, 8 * sizeof(unsigned long long), cudaHostAllocMapped);
    if (err == cudaSuccess) {
        memset(e->tick_flag_host, 0, 8 * sizeof(unsigned long long));
        err = cudaHostGetDevicePointer(&e->tick_flag_dev, e->tick_flag_host, 0);
    }
    if (err == cudaSuccess) err = cudaMalloc(&e->tick_count_dev, 8 * sizeof(unsigned));
    if (err == cudaSuccess) err = cudaMemset(e->tick_count_dev, 0, 8 * sizeof(unsigned));
    for (int k = 0; k < 3 && err == cudaSuccess; ++k) err = cudaEventCreate(&e->ev[k]);
    for (int k = 0; k < ahk_env::kHostDepth && err == cudaSuccess; ++k) err = cudaEventCreateWithFlags(&e->host_ev[k], cudaEventDisableTiming);
    if (err == cudaSuccess) err = ahk::step_prepare();
    if (err == cudaSuccess) err = ahk::raster_staged_prepare();
    if (err != cudaSuccess) {
        ahk_destroy(e);
        return fail(AHK_ERR_CUDA, "ahk_create: %s", cudaGetErrorString(err));
    }
    *out = e;
    return AHK_OK;
}

void ahk_destroy(ahk_env *e)
{
    if (!e) return;
    forget_engine(e);
    DeviceGuard guard(e->cfg.device);
    if (e->table_dev) cudaFree(e->table_dev);
    if (e->action_stage) cudaFree(e->action_stage);
    for (int k = 0; k < 2; ++k) if (e->bounds_dev[k]) cudaFree(e->bounds_dev[k]);
    if (e->end_dev) cudaFree(e->end_dev);
    if (e->done_dev) cudaFree(e->done_dev);
    if (e->host_stream) cudaStreamDestroy(e->host_stream);
    if (e->tick_flag_host) cudaFreeHost(e->tick_flag_host);
    if (e->tick_count_dev) cudaFree(e->tick_count_dev);
    for (int k = 0; k < 3; ++k) if (e->ev[k]) cudaEventDestroy(e->ev[k]);
    for (int k = 0; k < ahk_env::kHostDepth; ++k) if (e->host_ev[k]) cudaEventDestroy(e->host_ev[k]);
    delete e;
}

int ahk_set_table_image(ahk_env *e, const uint8_t *rgb, int height, int width)
{
    if (!e || !rgb) return fail(AHK_ERR_ARG, "env/rgb is NULL");
    if (height != AHK_TABLE_H || width != AHK_TABLE_W)
        return fail(AHK_ERR_ARG, "table image must be %dx%d (HxW), got %dx%d", AHK_TABLE_H, AHK_TABLE_W, height, width);
    // Image.resize((128,128), NEAREST) + transpose(2,0,1) of the static layer (processor.py:34-36)
    std::vector<uint8_t> chw(ahk::kFrameBytes);
    for (int i = 0; i < ahk::kDim; ++i) {
        const int y = ((2 * i + 1) * height) >> 8;
        for (int j = 0; j < ahk::kDim; ++j) {
            const int x = ((2 * j + 1) * width) >> 8;
            const uint8_t *p = rgb + (static_cast<size_t>(y) * width + x) * 3;
            for (int c = 0; c < 3; ++c) chw[(c * ahk::kDim + i) * ahk::kDim + j] = p[c];
        }
    }
    AHK_ON_DEVICE(e->cfg.device);
    AHK_CUDA(cudaMemcpy(e->table_dev, chw.data(), chw.size(), cudaMemcpyHostToDevice));
    e->table_set = true;
    return AHK_OK;
}

int ahk_bind_buffers(ahk_env *e, const ahk_buffers *b)
{
    if (!e || !b) return fail(AHK_ERR_ARG, "env/bufs is NULL");
    if (!b->state || !b->frames || !b->reward || !b->done || !b->robot_label || !b->human_label || !b->info || !b->drawlist)
        return fail(AHK_ERR_ARG, "every buffer except stats and warm must be non-NULL");
    if (e->depth > 3 && !b->warm) return fail(AHK_ERR_ARG, "ring_depth %d > 3 needs the warm buffer [3][n][3][128][128]", e->depth);
    if ((reinterpret_cast<uintptr_t>(b->frames) | reinterpret_cast<uintptr_t>(b->drawlist) | reinterpret_cast<uintptr_t>(b->warm)) & 15u)
        return fail(AHK_ERR_ARG, "frames, warm and drawlist must be 16-byte aligned");
    e->buf = *b;
    e->bound = true;
    return AHK_OK;
}

int ahk_reset(ahk_env *e, const uint8_t *mask, const uint8_t *episode, const uint8_t *use, void *stream)
{
    if (int rc = check_ready(e)) return rc;
    AHK_ON_DEVICE(e->cfg.device);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    ahk::EnvArgs a = env_args(e);
    a.ext_inputs = 1;
    a.reset_mask = mask;
    a.episode_override = episode;
    a.use_override = use;
    return run(e, a, 0, s);
}

int ahk_step(ahk_env *e, const uint8_t *robot, const uint8_t *human, void *stream)
{
    if (int rc = check_ready(e)) return rc;
    AHK_ON_DEVICE(e->cfg.device);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    ahk::EnvArgs a = env_args(e);
    a.robot = robot;
    a.human = human;
    a.ext_inputs = robot != nullptr || human != nullptr;
    e->head = (e->head + 1) % e->depth;
    return run(e, a, 1, s);
}

int ahk_step_many(ahk_env *e, int steps, const ahk_step_log *log, void *stream)
{
    if (int rc = check_ready(e)) return rc;
    if (steps < 0) return fail(AHK_ERR_ARG, "steps must be >= 0, got %d", steps);
    if (steps == 0) return AHK_OK;
    AHK_ON_DEVICE(e->cfg.device);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const size_t n = static_cast<size_t>(e->cfg.n_envs);
    const bool logging = log && (log->robot_label || log->human_label || log->done || log->reward);
    // The persistent kernel keeps at least two record buffers of 3 * kManyEnvs records per CTA: it takes the call when no CTA can own
    // more envs than that (the adaptive partition moves shares within +-50 % of even).  Larger batches — where a launch
    // is long and its fixed cost small — go step by step.  A/B switch AHK_MANY=loop.
    const int even = (e->cfg.n_envs + e->step_grid - 1) / e->step_grid;
    const bool fits = (e->balance ? even + (even + 1) / 2 + 1 : e->envs_per_cta) <= ahk::kManyEnvs;
    if (fits && !e->split && e->many_kernel) {
        if (e->host_early_ticket >= 0 && s != e->host_stream) AHK_CUDA(cudaStreamWaitEvent(s, e->host_ev[e->host_early_ticket], 0));
        ahk::EnvArgs a = env_args(e);
        ahk::StepArgs st = step_args(e, a, 3, s);
        st.pdl = 0;
        st.wait_first = 0;
        st.ras.head = (e->head + 1) % e->depth; // slot of the window's first step
        ahk::StepLogArgs lg{};
        if (logging) { lg.robot_label = log->robot_label; lg.human_label = log->human_label; lg.done = log->done; lg.reward = log->reward; }
        AHK_CUDA(ahk::launch_step_many(st, steps, lg, e->step_grid, s));
        e->head = (e->head + steps) % e->depth;
        e->launches += ahk::many_launches(steps);
        return AHK_OK;
    }
    for (int t = 0; t < steps; ++t) {
        ahk::EnvArgs a = env_args(e);
        e->head = (e->head + 1) % e->depth;
        if (int rc = run(e, a, 1, s)) return rc;
        if (logging) {
            if (log->robot_label) AHK_CUDA(cudaMemcpyAsync(log->robot_label + t * n, e->buf.robot_label, n, cudaMemcpyDeviceToDevice, s));
            if (log->human_label) AHK_CUDA(cudaMemcpyAsync(log->human_label + t * n, e->buf.human_label, n, cudaMemcpyDeviceToDevice, s));
            if (log->done) AHK_CUDA(cudaMemcpyAsync(log->done + t * n, e->buf.done, n, cudaMemcpyDeviceToDevice, s));
            if (log->reward) AHK_CUDA(cudaMemcpyAsync(log->reward + t * n, e->buf.reward, n * sizeof(float), cudaMemcpyDeviceToDevice, s));
        }
    }
    return AHK_OK;
}

int ahk_step_profiled(ahk_env *e, const uint8_t *robot, const uint8_t *human, void *stream, float *env_ms, float *raster_ms)
{
    if (int rc = check_ready(e)) return rc;
    if (!env_ms || !raster_ms) return fail(AHK_ERR_ARG, "env_ms/raster_ms is NULL");
    AHK_ON_DEVICE(e->cfg.device);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    ahk::EnvArgs a = env_args(e);
    a.robot = robot;
    a.human = human;
    a.ext_inputs = robot != nullptr || human != nullptr;
    e->head = (e->head + 1) % e->depth;
    AHK_CUDA(cudaEventRecord(e->ev[0], s));
    if (!e->split) { // the fused kernel is the whole step: reported as the second figure, the first is 0
        if (int rc = run(e, a, 1, s)) return rc;
        AHK_CUDA(cudaEventRecord(e->ev[2], s));
        AHK_CUDA(cudaEventSynchronize(e->ev[2]));
        *env_ms = 0.0f;
        AHK_CUDA(cudaEventElapsedTime(raster_ms, e->ev[0], e->ev[2]));
        return AHK_OK;
    }
    ahk::StepArgs st = step_args(e, a, 1, s);
    AHK_CUDA(ahk::launch_env(a, true, s));
    AHK_CUDA(cudaEventRecord(e->ev[1], s));
    if (e->staged) AHK_CUDA(ahk::launch_raster_staged(st.ras, e->raster_grid, s));
    else AHK_CUDA(ahk::launch_step(st, 2, e->step_grid, s));
    e->launches += 2;
    AHK_CUDA(cudaEventRecord(e->ev[2], s));
    AHK_CUDA(cudaEventSynchronize(e->ev[2]));
    AHK_CUDA(cudaEventElapsedTime(env_ms, e->ev[0], e->ev[1]));
    AHK_CUDA(cudaEventElapsedTime(raster_ms, e->ev[1], e->ev[2]));
    return AHK_OK;
}

// early != nullptr: ask the kernel to signal the end of its game ticks (ahk_step_host_results); *early = the sequence
// number to wait for, or 0 when some output buffer is not mapped and the caller has to wait for the whole step.
static int step_host_enqueue(ahk_env *e, const uint8_t *robot, const uint8_t *human, float *reward, uint8_t *done,
                             uint8_t *robot_label, uint8_t *human_label, uint8_t *frame, uint64_t *early)
{
    if (int rc = check_ready(e)) return rc;
    AHK_ON_DEVICE(e->cfg.device);
    cudaStream_t s = e->host_stream;
    const size_t n = static_cast<size_t>(e->cfg.n_envs);
    // Pinned buffers are read and written by the step kernel itself (mapped host memory, 2 + 7 bytes per env over PCIe):
    // the whole call is ONE kernel launch.  Pageable buffers go through the copy engine.
    const bool zc = !e->split && e->zero_copy;
    const uint8_t *robot_dev = zc ? mapped(robot) : nullptr, *human_dev = zc ? mapped(human) : nullptr;
    if (robot && !robot_dev) { robot_dev = e->action_stage; AHK_CUDA(cudaMemcpyAsync(e->action_stage, robot, n, cudaMemcpyHostToDevice, s)); }
    if (human && !human_dev) { human_dev = e->action_stage + n; AHK_CUDA(cudaMemcpyAsync(e->action_stage + n, human, n, cudaMemcpyHostToDevice, s)); }
    ahk::EnvArgs a = env_args(e);
    a.robot = robot_dev;
    a.human = human_dev;
    if (zc) {
        a.h_reward = mapped(reward);
        a.h_done = mapped(done);
        a.h_robot_label = mapped(robot_label);
        a.h_human_label = mapped(human_label);
    }
    if (early) {
        const bool all_mapped = zc && (!reward || a.h_reward) && (!done || a.h_done) && (!robot_label || a.h_robot_label) &&
                                (!human_label || a.h_human_label) && !frame;
        *early = 0;
        if (all_mapped) {
            a.h_tick_flag = e->tick_flag_dev;
            a.tick_count = e->tick_count_dev;
            a.tick_seq = *early = ++e->tick_seq;
        }
    }
    e->head = (e->head + 1) % e->depth;
    if (int rc = run(e, a, 1, s)) return rc;
    if (reward && !a.h_reward) AHK_CUDA(cudaMemcpyAsync(reward, e->buf.reward, n * sizeof(float), cudaMemcpyDeviceToHost, s));
    if (done && !a.h_done) AHK_CUDA(cudaMemcpyAsync(done, e->buf.done, n, cudaMemcpyDeviceToHost, s));
    if (robot_label && !a.h_robot_label) AHK_CUDA(cudaMemcpyAsync(robot_label, e->buf.robot_label, n, cudaMemcpyDeviceToHost, s));
    if (human_label && !a.h_human_label) AHK_CUDA(cudaMemcpyAsync(human_label, e->buf.human_label, n, cudaMemcpyDeviceToHost, s));
    if (frame)
        AHK_CUDA(cudaMemcpyAsync(frame, e->buf.frames + static_cast<size_t>(e->head) * n * ahk::kFrameBytes,
                                 n * ahk::kFrameBytes, cudaMemcpyDeviceToHost, s));
    const int ticket = static_cast<int>(e->host_ticket++ % ahk_env::kHostDepth);
    AHK_CUDA(cudaEventRecord(e->host_ev[ticket], s));
    return ticket;
}

int ahk_step_host_begin(ahk_env *e, const uint8_t *robot, const uint8_t *human, float *reward, uint8_t *done,
                        uint8_t *robot_label, uint8_t *human_label, uint8_t *frame)
{
    return step_host_enqueue(e, robot, human, reward, done, robot_label, human_label, frame, nullptr);
}

int ahk_step_host_results(ahk_env *e, const uint8_t *robot, const uint8_t *human, float *reward, uint8_t *done,
                          uint8_t *robot_label, uint8_t *human_label)
{
    uint64_t seq = 0;
    const int ticket = step_host_enqueue(e, robot, human, reward, done, robot_label, human_label, nullptr, &seq);
    if (ticket < 0) return ticket;
    e->host_early_ticket = ticket;
    if (seq == 0) return ahk_step_host_wait(e, ticket); // pageable buffers: results arrive with the copies, after the kernel
    const volatile unsigned long long *flag = e->tick_flag_host + (seq & 7u);
    for (unsigned spins = 1; *flag != seq; ++spins) {
        if ((spins & 0xFFFu) == 0u) { // a failed launch never signals: fall back to the event every 4096 polls
            const cudaError_t st = cudaEventQuery(e->host_ev[ticket]);
            if (st == cudaSuccess) break;
            if (st != cudaErrorNotReady) return fail(AHK_ERR_CUDA, "ahk_step_host_results: %s", cudaGetErrorString(st));
        }
#if defined(__x86_64__)
        __builtin_ia32_pause();
#endif
    }
    return AHK_OK;
}

int ahk_host_sync(ahk_env *e)
{
    if (!e) return fail(AHK_ERR_ARG, "env is NULL");
    AHK_ON_DEVICE(e->cfg.device);
    AHK_CUDA(cudaStreamSynchronize(e->host_stream));
    e->host_early_ticket = -1;
    return AHK_OK;
}

int ahk_step_host_wait(ahk_env *e, int ticket)
{
    if (!e) return fail(AHK_ERR_ARG, "env is NULL");
    if (ticket < 0 || ticket >= ahk_env::kHostDepth) return fail(AHK_ERR_ARG, "ticket %d out of range", ticket);
    AHK_ON_DEVICE(e->cfg.device);
    AHK_CUDA(cudaEventSynchronize(e->host_ev[ticket]));
    if (ticket == e->host_early_ticket) e->host_early_ticket = -1;
    return AHK_OK;
}

int ahk_step_host(ahk_env *e, const uint8_t *robot, const uint8_t *human, float *reward, uint8_t *done,
                  uint8_t *robot_label, uint8_t *human_label, uint8_t *frame)
{
    const int ticket = ahk_step_host_begin(e, robot, human, reward, done, robot_label, human_label, frame);
    if (ticket < 0) return ticket;
    return ahk_step_host_wait(e, ticket);
}

int ahk_observe(ahk_env *e, int format, void *out, void *stream)
{
    if (int rc = check_ready(e)) return rc;
    if (!out) return fail(AHK_ERR_ARG, "out is NULL");
    if (format < AHK_OBS_U8 || format > AHK_OBS_GRAY_DIFF_F64) return fail(AHK_ERR_ARG, "Processing mode is not supported");
    AHK_ON_DEVICE(e->cfg.device);
    if (e->host_early_ticket >= 0 && static_cast<cudaStream_t>(stream) != e->host_stream)
        AHK_CUDA(cudaStreamWaitEvent(static_cast<cudaStream_t>(stream), e->host_ev[e->host_early_ticket], 0));
    ahk::ObserveArgs a{};
    a.frames = e->buf.frames;
    a.warm = e->depth > 3 ? e->buf.warm : nullptr;
    a.meta = e->buf.state + static_cast<size_t>(15) * e->cfg.n_envs;
    a.out = out;
    a.n = e->cfg.n_envs;
    a.head = e->head;
    a.format = format;
    a.depth = e->depth;
    const size_t work = static_cast<size_t>(e->cfg.n_envs) * (format == AHK_OBS_GRAY_DIFF_F64 ? 1024 : 3 * ahk::kFrameVecs);
    size_t grid = (work + 255) / 256;
    if (grid > static_cast<size_t>(e->raster_grid) * 16) grid = static_cast<size_t>(e->raster_grid) * 16;
    AHK_CUDA(ahk::launch_observe(a, static_cast<int>(grid), static_cast<cudaStream_t>(stream)));
    e->launches += 1;
    return AHK_OK;
}

int ahk_head(const ahk_env *e) { return e ? e->head : fail(AHK_ERR_ARG, "env is NULL"); }
int ahk_ring_depth(const ahk_env *e) { return e ? e->depth : fail(AHK_ERR_ARG, "env is NULL"); }

int ahk_stack_view(const ahk_env *e, ahk_stack *out)
{
    if (!e || !out) return fail(AHK_ERR_ARG, "env/out is NULL");
    if (!e->bound) return fail(AHK_ERR_STATE, "ahk_bind_buffers has not been called");
    out->frames = e->buf.frames;
    out->warm = e->depth > 3 ? e->buf.warm : nullptr;
    out->meta = e->buf.state + static_cast<size_t>(15) * e->cfg.n_envs;
    out->n = e->cfg.n_envs;
    out->depth = e->depth;
    out->head = e->head;
    return AHK_OK;
}

int ahk_set_head(ahk_env *e, int head)
{
    if (!e || head < 0 || head >= e->depth) return fail(AHK_ERR_ARG, "head must be in [0, ring_depth)");
    e->head = head;
    return AHK_OK;
}

int ahk_set_pending(ahk_env *e, const uint8_t *pending, void *stream)
{
    if (int rc = check_ready(e)) return rc;
    if (!pending) return fail(AHK_ERR_ARG, "pending is NULL");
    AHK_ON_DEVICE(e->cfg.device);
    AHK_CUDA(ahk::launch_set_pending(e->buf.state, pending, e->cfg.n_envs, static_cast<cudaStream_t>(stream)));
    e->last_mode = -1; // the state was written by another kernel: the next step waits before it loads it
    e->launches += 1;
    return AHK_OK;
}

int ahk_set_trace(ahk_env *e, uint64_t *trace_dev)
{
    if (!e) return fail(AHK_ERR_ARG, "env is NULL");
    e->trace = reinterpret_cast<unsigned long long *>(trace_dev);
    return AHK_OK;
}

int ahk_trace_shape(const ahk_env *e, int *ctas, int *warps)
{
    if (!e || !ctas || !warps) return fail(AHK_ERR_ARG, "env/ctas/warps is NULL");
    *ctas = e->step_grid;
    *warps = ahk::kStepThreads / 32;
    return AHK_OK;
}

int ahk_get_partition(ahk_env *e, int32_t *bounds_host, int capacity)
{
    if (!e || !bounds_host) return fail(AHK_ERR_ARG, "env/bounds_host is NULL");
    const int g = e->step_grid;
    if (capacity < g + 1) return fail(AHK_ERR_ARG, "capacity %d < %d", capacity, g + 1);
    if (!e->balance) {
        for (int b = 0; b <= g; ++b) bounds_host[b] = b * e->envs_per_cta < e->cfg.n_envs ? b * e->envs_per_cta : e->cfg.n_envs;
        return g;
    }
    AHK_ON_DEVICE(e->cfg.device);
    AHK_CUDA(cudaDeviceSynchronize());
    AHK_CUDA(cudaMemcpy(bounds_host, e->bounds_dev[e->bounds_cur], (g + 1) * sizeof(int), cudaMemcpyDeviceToHost));
    return g;
}

int ahk_launch_count(const ahk_env *e, uint64_t *kernels)
{
    if (!e || !kernels) return fail(AHK_ERR_ARG, "env/kernels is NULL");
    *kernels = e->launches;
    return AHK_OK;
}

} // extern "C"
