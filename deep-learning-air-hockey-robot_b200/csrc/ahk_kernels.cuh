// ahk_kernels.cuh — kernel argument blocks and launchers shared by ahk_kernels.cu and ahk_api.cu.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace ahk {

constexpr int kTableW = 450, kTableH = 700; // native frame (SPEC.md §2)
constexpr int kDim = 128;                   // processor.py:5
constexpr int kFrameBytes = 3 * kDim * kDim;
constexpr int kFrameVecs = kFrameBytes / 16; // 3072 128-bit vectors per frame
constexpr int kEnvThreads = 128;
constexpr int kStepThreads = 1024;                                             // fused step kernel: 32 warps, one CTA per SM
constexpr int kManyEnvs = 512;                                                 // ahk_step_many's persistent kernel: envs per CTA (two record buffers)
constexpr int kRasterThreads = 256;
constexpr int kRasterGroups = 3;                                               // independent 256-thread groups per CTA
constexpr int kRasterGroupTail = kDim * 8 + kRasterThreads * 20 + 16;          // per-row spans + record batch + counter
constexpr int kRasterStagedSmem = (1 + kRasterGroups) * kFrameBytes + kRasterGroups * kRasterGroupTail; // table + stagings

struct EnvArgs {
    uint32_t *state;   // [16][n]
    uint32_t *drawlist; // [3][n][4]
    float *reward;
    uint8_t *done, *robot_label, *human_label, *info;
    uint64_t *stats;
    const uint8_t *robot, *human;                          // step: action labels or nullptr
    // ahk_step_host with pinned host buffers: zero-copy mirrors of reward/done/labels in mapped host memory, written by
    // the kernel itself over PCIe next to the device copies (nullptr: none) — no copy engine on the e2e path
    float *h_reward;
    uint8_t *h_done, *h_robot_label, *h_human_label;
    // ahk_step_host_results: once every CTA has ticked, the last one stores tick_seq into the mapped host word
    // h_tick_flag[tick_seq & 7]; tick_count[tick_seq & 7] counts the CTAs (device memory, left at 0 again)
    unsigned long long *h_tick_flag;
    unsigned *tick_count;
    unsigned long long tick_seq;
    const uint8_t *reset_mask, *episode_override, *use_override; // reset
    int ext_inputs; // robot / human / reset inputs may have been produced by earlier work on the stream (device pointers)
    uint2 key;
    uint32_t env_id_base;
    int n, auto_reset, max_episode_steps;
};

struct RasterArgs {
    const uint32_t *drawlist;
    const uint4 *table; // [3][128][8] vectors: the table image at the 128x128 sample points, CHW
    uint8_t *frames;    // [depth][n][3][128][128]
    uint8_t *warm;      // deep ring (depth > 3) only: [3][n][3][128][128] warm-up frames of each env's latest reset; else nullptr
    uint2 key;
    uint32_t env_id_base;
    int n, head, depth;
};

// per-step outputs kept by ahk_step_many: each [steps][n] or nullptr
struct StepLogArgs {
    uint8_t *robot_label, *human_label, *done;
    float *reward;
};

// Where frame `age` (0 = newest) of env's 3-frame stack lives.  3-slot ring: slot (head - age) mod 3, always.  Deep ring:
// a reset leaves its three warm-up frames (air_hockey_env.py:69-72) in the side buffer `warm` instead of rewriting older
// ring slots (which hold the previous episode's last frames — replay memory), so a frame older than the env's episode
// (age >= episode_steps, the low 16 bits of the meta word) is warm-up frame 3 - (age - episode_steps), stored at index
// 2 - (age - episode_steps).
__device__ __forceinline__ const uint8_t *stack_frame(const uint8_t *frames, const uint8_t *warm, const uint32_t *meta, int depth,
                                                      size_t n, int head, size_t env, int age)
{
    if (warm != nullptr) {
        const int steps = static_cast<int>(meta[env] & 0xFFFFu);
        if (age >= steps) return warm + (static_cast<size_t>(2 - (age - steps)) * n + env) * kFrameBytes;
    }
    return frames + (static_cast<size_t>((head + depth - age) % depth) * n + env) * kFrameBytes;
}

struct StepArgs { // the fused kernel: env block b of envs_per_cta environments belongs to CTA b
    EnvArgs env;
    RasterArgs ras;
    int envs_per_cta;
    // Adaptive partition (nullptr: CTA b owns [b * envs_per_cta, ...)).  SMs do not drain their stores equally fast (their
    // place in the memory system), and the difference is stable from launch to launch: every CTA stamps its end time,
    // the last CTA to finish moves the env boundaries against each CTA's lateness (ahk_step.cu, rebalance).
    const int *bounds;       // [grid + 1] env boundaries for THIS launch
    int *bounds_next;        // [grid + 1] written by the last CTA (nullptr: do not adapt)
    unsigned long long *end_ns; // [grid] %globaltimer at each CTA's end, this launch
    unsigned *done_count;    // CTAs finished in this launch
    int dyn_units;           // 1: warps claim raster units from a shared-memory counter instead of round-robin (default; AHK_UNITS=static for A/B)
    int wait_first;          // griddepcontrol.wait before the first load of caller-produced inputs (actions, masks, overrides)
    int pdl;                 // launched with programmatic stream serialization: griddepcontrol.* are live; 1: wait before the first frame store, 3: wait at the end (2: experiment, no wait)
    unsigned long long *trace; // measurement aid (ahk_set_trace): [grid][32 warps][4] %globaltimer stamps; selects the traced kernel
};

struct ObserveArgs {
    const uint8_t *frames;
    const uint8_t *warm;   // deep ring only
    const uint32_t *meta;  // state word 15 of every env (deep ring only)
    void *out;
    int n, head, format, depth;
};

void set_error_text(const char *text); // ahk_api.cu: fills the buffer behind ahk_last_error()

cudaError_t launch_env(const EnvArgs &a, bool step, cudaStream_t stream);
cudaError_t step_prepare();
cudaError_t launch_step(const StepArgs &a, int mode, int grid, cudaStream_t stream); // mode 0 reset, 1 step, 2 draw only
cudaError_t launch_step_many(const StepArgs &a, int steps, const StepLogArgs &log, int grid, cudaStream_t stream);
int many_launches(int steps); // kernel launches launch_step_many makes for a window of `steps` steps
cudaError_t raster_staged_prepare();
cudaError_t launch_raster_staged(const RasterArgs &a, int grid, cudaStream_t stream);
cudaError_t launch_observe(const ObserveArgs &a, int grid, cudaStream_t stream);
cudaError_t launch_set_pending(uint32_t *state, const uint8_t *pending, int n, cudaStream_t stream);

} // namespace ahk
