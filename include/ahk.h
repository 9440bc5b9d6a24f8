/*
 * ahk.h — C ABI of the B200 batched air-hockey environment engine (libahk.so).
 *
 * The reference has no native boundary on this path: `ai/supervised/generate_data.py:35-51` and
 * `ai/reinforcement/train.py:57-96` call the pure-Python gym wrapper
 * `gym-air-hockey/gym_air_hockey/envs/air_hockey_env.py`, which calls the (missing) pure-Python simulator.
 * Each entry point below names the reference interface it replaces; INTEGRATION.md shows the ctypes stub a
 * maintainer of the reference would add.  Plain pointers and sizes only; no exceptions cross this boundary.
 *
 * Conventions
 *   - every function returns AHK_OK (0) or a negative ahk_status; ahk_last_error() gives the text (thread-local).
 *   - "dev" pointers are CUDA device pointers owned by the caller (PyTorch allocates them), "host" pointers are
 *     ordinary (ideally pinned) host memory.  The engine allocates nothing on the step path.
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).  Calls are stream-ordered and
 *     asynchronous unless the name ends in _host.
 *   - action labels are the reference's 0..9 (`processor.py:7-18`); AHK_ACT_NONE (255) = None = the programmed AI
 *     plays that side (`air_hockey_env.py:86-89`, `processor.py:74-76`).
 */
#ifndef AHK_H
#define AHK_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AHK_DIM 128                       /* processor.py:5 */
#define AHK_FRAME_BYTES (3 * 128 * 128)   /* one CHW uint8 frame */
#define AHK_TABLE_W 450
#define AHK_TABLE_H 700
#define AHK_STATE_WORDS 16                /* SoA words per env, see DESIGN.md "HBM layout" */
#define AHK_N_ACTIONS 10                  /* air_hockey_env.py:16 */
#define AHK_ACT_NONE 255
#define AHK_STATS_WORDS 32

typedef enum {
    AHK_OK = 0,
    AHK_ERR_ARG = -1,      /* bad argument (NULL, size, mode) */
    AHK_ERR_CUDA = -2,     /* a CUDA call failed; text in ahk_last_error() */
    AHK_ERR_STATE = -3,    /* call order: buffers not bound / table not set */
    AHK_ERR_NO_DEVICE = -4 /* no CUDA device: there is NO CPU fallback */
} ahk_status;

enum { AHK_EP_NORMAL = 0, AHK_EP_MISSING = 1, AHK_EP_RANDOM = 2, AHK_EP_SAMPLE = 255 }; /* air_hockey_env.py:28 */
enum { AHK_USE_ARM = 1, AHK_USE_PUCK = 2, AHK_USE_TOP_MALLET = 4, AHK_USE_ALL = 7 };     /* air_hockey_env.py:32-34 */
enum { AHK_OBS_U8 = 0, AHK_OBS_RGB_F32 = 1, AHK_OBS_GRAY_DIFF_F64 = 2 };                /* processor.py:52,53-64 */

typedef struct ahk_env ahk_env; /* opaque; one per GPU (one per process) */

typedef struct {
    int32_t n_envs;            /* environments on THIS device */
    int32_t device;            /* CUDA ordinal */
    uint64_t seed;             /* master seed (SPEC.md §4) */
    uint32_t env_id_base;      /* global id of local env 0: results are independent of sharding */
    int32_t auto_reset;        /* 1: reset an env at the start of the step after a terminal / max_episode_steps */
    int32_t max_episode_steps; /* 0 = unlimited; generate_data.py:57 uses 200 */
    int32_t ring_depth;        /* frame ring slots: 0 or 3 = the 3-frame stack of processor.py:23; > 3 = a DEEP ring that doubles as
                                  replay memory / dataset staging (rl_agents.py:59-90: experience_buffer): every frame is written
                                  to HBM once, slot = step mod ring_depth, and a reset files its warm-up frames in ahk_buffers.warm
                                  instead of rewriting older slots */
    int32_t reserved[4];
} ahk_config;

/* Device memory the engine works in; all caller-owned (SURVEY.md §8b "Ownership"). */
typedef struct {
    uint32_t *state;      /* [AHK_STATE_WORDS][n] SoA: f32 puck x,y,vx,vy | bottom | top | serve,reset,noise counters | meta */
    uint8_t *frames;      /* [ring_depth][n][3][128][128] ring of CHW uint8 frames; slot `head` is the newest */
    float *reward;        /* [n]  air_hockey_env.py:95-145 */
    uint8_t *done;        /* [n]  terminal, air_hockey_env.py:90,147-148 */
    uint8_t *robot_label; /* [n]  info['robot_action'], air_hockey_env.py:92,155 */
    uint8_t *human_label; /* [n]  info['human_action'] */
    uint8_t *info;        /* [n]  bit0 puck_was_hit 1 hit_the_border 2 puck_is_at_the_bottom 3 distance_decreased 4 in_the_target, bits5-6 scored (0 none,1 top,2 bottom) */
    uint64_t *stats;      /* [AHK_STATS_WORDS] running sums: 0 steps, 1 episodes, 2 goals_top, 3 goals_bottom, 4 reward*2 (int64), 5.. robot label hist[10], 15.. human label hist[10] */
    uint32_t *drawlist;   /* [3][n][4] scratch between the env and raster kernels (split path only; the fused kernel keeps it in shared memory) */
    uint8_t *warm;        /* ring_depth > 3 only (else NULL): [3][n][3][128][128] the warm-up frames w1, w2, w3 of each env's latest reset
                             (air_hockey_env.py:69-72).  Frame `age` of an env's stack is ring slot (head - age) mod ring_depth when
                             age < episode_steps (state word 15, low 16 bits), else warm[2 - (age - episode_steps)] */
} ahk_buffers;

/* ≙ AirHockeyEnv.__init__ (air_hockey_env.py:9-24) minus the first reset. */
int ahk_create(const ahk_config *cfg, ahk_env **out);
void ahk_destroy(ahk_env *env);

/* ≙ the table sprite the missing simulator loaded.  rgb: host uint8[700][450][3].  The engine keeps the 128x128
 * NEAREST samples (processor.py:34-35) in device memory and stages them in shared memory per CTA. */
int ahk_set_table_image(ahk_env *env, const uint8_t *rgb_host, int height, int width);

int ahk_bind_buffers(ahk_env *env, const ahk_buffers *bufs);

/* ≙ AirHockeyEnv.reset (air_hockey_env.py:26-76), batched.  mask_dev [n] (NULL = every env): nonzero = reset.
 * episode_dev/use_dev [n] (NULL = sample per SPEC.md §4): force the episode type / object mask, AHK_EP_SAMPLE = sample. */
int ahk_reset(ahk_env *env, const uint8_t *mask_dev, const uint8_t *episode_dev, const uint8_t *use_dev, void *stream);

/* ≙ AirHockeyEnv.step (air_hockey_env.py:84-155), batched, ONE kernel: physics + programmed AI + reward/terminal/labels,
 * then the new frame of every env rasterised into ring slot `head` (advanced first).  robot_dev/human_dev [n] labels or NULL. */
int ahk_step(ahk_env *env, const uint8_t *robot_dev, const uint8_t *human_dev, void *stream);

/* T consecutive ACTION-FREE steps in one call — the data-generation loop, where the programmed AI plays both sides and
 * nothing outside the engine feeds a step (generate_data.py:44-58: `env.step()` with no arguments, reset on a goal or at
 * the step limit = auto_reset + max_episode_steps).  Same results as `steps` calls of ahk_step(env, NULL, NULL, stream);
 * the outputs of every step, which the loop appends to its dataset (generate_data.py:53-55), are kept when `log_dev` is
 * given: each member is a device array [steps][n] or NULL. */
typedef struct {
    uint8_t *robot_label; /* info['robot_action'] per step */
    uint8_t *human_label; /* info['human_action'] per step */
    uint8_t *done;        /* terminal per step */
    float *reward;        /* reward per step */
} ahk_step_log;
int ahk_step_many(ahk_env *env, int steps, const ahk_step_log *log_dev, void *stream);

/* The same call with HOST buffers, for callers that live on the CPU like the reference's loops do
 * (generate_data.py:44-58, reinforcement/train.py:87-96): actions in, step, reward/done/labels (and, if
 * frame_host != NULL, the new uint8 frame [n][3][128][128]) out, then waits.  Any pointer may be NULL.
 * PINNED buffers (cudaHostAlloc / cudaHostRegister, e.g. torch's pin_memory) are read and written by the step kernel
 * itself through their device mapping — 2 bytes in, 7 bytes out per env over PCIe, ONE kernel launch, no copy-engine
 * transfers; pageable buffers go through cudaMemcpyAsync.  The frame, when asked for, is always a copy (49 152 B/env).
 * Stream rule for all ahk_step_host* calls: they run on a stream the engine owns.  Work the engine itself enqueues later on
 * another stream (ahk_step / ahk_reset / ahk_observe) is ordered behind them; work enqueued EARLIER on another stream is
 * not waited for — synchronise that stream before switching an env from the device path to the host path. */
int ahk_step_host(ahk_env *env, const uint8_t *robot_host, const uint8_t *human_host, float *reward_host,
                  uint8_t *done_host, uint8_t *robot_label_host, uint8_t *human_label_host, uint8_t *frame_host);

/* ahk_step_host split in two so that a caller whose next actions do not depend on this step's results — the data
 * generation loop, where the programmed AI plays both sides (generate_data.py:51) — can keep up to 8 steps in flight
 * (each with its own output buffers): _begin enqueues and returns a ticket (>= 0) or a negative status, _wait blocks
 * until that step's results are in the host buffers.  ahk_step_host == _begin + _wait. */
int ahk_step_host_begin(ahk_env *env, const uint8_t *robot_host, const uint8_t *human_host, float *reward_host,
                        uint8_t *done_host, uint8_t *robot_label_host, uint8_t *human_label_host, uint8_t *frame_host);
int ahk_step_host_wait(ahk_env *env, int ticket);

/* ahk_step_host for a loop that needs each step's RESULTS before it chooses the next actions (reinforcement/train.py:87-96,
 * keras/test.py:42-51) but leaves the frames on the device: returns as soon as reward / done / labels of this step are in
 * the host buffers — the step kernel signals the end of its game ticks through mapped host memory — while the frames are
 * still being rasterised on the engine's own stream.  The next ahk_step_host* call is ordered behind it as usual;
 * ahk_host_sync waits for everything enqueued by the _host calls (call it before touching the ring from another stream).
 * Needs pinned output buffers; with pageable ones it behaves like ahk_step_host. */
int ahk_step_host_results(ahk_env *env, const uint8_t *robot_host, const uint8_t *human_host, float *reward_host,
                          uint8_t *done_host, uint8_t *robot_label_host, uint8_t *human_label_host);
int ahk_host_sync(ahk_env *env);

/* ahk_step with CUDA events recorded on `stream` around its kernel(s); waits, then reports the durations.  Default
 * (fused) path: *env_ms = 0 and *raster_ms = the one fused kernel (physics + AI + labels + raster).  With the A/B switch
 * AHK_PATH=split in the environment: the env kernel and the raster kernel separately.  Measurement aid for bench.py's
 * roofline (not a hot-path call). */
int ahk_step_profiled(ahk_env *env, const uint8_t *robot_dev, const uint8_t *human_dev, void *stream, float *env_ms,
                      float *raster_ms);

/* ≙ DataProcessor.process_observation's return value (processor.py:52-64) materialised from the ring:
 * AHK_OBS_U8 uint8[n][9][128][128] newest first; AHK_OBS_RGB_F32 float[n][9][128][128] = (x-128)/128;
 * AHK_OBS_GRAY_DIFF_F64 double[n][1][128][128]. */
int ahk_observe(ahk_env *env, int format, void *out_dev, void *stream);

int ahk_head(const ahk_env *env);                 /* ring slot holding the newest frame */
int ahk_ring_depth(const ahk_env *env);
/* Everything a consumer needs to find the 3-frame stack of every env (see ahk_buffers.warm): device pointers + head. */
typedef struct {
    const uint8_t *frames;  /* [depth][n][3][128][128] */
    const uint8_t *warm;    /* NULL for the 3-slot ring */
    const uint32_t *meta;   /* [n] state word 15 (episode_steps in the low 16 bits) */
    int32_t n, depth, head;
} ahk_stack;
int ahk_stack_view(const ahk_env *env, ahk_stack *out);
int ahk_set_head(ahk_env *env, int head);
int ahk_set_pending(ahk_env *env, const uint8_t *pending_dev, void *stream); /* arm/disarm the auto-reset per env */
int ahk_launch_count(const ahk_env *env, uint64_t *kernels); /* kernels launched so far by this handle */

/* Measurement aid (tools/phase_trace.py -> profiles/): when trace_dev != NULL the fused step kernel stores, per CTA and
 * warp, four %globaltimer stamps (ns) — CTA start, game tick done, records + table staged, last store issued — into
 * trace_dev[2][ctas][warps][4] (slot = launch count & 1, so consecutive overlapped launches can be compared);
 * ahk_trace_shape gives ctas and warps.  NULL switches it off (the default). */
int ahk_set_trace(ahk_env *env, uint64_t *trace_dev);
int ahk_trace_shape(const ahk_env *env, int *ctas, int *warps);
/* The env boundaries of the step kernel's CTAs right now (adaptive partition): bounds_host[0..ctas], returns
 * ctas or a negative status.  Synchronises the device.  Diagnostics and tests only. */
int ahk_get_partition(ahk_env *env, int32_t *bounds_host, int capacity);

const char *ahk_last_error(void);
const char *ahk_version(void);

/* ---- CNN policy forward (ai/supervised/keras/models.py:20-65, the 'rgb' network; rl_agents.py:78-98) ---- */
typedef struct {
    /* BatchNorm (axis=1, eps 1e-3, AFTER the ReLU: models.py:26-29) is passed raw; the engine folds it. All float32, Keras layouts:
     * conv kernels [kh][kw][cin][cout], dense [in][out]. */
    const float *conv1_w, *conv1_b, *bn1_gamma, *bn1_beta, *bn1_mean, *bn1_var; /* 8x8x9x5   */
    const float *conv2_w, *conv2_b, *bn2_gamma, *bn2_beta, *bn2_mean, *bn2_var; /* 4x4x5x32  */
    const float *conv3_w, *conv3_b, *bn3_gamma, *bn3_beta, *bn3_mean, *bn3_var; /* 4x4x32x64 */
    const float *dense1_w, *dense1_b; /* 4096x512 */
    const float *dense2_w, *dense2_b; /* 512x512  */
    const float *out_w, *out_b;       /* 512x10   */
} ahk_policy_weights; /* host pointers */

typedef struct ahk_policy ahk_policy;
int ahk_policy_create(int device, int max_batch, const ahk_policy_weights *w, ahk_policy **out);
void ahk_policy_destroy(ahk_policy *p);
/* frames_dev: the ring [3][n][3][128][128]; head: newest slot.  logits_dev float[n][10] (pre-softmax, i.e. the DDQN
 * linear head of rl_agents.py:160-172; softmax of it is the supervised model's output), action_dev uint8[n] argmax
 * (either may be NULL). */
int ahk_policy_forward(ahk_policy *p, const uint8_t *frames_dev, int n, int head, float *logits_dev,
                       uint8_t *action_dev, void *stream);
/* The same forward pass on an engine's current observation, whatever its ring depth (ahk_stack_view): the act() of the RL
 * loop (rl_agents.py:92-101) reading the deep ring that doubles as replay memory. */
int ahk_policy_forward_stack(ahk_policy *p, const ahk_stack *stack, float *logits_dev, uint8_t *action_dev, void *stream);
int ahk_policy_forward_stack_profiled(ahk_policy *p, const ahk_stack *stack, float *logits_dev, uint8_t *action_dev, void *stream,
                                      float *layer_ms);
/* The same forward pass with CUDA events recorded on `stream` around each of its six kernels (conv1, conv2, conv3,
 * dense1, dense2, out); waits, then reports their durations in layer_ms[6].  Measurement aid for bench.py's roofline. */
int ahk_policy_forward_profiled(ahk_policy *p, const uint8_t *frames_dev, int n, int head, float *logits_dev,
                                uint8_t *action_dev, void *stream, float *layer_ms);
/* NO COUNTERPART IN THE REFERENCE — added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  ahk_policy_forward for n <=
 * AHK_POLICY_STATIC_MAX_BATCH samples of an 'rgb' handle, with the activations between the layer kernels (the forward's
 * scratch memory) in a fixed array of the library's device image instead of the handle's buffers: the call allocates nothing
 * (no cudaMalloc / cudaFree / allocating library routine), the caller passes inputs and outputs as before (both required).
 * Same kernels, same results as ahk_policy_forward.  The array — 3 855 616 bytes — is held by every GPU that loads the
 * library until the process exits, is shared by every call on that GPU and keeps the last call's contents (nothing in it
 * needs clearing: each layer writes the rows the next one reads): calls that OVERLAP on the GPU — other streams, other
 * handles — race on it; issue them on one stream.  n outside [1, AHK_POLICY_STATIC_MAX_BATCH] and 'gray-diff' handles are
 * rejected with AHK_ERR_ARG.  The first call on a device resolves the array's device address and encodes four tensor maps
 * on the host (make it once before capturing a graph). */
#define AHK_POLICY_STATIC_MAX_BATCH 64
int ahk_policy_forward_static(ahk_policy *p, const uint8_t *frames_dev, int n, int head, float *logits_dev,
                              uint8_t *action_dev, void *stream);

/* Replace the weights of an 'rgb' policy handle in place (same folding and packing as ahk_policy_create): ≙
 * model.set_weights / the weights after model.fit (rl_agents.py:100-104,189) — the trainer pushes the updated network to
 * the forward kernels every training interval.  Waits for the device first (forward passes in flight read the old ones). */
int ahk_policy_set_weights(ahk_policy *p, const ahk_policy_weights *w);

/* The same with float32 Keras-layout weights that already live in DEVICE memory (w_dev holds device pointers): folded and
 * packed by kernels on `stream`, stream-ordered with the forward passes around it — after an optimizer step on the device
 * nothing crosses PCIe. */
int ahk_policy_set_weights_dev(ahk_policy *p, const ahk_policy_weights *w_dev, void *stream);

/* ---- model.fit on the device (rl_agents.py:189,276; ai/supervised/keras/train.py): training-mode forward, backward, optimizer ----
 * Training-mode forward pass of the 'rgb' network: BatchNormalization normalises each layer's post-ReLU output with the
 * BATCH's own per-channel mean and biased variance (Keras `fit`), which this call computes layer by layer and folds into
 * the next layer's weights exactly as the inference path folds the moving statistics — so the forward kernels are the
 * same tcgen05 kernels.  w_dev: float32 Keras-layout weights in device memory (its bn*_mean / bn*_var are not read).
 * frames_dev: [3][n][3][128][128] ring layout (ahk_replay_gather output), head = newest slot.  logits_dev float[n][10].
 * batch_stats_dev float[2][AHK_BN_STATS]: mean then variance, channels of bn1 at 0 (5 used of 8), bn2 at 8, bn3 at 40.
 * The layer outputs stay in the handle's activation buffers (ahk_policy_get_activations) for ahk_train_backward.
 * The handle's packed weights are left folded with the batch statistics: push the real ones (ahk_policy_set_weights_dev)
 * before the next inference call. */
#define AHK_BN_STATS 104
int ahk_policy_forward_train(ahk_policy *p, const ahk_policy_weights *w_dev, const uint8_t *frames_dev, int n, int head,
                             float *logits_dev, float *batch_stats_dev, void *stream);
typedef struct {
    const void *act1; /* bf16 [n][41*41][8]  relu(conv1), channels 5..7 zero */
    const void *act2; /* bf16 [n][19*19][32] relu(conv2) */
    const void *act3; /* bf16 [n][8*8][64]   relu(conv3): the dense1 input in (pixel, channel) order */
    const void *h1;   /* bf16 [n][512] */
    const void *h2;   /* bf16 [n][512] */
    int max_batch;
} ahk_policy_acts;
int ahk_policy_get_activations(ahk_policy *p, ahk_policy_acts *out);
/* Diagnostics and tests: activation buffer `which` (0 act1 .. 4 h2, layouts above) converted to float32 into dst_dev. */
int ahk_policy_copy_activation(ahk_policy *p, int which, float *dst_dev, size_t count, void *stream);

/* Backward pass of the forward pass ahk_policy_forward_train has just run on `p` (same stream, same frames / n / head /
 * batch_stats): given dlogits_dev float[n][10] = d loss / d logits (ahk_rl_ddqn_targets, ahk_rl_pg_targets), writes
 * d loss / d weight for the 18 trainable tensors into grads_dev (device pointers, Keras layouts, overwritten; its
 * bn*_mean / bn*_var slots are not touched).  ≙ the backward half of model.fit (rl_agents.py:189,276).  Hand-written
 * kernels only (csrc/ahk_train.cu); float32 accumulation over the bf16 activations the forward kernels produced. */
typedef struct ahk_trainer ahk_trainer;
int ahk_trainer_create(int device, int max_batch, ahk_trainer **out);
void ahk_trainer_destroy(ahk_trainer *t);
int ahk_train_backward(ahk_trainer *t, ahk_policy *p, const ahk_policy_weights *w_dev, const uint8_t *frames_dev, int n, int head,
                       const float *batch_stats_dev, const float *dlogits_dev, const ahk_policy_weights *grads_dev, void *stream);
/* One optimizer step over a FLAT float32 parameter vector (params_dev / grads_dev / state share the indexing), described
 * by up to 24 segments.  Gradients are taken as g + 2 * l2 * w on AHK_SEG_L2 segments (kernel_regularizer, models.py:17);
 * *loss_dev (nullable) += l2 * sum w^2.  AHK_SEG_BN_MEAN / _VAR segments are the BatchNorm moving statistics:
 * moving = bn_momentum * moving + (1 - bn_momentum) * batch_stats_dev[stat_offset + i] (Keras momentum 0.99;
 * batch_stats_dev nullable: left alone).
 * sgd:  keras.optimizers.SGD(lr, momentum, decay, clipnorm) as compiled at rl_agents.py:111,170: g *= clipnorm /
 *       max(||g||, clipnorm) over ALL trainable segments; v = momentum * v - lr_t * g; w += v.  lr_t = lr / (1 + decay * iterations)
 *       is the caller's.  clipnorm <= 0: no clipping.
 * adam: keras.optimizers.Adam (models.py:119): m, v moments; w -= lr_t * m / (sqrt(v) + eps) with
 *       lr_t = lr * sqrt(1 - beta2^t) / (1 - beta1^t) computed by the caller. */
#define AHK_SEG_TRAINABLE 1
#define AHK_SEG_L2 2
#define AHK_SEG_BN_MEAN 4
#define AHK_SEG_BN_VAR 8
typedef struct { int32_t offset, count, flags, stat_offset; } ahk_param_segment;
int ahk_train_sgd_step(ahk_trainer *t, float *params_dev, const float *grads_dev, float *velocity_dev, const ahk_param_segment *segs, int n_segs,
                       float lr_t, float momentum, float clipnorm, float l2, const float *batch_stats_dev, float bn_momentum, float *loss_dev,
                       void *stream);
int ahk_train_adam_step(ahk_trainer *t, float *params_dev, const float *grads_dev, float *m_dev, float *v_dev, const ahk_param_segment *segs, int n_segs,
                        float lr_t, float beta1, float beta2, float eps, float l2, const float *batch_stats_dev, float bn_momentum, float *loss_dev,
                        void *stream);
uint64_t ahk_trainer_launches(const ahk_trainer *t);
/* Diagnostics and tests: copy one of the backward pass's intermediate gradients (float32) to dst_dev: 0 dz5 [n][512],
 * 1 dz4 [n][512], 2 da3 [n][64 px][64], 3 dz3, 4 da2 [n][361][32], 5 dz2, 6 da1 [n][1681][8], 7 dz1. */
int ahk_trainer_copy_buffer(ahk_trainer *t, int which, float *dst_dev, size_t count, void *stream);

/* ---- batched RL step: ai/reinforcement/rl_agents.py around the CNN, on the device ----
 * Replay memory ≙ Agent.experience_buffer (rl_agents.py:61-62,88-90), kept as a time-ordered store of single uint8
 * frames, store_dev [capacity][n][3][128][128] (slot = step index mod capacity; the newest ring slot of every env is
 * copied in after each step), instead of two float32 (9,128,128) stacks per transition.
 * ≙ Agent._sample_experiences (rl_agents.py:64-77): for sample b, frames t-2, t-1, t, t+1 (t = t_dev[b], slots mod
 * capacity) of env env_dev[b] -> out_dev [4][batch][3][128][128].  out_dev[0:3] with head 2 is the stacked state s_t and
 * out_dev + batch*AHK_FRAME_BYTES (slots 1..3) with head 2 is s_{t+1}: both go to ahk_policy_forward as they are. */
int ahk_replay_gather(const uint8_t *store_dev, int capacity, int n, const int32_t *t_dev, const int32_t *env_dev, int batch,
                      uint8_t *out_dev, void *stream);
/* ≙ Agent.store_experience (rl_agents.py:88-90) for all n envs in one launch: files the step's action / reward / terminal flag
 * (each nullable: zeros — the observation an episode starts from) into row t of the replay arrays, together with the two
 * flags sampling needs from the engine's state word 15 (meta_dev = state_dev + 15 * n): pending_row = auto-reset armed (the
 * next step starts a new episode), stale_row (nullable) = no step since an explicit reset (frames are in ahk_buffers.warm). */
int ahk_replay_record(const uint32_t *meta_dev, const uint8_t *action_dev, const float *reward_dev, const uint8_t *done_dev, int n,
                      uint8_t *action_row, float *reward_row, uint8_t *done_row, uint8_t *pending_row, uint8_t *stale_row, void *stream);
/* ≙ the epsilon_greedy policy of Agent.__init__ / Agent.act (rl_agents.py:48-56,92-101) for n envs: action = a uniform draw
 * in [0, 10) with probability epsilon, else greedy_dev[i] (the argmax ahk_policy_forward wrote).  Draws are Philox4x32-10
 * keyed by `seed` with counter (iteration, GLOBAL env id), so they do not depend on the sharding.  hist_dev uint64[10]
 * (nullable) accumulates Agent.action_distribution. */
int ahk_rl_epsilon_greedy(const uint8_t *greedy_dev, int n, float epsilon, uint64_t seed, uint32_t env_id_base, uint32_t iteration,
                          uint8_t *action_dev, uint64_t *hist_dev, void *stream);
/* ≙ DDQNAgent.train (rl_agents.py:177-191), all float[n][10]: q = Q(s), q_next_online = Q(s'), q_next_target =
 * Q_target(s').  targets = q with targets[a] = reward + (1 - done) * discount * q_next_target[argmax q_next_online],
 * clipped to [-1, 1].  Also the Keras mean_squared_error loss of q_fit against targets (loss_dev, one float) and its
 * gradient dlogits = 2 (q_fit - targets) / (10 n); q_fit = output of the forward pass being fitted, NULL = q.
 * targets_dev / dlogits_dev / loss_dev may be NULL. */
int ahk_rl_ddqn_targets(const float *q_dev, const float *q_next_online_dev, const float *q_next_target_dev, const uint8_t *action_dev,
                        const float *reward_dev, const uint8_t *done_dev, int n, double discount, const float *q_fit_dev,
                        float *targets_dev, float *dlogits_dev, float *loss_dev, void *stream);
/* ≙ PGAgent.train::discount_rewards (rl_agents.py:257-268) for n envs at once: reward_dev/done_dev/returns_dev are
 * [steps][n]; per episode (a run of steps of one env ending at done != 0 or at the last step) G[t] = r[t] + discount *
 * G[t+1], then standardised by the episode's own mean and population std (zeros when the std is 0), in float64. */
int ahk_rl_pg_returns(const float *reward_dev, const uint8_t *done_dev, int steps, int n, double discount, float *returns_dev, void *stream);
/* ≙ PGAgent.store_experience + train (rl_agents.py:246-253,270-274) for m = steps*n samples: prob_dev float[m][10] = the
 * softmax output at act() time, targets = prob + learning_rate * (onehot(action) - prob) * returns; plus the categorical
 * cross-entropy (models.py:119) of prob_fit (NULL = prob) against targets and its gradient w.r.t. the logits. */
int ahk_rl_pg_targets(const float *prob_dev, const uint8_t *action_dev, const float *returns_dev, int m, float learning_rate,
                      const float *prob_fit_dev, float *targets_dev, float *dlogits_dev, float *loss_dev, void *stream);

/* ---- the 'gray-diff' network (ai/supervised/keras/models.py:67-116; ai/config.py:3 lists the mode) ----
 * (1,128,128) gray-diff observation (processor.py:53-64, computed from the ring by the engine) -> conv 16 k8 s4 ->
 * conv 24 k4 s2 -> conv 32 k4 s1 -> conv 48 k4 s1 (each ReLU then BatchNorm) -> flatten 3072 -> dense 512 -> dense 10.
 * Same handle type and the same ahk_policy_forward / _forward_profiled / _destroy calls as the 'rgb' network
 * (layer_ms = input + conv1, conv2, conv3, conv4, dense1, out). */
typedef struct {
    const float *conv1_w, *conv1_b, *bn1_gamma, *bn1_beta, *bn1_mean, *bn1_var; /* 8x8x1x16  */
    const float *conv2_w, *conv2_b, *bn2_gamma, *bn2_beta, *bn2_mean, *bn2_var; /* 4x4x16x24 */
    const float *conv3_w, *conv3_b, *bn3_gamma, *bn3_beta, *bn3_mean, *bn3_var; /* 4x4x24x32 */
    const float *conv4_w, *conv4_b, *bn4_gamma, *bn4_beta, *bn4_mean, *bn4_var; /* 4x4x32x48 */
    const float *dense1_w, *dense1_b; /* 3072x512 */
    const float *out_w, *out_b;       /* 512x10   */
} ahk_policy_gray_weights; /* host pointers, Keras layouts */
int ahk_policy_gray_create(int device, int max_batch, const ahk_policy_gray_weights *w, ahk_policy **out);

#ifdef __cplusplus
}
#endif
#endif /* AHK_H */
