// Batched RL step: the arithmetic of ai/reinforcement/rl_agents.py around the CNN, on the device.
//   replay_gather_kernel   ≙ Agent._sample_experiences (rl_agents.py:64-77) on a time-ordered uint8 frame store
//   ddqn_targets_kernel    ≙ DDQNAgent.train (rl_agents.py:177-191): double-Q target, clip, MSE and its gradient
//   pg_returns_kernel      ≙ PGAgent.train::discount_rewards (rl_agents.py:257-268), per episode segment, float64
//   pg_targets_kernel      ≙ PGAgent.store_experience / train (rl_agents.py:246-253, 270-274) + cross-entropy gradient
// All HBM-bound byte/float work: coalesced 128-bit accesses, no tensor cores.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/ahk.h"
#include "ahk_kernels.cuh"
#include "ahk_sim.cuh"

namespace ahk {
namespace rl {

constexpr int kFrameVec = AHK_FRAME_BYTES / 16; // 3072 uint4 per frame

// out[j][b] = store[(t[b] + cap - 2 + j) % cap][env[b]], j = 0..3: frames t-2, t-1, t, t+1 of the sampled env.
// out[0:3] with head 2 is the stacked state s_t, out[1:4] with head 2 is s_{t+1} (slot of age a = (head + 3 - a) % 3),
// so both feed ahk_policy_forward unchanged.  One CTA per (j, b) frame, 256 threads x 12 uint4.
__global__ void __launch_bounds__(256) replay_gather_kernel(const uint4 *__restrict__ store, int cap, int n,
                                                            const int32_t *__restrict__ t_idx, const int32_t *__restrict__ env_idx,
                                                            int batch, uint4 *__restrict__ out)
{
    for (int unit = blockIdx.x; unit < 4 * batch; unit += gridDim.x) {
        const int j = unit / batch, b = unit - j * batch;
        const int t = (t_idx[b] + cap - 2 + j) % cap;
        const uint4 *src = store + (static_cast<size_t>(t) * n + env_idx[b]) * kFrameVec;
        uint4 *dst = out + (static_cast<size_t>(j) * batch + b) * kFrameVec;
        uint4 v[12];
#pragma unroll
        for (int i = 0; i < 12; ++i) v[i] = __ldg(src + threadIdx.x + i * 256);
#pragma unroll
        for (int i = 0; i < 12; ++i) __stcs(dst + threadIdx.x + i * 256, v[i]);
    }
}

// rl_agents.py:177-191.  targets = Q(s); targets[a] = r + (1 - done) * discount * Q_target(s')[argmax Q(s')];
// clip to [-1, 1]; loss = mean over n x 10 of (Q(s) - targets)^2 (Keras mean_squared_error, mean over the batch);
// dlogits = d loss / d Q_fit = 2 (Q_fit - targets) / (10 n), Q_fit = the output of the forward pass being fitted (Keras
// `fit` runs BatchNorm on batch statistics, `predict` on the moving ones; NULL = Q(s) itself).  One thread per sample, float32 like numpy on float32 predictions
// (reward is float64 in the reference; the sum is formed in double and rounded once, as numpy's assignment does).
__global__ void __launch_bounds__(256) ddqn_targets_kernel(const float *__restrict__ q, const float *__restrict__ qn_online,
                                                           const float *__restrict__ qn_target, const uint8_t *__restrict__ action,
                                                           const float *__restrict__ reward, const uint8_t *__restrict__ done, int n,
                                                           double discount, const float *__restrict__ q_fit,
                                                           float *__restrict__ targets, float *__restrict__ dlogits,
                                                           float *__restrict__ loss)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float err2 = 0.0f;
    if (i < n) {
        float row[AHK_N_ACTIONS], on[AHK_N_ACTIONS];
#pragma unroll
        for (int k = 0; k < AHK_N_ACTIONS; ++k) { row[k] = q[i * AHK_N_ACTIONS + k]; on[k] = qn_online[i * AHK_N_ACTIONS + k]; }
        int best = 0; // np.argmax: first maximum
#pragma unroll
        for (int k = 1; k < AHK_N_ACTIONS; ++k) if (on[k] > on[best]) best = k;
        const double y = static_cast<double>(reward[i]) +
                         static_cast<double>(1 - (done[i] ? 1 : 0)) * discount * static_cast<double>(qn_target[i * AHK_N_ACTIONS + best]);
        const int a = action[i] < AHK_N_ACTIONS ? action[i] : 0;
        const float scale = 2.0f / (static_cast<float>(n) * AHK_N_ACTIONS);
#pragma unroll
        for (int k = 0; k < AHK_N_ACTIONS; ++k) {
            float t = (k == a) ? static_cast<float>(y) : row[k];
            t = fminf(fmaxf(t, -1.0f), 1.0f);
            const float d = (q_fit ? q_fit[i * AHK_N_ACTIONS + k] : row[k]) - t;
            if (targets) targets[i * AHK_N_ACTIONS + k] = t;
            if (dlogits) dlogits[i * AHK_N_ACTIONS + k] = scale * d;
            err2 += d * d;
        }
    }
    if (loss) {
        __shared__ float part[8];
        for (int o = 16; o > 0; o >>= 1) err2 += __shfl_xor_sync(0xffffffffu, err2, o);
        if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = err2;
        __syncthreads();
        if (threadIdx.x == 0) {
            float s = 0.0f;
            for (int w = 0; w < 8; ++w) s += part[w];
            atomicAdd(loss, s / (static_cast<float>(n) * AHK_N_ACTIONS));
        }
    }
}

// rl_agents.py:257-268 per episode.  reward/done are [T][n] (step-major, so consecutive threads read consecutive
// addresses); an episode of env e is a maximal run of steps ending at a step with done != 0 or at T-1.  Within it:
// G[t] = r[t] + discount * G[t+1] (the in-place reverse loop), then G -= mean(G), then G /= std(G) (population std; all
// zeros when std == 0: np.divide(..., where=std != 0) onto zeros).  float64 like the reference's np.array of floats.
__global__ void __launch_bounds__(128) pg_returns_kernel(const float *__restrict__ reward, const uint8_t *__restrict__ done, int T,
                                                         int n, double discount, float *__restrict__ returns)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    int end = T - 1;
    while (end >= 0) {
        double g = 0.0, sum = 0.0;
        int t = end;
        for (;; --t) { // pass 1: the episode's first step and the mean of its discounted returns
            g = static_cast<double>(reward[static_cast<size_t>(t) * n + e]) + discount * g;
            sum += g;
            if (t == 0 || done[static_cast<size_t>(t - 1) * n + e]) break;
        }
        const int start = t, len = end - start + 1;
        // passes 2 and 3 rebuild the returns from the rewards (T floats per env, L1/L2 hits) rather than keep them in float32
        const double mean = sum / len;
        double var = 0.0;
        g = 0.0;
        for (int u = end; u >= start; --u) {
            g = static_cast<double>(reward[static_cast<size_t>(u) * n + e]) + discount * g;
            var += (g - mean) * (g - mean);
        }
        const double sd = sqrt(var / len);
        g = 0.0;
        for (int u = end; u >= start; --u) {
            g = static_cast<double>(reward[static_cast<size_t>(u) * n + e]) + discount * g;
            returns[static_cast<size_t>(u) * n + e] = sd != 0.0 ? static_cast<float>((g - mean) / sd) : 0.0f;
        }
        end = start - 1;
    }
}

// rl_agents.py:246-253,270-274.  prob = the model's softmax output at act() time ([m][10], m = T*n), y = one-hot(action):
// targets = prob + learning_rate * (y - prob) * G.  The model is fitted with categorical cross-entropy (models.py:119):
// loss = -mean_m sum_k targets * log(prob_fit); with a softmax output d loss / d logits = (prob_fit * sum_k targets - targets) / m,
// prob_fit = the softmax output of the forward pass being fitted (NULL = prob).
__global__ void __launch_bounds__(256) pg_targets_kernel(const float *__restrict__ prob, const uint8_t *__restrict__ action,
                                                         const float *__restrict__ returns, int m, float learning_rate,
                                                         const float *__restrict__ prob_fit, float *__restrict__ targets, float *__restrict__ dlogits, float *__restrict__ loss)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float ce = 0.0f;
    if (i < m) {
        const int a = action[i];
        const float g = returns[i];
        float p[AHK_N_ACTIONS], t[AHK_N_ACTIONS], tsum = 0.0f;
#pragma unroll
        for (int k = 0; k < AHK_N_ACTIONS; ++k) {
            p[k] = prob[i * AHK_N_ACTIONS + k];
            const float grad = ((k == a) ? 1.0f : 0.0f) - p[k];
            t[k] = p[k] + learning_rate * (grad * g);
            tsum += t[k];
        }
#pragma unroll
        for (int k = 0; k < AHK_N_ACTIONS; ++k) {
            if (targets) targets[i * AHK_N_ACTIONS + k] = t[k];
            const float pf = prob_fit ? prob_fit[i * AHK_N_ACTIONS + k] : p[k];
            if (dlogits) dlogits[i * AHK_N_ACTIONS + k] = (pf * tsum - t[k]) / static_cast<float>(m);
            ce -= t[k] * logf(fminf(fmaxf(pf, 1e-7f), 1.0f - 1e-7f)); // Keras clips to [eps, 1 - eps] before the log
        }
    }
    if (loss) {
        __shared__ float part[8];
        for (int o = 16; o > 0; o >>= 1) ce += __shfl_xor_sync(0xffffffffu, ce, o);
        if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = ce;
        __syncthreads();
        if (threadIdx.x == 0) {
            float s = 0.0f;
            for (int w = 0; w < 8; ++w) s += part[w];
            atomicAdd(loss, s / static_cast<float>(m));
        }
    }
}

// rl_agents.py:48-56, one draw per env: with probability epsilon a uniform action in [0, 10), else the greedy one.
// Counter-based like every draw of the engine (SPEC.md §4): Philox keyed by the seed, counter (iteration, 0, GLOBAL env id,
// 'EPSG'), so the actions do not depend on how the envs are sharded over GPUs.  hist[10] += the actions taken
// (Agent.action_distribution, rl_agents.py:99).
__global__ void __launch_bounds__(256) epsilon_greedy_kernel(const uint8_t *__restrict__ greedy, int n, float epsilon, uint2 key,
                                                             uint32_t env_id_base, uint32_t iteration, uint8_t *__restrict__ action,
                                                             unsigned long long *__restrict__ hist)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    int a = -1;
    if (i < n) {
        const uint4 r = philox(iteration, 0u, env_id_base + static_cast<uint32_t>(i), 0x45505347u, key);
        a = unit(r.x) < epsilon ? static_cast<int>(__umulhi(r.y, 10u)) : static_cast<int>(greedy[i]);
        action[i] = static_cast<uint8_t>(a);
    }
    if (hist) {
        __shared__ unsigned s_hist[AHK_N_ACTIONS];
        if (threadIdx.x < AHK_N_ACTIONS) s_hist[threadIdx.x] = 0u;
        __syncthreads();
#pragma unroll
        for (int k = 0; k < AHK_N_ACTIONS; ++k) {
            const unsigned votes = __ballot_sync(0xffffffffu, a == k);
            if ((threadIdx.x & 31) == 0 && votes) atomicAdd(&s_hist[k], __popc(votes));
        }
        __syncthreads();
        if (threadIdx.x < AHK_N_ACTIONS && s_hist[threadIdx.x]) atomicAdd(hist + threadIdx.x, static_cast<unsigned long long>(s_hist[threadIdx.x]));
    }
}

static int rfail(int code, const char *text)
{
    set_error_text(text);
    return code;
}

static int launched(const char *what)
{
    const cudaError_t err = cudaGetLastError();
    if (err == cudaSuccess) return AHK_OK;
    char text[256];
    snprintf(text, sizeof text, "%s: %s", what, cudaGetErrorString(err));
    return rfail(AHK_ERR_CUDA, text);
}


// Agent.store_experience (rl_agents.py:88-90) for every env at once: the step's action, reward, terminal flag, and the two
// flags replay sampling needs from the engine's state word 15 (meta): auto-reset armed (bit 21: the NEXT step starts a new
// episode) and "no step since an explicit reset" (episode_steps == 0: the env's frames sit in the warm-up side buffer).
// One launch instead of eight small tensor ops per step.
__global__ void __launch_bounds__(256) replay_record_kernel(const uint32_t *__restrict__ meta, const uint8_t *__restrict__ action, const float *__restrict__ reward,
                                                            const uint8_t *__restrict__ done, int n, uint8_t *__restrict__ action_row,
                                                            float *__restrict__ reward_row, uint8_t *__restrict__ done_row, uint8_t *__restrict__ pending_row,
                                                            uint8_t *__restrict__ stale_row)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const uint32_t m = meta[i];
    action_row[i] = action ? action[i] : 0;
    reward_row[i] = reward ? reward[i] : 0.0f;
    done_row[i] = done ? (done[i] != 0) : 0;
    pending_row[i] = (m >> 21) & 1u;
    if (stale_row) stale_row[i] = (m & 0xFFFFu) == 0u;
}

} // namespace rl
} // namespace ahk

using namespace ahk::rl;

extern "C" {

int ahk_replay_gather(const uint8_t *store_dev, int capacity, int n, const int32_t *t_dev, const int32_t *env_dev, int batch,
                      uint8_t *out_dev, void *stream)
{
    if (!store_dev || !t_dev || !env_dev || !out_dev || capacity < 4 || n <= 0 || batch <= 0)
        return rfail(AHK_ERR_ARG, "ahk_replay_gather: bad argument (capacity must be >= 4, pointers non-NULL)");
    if ((reinterpret_cast<uintptr_t>(store_dev) | reinterpret_cast<uintptr_t>(out_dev)) & 15)
        return rfail(AHK_ERR_ARG, "ahk_replay_gather: frame buffers must be 16-byte aligned");
    // persistent over the frames: exactly as many CTAs as are resident at once (a second, partial wave would start its
    // whole share late)
    static int resident = 0;
    if (resident == 0) {
        int dev = 0, sms = 0, per_sm = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, replay_gather_kernel, 256, 0) != cudaSuccess || per_sm < 1)
            return launched("ahk_replay_gather");
        resident = sms * per_sm;
    }
    const long units = 4L * batch;
    const int grid = static_cast<int>(units < resident ? units : resident);
    replay_gather_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(reinterpret_cast<const uint4 *>(store_dev), capacity, n, t_dev,
                                                                            env_dev, batch, reinterpret_cast<uint4 *>(out_dev));
    return launched("ahk_replay_gather");
}

int ahk_rl_epsilon_greedy(const uint8_t *greedy_dev, int n, float epsilon, uint64_t seed, uint32_t env_id_base, uint32_t iteration,
                          uint8_t *action_dev, uint64_t *hist_dev, void *stream)
{
    if (!greedy_dev || !action_dev || n <= 0) return rfail(AHK_ERR_ARG, "ahk_rl_epsilon_greedy: bad argument");
    const uint2 key = make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
    epsilon_greedy_kernel<<<(n + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(greedy_dev, n, epsilon, key, env_id_base, iteration, action_dev,
                                                                                        reinterpret_cast<unsigned long long *>(hist_dev));
    return launched("ahk_rl_epsilon_greedy");
}

int ahk_rl_ddqn_targets(const float *q_dev, const float *q_next_online_dev, const float *q_next_target_dev, const uint8_t *action_dev,
                        const float *reward_dev, const uint8_t *done_dev, int n, double discount, const float *q_fit_dev,
                        float *targets_dev, float *dlogits_dev, float *loss_dev, void *stream)
{
    if (!q_dev || !q_next_online_dev || !q_next_target_dev || !action_dev || !reward_dev || !done_dev || n <= 0)
        return rfail(AHK_ERR_ARG, "ahk_rl_ddqn_targets: bad argument");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (loss_dev && cudaMemsetAsync(loss_dev, 0, sizeof(float), s) != cudaSuccess) return launched("ahk_rl_ddqn_targets");
    ddqn_targets_kernel<<<(n + 255) / 256, 256, 0, s>>>(q_dev, q_next_online_dev, q_next_target_dev, action_dev, reward_dev, done_dev, n,
                                                       discount, q_fit_dev, targets_dev, dlogits_dev, loss_dev);
    return launched("ahk_rl_ddqn_targets");
}

int ahk_replay_record(const uint32_t *meta_dev, const uint8_t *action_dev, const float *reward_dev, const uint8_t *done_dev, int n,
                      uint8_t *action_row, float *reward_row, uint8_t *done_row, uint8_t *pending_row, uint8_t *stale_row, void *stream)
{
    if (!meta_dev || !action_row || !reward_row || !done_row || !pending_row || n <= 0) return rfail(AHK_ERR_ARG, "ahk_replay_record: bad argument");
    replay_record_kernel<<<(n + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(meta_dev, action_dev, reward_dev, done_dev, n, action_row, reward_row,
                                                                                         done_row, pending_row, stale_row);
    return launched("ahk_replay_record");
}

int ahk_rl_pg_returns(const float *reward_dev, const uint8_t *done_dev, int steps, int n, double discount, float *returns_dev, void *stream)
{
    if (!reward_dev || !done_dev || !returns_dev || steps <= 0 || n <= 0) return rfail(AHK_ERR_ARG, "ahk_rl_pg_returns: bad argument");
    pg_returns_kernel<<<(n + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(reward_dev, done_dev, steps, n,
                                                                                    discount, returns_dev);
    return launched("ahk_rl_pg_returns");
}

int ahk_rl_pg_targets(const float *prob_dev, const uint8_t *action_dev, const float *returns_dev, int m, float learning_rate,
                      const float *prob_fit_dev, float *targets_dev, float *dlogits_dev, float *loss_dev, void *stream)
{
    if (!prob_dev || !action_dev || !returns_dev || m <= 0) return rfail(AHK_ERR_ARG, "ahk_rl_pg_targets: bad argument");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (loss_dev && cudaMemsetAsync(loss_dev, 0, sizeof(float), s) != cudaSuccess) return launched("ahk_rl_pg_targets");
    pg_targets_kernel<<<(m + 255) / 256, 256, 0, s>>>(prob_dev, action_dev, returns_dev, m, learning_rate, prob_fit_dev, targets_dev, dlogits_dev, loss_dev);
    return launched("ahk_rl_pg_targets");
}

} // extern "C"
