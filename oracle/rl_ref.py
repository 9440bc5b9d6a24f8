"""numpy restatement of the arithmetic in ai/reinforcement/rl_agents.py — TEST INFRASTRUCTURE, never on the product path.

Each function follows the cited lines of the reference with the dtypes numpy gives them there (Keras predictions are
float32; rewards collected from env.step are Python floats, i.e. float64 arrays).  Keras itself is absent, so the loss and
optimizer formulas restate Keras 2.x (`mean_squared_error`, `categorical_crossentropy`, `SGD.get_updates`, `clip_norm`).
"""
import numpy as np


def ddqn_targets(q, q_next_online, q_next_target, actions, rewards, done, discount_rate):
    """rl_agents.py:179-187.  q, q_next_*: float32 [B, 10]; rewards float64 [B]; done bool [B]."""
    targets = np.array(q, dtype=np.float32, copy=True)
    next_actions = np.argmax(q_next_online, axis=1)
    rows = np.arange(targets.shape[0])
    targets[rows, actions] = np.asarray(rewards, dtype=np.float64) + (1 - np.asarray(done, dtype=bool)) * discount_rate * q_next_target[rows, next_actions]
    return np.clip(targets, -1.0, 1.0)


def mse_loss_and_grad(q_fit, targets):
    """Keras mean_squared_error averaged over the batch, and d loss / d q_fit."""
    d = q_fit.astype(np.float64) - targets.astype(np.float64)
    return float(np.mean(d * d)), (2.0 * d / d.size).astype(np.float32)


def discount_rewards(rewards, discount_rate):
    """rl_agents.py:257-268 for ONE episode's rewards."""
    rewards = np.array(rewards, dtype=np.float64)
    for i in reversed(range(1, len(rewards))):
        rewards[i - 1] += discount_rate * rewards[i]
    rewards -= np.mean(rewards)
    std = np.std(rewards)
    return np.divide(rewards, std, out=np.zeros_like(rewards), where=std != 0)


def discount_rewards_batched(reward, ends, discount_rate):
    """discount_rewards applied to every episode of every env: reward/ends [T, n]; an episode is a run of steps ending
    where ends != 0 or at T-1."""
    T, n = reward.shape
    out = np.zeros((T, n), dtype=np.float64)
    for e in range(n):
        start = 0
        for t in range(T):
            if ends[t, e] or t == T - 1:
                out[start:t + 1, e] = discount_rewards(reward[start:t + 1, e], discount_rate)
                start = t + 1
    return out


def pg_targets(prob, actions, returns, learning_rate):
    """rl_agents.py:246-253,270-274: y one-hot float32; gradients = (y - prediction) * reward; targets = predictions + lr * gradients."""
    y = np.zeros_like(prob, dtype=np.float32)
    y[np.arange(len(actions)), actions] = 1.0
    gradients = (y - prob) * np.asarray(returns).reshape((-1, 1))
    return prob + learning_rate * gradients


def crossentropy_loss_and_grad(prob_fit, targets):
    """Keras categorical_crossentropy on a softmax output (clipped to [1e-7, 1 - 1e-7]) and d loss / d logits."""
    p = np.clip(prob_fit.astype(np.float64), 1e-7, 1 - 1e-7)
    t = targets.astype(np.float64)
    loss = float(np.mean(-np.sum(t * np.log(p), axis=1)))
    grad = (prob_fit.astype(np.float64) * t.sum(axis=1, keepdims=True) - t) / len(t)
    return loss, grad.astype(np.float32)


def keras_sgd_steps(params, grads_per_step, lr=0.0075, momentum=0.5, decay=1e-6, clipnorm=2.0):
    """keras.optimizers.SGD.get_updates (Keras 2.x) with clipnorm on the global norm, as compiled at rl_agents.py:111,170."""
    params = [np.array(p, dtype=np.float64) for p in params]
    vel = [np.zeros_like(p) for p in params]
    for it, grads in enumerate(grads_per_step):
        grads = [np.asarray(g, dtype=np.float64) for g in grads]
        norm = np.sqrt(sum(np.sum(g * g) for g in grads))
        if clipnorm and norm >= clipnorm:
            grads = [g * clipnorm / norm for g in grads]
        lr_t = lr * (1.0 / (1.0 + decay * it))
        for p, v, g in zip(params, vel, grads):
            v *= momentum
            v -= lr_t * g
            p += v
    return params


def l2_penalty_and_grads(kernels, l2=1e-4):
    """kernel_regularizer=regularizers.l2(l2) (models.py:17,22-62): what model.fit adds to the loss, l2 * sum ||W||^2, and to
    the gradient of every regularised kernel, 2 * l2 * W."""
    penalty = float(sum(l2 * np.sum(np.asarray(k, dtype=np.float64) ** 2) for k in kernels))
    return penalty, [(2.0 * l2 * np.asarray(k, dtype=np.float64)).astype(np.float32) for k in kernels]
