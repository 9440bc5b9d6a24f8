"""Batched RL step: ai/reinforcement/rl_agents.py + train.py:44-118 for N environments on a GPU.

What runs where
* acting (`Agent.act`, rl_agents.py:92-101) and the three Q evaluations of `DDQNAgent.train` (rl_agents.py:179-181):
  the tcgen05 forward kernels on uint8 frames (`PolicyNetwork`, csrc/ahk_policy.cu);
* replay memory (`experience_buffer`, rl_agents.py:61-62,88-90) and sampling (rl_agents.py:64-77): `ReplayStore`, a
  time-ordered uint8 frame store in HBM + `ahk_replay_gather` (csrc/ahk_rl.cu);
* targets, loss and the gradient at the network output (rl_agents.py:183-187, 246-253, 257-274): `ahk_rl_ddqn_targets`,
  `ahk_rl_pg_returns`, `ahk_rl_pg_targets` (csrc/ahk_rl.cu);
* the training-mode forward, the backward pass and the optimizer step of `model.fit` (rl_agents.py:189, 276):
  `train.NativeTrainer` — hand-written kernels (csrc/ahk_train.cu, ahk_policy_forward_train): no autograd, no cuDNN, no
  cuBLAS; the flat float32 gradient is averaged over ranks with one NCCL all-reduce, the optimizers restate the compiled
  Keras ones (SGD lr 0.0075 momentum 0.5 decay 1e-6 clipnorm 2: rl_agents.py:111,170; Adam: models.py:119);
* after every fit the new weights (and BatchNorm moving statistics) go back to the forward kernels with
  `ahk_policy_set_weights`.
"""
import ctypes as C

import torch

from . import _lib
from .policy import PolicyNetwork, keras_default_init
from .train import L2_REGULARIZER, NativeTrainer  # noqa: F401  (L2_REGULARIZER is part of this module's interface)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _stream(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


# ---------------------------------------------------------------------------------------------- replay memory
def valid_transitions(pending, cursor, filled, need_next=True, stale=None):
    """Which (slot, env) pairs of a ReplayStore can be sampled.  pending: uint8 [capacity, n], 1 where the env was armed
    for an auto-reset after that step (so the NEXT step rewrote all three ring slots with warm-up frames,
    air_hockey_env.py:69-72).  A transition at slot t uses frames t-2..t+1 (state = t, t-1, t-2; next state = t+1, t,
    t-1): all four must be recorded, consecutive, and no reset may fall on steps t-1, t, t+1 (pending at t-2, t-1, t).
    `cursor` = slot the next step will be written to, `filled` = recorded steps (<= capacity).  stale: optional uint8
    [capacity, n], 1 where the slot holds no frame of that env (a deep ring right after an explicit reset, whose frames sit
    in the warm-up side buffer): no window may contain such a slot.  Returns bool [capacity, n]."""
    cap = pending.shape[0]
    age = (cursor - 1 - torch.arange(cap, device=pending.device)) % cap  # 0 = newest recorded slot
    lo = 1 if need_next else 0                                              # t+1 must exist
    ok_t = (age >= lo) & (age <= filled - 3)                                # t-2 must exist
    clear = pending == 0
    ok = clear.roll(2, 0) & clear.roll(1, 0)                                # steps t-1, t were not resets
    if need_next:
        ok = ok & clear                                                     # step t+1 was not a reset
    if stale is not None:
        good = stale == 0
        ok = ok & good & good.roll(1, 0) & good.roll(2, 0)                  # frames t, t-1, t-2
        if need_next:
            ok = ok & good.roll(-1, 0)                                      # frame t+1
    return ok & ok_t[:, None]


class ReplayStore(object):
    """≙ Agent.experience_buffer (rl_agents.py:61-62,88-90): `capacity` steps of all n envs.  Frames are kept once, as
    uint8 (49 152 B per env-step); the reference keeps two float32 (9,128,128) stacks per transition (1.18 MB).

    Two forms.  With the engine's 3-slot ring the store is its own uint8 [capacity, n, 3, 128, 128] buffer and every step's
    newest frame is copied in (one device-to-device copy per step).  With a DEEP ring (`BatchedAirHockeyEnv(ring_depth=D)`,
    D > 3) the store is a VIEW of the ring: the rasteriser wrote each frame where replay sampling reads it, nothing is
    copied, capacity = D and slot = the ring slot of the step (`env.head`)."""

    def __init__(self, env, capacity=None):
        self.lib = _lib.load()
        self.view = getattr(env, 'ring_depth', 3) > 3
        capacity = env.ring_depth if self.view else capacity
        if capacity is None or capacity < 4:
            raise ValueError('capacity must be at least 4 steps')
        self.n, self.capacity, self.device = env.n_envs, int(capacity), env.device
        n, dev = self.n, self.device
        self.frames = env.frames if self.view else torch.zeros((self.capacity, n, 3, 128, 128), dtype=torch.uint8, device=dev)
        self.action = torch.zeros((self.capacity, n), dtype=torch.uint8, device=dev)
        self.reward = torch.zeros((self.capacity, n), dtype=torch.float32, device=dev)
        self.done = torch.zeros((self.capacity, n), dtype=torch.uint8, device=dev)
        self.pending = torch.zeros((self.capacity, n), dtype=torch.uint8, device=dev)
        self.stale = torch.zeros((self.capacity, n), dtype=torch.uint8, device=dev) if self.view else None
        self.cursor, self.filled = (env.head + 1) % self.capacity if self.view else 0, 0

    def reset(self):
        """≙ Agent._reset_experience_buffer (rl_agents.py:61-62).  (A view keeps following the ring's slots.)"""
        self.filled = 0
        if not self.view:
            self.cursor = 0

    def next_slot(self, env):
        """The slot the next record() will fill: the ring slot of the coming step (view) or the store's own cursor."""
        return (env.head + 1) % self.capacity if self.view else self.cursor

    def record_observation(self, env):
        """Store the newest frame of every env with no transition attached (the observation an episode starts from)."""
        self._put(env, None, None, None)

    def record(self, env, action, reward, done):
        """≙ Agent.store_experience (rl_agents.py:88-90) after env.step(action): the frame the step produced, the action
        that led to it, its reward and terminal flag.  (state, next_state) are implied by the neighbouring slots."""
        self._put(env, action, reward, done)

    def _put(self, env, action, reward, done):
        if self.view:
            t = env.head                                 # the rasteriser already wrote the frame here
        else:
            t = self.cursor
            self.frames[t].copy_(env.frames[env.head])  # 49 152 B per env, device to device
        if done is not None and done.dtype == torch.bool:
            done = done.view(torch.uint8)
        if action is not None and (action.dtype != torch.uint8 or not action.is_contiguous()):
            action = action.to(torch.uint8).contiguous()
        if reward is not None and (reward.dtype != torch.float32 or not reward.is_contiguous()):
            reward = reward.to(torch.float32).contiguous()
        # one kernel: action / reward / done rows, pending (auto-reset armed: the next step starts a new episode) and stale (no
        # step since an explicit reset: the env's frames are in env.warm) from the engine's state word 15
        _lib.check(self.lib.ahk_replay_record(_ptr(env.state[15]), _ptr(action), _ptr(reward), _ptr(done), self.n, _ptr(self.action[t]),
                                              _ptr(self.reward[t]), _ptr(self.done[t]), _ptr(self.pending[t]),
                                              _ptr(self.stale[t]) if self.view else None, _stream(self.device)))
        self.cursor = (t + 1) % self.capacity
        self.filled = min(self.filled + 1, self.capacity)

    def gather(self, t_idx, env_idx):
        """-> uint8 [4, B, 3, 128, 128]: frames t-2, t-1, t, t+1 of each sampled env (ahk_replay_gather)."""
        b = int(t_idx.numel())
        out = torch.empty((4, b, 3, 128, 128), dtype=torch.uint8, device=self.device)
        t32, e32 = t_idx.to(torch.int32).contiguous(), env_idx.to(torch.int32).contiguous()
        _lib.check(self.lib.ahk_replay_gather(_ptr(self.frames), self.capacity, self.n, _ptr(t32), _ptr(e32), b, _ptr(out),
                                              _stream(self.device)))
        return out

    def sample(self, batch_size, generator=None, need_next=True):
        """≙ Agent._sample_experiences (rl_agents.py:64-77): up to batch_size distinct transitions, uniformly.
        -> dict(frames [4,B,...], action = the action taken in s_t, reward, done of s_t -> s_{t+1}, t, env) or None."""
        ok = valid_transitions(self.pending, self.cursor, self.filled, need_next, self.stale)
        flat = ok.reshape(-1).nonzero().squeeze(1)
        if flat.numel() == 0:
            return None
        k = min(int(batch_size), int(flat.numel()))
        pick = flat[torch.randperm(flat.numel(), device=self.device, generator=generator)[:k]]
        t_idx, env_idx = pick // self.n, pick % self.n
        nxt = (t_idx + 1) % self.capacity
        return {'frames': self.gather(t_idx, env_idx), 't': t_idx, 'env': env_idx,
                'action': self.action[nxt, env_idx].contiguous(), 'reward': self.reward[nxt, env_idx].contiguous(),
                'done': self.done[nxt, env_idx].contiguous()}


def stacked(frames3):
    """uint8 [3, B, 3, 128, 128] oldest..newest -> the (9,128,128) stack per sample, newest first (processor.py:47-49)."""
    return torch.cat((frames3[2], frames3[1], frames3[0]), dim=1)


# ---------------------------------------------------------------------------------------------- agents
class Agent(object):
    """≙ rl_agents.Agent (rl_agents.py:19-127) for n envs at once.  `act` is ε-greedy per env with the reference's linear
    schedule (rl_agents.py:48-56); `iteration` counts calls, as there."""

    def __init__(self, env, weights=None, buffer_steps=64, batch_size=128, discount_rate=0.99, eps_min=0.0, eps_max=0.9,
                 eps_decay_steps=1000, seed=0, group=None, optimizer='sgd', l2_exclude=(), max_fit_batch=None,
                 pretrained_model_file=None, model_file=None):
        """pretrained_model_file: a Keras `.h5` of the reference (`load_model`, rl_agents.py:153-154) or an .npz —
        read by policy.load_weights; model_file: where `save_model` / `update_target_weights` write the network
        (`self.model.save(self.model_file)`, rl_agents.py:103), None = do not write."""
        self.lib = _lib.load()
        self.model_file, self.head_name = model_file, 'out'
        if pretrained_model_file is not None:
            from .policy import load_weights
            weights = load_weights(pretrained_model_file)
        self.env, self.device, self.n = env, env.device, env.n_envs
        self.n_actions = env.n_actions
        self.batch_size, self.discount_rate = int(batch_size), float(discount_rate)
        self.eps_min, self.eps_max, self.eps_decay_steps = eps_min, eps_max, eps_decay_steps
        self.iteration = 0
        self.group = group
        weights = keras_default_init() if weights is None else weights
        # model.fit on the device: flat float32 weights in Keras layouts + hand-written forward / backward / optimizer kernels;
        # kernel_regularizer=regularizers.l2(0.0001) on conv1..dense2 and 'out' (models.py:17,22-62) is part of its step
        self.trainer = NativeTrainer(weights, max_batch=max(self.batch_size, int(max_fit_batch or 0)), device=self.device.index,
                                     optimizer=optimizer, l2=L2_REGULARIZER, l2_exclude=l2_exclude)
        self.policy = PolicyNetwork(weights, max_batch=max(self.n, self.batch_size), device=self.device.index)
        self.replay = ReplayStore(env, buffer_steps)  # a view of the ring when the env has a deep one (then capacity = its depth)
        self.gen = torch.Generator(device=self.device)
        self.gen.manual_seed(int(seed))
        self.seed, self.env_id_base = int(seed), int(getattr(env, 'env_id_base', 0))
        self.action_distribution = torch.zeros(self.n_actions, dtype=torch.int64, device=self.device)
        self.loss_buffer = []
        self.logits = torch.zeros((self.n, 10), dtype=torch.float32, device=self.device)
        self.greedy = torch.zeros(self.n, dtype=torch.uint8, device=self.device)

    def _agree(self, value, op='min'):
        """An integer every rank must agree on before issuing collectives (how many fit batches, whether to fit at all):
        the MIN (or MAX) over the ranks of `group`; a no-op on one rank."""
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(self.group) == 1:
            return int(value)
        t = torch.tensor([int(value)], dtype=torch.int64, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.MIN if op == 'min' else dist.ReduceOp.MAX, group=self.group)
        return int(t.item())

    def epsilon(self):
        return max(self.eps_min, self.eps_max - (self.eps_max - self.eps_min) * self.iteration / self.eps_decay_steps)

    def act(self):
        """≙ Agent.act (rl_agents.py:92-101) on the env's current ring: forward kernels, then ε-greedy per env."""
        self.iteration += 1
        self.policy.forward(self.env, self.logits, self.greedy)
        action = torch.empty_like(self.greedy)
        _lib.check(self.lib.ahk_rl_epsilon_greedy(_ptr(self.greedy), self.n, float(self.epsilon()), self.seed, self.env_id_base,
                                                  self.iteration & 0xFFFFFFFF, _ptr(action), _ptr(self.action_distribution),
                                                  _stream(self.device)))
        return action

    def save_model(self, path=None):
        """≙ `self.model.save(self.model_file)` (rl_agents.py:103): the current weights and BatchNorm moving statistics in
        Keras' `save_weights` HDF5 layout (policy.save_weights_h5), which `model.load_weights` reads on the reference side."""
        from .policy import save_weights_h5, save_weights_npz
        path = path or self.model_file
        if path is None:
            raise ValueError('no model_file given')
        weights = self.trainer.keras_weights()
        if path.endswith('.npz'):
            save_weights_npz(path, weights)
        else:
            save_weights_h5(path, weights, head=self.head_name)
        return path

    def push_weights(self, policy=None):
        """The fitted network (BatchNorm moving statistics included) -> the forward kernels: folded and packed on the device
        from the trainer's flat weight vector (ahk_policy_set_weights_dev), stream-ordered, nothing crosses PCIe."""
        self.trainer.push(policy or self.policy)

    def _fit(self, frames3, backward_from):
        """One `model.fit` batch (rl_agents.py:189,276) on uint8 frames [3, B, 3, 128, 128] (ring layout, head 2): training-mode
        forward, `backward_from(logits) -> (dlogits, loss)`, backward, gradient all-reduce over ranks, optimizer step (L2
        regulariser, clipnorm, BatchNorm moving statistics), weights pushed to the forward kernels."""
        logits = self.trainer.forward(frames3, 2)
        dlogits, loss = backward_from(logits)
        self.trainer.backward(dlogits)
        self.trainer.all_reduce(self.group)
        self.trainer.step(loss)
        self.push_weights()
        return loss


class DDQNAgent(Agent):
    """≙ rl_agents.DDQNAgent (rl_agents.py:129-197): linear Dense(10) head, mean-squared error, SGD(lr 0.0075,
    momentum 0.5, decay 1e-6, clipnorm 2); target network = a second set of forward-kernel weights."""

    def __init__(self, env, weights=None, replace_softmax=True, **kw):
        if kw.get('pretrained_model_file') is not None:
            from .policy import load_weights
            weights = load_weights(kw.pop('pretrained_model_file'))
        if replace_softmax:  # rl_agents.py:160-172: a freshly initialised linear output layer (Glorot uniform, zero bias)
            weights = dict(keras_default_init() if weights is None else weights)
            fresh = keras_default_init(seed=kw.get('seed', 0) + 1)
            weights['out_w'], weights['out_b'] = fresh['out_w'], fresh['out_b']
        # the new Dense(10, activation='linear') is built without a kernel_regularizer (rl_agents.py:166)
        super().__init__(env, weights, optimizer='sgd', l2_exclude=('out_w',) if replace_softmax else (), **kw)
        self.target_policy = PolicyNetwork(self.trainer.keras_weights(), max_batch=self.batch_size, device=self.device.index)
        self.use_target = False
        if replace_softmax:
            self.head_name = 'output'  # the name of the new layer (rl_agents.py:166)

    def store_experience(self, action, reward, done):
        self.replay.record(self.env, action, reward, done)

    def update_target_weights(self):
        """≙ Agent.update_target_weights (rl_agents.py:103-107)."""
        self.push_weights(self.target_policy)
        self.use_target = True
        if self.model_file is not None and self.env_id_base == 0:  # one writer: the rank that owns env 0
            self.save_model()

    def train(self):
        """≙ DDQNAgent.train (rl_agents.py:177-191).  -> loss (float) or None if nothing can be sampled yet."""
        batch = self.replay.sample(self.batch_size, self.gen)
        if not self._agree(batch is not None):  # every rank fits or none does: a fit issues collectives
            return None
        g, b = batch['frames'], int(batch['t'].numel())
        q = self.policy.forward_ring(g[0:3], 2)[0]
        q_next = self.policy.forward_ring(g[1:4], 2)[0]
        q_next_target = (self.target_policy if self.use_target else self.policy).forward_ring(g[1:4], 2)[0]
        targets = torch.empty_like(q)
        loss = torch.zeros(1, dtype=torch.float32, device=self.device)

        def backward_from(q_fit):
            d = torch.empty_like(q_fit)
            _lib.check(self.lib.ahk_rl_ddqn_targets(_ptr(q), _ptr(q_next), _ptr(q_next_target), _ptr(batch['action']),
                                                    _ptr(batch['reward']), _ptr(batch['done']), b, self.discount_rate,
                                                    _ptr(q_fit), _ptr(targets), _ptr(d), _ptr(loss), _stream(self.device)))
            return d, loss

        self._fit(g[0:3], backward_from)
        self.targets = targets
        self.loss_buffer.append(float(loss.item()))
        return self.loss_buffer[-1]


class PGAgent(Agent):
    """≙ rl_agents.PGAgent (rl_agents.py:199-286): softmax output, categorical cross-entropy against
    prediction + learning_rate * (onehot - prediction) * standardised discounted return; the pretrained model keeps its
    compiled optimizer, Adam with the Keras defaults (models.py:119)."""

    def __init__(self, env, weights=None, learning_rate=0.01, episode_length=256, fit_batch=32, **kw):
        kw.setdefault('buffer_steps', episode_length + 3)
        if getattr(env, 'ring_depth', 3) > 3 and env.ring_depth < episode_length + 3:
            raise ValueError('PGAgent needs ring_depth >= episode_length + 3 to keep an episode in a deep ring')
        super().__init__(env, weights, optimizer='adam', max_fit_batch=fit_batch, **kw)
        self.learning_rate, self.episode_length, self.fit_batch = float(learning_rate), int(episode_length), int(fit_batch)
        self.prob = torch.zeros((self.replay.capacity, self.n, 10), dtype=torch.float32, device=self.device)

    def act(self):
        action = super().act()
        self._prediction = torch.softmax(self.logits, dim=1)  # self.prediction, rl_agents.py:99
        return action

    def store_experience(self, action, reward, done):
        """≙ PGAgent.store_experience (rl_agents.py:246-253); the prediction of the state the action was taken in is
        filed with the step's slot."""
        self.prob[self.env.head if self.replay.view else self.replay.cursor].copy_(self._prediction)  # the slot record() fills
        self.replay.record(self.env, action, reward, done)

    def train(self):
        """≙ PGAgent.train (rl_agents.py:255-282) over everything recorded since the last call; -> mean loss."""
        rp = self.replay
        steps = rp.filled
        if steps < 4:
            return None
        order = (rp.cursor - steps + torch.arange(steps, device=self.device)) % rp.capacity  # slots, oldest first
        # rewards of steps 1..steps-1 (slot 0 is the observation the rollout started from); an episode also ends where the
        # engine armed an auto-reset without a goal (max_episode_steps ≙ train.py:86 `episode_step < episode_length`)
        reward = rp.reward[order[1:]].contiguous()
        ends = (rp.done[order[1:]] | rp.pending[order[1:]]).contiguous()
        returns = torch.empty_like(reward)
        _lib.check(self.lib.ahk_rl_pg_returns(_ptr(reward), _ptr(ends), steps - 1, self.n, self.discount_rate, _ptr(returns),
                                              _stream(self.device)))
        # sample (k, e): the action of step k was taken in the state whose newest frame sits one slot earlier
        ok = valid_transitions(rp.pending, rp.cursor, rp.filled, need_next=True, stale=rp.stale)[order]
        k_idx, e_idx = ok[:-1].nonzero(as_tuple=True)  # state slot = order[k]; its action/reward/prob are at order[k+1]
        perm = torch.randperm(k_idx.numel(), device=self.device, generator=self.gen)  # fit(shuffle=True)
        # every fit batch issues two collectives (gradients, BatchNorm statistics) and the number of valid transitions
        # depends on each rank's own resets and goals: all ranks run the same number of batches, the smallest any of them
        # has (a rank's surplus samples — less than one episode's worth — are dropped)
        n_batches = self._agree(-(-int(perm.numel()) // self.fit_batch))
        losses = []
        for lo in range(0, n_batches * self.fit_batch, self.fit_batch):
            sel = perm[lo:lo + self.fit_batch]
            k, e = k_idx[sel], e_idx[sel]
            g = rp.gather(order[k], e)
            prob = self.prob[order[k + 1], e].contiguous()
            act = rp.action[order[k + 1], e].contiguous()
            ret = returns[k, e].contiguous()  # returns[k] belongs to step k+1, the one taken from state k
            loss = torch.zeros(1, dtype=torch.float32, device=self.device)

            def backward_from(logits):
                p_fit = torch.softmax(logits, dim=1).contiguous()
                d = torch.empty_like(p_fit)
                _lib.check(self.lib.ahk_rl_pg_targets(_ptr(prob), _ptr(act), _ptr(ret), int(sel.numel()), self.learning_rate,
                                                      _ptr(p_fit), None, _ptr(d), _ptr(loss), _stream(self.device)))
                return d, loss

            self._fit(g[0:3], backward_from)
            losses.append(loss)
        rp.reset()  # rl_agents.py:280
        if not losses:
            return None
        self.loss_buffer.append(float(torch.cat(losses).mean().item()))
        return self.loss_buffer[-1]


def run_training(agent, steps, training_start=100, training_interval=25, copy_interval=2500):
    """≙ the loop of ai/reinforcement/train.py:76-118 with every env stepping in lockstep: act, step, store; a DDQN agent
    trains every `training_interval` iterations after `training_start` and copies to the target network every
    `copy_interval`; a PG agent trains every `episode_length` steps.  -> list of losses."""
    env, losses = agent.env, []
    agent.replay.record_observation(env)
    for _ in range(int(steps)):
        action = agent.act()
        _, reward, done, _ = env.step(action)
        agent.store_experience(action, reward, done)
        if isinstance(agent, DDQNAgent):
            if agent.iteration > training_start and agent.iteration % training_interval == 0:
                losses.append(agent.train())
            if agent.iteration > copy_interval and agent.iteration % copy_interval == 0:
                agent.update_target_weights()
        elif agent.replay.filled >= agent.episode_length + 1:
            losses.append(agent.train())
            agent.replay.record_observation(env)
    return [l for l in losses if l is not None]
