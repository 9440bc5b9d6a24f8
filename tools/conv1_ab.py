import os, sys, numpy as np
sys.path.insert(0, 'deep-learning-air-hockey-robot_b200'); sys.path.insert(0, 'oracle')
import torch, cnn_ref
from gym_air_hockey_b200 import BatchedAirHockeyEnv
from gym_air_hockey_b200.policy import PolicyNetwork, keras_default_init
from gym_air_hockey_b200.assets import make_table_image
w = keras_default_init(seed=5, randomize_bn=True)
n = 400
env = BatchedAirHockeyEnv(n, table=make_table_image(), seed=9, max_episode_steps=50)
env.reset()
for _ in range(7): env.step()
fr = env.frames.cpu().numpy(); head = env.head
obs = np.concatenate([fr[(head + 3 - age) % 3] for age in range(3)], axis=1)
ref32 = cnn_ref.forward(w, obs); ref16 = cnn_ref.forward(w, obs, emulate_bf16=True)
scale = np.abs(ref32).max()
out = {}
for mode in ('tmem', 'normal', 'smem'):
    os.environ['AHK_CONV1'] = mode
    net = PolicyNetwork(w, max_batch=512)
    logits, action = net.forward_ring(env.frames, head)
    torch.cuda.synchronize()
    l = logits.cpu().numpy(); out[mode] = l
    print(mode, 'max|l-ref32|/scale %.5f  max|l-ref16|/scale %.5f  mean|l-ref32|/scale %.6f' % (np.abs(l - ref32).max() / scale, np.abs(l - ref16).max() / scale, np.abs(l - ref32).mean() / scale))
print('tmem vs normal max diff / scale', np.abs(out['tmem'] - out['normal']).max() / scale, 'normal vs smem', np.abs(out['normal'] - out['smem']).max() / scale)

# --- the same at the benchmarked size: 8192 samples (110 work items per CTA), bitwise across kernels and across repeats
n = 8192
env = BatchedAirHockeyEnv(n, seed=3, max_episode_steps=50)
env.reset()
for _ in range(9): env.step()
big = {}
for mode in ('tmem', 'normal', 'smem'):
    os.environ['AHK_CONV1'] = mode
    net = PolicyNetwork(w, max_batch=n)
    runs = []
    for rep in range(6):
        logits, action = net.forward_ring(env.frames, env.head)
        torch.cuda.synchronize()
        runs.append(logits.clone())
    big[mode] = runs[0]
    print(mode, 'repeat-to-repeat max diff', max(float((r - runs[0]).abs().max()) for r in runs[1:]), 'rows differing', int(((runs[-1] != runs[0]).any(1)).sum()))
for a, b in (('tmem', 'normal'), ('normal', 'smem')):
    d = (big[a] - big[b]).abs()
    print(a, 'vs', b, 'max diff', float(d.max()), 'rows differing', int((d > 0).any(1).sum()))
