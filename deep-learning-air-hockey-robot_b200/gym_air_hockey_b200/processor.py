"""Host-side mirror of the reference's `DataProcessor` (gym-air-hockey/gym_air_hockey/processor.py:4-85).

The heavy half of the reference class — resize, transpose, 3-frame shift register, normalisation
(`process_observation`, processor.py:31-66) — is fused into the CUDA rasteriser, which writes uint8 CHW frames
straight into a ring in HBM; `BatchedAirHockeyEnv.observation()` materialises the reference's float form on
request.  What stays on the host is the small integer bookkeeping: the action table and the label formula.
"""
import numpy as np


class DataProcessor(object):
    MODES = ('rgb', 'gray-diff')  # ai/config.py:1

    def __init__(self, dim=128, mode='rgb'):
        if dim != 128:
            raise ValueError('the B200 engine renders 128x128 observations only (processor.py:5)')
        self.dim = dim
        # processor.py:7-18 — label -> (x, y) move; row 9 ("Undefined") is also "stand"
        self.actions = np.array([[-1, -1], [-1, 0], [-1, 1], [0, -1], [0, 0], [0, 1], [1, -1], [1, 0], [1, 1], [0, 0]],
                                dtype=np.int8)
        self.metrics = []
        self.metrics_names = []
        self.mode = mode

    def check_mode(self):
        if self.mode not in self.MODES:
            raise ValueError('Processing mode is not supported')  # processor.py:66

    def process_step(self, observation, reward, done, info):
        return observation, self.process_reward(reward), done, self.process_info(info)

    def process_reward(self, reward):
        return np.clip(reward, -1.0, 1.0)  # processor.py:68-69

    def process_info(self, info):
        return info

    def process_action(self, label):
        if label is None:
            return None  # processor.py:74-76: None = the programmed AI plays
        return self.actions[label]

    def process_state_batch(self, batch):
        _, depth, height, width = batch[0].shape
        return np.array(batch).reshape(len(batch), depth, height, width)

    def action_to_label(self, action):
        return int(action[0] * 3 + action[1] + 4)  # processor.py:84-85
