"""Stand-in for the reference's MISSING simulator module `air_hockey` — TEST INFRASTRUCTURE.

Implements exactly the contract the reference wrapper relies on (call sites
gym-air-hockey/gym_air_hockey/envs/air_hockey_env.py:5,11,67,71,80,89-93,98-118,153):
`AirHockey(video_file=None)`, `.reset(use_object=dict)`, `.step(robot_action, human_action, use_object) -> GameInfo`.
The game itself is oracle/ahk_oracle.c (SPEC.md).  With this on sys.path plus oracle/gym_shim, the reference's
own AirHockeyEnv + DataProcessor run unmodified; tests/golden/make_golden.py uses that to make the fixtures.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import orc  # noqa: E402

SEED = 0      # module-level so the golden script can choose before the reference wrapper constructs AirHockey()
ENV_ID = 0
TABLE = None  # uint8[700][450][3]; must be set before AirHockey() is constructed


class GameInfo(object):
    def __init__(self, frame, info):
        self.frame = frame
        self.scored = orc.SCORED[info.scored]
        self.robot_action = [int(info.robot_action[0]), int(info.robot_action[1])]
        self.human_action = [int(info.human_action[0]), int(info.human_action[1])]
        self.puck_was_hit = bool(info.puck_was_hit)
        self.hit_the_border = bool(info.hit_the_border)
        self.puck_is_at_the_bottom = bool(info.puck_is_at_the_bottom)
        self.distance_decreased = bool(info.distance_decreased)
        self.in_the_target = bool(info.in_the_target)


def _mask(use_object):
    if use_object is None:
        return orc.USE_ALL
    return (orc.USE_ARM if use_object['arm'] else 0) | (orc.USE_PUCK if use_object['puck'] else 0) | \
           (orc.USE_TOP if use_object['top_mallet'] else 0)


class AirHockey(object):
    def __init__(self, video_file=None):
        if TABLE is None:
            raise RuntimeError("set air_hockey.TABLE first")
        self.sim = orc.Sim(seed=SEED, env_id=ENV_ID, table=TABLE)
        self.sim.reset()
        self.video_file = video_file
        self.trace = []  # (state float32[12], GameInfo-ish) per step, for the fixtures

    def reset(self, use_object=None):
        self.sim.reset()

    def step(self, robot_action=None, human_action=None, use_object=None):
        use = _mask(use_object)
        info = self.sim.step(robot_action, human_action, use)
        frame = self.sim.draw(use)
        game_info = GameInfo(frame, info)
        self.trace.append((self.sim.state.copy(), use, info.scored))
        return game_info
