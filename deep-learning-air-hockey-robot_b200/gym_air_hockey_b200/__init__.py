"""gym_air_hockey_b200 — the reference's `gym_air_hockey` package (gym-air-hockey/gym_air_hockey/__init__.py:1-8)
re-hosted on a B200: same Env API, the step path in hand-written sm_100a CUDA behind a C ABI (include/ahk.h)."""
from .processor import DataProcessor  # noqa: F401
from .env import AirHockeyEnv, BatchedAirHockeyEnv, make  # noqa: F401
