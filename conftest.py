import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(ROOT, "deep-learning-air-hockey-robot_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)
