"""sspp_b200 — B200-native geometric-query engine behind the sspp planner API.

Only what the hot path needs lives here:
  csrc/      CUDA kernels (sm_100a) + the C ABI declared in include/sspp_gpu.h
  host/      C++ mirror of the reference's planner interface (PathPlanner / SearchSpace / Heuristic ...)
  engine.py  ctypes binding of the C ABI used by tests/ and bench.py
  scenes.py  synthetic scenes of SURVEY.md §8(d)
There is no CPU fallback: importing `engine` without the built CUDA library raises.
"""
from .build import build_all, LIB_GPU, LIB_HOST  # noqa: F401
