"""B200-native cost-propagation core for the FD*, MS-DFM and MFD* grid replanners.

The product is the CUDA library behind include/tasi_gpu.h (csrc/, built by build.py); planner.py mirrors the
reference's planner interface on top of it. Importing the package does not load the library; creating a
planner does, and raises if it (or a CUDA device) is missing.
"""
from .planner import (DFM, FD, MFD, DFMPlanner, FieldDPlanner, FieldView, GpuPlanner, ShiftedGridPlanner,  # noqa: F401
                      ShardedPlanner, TasiError, PLANNERS, LOOP_OK, LOOP_FAILURE_NO_GRAPH, LOOP_FAILURE_NO_GOAL)
