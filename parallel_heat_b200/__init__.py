"""parallel_heat_b200 -- B200-native solver loop for manospits/parallel_heat.

csrc/     hand-written sm_100a CUDA kernels + the C ABI (include/heat.h)
api.py    ctypes mirror of that C ABI (reference parameter names)
strips.py one-process-per-GPU row-strip driver over torch.distributed
build.py  in-tree nvcc build of libheat_b200.so
"""
from .api import (F32_FMA, F32_NOFMA, GRID_F32, GRID_F64, MPI_F64MIX, PHASE_CUDA, PHASE_MPI, STEPS_CUDA, STEPS_MPI,  # noqa: F401
                  HeatError, HeatParams, HeatSolver, default_params, device_count, group_enqueue, group_run,
                  inidat, inidat3d, inidat_f64, load, prtdat, prtdat_bin, readdat_bin, solve, solve_gpus, step_device, step_residual_device,
                  strip_rows)

__all__ = ["HeatSolver", "HeatParams", "HeatError", "solve", "solve_gpus", "inidat", "prtdat", "prtdat_bin", "readdat_bin",
           "strip_rows", "default_params", "device_count", "load", "group_run", "group_enqueue",
           "step_device", "step_residual_device",
           "F32_FMA", "F32_NOFMA", "MPI_F64MIX", "GRID_F32", "GRID_F64", "inidat_f64", "inidat3d", "STEPS_CUDA", "STEPS_MPI", "PHASE_CUDA", "PHASE_MPI"]
