"""mpcgpu_b200 -- B200-native (sm_100a) linear-system hot path of MPCGPU:
Schur-complement formation (S, symmetric-stair preconditioner, gamma) and the
block-tridiagonal PCG solve, behind the reference's pluggable linear-solver interface.

Layout: csrc/ (CUDA kernels + C ABI -> libmpcgpu_b200.so), api.py (host-side mirror of the
reference interface over that ABI), shim/ (drop-in headers for the reference's own
include/pcg/sqp.cuh), synth.py (seeded synthetic KKT blocks).
"""
from . import synth  # noqa: F401
from .api import (HostSolver, KnotSplitGroup, MpcGpuError, PcgConfig, epoch_stride,  # noqa: F401
                  compute_dz, form_schur_system, kkt_solve, kkt_solve_dz, knot_range, knotsplit_pcg, launch_count, lib,
                  mailbox_bytes, pcg, schur_rows_local, schur_slice, sizes, knotsplit_last_timeout)

__all__ = ["HostSolver", "MpcGpuError", "PcgConfig", "form_schur_system", "kkt_solve", "kkt_solve_dz", "compute_dz", "pcg",
           "sizes", "synth", "lib", "launch_count"]
