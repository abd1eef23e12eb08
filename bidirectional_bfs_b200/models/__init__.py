"""Search variants ("models"): one class per way of answering an s-t query.

  serial      SerialSolver     v1 equivalent       (v1/main-v1.cpp)
  cpu-ranks   CpuRanksSolver   v2 equivalent       (v2/second_try.cpp)
  gpu         GpuSolver        v3 replacement      (v3/bibfs_cuda_only.cu)
  multi-gpu   MultiGpuSolver   v4 replacement      (v4/mpi_bas.cpp + v4/comp.cu)
  gpu-resident ResidentSolver  one resident kernel fed through a pinned mailbox: no launch per query (no counterpart)
  batch       BatchSolver      many queries in one launch (no counterpart in the reference)
  naive-gpu   NaiveGpuSolver   reference-style level-synchronous baseline, for comparisons only
"""
from .solvers import (BatchSolver, CpuRanksSolver, GpuSolver, MultiGpuSolver, NaiveGpuSolver,  # noqa: F401
                      ResidentSolver, SerialSolver, make_solver, VARIANTS)
