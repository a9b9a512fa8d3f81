"""defect-gnn-cpp_b200 -- B200-native crystal-graph construction (periodic radius-neighbour search + Gaussian RBF
edge featurisation), a drop-in for the graph layer of bamarler/Defect-GNN-cpp.

Layout
  csrc/      CUDA kernels (sm_100a) + the C ABI (include/dgnn_cabi.h)  -> libdgnn_b200.so
  host/      C++17 facade keeping the reference's class API (NeighborList, gaussian_rbf, CrystalGraph)
  binding.py ctypes plumbing over the C ABI for tests / bench (no compute in Python, no CPU fallback)
  sharding.py structure-sharded multi-GPU helpers (no collective on the data path)
  synthetic.py seeded workload generators for the BASELINE configs

The directory name contains '-' so it is loaded by path; use `__graft_entry__.load_package()`.
"""
from . import binding, sharding, synthetic  # noqa: F401
from .binding import GraphBuilder, DgnnError, UNCAPPED  # noqa: F401
from .build import build_library, build_host_facade  # noqa: F401
