"""Build the native pieces in-tree (nvcc -gencode arch=compute_100a,code=sm_100a; g++ for the host facade)."""
from __future__ import annotations

import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))


def _run(cmd, cwd=None):
    p = subprocess.run(cmd, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if p.returncode != 0:
        raise RuntimeError(f"command failed: {' '.join(cmd)}\n{p.stdout}")
    return p.stdout


def build_library(force: bool = False) -> str:
    """Compile csrc/*.cu into defect-gnn-cpp_b200/libdgnn_b200.so (cross-compiles without a GPU)."""
    csrc = os.path.join(HERE, "csrc")
    if force:
        _run(["make", "-C", csrc, "clean"])
    _run(["make", "-C", csrc, f"-j{max(1, min(8, os.cpu_count() or 1))}", "all"])
    out = os.path.join(HERE, "libdgnn_b200.so")
    if not os.path.exists(out):
        raise RuntimeError("libdgnn_b200.so was not produced")
    return out


def build_host_facade(force: bool = False) -> str:
    """Compile the C++ facade (host/) into libdefect_gnn_graph.so, linked against libdgnn_b200.so."""
    host = os.path.join(HERE, "host")
    if not os.path.exists(os.path.join(host, "Makefile")):
        return ""
    if force:
        _run(["make", "-C", host, "clean"])
    _run(["make", "-C", host, "all"])
    return os.path.join(HERE, "libdefect_gnn_graph.so")
