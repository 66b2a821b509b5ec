"""The decoder kernel's schedule tables (tm_build_tables, csrc/ldpc_kernels_tmem.cuh), checked on the host for every base
graph, lifting size and TMEM budget: every row exactly once and in order, pairs adjacent / orthogonal / of a compiled shape,
edge records equal to the base graph's adjacency, message words disjoint and inside the budget, layer limits mapped to the
shortest cover of whole steps - and above all that the builder accepts every graph, so that no (BG, Z) can fall back to the
general kernel silently. Needs nvcc (host compilation only), no GPU."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_schedule_tables_of_all_graphs(tmp_path):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not found")
    exe = tmp_path / "tm_schedule_check"
    subprocess.run([nvcc, "-std=c++17", "--expt-relaxed-constexpr", "-gencode", "arch=compute_100a,code=sm_100a",
                    "-I", os.path.join(ROOT, "srsran_project_23.5_b200", "csrc"), "-o", str(exe),
                    os.path.join(ROOT, "tests", "tm_schedule_check.cu")], check=True, capture_output=True, timeout=300)
    out = subprocess.run([str(exe)], capture_output=True, text=True, timeout=120).stdout.strip()
    assert out == "ok 204", out  # 2 base graphs x 51 lifting sizes x 2 TMEM budgets
