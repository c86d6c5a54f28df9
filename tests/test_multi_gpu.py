"""Partitioned run over NCCL (2 ranks) against the global-mesh oracle; needs two B200s."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("dim,n", [(3, 8), (2, 16)])
def test_two_rank_relaxation_matches_oracle(dim, n):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29517",
                          os.path.join(ROOT, "tests", "dist", "run_ma_2rank.py"), str(dim), str(n)],
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "DIST-OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]


@pytest.mark.parametrize("split_min", ["200000", "1"])
@pytest.mark.parametrize("dim,n", [(3, 20), (2, 96)])
def test_two_rank_multigrid_poisson(dim, n, split_min):
    """split_min = 1: the replicated level 1 of these small hierarchies is work-split over the ranks
    (row chunks + in-place allgather, csrc/mv_amg_cycle.cu), as happens by default from 200 000 rows on."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29519",
                          os.path.join(ROOT, "tests", "dist", "run_amg_2rank.py"), str(dim), str(n)],
                         capture_output=True, text=True, timeout=600, env=dict(os.environ, MV_AMG_SPLIT_MIN=split_min))
    assert out.returncode == 0 and "DIST-AMG-OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]


def test_two_rank_laplacian_spring_quasi_newton_match_oracle():
    """LaplacianSmoother, SpringMover ('+' cell by global cell number) and the quasi-Newton mover under a 2-way
    partition against the global-mesh oracle: coordinates within 1e-10, Newton iteration count +-1."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29521",
                          os.path.join(ROOT, "tests", "dist", "run_movers_2rank.py")],
                         capture_output=True, text=True, timeout=900)
    assert out.returncode == 0 and "DIST-MOVERS-OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]


def test_two_rank_peer_memory_timeout_is_reported_once_and_recovered_from():
    """A neighbour that is late by more than MV_P2P_TIMEOUT_S fails ONE exchange with MV_ERR_COMM on the waiting rank;
    the next exchanges succeed (csrc/mv_p2p.cu: wall-clock time-out, status cleared when reported)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29523",
                          os.path.join(ROOT, "tests", "dist", "run_p2p_timeout_2rank.py")],
                         capture_output=True, text=True, timeout=600, env=dict(os.environ, MV_P2P_TIMEOUT_S="2"))
    assert out.returncode == 0 and "DIST-P2P-TIMEOUT-OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]
