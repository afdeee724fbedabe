"""Multi-GPU parity: runs tests/dist_check.py under torchrun when >= 2 GPUs are visible."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.parametrize("case,side", [("sedov", 40), ("noh", 40), ("turbulence", 36)])
def test_two_ranks_nccl_domain_equals_reference(case, side):
    """the C++ multi-rank domain over NCCL (sphb200_domain_*): every rank's owned particles == the reference's
    single-rank CPU run (tests/domain_check.py); the same code path runs on one GPU through the loopback transport in
    tests/test_gpu_domain_dist.py"""
    if _ngpu() < 2:
        pytest.skip("needs >= 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29541", os.path.join(ROOT, "tests", "domain_check.py"), "--case", case,
           "--side", str(side), "--steps", "3"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert res.returncode == 0 and "DOMAIN_CHECK PASS" in res.stdout, res.stdout[-3000:] + res.stderr[-3000:]


@pytest.mark.parametrize("case,side", [("sedov", 40), ("noh", 40), ("kh", 64)])
def test_two_ranks_equal_one_rank(case, side):
    if _ngpu() < 2:
        pytest.skip("needs >= 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tests", "dist_check.py"), "--case", case,
           "--side", str(side), "--steps", "3"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert res.returncode == 0 and "DIST_CHECK PASS" in res.stdout, res.stdout[-3000:] + res.stderr[-3000:]


def test_two_ranks_block_time_steps():
    """DistHydroVeBdtProp on 2 ranks == the one-rank tier-2 block-time-step propagator, bit for bit, over two
    hierarchies (full syncs with migration-capable reorder of the rung field, partial syncs with halo-only exchanges)"""
    if _ngpu() < 2:
        pytest.skip("needs >= 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29534", os.path.join(ROOT, "tests", "dist_check.py"), "--case", "sedov",
           "--side", "40", "--steps", "10", "--bdt"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert res.returncode == 0 and "DIST_CHECK PASS" in res.stdout, res.stdout[-3000:] + res.stderr[-3000:]


def test_two_ranks_stirred_turbulence():
    """turbulence-ve on 2 ranks (every rank advances the same OU sequence and stirs its own particles) == 1 rank;
    step_host with the stirring kernel per range == step; restart carries the driver state"""
    if _ngpu() < 2:
        pytest.skip("needs >= 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29535", os.path.join(ROOT, "tests", "dist_check.py"), "--case", "turbulence",
           "--side", "40", "--steps", "3", "--stir"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert res.returncode == 0 and "DIST_CHECK PASS" in res.stdout, res.stdout[-3000:] + res.stderr[-3000:]


def test_two_ranks_block_time_steps_with_stirring():
    """TurbVeBdtProp on 2 ranks == the one-rank tier-2 propagator with a driver attached (the active rungs are stirred
    after the acceleration time step, one OU step per substep on every rank); restart carries the driver state"""
    if _ngpu() < 2:
        pytest.skip("needs >= 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29536", os.path.join(ROOT, "tests", "dist_check.py"), "--case", "sedov",
           "--side", "40", "--steps", "10", "--bdt", "--stir"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert res.returncode == 0 and "DIST_CHECK PASS" in res.stdout, res.stdout[-3000:] + res.stderr[-3000:]
