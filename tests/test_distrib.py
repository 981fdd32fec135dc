"""Distributed sort: host-side plan on CPU (gloo, world size 2 and 3) and the
full NCCL/NVLink path on GPUs (needs >= 2 devices)."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "dist_worker.py")


def torchrun(nproc, port, *args, timeout=300, env=None):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc),
           "--master-addr", "127.0.0.1", "--master-port", str(port), WORKER] + [str(a) for a in args]
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT, env=e)


def test_plan_functions(pkg):
    import importlib
    plan = importlib.import_module(pkg.__name__ + ".distrib_plan")
    s = np.arange(0, 1000, dtype=np.uint32)
    spl = plan.choose_splitters(s, 4)
    assert spl.tolist() == [250, 500, 750]
    d = plan.destination(np.array([0, 249, 250, 999, 500], np.uint32), spl)
    assert d.tolist() == [0, 0, 1, 3, 2]          # a key equal to a splitter goes right (SplitBinner::bin)
    # ties: a value spanning several quantiles is cut by position; all-equal input balances
    same = np.full(4000, 7, np.uint32)
    spl, thr = plan.choose_splitters(same, 4, with_ties=True)
    assert spl.tolist() == [7, 7, 7] and thr.tolist() == [1 << 30, 1 << 31, 3 << 30]
    d = plan.destination(same, spl, thr)
    assert np.abs(np.bincount(d, minlength=4) - 1000).max() <= 2          # low-discrepancy position hash
    mixed = np.sort(np.concatenate([np.arange(100, dtype=np.uint32), np.full(700, 100, np.uint32),
                                    np.arange(101, 301, dtype=np.uint32)]))
    spl, thr = plan.choose_splitters(mixed, 4, with_ties=True)
    assert spl.tolist() == [100, 100, 100] and thr[0] > 0
    d = plan.destination(mixed, spl, thr)
    assert np.all(np.diff(np.array([mixed[d == r].max() for r in range(4)]).astype(np.int64)) >= 0)   # ranges stay ordered
    assert np.bincount(d, minlength=4).max() <= 260                                                   # and balanced
    spl, thr = plan.choose_splitters(np.arange(1000, dtype=np.uint32), 4, with_ties=True)
    assert thr.tolist() == [0, 0, 0]                                                                  # no ties, no thresholds
    # step-wise mode: bases are bucket-major, rank-minor; positions keep the slice's order inside a digit
    cm = np.zeros((2, 256), dtype=np.int64)
    cm[0, 1], cm[0, 3], cm[1, 1], cm[1, 2] = 2, 1, 1, 4
    assert plan.step_bases(cm, 0)[[1, 2, 3]].tolist() == [0, 3, 7] and plan.step_bases(cm, 1)[[1, 2, 3]].tolist() == [2, 3, 8]
    pos = plan.step_positions(np.array([3, 1, 1], np.uint32), 0, 8, plan.step_bases(cm, 0))
    assert pos.tolist() == [7, 0, 1]
    owner, off = plan.slice_of(np.array([0, 3, 4, 7]), [0, 4, 8])
    assert owner.tolist() == [0, 0, 1, 1] and off.tolist() == [0, 3, 0, 3]
    off, tot = plan.recv_layout([[1, 2], [3, 4]])
    assert off.tolist() == [[0, 0], [1, 2]] and tot.tolist() == [4, 6]
    assert plan.imbalance([4, 6]) == pytest.approx(1.2)
    with pytest.raises(ValueError):
        plan.choose_splitters(s, 17)


@pytest.mark.parametrize("world", [2, 3])
def test_plan_over_gloo(world):
    r = torchrun(world, 29731 + world, "gloo", 20000)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert r.stdout.count("gloo plan ok") == world


@pytest.mark.gpu
@pytest.mark.parametrize("plan", ["auto", "counts", "cursor"])
def test_distributed_sort_on_gpus(plan):
    """Every plan of the NVLink exchange: histogram splitters with automatic fall-back to sampled ones,
    sampled splitters + counted layout, remote cursor atomics."""
    import torch
    ndev = torch.cuda.device_count()
    if ndev < 2:
        pytest.skip("needs >= 2 GPUs")
    world = 2 if ndev < 4 else 4
    r = torchrun(world, {"auto": 29741, "counts": 29743, "cursor": 29745}[plan], "nccl", 300000, timeout=600, env={"LIBSORT_EXCHANGE_PLAN": plan})
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert r.stdout.count(" ok; keys/rank") == 10
    assert r.stdout.count("stepwise") == 3
