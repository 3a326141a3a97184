"""The batching controller: unchanged per-structure tasks, one launch for all of them."""
import threading
import time

import pytest

from igm_b200 import kernel
from igm_b200.controller import B200Controller, PopulationBatcher


class _M(object):
    def __init__(self, uid):
        self.id = uid


@pytest.fixture
def fake_gpu(monkeypatch):
    calls = []

    def fake(models, cfg, device=None):
        calls.append([m.id for m in models])
        time.sleep(0.01)
        return [{"final-energy": float(m.id)} for m in models]

    monkeypatch.setattr(kernel, "optimize_population", fake)
    return calls


def _task(i, skip=(), fail=()):
    time.sleep(0.001 * (i % 5))                 # uneven host-side preparation
    if i in skip:
        return ("skipped", i)                    # e.g. the .ready file already exists (ModelingStep.py:167-169)
    if i in fail:
        raise ValueError("task %d" % i)
    info = kernel.optimize(_M(i), {"optimization": {}})
    return info["final-energy"]


def test_all_tasks_share_one_launch(fake_gpu):
    ctl = B200Controller(max_threads=64)
    out = ctl.map(_task, range(40))
    assert out == [float(i) for i in range(40)]
    assert len(fake_gpu) == 1 and sorted(fake_gpu[0]) == list(range(40))


def test_more_tasks_than_threads_and_skipped_tasks(fake_gpu):
    ctl = B200Controller(max_threads=8)
    skip = {3, 4, 17}
    out = ctl.map(lambda i: _task(i, skip=skip), range(30))
    assert [o for k, o in enumerate(out) if k not in skip] == [float(i) for i in range(30) if i not in skip]
    assert all(out[k] == ("skipped", k) for k in skip)
    assert sorted(i for b in fake_gpu for i in b) == [i for i in range(30) if i not in skip]
    assert max(len(b) for b in fake_gpu) <= 8 and len(fake_gpu) <= 8


def test_batch_size_limit(fake_gpu):
    ctl = B200Controller(max_threads=32, batch_size=10)
    ctl.map(_task, range(32))
    assert sorted(i for b in fake_gpu for i in b) == list(range(32))
    assert all(len(b) <= 32 for b in fake_gpu) and len(fake_gpu) >= 2


def test_kernel_failure_reaches_every_caller_and_task_errors_propagate(monkeypatch):
    def boom(models, cfg, device=None):
        raise RuntimeError("b200 kernel produced non-finite coordinates or energy (modelid: 1)")
    monkeypatch.setattr(kernel, "optimize_population", boom)
    with pytest.raises(RuntimeError, match="modelid"):
        B200Controller(max_threads=4).map(_task, range(4))
    with pytest.raises(ValueError, match="task 2"):
        B200Controller(max_threads=4).map(lambda i: _task(i, fail={2}), range(3))


def test_without_the_controller_optimize_is_a_batch_of_one(fake_gpu):
    assert kernel.optimize(_M(5), {"optimization": {}}) == {"final-energy": 5.0}
    assert fake_gpu == [[5]]


def test_batcher_alone():
    got = []
    b = PopulationBatcher(run=lambda models, cfg: [got.append(len(models)) or {"n": len(models)} for _ in models])
    b.set_workers(3)
    res = [None] * 3

    def w(k):
        res[k] = b.optimize(_M(k), {})
        b.worker_done()
    ts = [threading.Thread(target=w, args=(k,)) for k in range(3)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert res == [{"n": 3}] * 3 and b.batches == [3]
