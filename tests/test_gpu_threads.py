"""The threaded deployment of the reference: one process, T trainer threads, each with its own tree maker
(J/worker/TrainWorker.java:167-200; INTEGRATION.md maps thread t to GPU t).  Handles must not share
hidden per-process state: per-device kernel attributes, the NCCL loader, error strings."""
import threading

import numpy as np
import pytest

import oracle_binding as ob
from gpu_common import assert_same_tree, maker, synth

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("gpu_required")]

KW = dict(max_depth=6, max_leaf_cnt=-1, grow_policy="level", min_child_hessian_sum=1.0,
          min_split_loss=0.0, l1=0.0, l2=1.0, learning_rate=0.1)


def _run(device, seed, out, idx):
    try:
        x, start, gp = synth(50000, 28, 255, seed, skew=True)
        tm = maker(KW, x, start, gp, device=device)
        trees = [tm.make(0) for _ in range(3)]
        pos = tm.get_positions()
        tm.close()
        out[idx] = (x, start, gp, trees, pos)
    except Exception as e:          # surfaced by the asserting thread
        out[idx] = e


def _check(res):
    for r in res:
        assert not isinstance(r, Exception), r
        x, start, gp, trees, pos = r
        o, opos, _ = ob.make_tree(ob.make_params(**KW), x, gp, 0, start)
        for t in trees:
            assert_same_tree(t, o)
        assert np.array_equal(pos, opos)


def test_two_threads_two_handles_on_one_device():
    res = [None, None]
    th = [threading.Thread(target=_run, args=(0, 11 + i, res, i)) for i in range(2)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    _check(res)


def test_one_process_one_thread_per_device():
    """The histogram kernel's shared-memory opt-in is per device: every device a handle is created on
    must be able to launch it."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    res = [None, None]
    th = [threading.Thread(target=_run, args=(i, 21 + i, res, i)) for i in range(2)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    _check(res)
