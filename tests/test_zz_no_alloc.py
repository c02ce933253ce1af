"""ytk_gbdt_partition_rows_static: the one entry point that allocates nothing.

It is for hosts that cannot call cudaMalloc where they need a row partition -- all device memory
belongs to their own pool, or they cannot take the device-wide stall an allocation implies -- so its
work space is a fixed array inside the library, shared by every call.  The test drives it through the
C ABI on caller-owned device buffers and compares it with the allocating entry point
(ytk_gbdt_partition_rows): the row lists and the left count are integers, so they must match exactly.
Two calls in a row, the second smaller than the first, show that nothing left in the shared work
space by one call leaks into the next.  (Collected last on purpose: a failure here must not hide any
other test of a run that stops at the first failure.)
"""
import ctypes as C

import numpy as np
import pytest

from gpu_common import maker, synth, yb

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("gpu_required")]

KW = dict(max_depth=6, max_leaf_cnt=-1, grow_policy="level", min_child_hessian_sum=1.0,
          min_split_loss=0.0, l1=0.0, l2=1.0, learning_rate=0.1)
MAX_ROWS = 1 << 22


def _static_partition(bins_dev, stride, rows, fid, cond):
    import torch
    rows_dev = torch.from_numpy(np.ascontiguousarray(rows, dtype=np.int32)).cuda()
    out_dev = torch.full_like(rows_dev, -1)
    left_dev = torch.full((1,), -1, dtype=torch.int64, device="cuda")
    stream = torch.cuda.current_stream()
    rc = yb.lib().ytk_gbdt_partition_rows_static(
        C.c_void_p(bins_dev.data_ptr()), stride, C.c_void_p(rows_dev.data_ptr()), len(rows), fid,
        C.c_float(cond), C.c_void_p(out_dev.data_ptr()), C.c_void_p(left_dev.data_ptr()),
        C.c_void_p(stream.cuda_stream))
    if rc != 0:
        raise yb.YtkError(rc, yb.lib().ytk_gbdt_last_error().decode())
    stream.synchronize()
    return out_dev.cpu().numpy(), int(left_dev.item())


def test_static_partition_matches_allocating_entry_point_twice():
    import torch
    n, F, stride = 300000, 28, 32
    x, start, gp = synth(n, F, 255, 17, skew=True)
    tm = maker(KW, x, start, gp)                         # the existing, allocating path
    padded = np.zeros((n, stride), dtype=np.uint8)      # the caller's own device copy of the slots
    padded[:, :F] = x[:, :F]
    bins_dev = torch.from_numpy(padded).cuda()
    rng = np.random.default_rng(5)
    calls = [(rng.permutation(n)[:250001].astype(np.int32), 3, 100.5),
             (np.sort(rng.permutation(n)[:4097]).astype(np.int32), 27, 0.5)]   # smaller: stale scratch must not matter
    for rows, fid, cond in calls:
        want_rows, want_left = tm.partition_rows(rows, fid, cond)
        got_rows, got_left = _static_partition(bins_dev, stride, rows, fid, cond)
        assert got_left == want_left
        assert np.array_equal(got_rows, want_rows)
    # degenerate sizes
    got_rows, got_left = _static_partition(bins_dev, stride, np.zeros(0, np.int32), 0, 1.0)
    assert got_left == 0 and got_rows.size == 0


def test_static_partition_refuses_more_than_its_fixed_work_space():
    import torch
    bins_dev = torch.zeros((8, 32), dtype=torch.uint8, device="cuda")
    dummy = torch.zeros(8, dtype=torch.int32, device="cuda")
    left = torch.zeros(1, dtype=torch.int64, device="cuda")
    rc = yb.lib().ytk_gbdt_partition_rows_static(
        C.c_void_p(bins_dev.data_ptr()), 32, C.c_void_p(dummy.data_ptr()), MAX_ROWS + 1, 0, C.c_float(1.0),
        C.c_void_p(dummy.data_ptr()), C.c_void_p(left.data_ptr()), None)
    assert rc == -3
    msg = yb.lib().ytk_gbdt_last_error().decode()
    assert "4194304" in msg and "exceed" in msg
