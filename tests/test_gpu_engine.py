"""The device-planned level loop (csrc/engine.cu): one enqueue and one host wait per tree, replayed from a
recorded CUDA graph on one GPU.  Its trees must be the host-planned path's, bit for bit (that path is
checked against the oracle everywhere else, and so is this one wherever it is the default)."""
import numpy as np
import pytest

from gpu_common import maker, synth, yb

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("gpu_required")]

KW = dict(max_depth=7, max_leaf_cnt=-1, grow_policy="level", min_child_hessian_sum=1.0,
          min_split_loss=0.0, l1=0.0, l2=1.0, learning_rate=0.1)


def _same(a, b):
    assert a.n_nodes == b.n_nodes
    for name, _ in yb.Tree.FIELDS:
        assert np.array_equal(getattr(a, name), getattr(b, name), equal_nan=True), name


@pytest.mark.parametrize("kw", [KW, dict(KW, max_depth=1), dict(KW, max_depth=10, min_split_loss=0.5),
                                dict(KW, max_depth=4, max_leaf_cnt=16)], ids=["depth7", "depth1", "depth10_minloss", "full_budget"])
def test_engine_graph_replay_and_host_planned_path_agree(kw, monkeypatch):
    n, F, K = 60000, 28, 2
    x, start, gp = synth(n, F, 255, 17, K=K, skew=True)
    a = maker(kw, x, start, gp, K=K)                      # device-planned, CUDA graph
    trees, pos = [], []
    for gid in (0, 1, 0, 0):                              # recorded once per class, then replayed
        trees.append(a.make(gid))
        pos.append(a.get_positions())
        assert a.last_stats.batches == 0                  # no host wait between the root and the last level
    assert a.last_stats.graph_replayed == 1
    _same(trees[0], trees[2]); _same(trees[0], trees[3])
    mask = (np.random.default_rng(1).random(n) < 0.6).astype(np.uint8)
    fea = np.array([1, 4, 9, 20, 27], np.int32)
    t_mask = a.make(1, fea_index=fea, row_mask=mask)      # a row mask is queued launch by launch
    p_mask = a.get_positions()
    assert a.last_stats.graph_replayed == 0
    _same(a.make(0), trees[0])                            # ... and the recorded sequence still replays afterwards
    monkeypatch.setenv("YTK_GBDT_NO_GRAPH", "1")
    b = maker(kw, x, start, gp, K=K)                      # device-planned, queued launch by launch
    monkeypatch.setenv("YTK_GBDT_NO_ENGINE", "1")
    c = maker(kw, x, start, gp, K=K)                      # host-planned
    for other in (b, c):
        for i, gid in enumerate((0, 1)):
            _same(other.make(gid), trees[i])
            assert np.array_equal(other.get_positions(), pos[i])
        _same(other.make(1, fea_index=fea, row_mask=mask), t_mask)
        assert np.array_equal(other.get_positions(), p_mask)
    assert b.last_stats.graph_replayed == 0 and c.last_stats.batches > 0


def test_kernel_time_spans_agree_with_cuda_events(monkeypatch):
    """gpu_hist_ms of the device-planned loop is the histogram kernels' own span on the GPU's global timer (the
    only clock a graph replay can be read with); queued launch by launch the same launches are also bracketed
    by CUDA events, and the two must tell the same time."""
    monkeypatch.setenv("YTK_GBDT_NO_GRAPH", "1")
    n, F = 2_000_000, 28
    rng = np.random.default_rng(3)
    x = rng.integers(0, 255, size=(n, F), dtype=np.uint8)
    start = (np.arange(F + 1) * 255).astype(np.int32)
    y = (rng.random(n) < 0.4)
    gp = np.stack([0.5 - y, np.full(n, 0.25)], 1).astype(np.float32).reshape(-1)
    tm = yb.GpuTreeMaker(yb.make_params(**dict(KW, max_depth=6, min_child_hessian_sum=10.0)), start, n)
    tm.upload_rows_u8(x)
    tm.set_grads(gp)
    for _ in range(3):
        tm.make(0)
    ts = tm.last_stats
    assert ts.stage_times_from_gpu == 1 and ts.hist_launches == 6
    assert ts.hist_events_ms > 0.05
    # the events also see what lies between a launch and its first CTA (a few us per launch), never less than the kernels
    assert ts.gpu_hist_ms <= ts.hist_events_ms + 0.005, (ts.gpu_hist_ms, ts.hist_events_ms)
    assert ts.hist_events_ms - ts.gpu_hist_ms <= 0.10 * ts.hist_events_ms + 0.010 * ts.hist_launches, \
        (ts.gpu_hist_ms, ts.hist_events_ms)
    assert ts.device_span_ms >= ts.gpu_hist_ms + ts.partition + ts.find_best_split
