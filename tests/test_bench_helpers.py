"""bench.py's workload helpers (no GPU): the frame order of a step and the config dict both arms print."""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def test_frame_order_walks_the_distinct_frames_back_and_forth():
    bench = importlib.import_module("bench")
    for F in (1, 4, 16, 31, 64, 128):
        order = bench.frame_order(F)
        n = min(F, bench.DISTINCT_FRAMES)
        assert len(order) == F and min(order) == 0 and max(order) == n - 1
        # every consecutive pair of a step is a real consecutive pair of the synthetic sequence (or the same frame)
        assert all(abs(a - b) == 1 for a, b in zip(order, order[1:])) or n == 1
        # a longer step is the same workload: every distinct frame is visited, interior frames about twice as often
        counts = [order.count(k) for k in range(n)]
        assert min(counts) >= 1 and max(counts) - min(counts) <= max(2, F // n)


def test_both_arms_describe_the_same_workload():
    bench = importlib.import_module("bench")
    a = bench.workload_config(1920, 1200, 64, 1)
    b = bench.workload_config(1920, 1200, 64, 8)
    assert a == b   # nothing rank- or arm-dependent in `config`; those details go to `run_detail`
    assert a["distinct_frames"] == 16 and a["frames_per_step_per_gpu"] == 64 and "workload" in a
