"""bench.py's layer split: the head GPU (which also runs the draft) takes the share of the target's layers that balances the work per
run; checked here against the splits the round-2 measurements ran with."""
import os
import sys

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)


def test_head_layers_of_the_headline_config():
    import bench
    t, d, _ = bench.CONFIGS["70b"]
    assert [bench.head_layers(n, t, d, 3)[0] for n in (2, 4, 8)] == [30, 5, 1]
    # a draft that costs more than the whole pipeline leaves the head one layer; a free draft gives the equal share
    t120, d120, _ = bench.CONFIGS["120b"]
    assert bench.head_layers(8, t120, d120, 8)[0] == 1
    assert bench.head_layers(4, t, d, 0)[0] == 20


def test_head_share_becomes_the_measured_ranges():
    import bench
    try:
        from pipeinfer_b200 import pipeline
        pipeline.split_layers(8, [0.5, 0.5])
    except Exception as e:                      # the library is not built
        pytest.skip(str(e))
    t, d, _ = bench.CONFIGS["70b"]
    want = {2: [(0, 30), (30, 80)], 4: [(0, 5), (5, 29), (29, 53), (53, 80)],
            8: [(0, 1), (1, 12), (12, 23), (23, 34), (34, 45), (45, 56), (56, 67), (67, 80)]}
    for n, ranges in want.items():
        l0, _ = bench.head_layers(n, t, d, 3)
        got = pipeline.split_layers(t[0], pipeline.default_weights(n, (l0 + 0.25) / t[0]))
        assert [tuple(r) for r in got] == ranges, (n, got)
