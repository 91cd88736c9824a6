"""GPU: the matching.py-shaped driver end to end (batched GPU matching of all view-graph edges + host
walk) against the restated reference loop driven by the cv2 oracle matcher, and the reference loop
itself driven by the product's LightGlue-shaped callable (the literal drop-in at matching.py:122)."""
import importlib
import os

import numpy as np
import pytest

from oracle import oracle_driver
from tests.scene import make_scene

pytestmark = pytest.mark.gpu


def _same(a, b):
    pa, ma = a
    pb, mb = b
    assert [tuple(map(int, p)) for p in pa] == [tuple(map(int, p)) for p in pb]
    assert len(ma) == len(mb)
    for x, y in zip(ma, mb):
        for u, v in zip(x, y):
            assert np.array_equal(np.asarray(u), np.asarray(v))


def test_driver_equals_reference_loop(pkg, tmp_path):
    drv = importlib.import_module("3d_reconstruction_b200.driver")
    desc, pts, size, k, codebook = make_scene()
    connection, start = drv.bow_view_graph(desc, k, codebook, cos_thr=0.0)
    ref = oracle_driver.reference_loop(desc, pts, size, connection, start, oracle_driver.cv2_matcher(), min_matches=60)
    assert len(ref[0]) >= 3
    # 1. batched driver
    got = drv.match_view_graph(desc, pts, connection, start, gates=drv.Gates(min_matches=60))
    _same(ref, got)
    # 2. the product matcher as the callable of the reference's own loop (matching.py:122 unchanged)
    via_callable = oracle_driver.reference_loop(desc, pts, size, connection, start, pkg.B200Matcher(), min_matches=60)
    _same(ref, via_callable)
    # 3. stage files for sfm.py
    drv.save_matches(*got, output_dir=str(tmp_path))
    all_matches = np.load(os.path.join(str(tmp_path), "all_matches.npy"), allow_pickle=True)
    assert len(all_matches) == len(ref[1]) and len(all_matches[0]) == 3
