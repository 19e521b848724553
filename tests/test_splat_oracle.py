"""CPU: the numpy restatement of the reference's input-heatmap generator (oracle/ref_heatmap.py)
against vectors produced by the reference's own method (tests/golden/make_golden_splat.py)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import ref_heatmap  # noqa: E402

G = np.load(os.path.join(ROOT, "tests", "golden", "golden_splat.npz"))


@pytest.mark.parametrize("name", ["panoptic", "shelf", "campus", "campus4"])
@pytest.mark.parametrize("k", [0, 1, 2, 3])
def test_restatement_equals_reference_output(name, k):
    iw, ih, hw, hh, J = (int(v) for v in G[f"{name}_sizes"])
    joints, vis = G[f"{name}_{k}_joints"], G[f"{name}_{k}_vis"]
    want = G[f"{name}_{k}_heatmap"]
    use_vis = bool(G[f"{name}_{k}_use_vis"])
    if len(joints) == 0:
        assert not want.any()
        return
    sigma = int(G[f"{name}_sigma"]) if f"{name}_sigma" in G else 3
    got = ref_heatmap.generate_input_heatmap(list(joints), list(vis) if use_vis else None, sigma, (iw, ih), (hw, hh))
    assert got.shape == want.shape == (J, hh, hw)
    assert np.array_equal(got, want), float(np.abs(got - want).max())
    assert want.max() > 0.5 or len(joints) <= 1      # the fixtures are not trivially empty
