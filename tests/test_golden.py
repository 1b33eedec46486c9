"""The oracle must keep reproducing the committed fixtures (tests/golden/*.npz, made by make_golden.py)."""
import os

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SCENES = ["book-one", "cornell-box", "book-two", "stress:5000", "checker-lists"]


def load(name):
    return np.load(os.path.join(GOLDEN, name.replace(":", "_") + ".npz"))


@pytest.mark.parametrize("name", SCENES)
def test_oracle_reproduces_golden_trace(orc, name):
    g = load(name)
    h = orc.Scene(name, 1).trace(g["rays"].astype(np.float64), seed=1)
    assert np.array_equal(h["hit"], g["hit"]) and np.array_equal(h["prim"], g["prim"]) and np.array_equal(h["stable"], g["stable"])
    m = g["hit"] == 1
    # same binary, same flags -> bit-identical; a different libm may move the last bits
    assert np.allclose(h["t"][m], g["t"][m], rtol=1e-12, atol=0)
    assert np.allclose(h["p"][m], g["p"][m], rtol=1e-12, atol=1e-12)
    assert np.allclose(h["n"][m], g["n"][m], rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("name", SCENES)
def test_oracle_reproduces_golden_paired_render(orc, name):
    g = load(name)
    w, h, spp, depth = [int(v) for v in g["paired_shape"]]
    img, _, st = orc.Scene(name, 1).render(w, h, spp, depth, seed=1, mode=orc.PAIRED)
    assert st.segments == int(g["paired_segments"][0])
    assert np.allclose(img, g["paired_sum"], rtol=1e-9, atol=1e-12)
