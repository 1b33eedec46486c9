"""rt_render_noalloc (include/rt_b200.h): the render entry point that allocates nothing — its wavefront scratch is a static device
array of RT_NOALLOC_SCRATCH_BYTES.  No counterpart in the reference.  Collected last
(test_zz_*).  Checked against rt_render, the existing entry point, on two calls in a row."""
import importlib

import numpy as np
import pytest


@pytest.fixture(scope="module")
def rt():
    return importlib.import_module("ray-tracer_b200")


def _device_used_bytes():
    """Device memory in use (cuda-python), only to observe that a call leaves the device's free memory alone; None when unavailable."""
    try:
        from cuda.bindings import runtime as cudart
    except ImportError:
        return None
    err, free, total = cudart.cudaMemGetInfo()
    return total - free if int(err) == 0 else None


@pytest.mark.gpu
def test_render_noalloc_matches_rt_render_twice_in_a_row(rt):
    dev = rt.Device(0)
    for name, w, h, spp, depth in [("cornell-box", 96, 96, 16, 50), ("book-two", 128, 64, 8, 50)]:
        sc = rt.Scene.example(dev, name, 1)
        cam = sc.flat.camera(w / h)
        ref, ref_q, ref_st = sc.render(cam, w, h, spp, depth, seed=7)  # rt_render: two lanes, 4 Mi-ray queues, cudaMalloc'ed state
        a, aq, sa = sc.render_noalloc(cam, w, h, spp, depth, seed=7)
        used_before = _device_used_bytes()  # (after the first call: a kernel's first launch may load its module lazily)
        b, bq, sb = sc.render_noalloc(cam, w, h, spp, depth, seed=7)  # second call: the array still holds the first call's contents
        assert _device_used_bytes() == used_before, "rt_render_noalloc changed the device's free memory"
        assert sa.paths == sb.paths == ref_st.paths == w * h * spp and sa.rays == sb.rays == ref_st.rays
        assert w * h * spp > rt.NOALLOC_QUEUE_RAYS  # the render takes several queue refills
        # same Philox keys, same hits: the sums differ only by the order of the f64 accumulations (the tolerance of
        # test_queue_capacity_does_not_change_the_image / test_wavefront_lanes_do_not_change_the_image)
        for got, got_q in ((a, aq), (b, bq)):
            assert np.allclose(got, ref, rtol=1e-5, atol=1e-5) and np.allclose(got_q, ref_q, rtol=1e-5, atol=1e-5)
        assert np.array_equal(a, b) or np.allclose(a, b, rtol=1e-6, atol=1e-6)
        assert ref.mean() > 0
        # a row shard, no sum of squares
        c, cq, _ = sc.render_noalloc(cam, w, h, spp, depth, seed=7, row_begin=1, row_stride=3, want_sumsq=False)
        assert cq is None and np.allclose(c, ref[1::3], rtol=1e-5, atol=1e-5)
        # and the context still renders through its own buffers afterwards
        again, _, _ = sc.render(cam, w, h, spp, depth, seed=7)
        assert np.allclose(again, ref, rtol=1e-5, atol=1e-5)
        sc.close()
    dev.close()


@pytest.mark.gpu
def test_render_noalloc_rejects_sizes_beyond_the_static_maximum(rt):
    dev = rt.Device(0)
    sc = rt.Scene.example(dev, "cornell-box", 1)
    cam = sc.flat.camera(1.0)
    with pytest.raises(rt.BackendError) as e:
        sc.render_noalloc(cam, 257, 256, 1, 4)  # 65 792 pixels > RT_NOALLOC_MAX_PIXELS
    assert e.value.code == rt.RT_ERR_INVALID and "RT_NOALLOC_MAX_PIXELS" in str(e.value) and "65536" in str(e.value)
    s, _, st = sc.render_noalloc(cam, 256, 256, 1, 4)  # exactly the maximum
    assert st.paths == 65536 and s.shape == (256, 256, 3)
    s2, _, _ = sc.render_noalloc(cam, 512, 256, 1, 4, row_begin=0, row_stride=2)  # the limit is on RENDERED rows
    assert s2.shape == (128, 512, 3)
    sc.close()
    dev.close()
