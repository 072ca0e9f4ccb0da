"""The two CPU restatements of CCICP::classifyPoints (oracle/classify_port.py) agree with each other and with
hand-computed answers, and both are pinned to the reference's own function compiled from icpTools.cpp:36-103."""
import numpy as np

from oracle import classify_port as cp


def synthetic_cloud(seed, n=4000):
    rng = np.random.default_rng(seed)
    walls = np.c_[rng.uniform(-40, 40, n // 2), np.round(rng.uniform(-3, 3, n // 2)) * 7.3 + rng.normal(0, 0.05, n // 2)]
    blobs = rng.normal(0, 1, (n // 4, 2)) * 3 + rng.uniform(-30, 30, (1, 2))
    far = rng.uniform(-320, 320, (n - n // 2 - n // 4, 2))              # some outside the 600 m grid / on the border ring
    xy = np.r_[walls, blobs, far]
    rng.shuffle(xy)
    return np.c_[xy, rng.uniform(-1, 2, len(xy))].astype(np.float32)


def test_literal_and_vectorised_restatements_agree():
    for seed in range(4):
        xyz = synthetic_cloud(seed)
        a, b = cp.classify_literal(xyz), cp.classify(xyz)
        for u, v in zip(a, b):
            assert np.array_equal(u, v)
        assert 0 < len(a[0]) < len(xyz) and 0 < a[1].sum() < len(a[1])


def test_known_answers():
    # a filled 3x3 block of bins around the origin: centre bin has 0 empty neighbours, edges 3, corners 5
    pts = [(0.5 * i + 0.25, 0.5 * j + 0.25, 0.0) for i in range(-1, 2) for j in range(-1, 2)]
    xyz, ga, src = cp.classify(np.array(pts, np.float32))
    assert len(xyz) == 9 and ga.tolist() == [1, 1, 1, 1, 0, 1, 1, 1, 1]
    assert src.tolist() == list(range(9))                                # x-major bin order = generation order here
    # exact bin edges: -300 is bin 0 (border, dropped), 299.5 is bin 1199 (border), 300 is outside
    e = np.array([[-300, 0, 0], [299.5, 0, 0], [300, 0, 0], [-299.5, 0.2, 0], [float("nan"), 0, 0]], np.float32)
    xyz, ga, src = cp.classify(e)
    assert src.tolist() == [3] and ga.tolist() == [1]
    # two points in one bin keep input order; bin order is x-major
    e = np.array([[1.1, 5.1, 1], [-2.0, 0.1, 2], [1.2, 5.2, 3], [1.1, -5.0, 4]], np.float32)
    assert cp.classify(e)[2].tolist() == [1, 3, 0, 2]


def test_port_is_pinned_to_the_compiled_reference():
    """oracle/_ref/libclassify_ref.so is CCICP::classifyPoints itself (icpTools.cpp:36-103 compiled unmodified against a
    pcl::PointCloud shim, oracle/Makefile): the numpy port must reproduce its output order, coordinates and flags."""
    import pytest
    from oracle import refbind
    if not refbind.classify_available():
        pytest.skip("oracle/_ref/libclassify_ref.so not built (needs /root/reference at build time)")
    clouds = [synthetic_cloud(s, n) for s, n in ((0, 4000), (1, 70000), (2, 257), (3, 12345))]
    clouds.append(np.array([[-300, 0, 0], [299.5, 0, 0], [300, 0, 0], [-299.5, 0.2, 0], [1.1, 5.1, 1], [-2.0, 0.1, 2],
                            [1.2, 5.2, 3], [1.1, -5.0, 4], [1.1, 5.1, 5]], np.float32))
    clouds.append(np.repeat(np.array([[1.1, 5.1, 1.0]], np.float32), 500, 0))
    for xyz in clouds:
        rx, rg = refbind.classify_ref(xyz)
        px, pg, ps = cp.classify(xyz)
        assert len(rx) == len(px)
        assert np.array_equal(rx.view(np.uint32), px.view(np.uint32)) and np.array_equal(rg, pg)
        assert np.array_equal(xyz[ps].view(np.uint32), rx.view(np.uint32))
