"""CPU statement of the curvature / normal features (oracle/curv_port.py; a north_star extra without reference
counterpart, parity unpinned by construction): known answers on simple shapes."""
import numpy as np

from oracle import curv_port as cv


def test_wall_corner_and_sparse_points():
    wall = np.c_[np.linspace(-2, 2, 41), np.full(41, 3.0)]                    # straight wall y = 3, 10 cm spacing
    corner = np.r_[np.c_[np.linspace(0, 1, 11), np.zeros(11)], np.c_[np.ones(10), np.linspace(0.1, 1, 10)]] + [4.0, 4.0]
    lonely = np.array([[10.0, 0], [20.0, 0], [30.0, 0]])                       # neighbours beyond the radius
    xy = np.r_[wall, corner, lonely]
    off = np.array([0, 41, 41, 41 + 21, 41 + 21, 41 + 21, 41 + 24], np.int32)   # three clouds, all points in their GA block
    nrm, cur = cv.features(xy, off, k=5, radius=0.5)
    assert np.nanmax(cur[:41]) < 1e-12                                         # a line has no curvature
    np.testing.assert_allclose(nrm[:41], np.tile([0.0, -1.0], (41, 1)), atol=1e-6)   # towards the origin
    c = cur[41:62]
    assert c[10] == np.nanmax(c) and c[10] > 0.1 and c[0] < 1e-12 and c[-1] < 1e-12   # the corner point stands out
    assert np.isnan(cur[62:]).all() and (nrm[62:] == 0).all()                  # fewer than 3 neighbours: no feature
    assert (cur[~np.isnan(cur)] >= 0).all() and (cur[~np.isnan(cur)] <= 0.5 + 1e-7).all()


def test_gate_zero_tolerance_still_matches_featureless_points():
    rng = np.random.default_rng(0)
    model = np.c_[np.linspace(-3, 3, 61), 2 + 0.02 * rng.standard_normal(61)]
    scene = model + [0.05, -0.03]
    I6 = np.array([1.0, 0, 0, 1, 0, 0])
    E = np.zeros((0, 2))
    free = cv.fit_curv(E, model, E, scene, I6, max_iter=5, tol=1.0)            # tolerance above 0.5: gate inactive
    tight = cv.fit_curv(E, model, E, scene, I6, max_iter=5, tol=1e-4)
    assert free["ncorr"][-1] == 61 and 0 < tight["ncorr"][-1] <= 61
    np.testing.assert_allclose(free["Rt"][4:], [-0.05, 0.03], atol=0.02)
