import numpy as np

from libgran_b200 import scenarios as S


def test_aggregate_fixture():
    a = S.aggregate_coordinates()
    assert a.shape == (150, 3)
    d = np.linalg.norm(a[:, None, :] - a[None, :, :], axis=-1) + np.eye(150) * 10
    assert 1.999 < d.min() < 2.001  # touching spheres, SURVEY.md section 8


def test_fractal_generator_touching_and_non_overlapping():
    x = S.fractal_aggregate_coordinates(400, seed=5)
    d = np.linalg.norm(x[:, None, :] - x[None, :, :], axis=-1) / S.R_PART + np.eye(400) * 10
    assert d.min() > 2.0 - 1e-6
    assert (d.min(axis=1) < 2.0 + 1e-6).all()  # every particle touches at least one other


def test_multi_aggregate_layout():
    scn = S.multi_aggregate_collision(n_copies=12)
    assert scn.cfg.n == 1800 and scn.x0.shape == (1800, 3)
    assert set(np.unique(scn.v0[:, 0])) == {-0.5, 0.5}


def test_random_packing_no_overlap():
    scn = S.random_packing(n=3000, seed=1)
    from scipy.spatial import cKDTree

    pairs = cKDTree(scn.x0).query_pairs(2.0 * S.R_PART * (1 - 1e-12))
    assert len(pairs) == 0
