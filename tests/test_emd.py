"""The EMD stage of side_effect_score (side_effects.py:12-56 -> pyemd.emd).  pyemd is not
installed and the reference ships no value for it: parity of this stage is UNPINNED.  What can be checked
without it are the properties of the transportation problem the reference hands to pyemd."""
import numpy as np

from safelife_k2_b200 import side_effects as se


def test_emd_known_answers():
    a = np.zeros((5, 5)); b = np.zeros((5, 5))
    assert se.earth_mover_distance(a, b) == 0.0
    a[0, 0] = 1.0; b[0, 2] = 1.0  # one unit moved two cells: tanh(2 / 5)
    assert abs(se.earth_mover_distance(a, b) - np.tanh(2 / 5)) < 1e-9
    # the reference wraps only positive coordinate differences (side_effects.py:46-50: min(dx, W - dx) with
    # dx = x_a - x_b signed), so the ground distance is not symmetric — restated as it is
    a[:] = 0; b[:] = 0; a[0, 4] = 1.0; b[0, 0] = 1.0   # dx = +4 -> wraps to 1
    assert abs(se.earth_mover_distance(a, b) - np.tanh(1 / 5)) < 1e-9
    assert abs(se.earth_mover_distance(b, a) - np.tanh(4 / 5)) < 1e-9  # dx = -4 stays 4
    b[0, 0] = 0.25                # unequal masses: the surplus pays extra_mass_penalty = 1 per unit
    assert abs(se.earth_mover_distance(a, b) - (0.25 * np.tanh(1 / 5) + 0.75)) < 1e-9


def test_emd_is_bounded():
    rng = np.random.RandomState(3)
    for _ in range(5):
        a = np.zeros((9, 11)); b = np.zeros((9, 11))
        a.flat[rng.choice(99, 12, replace=False)] = rng.randint(1, 1000, 12) / 1000
        b.flat[rng.choice(99, 12, replace=False)] = rng.randint(1, 1000, 12) / 1000
        d1 = se.earth_mover_distance(a, b)
        # every unit of flow costs less than 1 (tanh), every unit of surplus exactly 1
        assert d1 <= max(a.sum(), b.sum()) + 1e-9 and d1 >= abs(a.sum() - b.sum()) - 1e-9


# ---------------------------------------------------------------------------------------------
# the two independent exact solvers of the oracle (oracle/emd.py), cross-checked
# ---------------------------------------------------------------------------------------------
def _random_maps(rng, h, w, cells, overlap=0.5, levels=1000):
    """two density maps (multiples of 1 / levels, like counts / num_samples) that differ on ~`cells` cells,
    partly overlapping so that the in-bin cancellation has something to cancel"""
    a = np.zeros((h, w)); b = np.zeros((h, w))
    ia = rng.choice(h * w, cells, replace=False)
    ib = np.where(rng.rand(cells) < overlap, ia, rng.choice(h * w, cells))
    a.flat[ia] = rng.randint(1, levels + 1, cells) / levels
    b.flat[ib] = rng.randint(1, levels + 1, cells) / levels
    return a, b


def test_oracle_solvers_agree_on_1000_random_cases():
    """successive shortest paths (C) vs HiGHS LP on the transportation problem EMD-hat leaves after the in-bin
    cancellation, with the reference's ASYMMETRIC ground distance (one-sided wrap, side_effects.py:44-49): 1e-12."""
    from oracle import emd

    rng = np.random.RandomState(11)
    worst, asym = 0.0, 0
    for k in range(1000):
        h, w = rng.randint(3, 27), rng.randint(3, 27)
        a, b = _random_maps(rng, h, w, rng.randint(1, min(h * w, 40)), overlap=rng.rand())
        ch, dist = emd.changed_cells_and_distance(a, b)
        if dist is None:
            continue
        asym += bool(np.abs(dist - dist.T).max() > 0)
        v1 = emd.emd_hat_gd_metric(a[ch], b[ch], dist, 1.0, emd.transport_ssp)
        v2 = emd.emd_hat_gd_metric(a[ch], b[ch], dist, 1.0, emd.transport_lp)
        worst = max(worst, abs(v1 - v2))
        assert abs(v1 - v2) <= 1e-12 * max(1.0, abs(v1)), (k, v1, v2)
        # the product's host function (scipy LP of the same reduced problem; pyemd absent) says the same
        assert abs(se.earth_mover_distance(a, b) - v1) <= 1e-12 * max(1.0, abs(v1)), k
    assert asym > 500  # the asymmetric case is the common one


def test_oracle_solvers_agree_on_the_golden_density_maps():
    """the density maps the unmodified reference produced (tests/golden/side_effects.npz): both solvers, the
    host function, and the size of the open risk — the plain LP of EMD-hat's definition WITHOUT the in-bin
    cancellation (what a non-metric ground distance could make different) — on real maps."""
    from oracle import emd
    from tests import golden_util as gu

    d = gu.load("side_effects.npz")
    i, checked, risk = 0, 0, 0.0
    while "keys_%d" % i in d:
        ns = int(d["num_samples_%d" % i])
        for s, k in enumerate(d["keys_%d" % i]):
            a = d["counts_%d" % i][0, s].astype(np.float64) / ns
            b = d["counts_%d" % i][1, s].astype(np.float64) / ns
            ch, dist = emd.changed_cells_and_distance(a, b)
            if dist is None:
                assert se.earth_mover_distance(a, b) == 0.0
                continue
            v1 = emd.earth_mover_distance(a, b)
            if ch.sum() <= 150:  # (the LP of the 420-cell map takes a while)
                v2 = emd.earth_mover_distance(a, b, solver=emd.transport_lp)
                assert abs(v1 - v2) <= 1e-12 * max(1.0, abs(v1)), (i, k)
                assert abs(se.earth_mover_distance(a, b) - v1) <= 1e-12 * max(1.0, abs(v1)), (i, k)
                risk = max(risk, abs(emd.emd_unreduced(a[ch], b[ch], dist) - v1))
            # pyemd's 1e-6 quantisation of masses and costs moves the value by about that much, no more
            assert abs(emd.earth_mover_distance(a, b, quantize=True) - v1) <= 1e-3 * max(1.0, abs(v1))
            checked += 1
        i += 1
    assert checked >= 15
    assert risk <= 1e-12  # on the reference's own maps the cancellation changes nothing


def test_in_bin_cancellation_matters_for_a_non_metric_distance():
    """why the open risk is stated: with the one-sided wrap a relay through a bin that holds mass on both sides
    can be cheaper than the direct arc, and emd_hat_gd_metric (pyemd's entry point) does not see it."""
    from oracle import emd

    w = 26
    a = np.zeros((1, w)); b = np.zeros((1, w))
    a[0, 0] = 1.0; b[0, 25] = 1.0          # direct arc 0 -> 25: dx = -25 does not wrap: tanh(25 / 5)
    a[0, 13] = 1.0; b[0, 13] = 0.5         # a bin with mass on both sides half way
    ch, dist = emd.changed_cells_and_distance(a, b)
    red = emd.emd_hat_gd_metric(a[ch], b[ch], dist)
    full = emd.emd_unreduced(a[ch], b[ch], dist)
    assert full <= red + 1e-15
    assert abs(se.earth_mover_distance(a, b) - red) < 1e-12  # the product follows pyemd's entry point
