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
