"""The oracle's analytic gradient and virial UNDER PERIODIC BOUNDARY CONDITIONS against central differences of its own
energy -- the reference's own consistency tests (test/unit/test_dftd4.f90:122-168 numgrad, :171-224 numsigma) with the
cutoffs of test/unit/test_periodic.f90:34-35 (disp3 = 15 bohr) and with all-default cutoffs, on a triclinic cell that
contains Si (the element of the headline workload whose EEQ row no reference golden touches).  CPU only.

This pins what no reference golden vector pins: the periodic gradient / virial of the oracle that the `-m gpu` parity
tests then compare the CUDA path with (Ewald dq/dr and dq/dL chain rule, image sums of the pair / ATM / CN terms).
"""
import numpy as np
import pytest

STEP = 1.0e-4   # bohr / strain: truncation error O(step^2) ~ 1e-10; rounding noise of the energy (1e-13) / step ~ 1e-9 stays below


def _cell(seed, n_side=(2, 2, 3)):
    rng = np.random.default_rng(seed)
    lat = np.array([[9.5, 0.0, 0.0], [1.8, 9.0, 0.0], [0.9, 1.4, 11.0]])
    g = np.stack(np.meshgrid(*[np.arange(k) for k in n_side], indexing="ij"), axis=-1).reshape(-1, 3)
    frac = (g + 0.5) / np.array(n_side) + rng.uniform(-0.07, 0.07, size=(len(g), 3))
    numbers = np.array([14, 8, 8, 6, 1, 1, 7, 1, 8, 14, 6, 1][: len(g)], dtype=np.int32)
    return numbers, frac @ lat, lat


def _energy(O, numbers, pos, lat, p, cutoff):
    return O.get_dispersion(numbers, pos, p, 0.0, lat, None, cutoff=cutoff, grad=False)["energy"]


CASES = [
    ("disp3-15", (60.0, 15.0, 30.0, 0.0, 0.0), "pbe", ((0, 0), (3, 1), (7, 2), (9, 0), (11, 1))),
    ("default", (60.0, 40.0, 30.0, 0.0, 0.0), "pbe", ((0, 2), (5, 0))),
    ("smooth", (40.0, 15.0, 30.0, 6.0, 4.0), "r2scan", ((1, 1), (10, 2))),
]


@pytest.mark.parametrize("name,cutoff,method,probes", CASES, ids=[c[0] for c in CASES])
def test_periodic_gradient_matches_central_differences(oracle, name, cutoff, method, probes):
    numbers, pos, lat = _cell(21)
    p = oracle.load_rational_damping(method, True)
    ref = oracle.get_dispersion(numbers, pos, p, 0.0, lat, None, cutoff=cutoff, grad=True)
    for i, c in probes:
        a, b = pos.copy(), pos.copy()
        a[i, c] += STEP
        b[i, c] -= STEP
        fd = (_energy(oracle, numbers, a, lat, p, cutoff) - _energy(oracle, numbers, b, lat, p, cutoff)) / (2 * STEP)
        assert abs(fd - ref["gradient"][i, c]) < 1.0e-9, (name, i, c, fd, ref["gradient"][i, c])


def _strain_fd(O, numbers, pos, lat, p, cutoff, a, b, step, five_point=False):
    """dE/d(eps_ab) by central differences; five_point: O(step^4) stencil (the energy goes like L^-6 in the strain, so
    the O(step^2) truncation error of the plain difference is 1e-7 at step 1e-4)."""
    def e_at(x):
        eps = np.eye(3)
        eps[a, b] += x
        return _energy(O, numbers, pos @ eps.T, lat @ eps.T, p, cutoff)
    if five_point:
        return (-e_at(2 * step) + 8 * e_at(step) - 8 * e_at(-step) + e_at(-2 * step)) / (12 * step)
    return (e_at(step) - e_at(-step)) / (2 * step)


def test_periodic_virial_matches_strain_differences_smooth(oracle):
    """Smooth cutoffs (cutoff.f90:74-114) make the energy differentiable in the strain (a sharp cutoff lets ~50 pair
    images cross 60 bohr per 1e-4 of strain), so the virial can be pinned tightly: all six independent components."""
    numbers, pos, lat = _cell(22)
    cutoff = (50.0, 15.0, 30.0, 8.0, 5.0)
    p = oracle.load_rational_damping("pbe", True)
    ref = oracle.get_dispersion(numbers, pos, p, 0.0, lat, None, cutoff=cutoff, grad=True)
    for a, b in ((0, 0), (1, 1), (2, 2), (0, 1), (1, 2), (2, 0)):
        fd = _strain_fd(oracle, numbers, pos, lat, p, cutoff, a, b, 2 * STEP, five_point=True)
        assert abs(fd - ref["sigma"][a, b]) < 1.0e-9, (a, b, fd, ref["sigma"][a, b])
    assert np.max(np.abs(ref["sigma"] - ref["sigma"].T)) < 1e-12


@pytest.mark.parametrize("name,cutoff,method,probes", CASES[:2], ids=[c[0] for c in CASES[:2]])
def test_periodic_virial_matches_strain_differences_sharp(oracle, name, cutoff, method, probes):
    """Sharp cutoffs, exactly the reference's numsigma: step 1e-7, tolerance 100 sqrt(eps) (test_periodic.f90:33,153-180)."""
    numbers, pos, lat = _cell(22)
    p = oracle.load_rational_damping(method, True)
    ref = oracle.get_dispersion(numbers, pos, p, 0.0, lat, None, cutoff=cutoff, grad=True)
    for a, b in ((0, 0), (1, 2)):
        fd = _strain_fd(oracle, numbers, pos, lat, p, cutoff, a, b, 1.0e-7)
        assert abs(fd - ref["sigma"][a, b]) < 1.5e-6, (name, a, b, fd, ref["sigma"][a, b])
    assert np.max(np.abs(ref["sigma"] - ref["sigma"].T)) < 1e-12
