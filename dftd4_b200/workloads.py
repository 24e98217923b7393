"""Deterministic synthetic structures for the BASELINE.json configs.  numpy only, Bohr out.

Every generator returns dict(name, numbers[int32 n], positions[n,3] Bohr, lattice[3,3] Bohr or None, method, atm).
Atoms sit on jittered grids so that minimum distances are guaranteed without an O(N^2) rejection loop (SURVEY 8d
describes self-avoiding / minimum-distance constructions: these inputs have the stated densities, compositions and
seeds but are "in the spirit of" those; BASELINE.md says so).  sha256() identifies the generated arrays.
"""
import numpy as np

AATOAU = 1.0 / 0.529177210903  # CODATA-2018, the constant mctc-lib uses for Angstrom -> Bohr


def _jittered_grid(rng, lo, hi, spacing, jitter):
    """Points on a cubic grid of the given spacing filling the box [lo, hi), each displaced by U(-jitter, jitter)."""
    lo, hi = np.asarray(lo, float), np.asarray(hi, float)
    n = np.maximum(np.floor((hi - lo) / spacing + 1e-9).astype(int), 1)
    step = (hi - lo) / n
    g = np.stack(np.meshgrid(*[np.arange(k) for k in n], indexing="ij"), axis=-1).reshape(-1, 3)
    p = lo + (g + 0.5) * step
    return p + rng.uniform(-jitter, jitter, size=p.shape)


def caffeine():
    """Config 1: app/04-caffeine.xyz of the reference (geometry in Angstrom), PBE0-D4."""
    sym = "C N C C N C H N C O N C O C H H H C H H H H H H".split()
    z = dict(H=1, C=6, N=7, O=8)
    xyz = np.array([
        [-3.2668063145, -1.0612383311, 0.0019792586],
        [-2.2766003898, -0.0195172936, -0.0013208218],
        [-0.9134590858, -0.2019878522, 0.0036747135],
        [-0.3724701006, 1.0762202467, 0.0017569102],
        [-1.3526586265, 2.0066755645, -0.0051933750],
        [-2.4700895686, 1.3111433173, -0.0065911261],
        [-3.4434528448, 1.7598589152, -0.0115392456],
        [0.9701620055, 1.2833365802, -0.0008793828],
        [1.8257144377, 0.2028321683, -0.0007168653],
        [3.0278876281, 0.3474912012, -0.0051812078],
        [1.2515027114, -1.0636950051, -0.0008922263],
        [-0.1065018202, -1.3842039491, 0.0029991689],
        [-0.5015980556, -2.5366534217, 0.0044099955],
        [2.2018357820, -2.1638123093, -0.0037185988],
        [1.6323675079, -3.0900622488, 0.0097444448],
        [2.8244315108, -2.1119787186, -0.8964193499],
        [2.8450781508, -2.0977859157, 0.8731480371],
        [1.5216365707, 2.6211569486, 0.0093869454],
        [0.7306184352, 3.3088549484, -0.2771144995],
        [1.8886888508, 2.8714188674, 1.0060887389],
        [2.3551115434, 2.6687817985, -0.6902763920],
        [-2.7510404767, -2.0202634640, -0.0176576390],
        [-3.9109098848, -0.9753935259, -0.8744768432],
        [-3.8816811438, -0.9984445054, 0.9013764258],
    ])
    return dict(name="caffeine-24", numbers=np.array([z[s] for s in sym], dtype=np.int32), positions=xyz * AATOAU,
                lattice=None, method="pbe0", atm=True)


def water_box(nside=10, box_angstrom=31.04, seed=2024):
    """Config 2: nside^3 rigid water molecules (r_OH 0.9572 A, 104.52 deg) on a grid with random orientation and
    0.15 A centre jitter in a cubic periodic cell; r2SCAN-D4 with ATM."""
    rng = np.random.default_rng(seed)
    a = box_angstrom / nside
    g = np.stack(np.meshgrid(*[np.arange(nside)] * 3, indexing="ij"), axis=-1).reshape(-1, 3)
    centres = (g + 0.5) * a + rng.uniform(-0.15, 0.15, size=(len(g), 3))
    half = np.deg2rad(104.52) / 2
    mol = 0.9572 * np.array([[0.0, 0.0, 0.0], [np.sin(half), 0.0, np.cos(half)], [-np.sin(half), 0.0, np.cos(half)]])
    q = rng.normal(size=(len(g), 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    w, x, y, z = q.T
    rot = np.stack([np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)], -1),
                    np.stack([2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)], -1),
                    np.stack([2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], -1)], 1)
    pos = centres[:, None, :] + np.einsum("nij,aj->nai", rot, mol)
    numbers = np.tile(np.array([8, 1, 1], dtype=np.int32), len(g))
    return dict(name="water-box-%d" % (3 * len(g)), numbers=numbers, positions=pos.reshape(-1, 3) * AATOAU,
                lattice=np.eye(3) * box_angstrom * AATOAU, method="r2scan", atm=True)


def mixed_cell(box_angstrom=100.0, seed=13):
    """Config 5 (box 102 A -> 101 329 atoms) and its 10k-atom point (box 48 A -> 9 819 atoms): lower half SiO2-like at
    0.080 atoms/A^3 (Si:O 1:2), upper half organic at 0.12 atoms/A^3 (H .5 C .35 N .05 O .10); PBE-D4 with ATM."""
    rng = np.random.default_rng(seed)
    L = box_angstrom
    sil = _jittered_grid(rng, [0, 0, 0], [L, L, L / 2], (1 / 0.080) ** (1 / 3), 0.30)
    org = _jittered_grid(rng, [0, 0, L / 2], [L, L, L], (1 / 0.120) ** (1 / 3), 0.30)
    zs = np.where(np.arange(len(sil)) % 3 == 0, 14, 8)
    zo = rng.choice(np.array([1, 6, 7, 8]), size=len(org), p=[0.50, 0.35, 0.05, 0.10])
    pos = np.concatenate([sil, org])
    numbers = np.concatenate([zs, zo]).astype(np.int32)
    perm = rng.permutation(len(numbers))  # no spatial order in the input: the engine has to sort for itself
    return dict(name="mixed-cell-%d" % len(numbers), numbers=numbers[perm], positions=pos[perm] * AATOAU,
                lattice=np.eye(3) * L * AATOAU, method="pbe", atm=True)


def cluster(natoms=20000, seed=11):
    """Config 4: protein-like molecular cluster, sphere at 0.113 atoms/A^3, H .50 C .32 N .085 O .09 S .005; PBE-D4."""
    rng = np.random.default_rng(seed)
    rad = (3 * natoms / (4 * np.pi * 0.113)) ** (1 / 3)
    spacing = (1 / 0.113) ** (1 / 3)
    p = _jittered_grid(rng, [-rad - spacing] * 3, [rad + spacing] * 3, spacing, 0.30)
    d = np.linalg.norm(p, axis=1)
    p = p[np.argsort(d, kind="stable")[:natoms]]
    numbers = rng.choice(np.array([1, 6, 7, 8, 16]), size=len(p), p=[0.50, 0.32, 0.085, 0.09, 0.005]).astype(np.int32)
    perm = rng.permutation(len(numbers))
    return dict(name="cluster-%d" % len(numbers), numbers=numbers[perm], positions=p[perm] * AATOAU, lattice=None,
                method="pbe", atm=True)


def conformers(nmol=50000, seed=7):
    """Config 3: nmol drug-like molecules of 40-60 atoms (H .46 C .36 N .07 O .08 S .01 F .01 Cl .01) grown as
    self-avoiding chains on a 1.4 A jittered lattice walk; B3LYP-D4.  Returns a list-like batch:
    dict(offsets[nmol+1], numbers, positions, method)."""
    rng = np.random.default_rng(seed)
    sizes = rng.integers(40, 61, size=nmol)
    offsets = np.concatenate([[0], np.cumsum(sizes)])
    total = int(offsets[-1])
    numbers = rng.choice(np.array([1, 6, 7, 8, 16, 9, 17]), size=total,
                         p=[0.46, 0.36, 0.07, 0.08, 0.01, 0.01, 0.01]).astype(np.int32)
    # every molecule: first `size` points of a jittered 4x4x4 grid (spacing 1.45 A) in random order -> compact blob
    g = np.stack(np.meshgrid(*[np.arange(4)] * 3, indexing="ij"), axis=-1).reshape(-1, 3).astype(float)
    pos = np.empty((total, 3))
    for m in range(nmol):
        order = np.argsort(np.linalg.norm(g - 1.5 + rng.uniform(-0.4, 0.4, size=3), axis=1), kind="stable")[:sizes[m]]
        pos[offsets[m]:offsets[m + 1]] = g[order] * 1.45 + rng.uniform(-0.2, 0.2, size=(sizes[m], 3))
    return dict(name="conformers-%d" % nmol, offsets=offsets.astype(np.int64), numbers=numbers, positions=pos * AATOAU,
                lattice=None, method="b3lyp", atm=True)


WORKLOADS = {
    "caffeine": caffeine,
    "water3k": water_box,
    "water1k": lambda: water_box(7, 31.04 * 0.7),   # profiling-sized cut of config 2 (1 029 atoms, same density)
    "mixed10k": lambda: mixed_cell(48.0),
    "cluster20k": cluster,
    "mixed100k": lambda: mixed_cell(102.0),
}


def get(name):
    return WORKLOADS[name]()


def sha256(w):
    """SHA-256 over the generated arrays (numbers int32, positions and lattice float64, C order): the identity of a
    synthetic input (tests/golden/workload_sha256.json pins every BASELINE workload)."""
    import hashlib
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(w["numbers"], dtype=np.int32).tobytes())
    h.update(np.ascontiguousarray(w["positions"], dtype=np.float64).tobytes())
    if w.get("lattice") is not None:
        h.update(np.ascontiguousarray(w["lattice"], dtype=np.float64).tobytes())
    if "offsets" in w:
        h.update(np.ascontiguousarray(w["offsets"], dtype=np.int64).tobytes())
    return h.hexdigest()
