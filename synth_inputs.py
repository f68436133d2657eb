"""Seeded synthetic inputs for the five BASELINE.json configs.

numpy-only (RandomState, stable across torch versions) and independent of both the product package
and the oracle, so bench.py, the tests and oracle/make_golden.py all draw identical inputs.
No dataset or reference file is read: there is no network and /root/reference does not exist on
the GPU box.  Everything is float32 / int64 numpy; callers move it to torch.
"""
import numpy as np

ELEMENTS = np.array([1, 6, 7, 8, 9], dtype=np.int64)   # as the reference's example batch (tests/utils.py:28)


def _lattice_sites(n, spacing, stretch=(1.0, 1.0, 1.0)):
    """The n simple-cubic sites nearest the origin (after an anisotropic stretch of the metric)."""
    m = int(np.ceil((n * max(stretch) ** 3) ** (1 / 3))) + 3
    g = np.arange(-m, m + 1)
    X = np.stack(np.meshgrid(g, g, g, indexing="ij"), -1).reshape(-1, 3).astype(np.float64)
    r2 = ((X / np.asarray(stretch)) ** 2).sum(1) + 1e-6 * (X * [1.0, 2.0, 3.0]).sum(1)  # deterministic tie-break
    return X[np.argsort(r2, kind="stable")[:n]] * spacing


def molecule(n, rng, spacing=1.5, jitter=0.2, stretch=(1.0, 1.0, 1.0)):
    """Lattice blob + U(-jitter, jitter): min distance >= spacing-2*sqrt(3)*jitter keeps forces O(1)."""
    p = _lattice_sites(n, spacing, stretch) + rng.uniform(-jitter, jitter, (n, 3))
    return p.astype(np.float32)


def qm9_like_batch(n_mol=128, seed=1, n_min=3, n_max=29):
    """cfg 1: 128 QM9-size molecules (3..29 atoms), random elements from {H,C,N,O,F}."""
    rng = np.random.RandomState(seed)
    sizes = rng.randint(n_min, n_max + 1, n_mol)
    pos = np.concatenate([molecule(int(n), rng) for n in sizes])
    z = ELEMENTS[rng.randint(0, len(ELEMENTS), pos.shape[0])]
    batch = np.repeat(np.arange(n_mol, dtype=np.int64), sizes)
    return z, pos, batch


def chignolin_like(seed=2, n=166):
    """cfg 2: one 166-atom peptide-sized system.  An elongated (2.6:1:1) 2.0 A lattice blob whose 10 A
    neighbourhood statistics are close to chignolin's (E = 19 878 directed pairs incl. self
    loops, mean 118.7 neighbours; this blob: 20 100 and 120.1); composition roughly peptide-like (H, C, N, O)."""
    rng = np.random.RandomState(seed)
    pos = molecule(n, rng, spacing=2.0, jitter=0.25, stretch=(2.6, 1.0, 1.0))
    z = rng.choice(np.array([1, 6, 7, 8], dtype=np.int64), size=n, p=[0.47, 0.33, 0.08, 0.12])
    return z, pos, np.zeros(n, dtype=np.int64)


def md_step_noise(pos, step, seed=2, sigma=0.05):
    """MD-step variant of cfg 2: N(0, 0.05 A) displacement per step (fresh positions each step)."""
    rng = np.random.RandomState(seed * 1000003 + step)
    return (pos + rng.normal(0.0, sigma, pos.shape)).astype(np.float32)


ASPIRIN_Z = np.array([6] * 9 + [1] * 8 + [8] * 4, dtype=np.int64)   # C9H8O4, 21 atoms (rMD17/MD17 aspirin-sized)


def md17_like_batch(n_mol=8192, seed=3):
    """cfg 3: n_mol copies of one 21-atom lattice molecule, each with its own jitter."""
    rng = np.random.RandomState(seed)
    base = _lattice_sites(21, 1.5)
    pos = (base[None] + rng.uniform(-0.2, 0.2, (n_mol, 21, 3))).reshape(-1, 3).astype(np.float32)
    z = np.tile(ASPIRIN_Z, n_mol)
    batch = np.repeat(np.arange(n_mol, dtype=np.int64), 21)
    return z, pos, batch


def water_box(n_side, seed=4, density=0.1, order="raster"):
    """cfg 4: n_side^3 water-like triplets (O + 2 H at 0.96 A, random orientation) on a jittered lattice in
    a periodic cubic box at `density` atoms/A^3.  n_side=70 -> 1 029 000 atoms (L = 217.5 A);
    n_side=19 -> 20 577 atoms (L = 59.0 A) is the CPU-parity size.  Returns (z, pos, batch, box[3,3])."""
    rng = np.random.RandomState(seed)
    nmol = n_side ** 3
    L = (3 * nmol / density) ** (1 / 3)
    a = L / n_side
    g = np.arange(n_side)
    O = (np.stack(np.meshgrid(g, g, g, indexing="ij"), -1).reshape(-1, 3) + 0.5) * a
    O = O + rng.uniform(-0.15, 0.15, O.shape) * a
    if order == "morton":      # molecules along a Z-order curve (spatially local atom indices)
        gi = np.stack(np.meshgrid(g, g, g, indexing="ij"), -1).reshape(-1, 3).astype(np.uint64)
        key = np.zeros(nmol, dtype=np.uint64)
        for b in range(10):
            for ax in range(3):
                key |= ((gi[:, ax] >> np.uint64(b)) & np.uint64(1)) << np.uint64(3 * b + ax)
        O = O[np.argsort(key, kind="stable")]
    elif order == "shuffle":   # worst case: no relation between index and position
        O = O[rng.permutation(nmol)]
    def unit(v):
        return v / np.linalg.norm(v, axis=1, keepdims=True)
    h1 = unit(rng.normal(size=(nmol, 3)))
    t = unit(np.cross(h1, rng.normal(size=(nmol, 3))))
    ang = np.deg2rad(104.5)
    h2 = np.cos(ang) * h1 + np.sin(ang) * t
    pos = np.stack([O, O + 0.96 * h1, O + 0.96 * h2], 1).reshape(-1, 3)
    pos = pos - L * np.floor(pos / L) - L / 2          # wrap into [-L/2, L/2)
    z = np.tile(np.array([8, 1, 1], dtype=np.int64), nmol)
    box = np.diag([L, L, L]).astype(np.float32)
    return z, pos.astype(np.float32), np.zeros(3 * nmol, dtype=np.int64), box


def uniform_gas(n, seed=5, density=0.1):
    """cfg 5: pos = rand(N,3)*L with L = (N/density)^(1/3) (benchmarks/neighbors.py:59-70 setup)."""
    rng = np.random.RandomState(seed)
    L = (n / density) ** (1 / 3)
    pos = (rng.rand(n, 3) * L - L / 2).astype(np.float32)
    return pos, np.diag([L, L, L]).astype(np.float32)
