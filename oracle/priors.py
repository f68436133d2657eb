"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy, fp64) of the reference's pairwise priors:

    ZBL.post_reduce      torchmdnet/priors/zbl.py:47-63
    Coulomb.post_reduce  torchmdnet/priors/coulomb.py:38-50
    D2.post_reduce       torchmdnet/priors/d2.py:163-193

on the pair set of the reference's legacy ``Distance`` module (models/utils.py:393-455: same molecule, no self pairs, both
directions, d < cutoff; open boundaries).  Energies per molecule; forces by central differences of these energies, so they do
not share a line of algebra with the analytic derivatives of csrc/pair_prior.cu.  Pinned by tests/test_oracle_pinned.py against
the fixtures the unmodified reference wrote (tests/golden/small_tensornet_{pairpriors,d2}.npz) and the Psi4 values of the
reference's tests/test_prior_d2.py."""
import math

import numpy as np

# Grimme 2006, Table 1 (as priors/d2.py:48-105): C_6 in J nm^6 / mol, R_0 in Angstrom, index = atomic number
D2_C6 = np.array([np.nan, 0.14, 0.08, 1.61, 1.61, 3.13, 1.75, 1.23, 0.70, 0.75, 0.63, 5.71, 5.71, 10.79, 9.23, 7.84, 5.57, 5.07,
                  4.61] + [10.80] * 12 + [16.99, 17.10, 16.37, 12.64, 12.47, 12.01] + [24.67] * 12
                 + [37.32, 38.71, 38.44, 31.74, 31.50, 29.99])
D2_R0 = np.array([np.nan, 1.001, 1.012, 0.825, 1.408, 1.485, 1.452, 1.397, 1.342, 1.287, 1.243, 1.144, 1.364, 1.639, 1.716,
                  1.705, 1.683, 1.639, 1.595, 1.485, 1.474] + [1.562] * 10
                 + [1.650, 1.727, 1.760, 1.771, 1.749, 1.727, 1.628, 1.606] + [1.639] * 10
                 + [1.672, 1.804, 1.881, 1.892, 1.892, 1.881])

_erf = np.vectorize(math.erf)


def _pairs(pos, batch, cutoff):
    """directed pairs (i, j), i != j, same molecule, d < cutoff -> (i, j, d)"""
    pos = np.asarray(pos, dtype=np.float64)
    batch = np.asarray(batch)
    d = np.linalg.norm(pos[:, None, :] - pos[None, :, :], axis=-1)
    keep = (batch[:, None] == batch[None, :]) & ~np.eye(len(pos), dtype=bool)
    if cutoff is not None:
        keep &= d < cutoff
    i, j = np.nonzero(keep)
    return i, j, d[i, j]


def _per_molecule(e_pair, i, batch, n_mol):
    out = np.zeros(n_mol)
    np.add.at(out, np.asarray(batch)[i], e_pair)
    return out


def zbl_energy(pos, batch, z, n_mol, atomic_number, cutoff_distance, distance_scale, energy_scale):
    """zbl.py:47-63 (cosine cutoff of models/utils.py:361-391 with cutoff_lower = 0)"""
    i, j, d = _pairs(pos, batch, cutoff_distance)
    an = np.asarray(atomic_number, dtype=np.float64)[np.asarray(z)]
    zi, zj = an[i], an[j]
    a = 0.8854 * 5.29177210903e-11 / (zi ** 0.23 + zj ** 0.23)
    x = d * distance_scale / a
    f = 0.1818 * np.exp(-3.2 * x) + 0.5099 * np.exp(-0.9423 * x) + 0.2802 * np.exp(-0.4029 * x) + 0.02817 * np.exp(-0.2016 * x)
    f = f * 0.5 * (np.cos(d * math.pi / cutoff_distance) + 1.0) * (d < cutoff_distance)
    e = f * zi * zj / d
    return 0.5 * (2.30707755e-28 / energy_scale / distance_scale) * _per_molecule(e, i, batch, n_mol)


def coulomb_energy(pos, batch, charges, n_mol, alpha, distance_scale, energy_scale):
    """coulomb.py:38-50: positions in nm, no cutoff"""
    x = 1e9 * distance_scale * np.asarray(pos, dtype=np.float64)
    a = alpha / (1e9 * distance_scale)
    i, j, d = _pairs(x, batch, None)
    q = np.asarray(charges, dtype=np.float64)
    e = _erf(a * d) * q[i] * q[j] / d if len(d) else np.zeros(0)
    return 0.5 * (2.30707e-28 / energy_scale / distance_scale) * _per_molecule(e, i, batch, n_mol)


def d2_energy(pos, batch, z, n_mol, atomic_number, cutoff_distance, distance_scale, energy_scale, s_6=1.0, d=20.0):
    """d2.py:163-193: nm and J/mol internally"""
    i, j, r = _pairs(pos, batch, cutoff_distance)
    zz = np.asarray(atomic_number)[np.asarray(z)]
    R = r * distance_scale * 1e9
    c6 = np.sqrt(D2_C6[zz[i]] * D2_C6[zz[j]])
    rr = (D2_R0[zz[i]] + D2_R0[zz[j]]) * 0.1
    e = c6 / R ** 6 / (1.0 + np.exp(-d * (R / rr - 1.0)))
    return -s_6 * 0.5 * _per_molecule(e, i, batch, n_mol) / (energy_scale * 6.02214076e23)


def forces_fd(energy_fn, pos, h=1e-5):
    """-dE_total/dpos by central differences (E_total = sum over molecules); fp64"""
    pos = np.array(pos, dtype=np.float64)
    f = np.zeros_like(pos)
    for a in range(pos.shape[0]):
        for c in range(3):
            p = pos.copy(); p[a, c] += h
            m = pos.copy(); m[a, c] -= h
            f[a, c] = -(energy_fn(p).sum() - energy_fn(m).sum()) / (2 * h)
    return f
