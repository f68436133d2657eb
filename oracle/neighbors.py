"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy) of the reference neighbour-pair op.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may import this.  The product path never does.

Pinned against: the reference's compiled CPU op (oracle/_ref, built from neighbors_cpu.cpp) in
tests/test_oracle_pinned.py when the reference is present, and against the committed golden pair
lists in tests/golden/neighbors_*.npz that were produced by that op (oracle/make_golden.py).

Algorithm followed (reference file:line):
  * all i>j pairs, same-batch mask                neighbors/neighbors_cpu.cpp:55-60
  * delta = pos[i]-pos[j]                          neighbors_cpu.cpp:61-62
  * triclinic minimum image, round(), z->y->x      neighbors_cpu.cpp:63-70  (= common.cuh:180-194)
  * rectangular minimum image floor(x/L+0.5)       common.cuh:142-148       (cell strategy only)
  * keep lo <= d < hi                              neighbors_cpu.cpp:71-75
    (the CUDA kernels test lo^2 <= d^2 < hi^2,      neighbors_cuda_brute.cuh:36; `d2_rule` selects it)
  * append transposed pairs, then self loops       neighbors_cpu.cpp:76-86
Pair order is unspecified in the reference (global atomics on CUDA); compare after `canonical_order`.
"""
import numpy as np


def min_image_triclinic(delta, box):
    """neighbors_cpu.cpp:63-70 / common.cuh:180-194.  `box` rows are a, b, c."""
    delta = delta.copy()
    for ax in (2, 1, 0):
        s = np.round(delta[:, ax] / box[ax, ax])
        delta -= s[:, None] * box[ax][None, :]
    return delta


def min_image_rect(delta, box):
    """common.cuh:142-148 (diagonal boxes, cell strategy)."""
    L = np.diag(box).astype(delta.dtype)
    return delta - np.floor(delta / L + delta.dtype.type(0.5)) * L


def neighbor_pairs(pos, batch=None, box=None, cutoff_lower=0.0, cutoff_upper=5.0, loop=False,
                   include_transpose=True, d2_rule=True, rect=False, chunk=4096):
    """Returns (neighbors int32[2,E], deltas[E,3], distances[E]) exactly as the reference CPU op would
    (unpadded), in the reference's emission order: lower-triangle pairs, transposes, self loops.

    Works in the dtype of `pos` (float32 or float64) so results are bit-comparable with a device
    kernel doing the same arithmetic.  Row-chunked so memory stays O(chunk*N).
    """
    pos = np.ascontiguousarray(pos)
    dt = pos.dtype
    n = pos.shape[0]
    if batch is None:
        batch = np.zeros(n, dtype=np.int64)
    batch = np.asarray(batch)
    if box is not None:
        box = np.asarray(box, dtype=dt)
    lo, hi = dt.type(cutoff_lower), dt.type(cutoff_upper)
    I, J, D, R = [], [], [], []
    for r0 in range(0, n, chunk):
        r1 = min(n, r0 + chunk)
        ii, jj = np.nonzero(np.arange(r0, r1)[:, None] > np.arange(n)[None, :])
        ii = ii + r0
        keep = batch[ii] == batch[jj]
        ii, jj = ii[keep], jj[keep]
        delta = pos[ii] - pos[jj]
        if box is not None:
            delta = min_image_rect(delta, box) if rect else min_image_triclinic(delta, box)
        d2 = (delta[:, 0] * delta[:, 0] + delta[:, 1] * delta[:, 1]) + delta[:, 2] * delta[:, 2]
        if d2_rule:
            keep = (d2 < hi * hi) & (d2 >= lo * lo)
        else:
            dd = np.sqrt(d2)
            keep = (dd < hi) & (dd >= lo)
        I.append(ii[keep]); J.append(jj[keep]); D.append(delta[keep]); R.append(np.sqrt(d2[keep]))
    I = np.concatenate(I) if I else np.zeros(0, np.int64)
    J = np.concatenate(J) if J else np.zeros(0, np.int64)
    D = np.concatenate(D) if D else np.zeros((0, 3), dt)
    R = np.concatenate(R) if R else np.zeros(0, dt)
    nb = np.stack([I, J]).astype(np.int32)
    if include_transpose:
        nb = np.concatenate([nb, nb[::-1]], axis=1)
        D = np.concatenate([D, -D], axis=0)
        R = np.concatenate([R, R])
    if loop:
        ar = np.arange(n, dtype=np.int32)
        nb = np.concatenate([nb, np.stack([ar, ar])], axis=1)
        D = np.concatenate([D, np.zeros((n, 3), dt)], axis=0)
        R = np.concatenate([R, np.zeros(n, dt)])
    return nb, D.astype(dt), R.astype(dt)


def canonical_order(neighbors, deltas=None, distances=None):
    """Sort pairs by (row0, row1) -- the comparison the reference's tests do (tests/test_neighbors.py:9-11).
    Drops -1 padding."""
    neighbors = np.asarray(neighbors)
    valid = neighbors[0] >= 0
    nb = neighbors[:, valid]
    order = np.lexsort((nb[1], nb[0]))
    out = [nb[:, order]]
    if deltas is not None:
        out.append(np.asarray(deltas)[valid][order])
    if distances is not None:
        out.append(np.asarray(distances)[valid][order])
    return out if len(out) > 1 else out[0]


def neighbor_backward(neighbors, deltas, distances, grad_deltas, grad_distances, n_atoms):
    """neighbors/neighbors_cuda.cu:43-71: g = grad_vec*[w!=0] + (vec/w)*grad_w*[w!=0];
    grad_pos[i] += g, grad_pos[j] -= g; rows with w == 0 (self loops, -1 padding) are skipped."""
    nz = distances != 0
    g = np.zeros_like(deltas)
    g[nz] = grad_deltas[nz] + deltas[nz] / distances[nz, None] * grad_distances[nz, None]
    out = np.zeros((n_atoms, 3), dtype=deltas.dtype)
    np.add.at(out, neighbors[0][nz], g[nz])
    np.add.at(out, neighbors[1][nz], -g[nz])
    return out
