"""CUDA neighbour list (through the torch op -> C ABI) against the oracle restatement of the reference op and
the reference's own golden pair lists.  Mirrors the reference's tests/test_neighbors.py: every list is sorted
before comparison (:9-11), max_num_pairs is the exact count so shapes are asserted too (:114,133-135)."""
import numpy as np
import pytest
import torch

from helpers import load_npz
from oracle import neighbors as onb

pytestmark = pytest.mark.gpu


def run_op(strategy, pos, batch, box, lo, hi, max_pairs, loop, transpose, periodic=None):
    from torchmdnet.neighbors import get_neighbor_pairs_kernel
    dev = "cuda"
    tb = torch.empty(0, 0) if box is None else torch.from_numpy(np.asarray(box)).to(torch.from_numpy(pos).dtype)
    nb, dl, ds, cnt = get_neighbor_pairs_kernel(strategy, torch.from_numpy(pos).to(dev),
                                                torch.from_numpy(batch).to(dev), tb,
                                                (box is not None) if periodic is None else periodic, lo, hi,
                                                max_pairs, loop, transpose)
    torch.cuda.synchronize()
    return nb.cpu().numpy(), dl.cpu().numpy(), ds.cpu().numpy(), int(cnt.item())


@pytest.mark.parametrize("i", range(8))
@pytest.mark.parametrize("strategy", ["brute", "shared", "cell"])
def test_reference_golden_pair_lists(i, strategy):
    g = load_npz("neighbors_%d.npz" % i)
    box = g["box"] if g["box"].size else None
    if strategy == "cell":
        if box is not None and np.abs(box - np.diag(np.diag(box))).max() > 0:
            pytest.skip("cell strategy: rectangular boxes only (neighbors_cuda_cell.cuh:349-352)")
        if box is None:
            L = float(np.abs(g["pos"]).max() * 2 + 1)
            box_arg, periodic = np.diag([L, L, L]), False
        else:
            box_arg, periodic = box, True
    else:
        box_arg, periodic = box, box is not None
    n_ref = int(g["count"][0])
    from torchmdnet.neighbors import get_neighbor_pairs_kernel
    pos, batch = torch.from_numpy(g["pos"]).cuda(), torch.from_numpy(g["batch"]).cuda()
    tb = torch.empty(0, 0) if box_arg is None else torch.from_numpy(np.asarray(box_arg)).to(pos.dtype)
    nb, dl, ds, cnt = get_neighbor_pairs_kernel(strategy, pos, batch, tb, periodic, 0.0, float(g["cutoff"]),
                                                max(n_ref, 1), bool(g["loop"]), bool(g["transpose"]))
    assert int(cnt.item()) == n_ref
    a = onb.canonical_order(nb.cpu().numpy(), dl.cpu().numpy(), ds.cpu().numpy())
    b = onb.canonical_order(g["neighbors"], g["deltas"], g["distances"])
    if strategy == "cell" and not bool(g["transpose"]):
        # the half list keeps (larger, smaller) index pairs in every strategy
        assert (a[0][0] >= a[0][1]).all()
    assert np.array_equal(a[0], b[0])
    np.testing.assert_allclose(a[1], b[1], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(a[2], b[2], rtol=1e-5, atol=1e-6)


CASES = []
for n_batches in (1, 2, 3, 4, 128):
    for cutoff in (0.1, 1.0, 3.0, 4.9):
        for loop in (False, True):
            for transpose in (False, True):
                for boxkind in (None, "triclinic", "rectangular"):
                    CASES.append((n_batches, cutoff, loop, transpose, boxkind))


@pytest.mark.parametrize("strategy", ["brute", "cell"])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_sweep_against_oracle(strategy, dtype):
    """The parameter grid of the reference's test_neighbors (tests/test_neighbors.py:69-139); pair sets and,
    because the kernels use the oracle's exact rounding sequence, deltas/distances must be bit-identical."""
    rng = np.random.RandomState(123)
    checked = 0
    for (n_batches, cutoff, loop, transpose, boxkind) in CASES:
        if strategy == "cell" and boxkind == "triclinic":
            continue
        n_per = rng.randint(3, 60)
        batch = np.repeat(np.arange(n_batches, dtype=np.int64), n_per)
        lbox = 10.0
        pos = (rng.rand(batch.size, 3) * 2 * lbox - lbox).astype(dtype)
        if boxkind == "triclinic":
            box = np.array([[lbox, 0, 0], [0.1 * lbox, lbox, 0], [0.3 * lbox, 0.2 * lbox, lbox]])
        elif boxkind == "rectangular":
            box = np.diag([lbox, lbox, lbox])
        else:
            box = None
        rect = strategy == "cell"
        if strategy == "cell" and box is None:
            pos = (pos / 2).astype(dtype)      # open boundaries: atoms must lie inside the binning box
            obox, periodic = np.diag([lbox + 1e-3] * 3), False
            ref = onb.neighbor_pairs(pos, batch, None, 0.0, cutoff, loop, transpose, d2_rule=True)
        else:
            obox, periodic = box, box is not None
            ref = onb.neighbor_pairs(pos, batch, None if box is None else box.astype(dtype), 0.0, cutoff, loop,
                                     transpose, d2_rule=True, rect=rect)
        n_ref = ref[0].shape[1]
        nb, dl, ds, cnt = run_op(strategy, pos, batch, obox, 0.0, cutoff, max(n_ref, 1), loop, transpose,
                                 periodic=periodic)
        assert cnt == n_ref, (n_batches, cutoff, loop, transpose, boxkind)
        a = onb.canonical_order(nb, dl, ds)
        b = onb.canonical_order(*ref)
        if strategy == "cell" and not transpose:
            # cell half list orders each pair (larger, smaller) exactly as brute: compare as is
            pass
        assert np.array_equal(a[0], b[0]), (n_batches, cutoff, loop, transpose, boxkind)
        assert np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2]), (n_batches, cutoff, boxkind)
        checked += 1
    assert checked > 100


def test_padding_and_overflow_semantics():
    """Slots beyond the found count hold -1/0/0; the count keeps counting past the capacity (common.cuh:70-116)."""
    rng = np.random.RandomState(5)
    pos = (rng.rand(200, 3) * 6).astype(np.float32)
    batch = np.zeros(200, dtype=np.int64)
    ref = onb.neighbor_pairs(pos, batch, None, 0.0, 2.0, False, True)
    n_ref = ref[0].shape[1]
    nb, dl, ds, cnt = run_op("brute", pos, batch, None, 0.0, 2.0, n_ref + 57, False, True)
    assert cnt == n_ref
    assert (nb[:, n_ref:] == -1).all() and (dl[n_ref:] == 0).all() and (ds[n_ref:] == 0).all()
    assert (nb[:, :n_ref] >= 0).all()
    nb, dl, ds, cnt = run_op("brute", pos, batch, None, 0.0, 2.0, n_ref // 2, False, True)
    assert cnt == n_ref and (nb >= 0).all()
    stored = set(map(tuple, nb.T.tolist()))
    assert stored <= set(map(tuple, ref[0].T.tolist()))


def test_lower_cutoff():
    rng = np.random.RandomState(6)
    pos = (rng.rand(300, 3) * 8).astype(np.float32)
    batch = np.sort(rng.randint(0, 3, 300)).astype(np.int64)
    ref = onb.neighbor_pairs(pos, batch, None, 1.5, 3.0, False, True)
    nb, dl, ds, cnt = run_op("brute", pos, batch, None, 1.5, 3.0, ref[0].shape[1], False, True)
    a, b = onb.canonical_order(nb, dl, ds), onb.canonical_order(*ref)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[2], b[2])
    assert ds.min() >= 1.5


@pytest.mark.parametrize("strategy", ["brute", "cell"])
def test_large_size(strategy):
    """32 000 atoms as the reference's test_large_size (tests/test_neighbors.py:407-469), and beyond the
    reference's brute limit of 32 767 atoms (neighbors_cuda_brute.cuh:75)."""
    import synth_inputs as S
    pos, box = S.uniform_gas(40000, seed=5)
    batch = np.zeros(pos.shape[0], dtype=np.int64)
    ref = onb.neighbor_pairs(pos, batch, box, 0.0, 5.0, False, True, rect=(strategy == "cell"))
    n_ref = ref[0].shape[1]
    nb, dl, ds, cnt = run_op(strategy, pos, batch, box, 0.0, 5.0, n_ref, False, True)
    assert cnt == n_ref
    a, b = onb.canonical_order(nb, dl, ds), onb.canonical_order(*ref)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])


def test_brute_and_cell_agree_at_scale():
    """Size-independent property at a size the O(N^2) oracle cannot reach: both strategies give the same sorted
    pair set on 500 000 atoms, every atom's neighbour count matches, and the list is symmetric."""
    import synth_inputs as S
    pos, box = S.uniform_gas(500000, seed=7)
    batch = np.zeros(pos.shape[0], dtype=np.int64)
    cap = int(pos.shape[0] * 52.4 * 1.3)
    nb_c, dl_c, ds_c, cnt_c = run_op("cell", pos, batch, box, 0.0, 5.0, cap, False, True)
    assert cnt_c <= cap
    v = nb_c[:, :cnt_c]
    key = v[0].astype(np.int64) * pos.shape[0] + v[1]
    rkey = v[1].astype(np.int64) * pos.shape[0] + v[0]
    assert np.array_equal(np.sort(key), np.sort(rkey))            # symmetric
    assert np.unique(key).size == key.size                          # no duplicates
    assert ds_c[:cnt_c].max() <= 5.0 and ds_c[:cnt_c].min() > 0   # d^2 < 25 may round to d == 5.0f
    # a 20k-atom slab of targets against the tiled brute kernel restricted through batch ids is not possible
    # (different batch ids change the pair set), so compare full brute at 100k atoms instead
    pos2, box2 = S.uniform_gas(100000, seed=8)
    batch2 = np.zeros(pos2.shape[0], dtype=np.int64)
    cap2 = int(pos2.shape[0] * 52.4 * 1.3)
    a = run_op("cell", pos2, batch2, box2, 0.0, 5.0, cap2, False, True)
    b = run_op("brute", pos2, batch2, box2, 0.0, 5.0, cap2, False, True)
    assert a[3] == b[3]
    ca, cb = onb.canonical_order(a[0], a[1], a[2]), onb.canonical_order(b[0], b[1], b[2])
    assert np.array_equal(ca[0], cb[0]) and np.array_equal(ca[2], cb[2])


def test_backward_matches_oracle():
    rng = np.random.RandomState(9)
    pos = (rng.rand(500, 3) * 9).astype(np.float64)
    from torchmdnet.neighbors import get_neighbor_pairs_kernel
    p = torch.from_numpy(pos).cuda().requires_grad_(True)
    batch = torch.zeros(500, dtype=torch.long, device="cuda")
    nb, dl, ds, cnt = get_neighbor_pairs_kernel("brute", p, batch, torch.empty(0, 0), False, 0.0, 3.0, 40000, True,
                                                True)
    gd = torch.from_numpy(rng.randn(*dl.shape)).cuda()
    gw = torch.from_numpy(rng.randn(*ds.shape)).cuda()
    (gp,) = torch.autograd.grad((dl * gd).sum() + (ds * gw).sum(), p)
    ref = onb.neighbor_backward(nb.cpu().numpy(), dl.detach().cpu().numpy(), ds.detach().cpu().numpy(),
                                gd.cpu().numpy(), gw.cpu().numpy(), 500)
    np.testing.assert_allclose(gp.cpu().numpy(), ref, rtol=1e-10, atol=1e-10)


def test_optimized_distance_module():
    from torchmdnet.models.utils import OptimizedDistance
    rng = np.random.RandomState(10)
    pos = torch.from_numpy((rng.rand(100, 3) * 6).astype(np.float32)).cuda()
    od = OptimizedDistance(0.0, 3.0, max_num_pairs=-64, return_vecs=True, loop=True)
    ei, ew, ev = od(pos)
    assert ei.dtype == torch.long and ei.shape[0] == 2 and ew.shape[0] == ei.shape[1] == ev.shape[0]
    assert (ei >= 0).all()
    ref = onb.neighbor_pairs(pos.cpu().numpy(), None, None, 0.0, 3.0, True, True)
    assert ei.shape[1] == ref[0].shape[1]
    od2 = OptimizedDistance(0.0, 3.0, max_num_pairs=10, check_errors=True)
    with pytest.raises(RuntimeError, match="max_num_pairs"):
        od2(pos)
    od3 = OptimizedDistance(0.0, 3.0, max_num_pairs=-64, resize_to_fit=False, check_errors=False,
                            long_edge_index=False)
    ei3, ew3, ev3 = od3(pos)
    assert ei3.dtype == torch.int32 and ei3.shape[1] == 6400 and ev3 is None


def test_cuda_graph_capture_of_the_op():
    """check_errors=False, resize_to_fit=False is capturable (tests/test_neighbors.py:540-693)."""
    from torchmdnet.models.utils import OptimizedDistance
    rng = np.random.RandomState(11)
    pos = torch.from_numpy((rng.rand(256, 3) * 8).astype(np.float32)).cuda()
    od = OptimizedDistance(0.0, 3.0, max_num_pairs=-64, return_vecs=True, loop=False, resize_to_fit=False,
                           check_errors=False)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3):
            od(pos)
    torch.cuda.current_stream().wait_stream(s)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        ei, ew, ev = od(pos)
    pos.copy_(torch.from_numpy((rng.rand(256, 3) * 8).astype(np.float32)))
    graph.replay()
    torch.cuda.synchronize()
    ref = onb.neighbor_pairs(pos.cpu().numpy(), None, None, 0.0, 3.0, False, True)
    a = onb.canonical_order(ei.cpu().numpy().astype(np.int32), ev.cpu().numpy(), ew.cpu().numpy())
    b = onb.canonical_order(*ref)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[2], b[2])


def _csr_reference(pos, batch, box, hi, rect):
    nb, dl, ds = onb.neighbor_pairs(pos, batch, box, 0.0, hi, loop=True, include_transpose=True, rect=rect)
    order = np.lexsort((nb[1], nb[0]))
    return nb[:, order], dl[order], ds[order]


@pytest.mark.parametrize("case", ["molecules", "single", "periodic_tric", "periodic_rect_cell", "open_cell"])
def test_direct_csr_build_matches_oracle(case):
    """tmdn_csr_build_f32 (the fused models' graph builder): same pair set as the reference op with loop=True and
    include_transpose=True, rows contiguous, geom = (unit vector target-source, distance)."""
    from torchmdnet import _native
    _native.load()
    rng = np.random.RandomState(21)
    strategy, box, periodic = "brute", None, False
    if case == "molecules":
        sizes = rng.randint(1, 40, 50)
        batch = np.repeat(np.arange(50, dtype=np.int64), sizes)
        pos = (rng.rand(batch.size, 3) * 6).astype(np.float32)
    elif case == "single":
        pos = (rng.rand(700, 3) * 12).astype(np.float32)
        batch = np.zeros(700, dtype=np.int64)
    elif case == "periodic_tric":
        pos = (rng.rand(500, 3) * 24 - 12).astype(np.float32)
        batch = np.zeros(500, dtype=np.int64)
        box, periodic = np.array([[12.0, 0, 0], [1.2, 12.0, 0], [3.6, 2.4, 12.0]], dtype=np.float32), True
    else:
        pos = (rng.rand(3000, 3) * 30 - 15).astype(np.float32)
        batch = np.sort(rng.randint(0, 2, 3000)).astype(np.int64)
        strategy = "cell"
        box = np.diag([30.0, 30.0, 30.0]).astype(np.float32)
        periodic = case == "periodic_rect_cell"
    hi = 3.0
    ref_nb, ref_dl, ref_ds = _csr_reference(pos, batch, box if periodic else None, hi, rect=(strategy == "cell"))
    E = ref_nb.shape[1]
    tb = torch.empty(0, 0) if box is None else torch.from_numpy(box)
    rp, col, geom, num = torch.ops.torchmdnet_b200.csr_build(strategy, torch.from_numpy(pos).cuda(),
                                                             torch.from_numpy(batch).cuda(), tb, periodic, 0.0, hi,
                                                             E + 13)
    assert int(num) == E
    rp, col, geom = rp.cpu().numpy(), col.cpu().numpy()[:E], geom.cpu().numpy()[:E]
    n = pos.shape[0]
    assert rp[0] == 0 and rp[-1] == E
    assert np.array_equal(np.diff(rp), np.bincount(ref_nb[0], minlength=n))
    tgt = np.repeat(np.arange(n), np.diff(rp))
    if strategy == "brute":
        assert np.array_equal(col, ref_nb[1])           # ascending source index inside every row
    order = np.lexsort((col, tgt))
    assert np.array_equal(col[order], ref_nb[1])
    g = geom[order]
    assert np.array_equal(g[:, 3], ref_ds)               # same rounding sequence -> identical distances
    nz = ref_ds > 0
    np.testing.assert_allclose(g[nz, :3], ref_dl[nz] / ref_ds[nz, None], rtol=0, atol=3e-7)
    assert (g[~nz] == 0).all()
    # determinism: a second build gives the identical structure (cells are internally re-ranked by atom index)
    rp2, col2, geom2, _ = torch.ops.torchmdnet_b200.csr_build(strategy, torch.from_numpy(pos).cuda(),
                                                              torch.from_numpy(batch).cuda(), tb, periodic, 0.0, hi,
                                                              E + 13)
    assert torch.equal(col2.cpu()[:E], torch.from_numpy(col)) and torch.equal(geom2.cpu()[:E], torch.from_numpy(geom))
    # overflow: the structure is truncated, the count is not
    rp3, col3, geom3, num3 = torch.ops.torchmdnet_b200.csr_build(strategy, torch.from_numpy(pos).cuda(),
                                                                 torch.from_numpy(batch).cuda(), tb, periodic, 0.0,
                                                                 hi, E // 2)
    assert int(num3) == E and int(rp3.max()) == E // 2


@pytest.mark.parametrize("case", ["one_molecule_166", "molecules_256", "periodic_tric_200", "single_atom", "lower_cutoff_97"])
def test_small_csr_single_kernel_equals_the_four_launch_build(case):
    """Systems of <= 256 atoms build their CSR in one single-CTA kernel (tuning knob "small_csr", csr_direct.cu): checked
    against the oracle's pair list and bit for bit against the four-launch build, truncated structures included."""
    from torchmdnet import _native
    _native.load()
    ops = torch.ops.torchmdnet_b200
    rng = np.random.RandomState(33)
    box, periodic, lo, hi = None, False, 0.0, 3.0
    if case == "one_molecule_166":
        pos, batch = (rng.rand(166, 3) * 9).astype(np.float32), np.zeros(166, dtype=np.int64)
    elif case == "molecules_256":
        sizes = [1, 40, 33, 2, 64, 17, 99]
        batch = np.repeat(np.arange(len(sizes), dtype=np.int64), sizes)
        pos = (rng.rand(256, 3) * 5).astype(np.float32)
    elif case == "periodic_tric_200":
        pos, batch = (rng.rand(200, 3) * 20 - 10).astype(np.float32), np.zeros(200, dtype=np.int64)
        box, periodic = np.array([[10.0, 0, 0], [1.0, 10.0, 0], [3.0, 2.0, 10.0]], dtype=np.float32), True
    elif case == "single_atom":
        pos, batch = np.zeros((1, 3), dtype=np.float32), np.zeros(1, dtype=np.int64)
    else:
        pos, batch, lo = (rng.rand(97, 3) * 6).astype(np.float32), np.zeros(97, dtype=np.int64), 1.5
    tb = torch.empty(0, 0) if box is None else torch.from_numpy(box)
    tp, tbatch = torch.from_numpy(pos).cuda(), torch.from_numpy(batch).cuda()
    if lo == 0.0:
        ref_nb, ref_dl, ref_ds = _csr_reference(pos, batch, box if periodic else None, hi, rect=False)
        E = ref_nb.shape[1]
    else:
        E = None
    old = ops.set_tuning("small_csr", -1)
    assert old == 1
    try:
        out = {}
        for knob in (1, 0):
            ops.set_tuning("small_csr", knob)
            n0 = ops.launch_count()
            full = ops.csr_build("brute", tp, tbatch, tb, periodic, lo, hi, 256 * 257)
            launches = ops.launch_count() - n0
            assert (launches == 1) if knob else (launches >= 4), launches
            found = int(full[3])
            cut = ops.csr_build("brute", tp, tbatch, tb, periodic, lo, hi, max(found // 2, 1))
            out[knob] = (full, cut)
        for a, b in zip(out[1], out[0]):
            found = int(a[3])
            assert found == int(b[3]) and torch.equal(a[0], b[0])
            k = int(a[0][-1])          # row pointers are clamped to the capacity
            assert torch.equal(a[1][:k], b[1][:k]) and torch.equal(a[2][:k], b[2][:k])
        rp, col, geom, num = out[1][0]
        if E is not None:
            assert int(num) == E and np.array_equal(col.cpu().numpy()[:E], ref_nb[1])
            assert np.array_equal(geom.cpu().numpy()[:E, 3], ref_ds)
            assert np.array_equal(np.diff(rp.cpu().numpy()), np.bincount(ref_nb[0], minlength=pos.shape[0]))
        else:
            d = geom.cpu().numpy()[:int(num), 3]
            assert ((d == 0) | ((d >= lo) & (d < hi))).all() and int(num) > pos.shape[0]
        assert int(out[1][1][0].max()) == max(int(num) // 2, 1)
    finally:
        ops.set_tuning("small_csr", old)


@pytest.mark.parametrize("strategy,box_type", [("brute", None), ("brute", "triclinic"), ("shared", "rectangular"),
                                               ("cell", "rectangular"), ("cell", None)])
@pytest.mark.parametrize("loop,include_transpose", [(True, True), (False, False)])
def test_jit_script_compatible(strategy, box_type, loop, include_transpose):
    """torch.jit.script(OptimizedDistance(...)) -- the reference's tests/test_neighbors.py:471-537: the scripted module gives
    the pair list of the oracle (and of the eager module, bit for bit), molecules of 3-99 atoms."""
    from torchmdnet.models.utils import OptimizedDistance
    rng = np.random.RandomState(4321)
    sizes = rng.randint(3, 100, 12)
    batch = np.repeat(np.arange(12, dtype=np.int64), sizes)
    lbox, cutoff = 10.0, 1.5
    pos = (rng.rand(batch.size, 3) * lbox).astype(np.float32)
    pos[0] = pos[1] = 0.0                                  # at least one pair (d = 0: kept by both, as in the reference's test)
    box = None
    if box_type == "rectangular":
        box = np.diag([lbox] * 3).astype(np.float32)
    elif box_type == "triclinic":
        box = np.array([[lbox, 0, 0], [0.1 * lbox, lbox, 0], [0.3 * lbox, 0.2 * lbox, lbox]], dtype=np.float32)
    ref = onb.neighbor_pairs(pos, batch, box, 0.0, cutoff, loop, include_transpose, rect=(strategy == "cell"))
    n_ref = ref[0].shape[1]
    nl = OptimizedDistance(cutoff_lower=0.0, loop=loop, cutoff_upper=cutoff, max_num_pairs=n_ref, strategy=strategy,
                           box=None if box is None else torch.from_numpy(box), return_vecs=True,
                           include_transpose=include_transpose)
    tp, tb = torch.from_numpy(pos).cuda().requires_grad_(True), torch.from_numpy(batch).cuda()
    e0 = nl(tp, tb)
    snl = torch.jit.script(nl)
    ei, ew, ev = snl(tp, tb)
    assert ei.shape == (2, n_ref) and ew.shape == (n_ref,) and ev.shape == (n_ref, 3)
    a = onb.canonical_order(ei.cpu().numpy().astype(np.int32), ev.detach().cpu().numpy(), ew.detach().cpu().numpy())
    b = onb.canonical_order(*ref)
    c = onb.canonical_order(e0[0].cpu().numpy().astype(np.int32), e0[2].detach().cpu().numpy(), e0[1].detach().cpu().numpy())
    assert np.array_equal(a[0], b[0]) and np.allclose(a[2], b[2], atol=1e-6) and np.allclose(a[1], b[1], atol=1e-6)
    assert np.array_equal(a[0], c[0]) and np.array_equal(a[1], c[1]) and np.array_equal(a[2], c[2])   # eager == scripted
    (g,) = torch.autograd.grad(ew.sum() + (ev * ev).sum(), tp)       # the scripted module differentiates like the eager one
    (g0,) = torch.autograd.grad(e0[1].sum() + (e0[2] * e0[2]).sum(), tp)
    torch.testing.assert_close(g, g0, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("strategy,box_type", [("brute", None), ("brute", "triclinic"), ("cell", "rectangular")])
@pytest.mark.parametrize("grad", ["deltas", "distances", "combined"])
def test_neighbor_autograds(strategy, box_type, grad):
    """gradcheck + gradgradcheck of the op in fp64 (reference: tests/test_neighbors.py:254-299, same tolerances): first and
    second derivatives of distances and displacement vectors with respect to the positions."""
    from torchmdnet.models.utils import OptimizedDistance
    torch.manual_seed(123)
    n, cutoff, lbox = 10, 4.999999, 10.0
    box = None
    if box_type == "rectangular":
        box = torch.eye(3, dtype=torch.float64) * lbox
    elif box_type == "triclinic":
        box = torch.tensor([[lbox, 0.0, 0.0], [0.1, lbox, 0.0], [0.3, 0.2, lbox]], dtype=torch.float64)
    nl = OptimizedDistance(cutoff_upper=cutoff, max_num_pairs=n * n, strategy=strategy, loop=False, include_transpose=True,
                           return_vecs=True, resize_to_fit=True, box=box)
    pos = (0.25 * lbox * torch.rand(n, 3, dtype=torch.float64)).cuda().requires_grad_(True)
    batch = torch.zeros(n, dtype=torch.long, device="cuda")

    def fn(p):
        ei, ew, ev = nl(p, batch)
        if grad == "deltas":
            return ev
        if grad == "distances":
            return ew
        return ev.sum(dim=-1) + ew

    assert torch.autograd.gradcheck(fn, pos, eps=1e-4, atol=1e-4, rtol=1e-4, nondet_tol=1e-4)
    assert torch.autograd.gradgradcheck(fn, pos, eps=1e-4, atol=1e-4, rtol=1e-4, nondet_tol=1e-4)


@pytest.mark.parametrize("strategy,box_type", [("brute", None), ("shared", "triclinic"), ("cell", "rectangular")])
def test_cuda_graph_compatible_backward(strategy, box_type):
    """The op and its backward inside one CUDA graph (reference: tests/test_neighbors.py:622-693): warm up, capture
    forward + backward, replay, compare the captured gradient with an eager evaluation."""
    from torchmdnet.models.utils import OptimizedDistance
    rng = np.random.RandomState(77)
    sizes = rng.randint(3, 100, 16)
    lbox, cutoff = 10.0, 1.5
    box = None
    if box_type == "rectangular":
        box = torch.eye(3) * lbox
    elif box_type == "triclinic":
        box = torch.tensor([[lbox, 0.0, 0.0], [1.0, lbox, 0.0], [3.0, 2.0, lbox]])
    nl = OptimizedDistance(cutoff_lower=0.0, loop=True, cutoff_upper=cutoff, max_num_pairs=-40, strategy=strategy, box=box,
                           return_vecs=True, include_transpose=True, check_errors=False, resize_to_fit=False)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):          # the leaf lives on the capture stream, so its gradient accumulates there as well
        batch = torch.from_numpy(np.repeat(np.arange(16, dtype=np.int64), sizes)).cuda()
        pos = torch.from_numpy((rng.rand(int(sizes.sum()), 3) * lbox).astype(np.float32)).cuda().requires_grad_(True)
        ei, ew, ev = nl(pos, batch)
        (g_ref,) = torch.autograd.grad(ew.sum() + (ev * ev).sum(), pos)
        for _ in range(3):
            ei, ew, ev = nl(pos, batch)
            (ew.sum() + (ev * ev).sum()).backward()
            pos.grad.data.zero_()
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            ei, ew, ev = nl(pos, batch)
            (ew.sum() + (ev * ev).sum()).backward()
        pos.grad.data.zero_()
        graph.replay()
        torch.cuda.synchronize()
        grad = pos.grad.clone()
    torch.cuda.current_stream().wait_stream(s)
    torch.testing.assert_close(grad, g_ref, rtol=1e-5, atol=1e-5)
    assert int((ei[0] >= 0).sum()) > int(sizes.sum())           # self loops + at least some pairs


def test_legacy_distance_module():
    """``Distance`` (the reference's radius_graph-based module, models/utils.py:393-455) as the reference tests it
    (tests/test_model_utils.py:9-84): pair counts for two atoms inside / below / above the cutoffs with and without self
    loops, all pairs of 100 atoms in a unit cube, and the AssertionError when an atom exceeds max_num_neighbors."""
    from torchmdnet.models.utils import Distance
    for cutoff_lower in (0, 2):
        for cutoff_upper in (5, 10):
            for return_vecs in (False, True):
                for loop in (False, True):
                    dist = Distance(cutoff_lower, cutoff_upper, max_num_neighbors=100, return_vecs=return_vecs, loop=loop)
                    batch = torch.tensor([0, 0]).cuda()
                    extra = 2 if loop else 0
                    pos = torch.tensor([[0, 0, 0], [(cutoff_lower + cutoff_upper) / 2, 0, 0]], dtype=torch.float).cuda()
                    ei, ew, ev = dist(pos, batch)
                    assert ei.size(1) == 2 + extra and ei.dtype == torch.long
                    assert (ev is not None) == return_vecs
                    if return_vecs:
                        torch.testing.assert_close(ev, pos[ei[0]] - pos[ei[1]])
                    torch.testing.assert_close(ew, (pos[ei[0]] - pos[ei[1]]).norm(dim=-1))
                    if cutoff_lower > 0:
                        pos = torch.tensor([[0, 0, 0], [cutoff_lower / 2, 0, 0]], dtype=torch.float).cuda()
                        assert dist(pos, batch)[0].size(1) == extra
                    pos = torch.tensor([[0, 0, 0], [cutoff_upper + 1, 0, 0]], dtype=torch.float).cuda()
                    assert dist(pos, batch)[0].size(1) == extra
                    torch.manual_seed(1)
                    pos = torch.rand(100, 3).cuda()
                    ei, ew, ev = dist(pos, torch.zeros(100, dtype=torch.long).cuda())
                    extra = 100 if loop else 0
                    assert ei.size(1) == (extra if cutoff_lower > 0 else 100 * 99 + extra)
    dist = Distance(0, 5, max_num_neighbors=32)
    pos = torch.rand(100, 3).cuda()
    with pytest.raises(AssertionError, match="max_num_neighbors"):
        dist(pos, torch.zeros(100, dtype=torch.long).cuda())
    with pytest.raises(AssertionError, match="max_num_neighbors"):
        dist(pos, torch.tensor([0] * 20 + [1] * 80).cuda())
    ei, ew, ev = dist(pos[:30], torch.zeros(30, dtype=torch.long).cuda())        # 29 neighbours each: within the budget
    assert ei.size(1) == 30 * 29
    p = pos[:10].clone().requires_grad_(True)                                    # differentiable like OptimizedDistance
    ei, ew, ev = Distance(0, 5, return_vecs=True)(p, None)
    (g,) = torch.autograd.grad(ew.sum(), p)
    u = (p[ei[0]] - p[ei[1]]) / ew.unsqueeze(1)
    ref = torch.zeros_like(p).index_add_(0, ei[0], u).index_add_(0, ei[1], -u)
    torch.testing.assert_close(g, ref.detach(), rtol=1e-5, atol=1e-5)
