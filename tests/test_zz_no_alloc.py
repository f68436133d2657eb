"""tmdn_csr_build_static_f32 (include/tmdn.h): the allocation-free variant of the direct CSR builder, scratch in a
file-scope static __device__ array.  Called through the C ABI (ctypes) on torch-owned inputs/outputs and compared
bit for bit with tmdn_csr_build_f32 (integer / index work and the same rounding sequence: the bar is equality), on two
calls in a row (the second one starts from the scratch contents the first one left).  Collected last on purpose."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

BRUTE, CELL = 0, 2
STATIC_BYTES = 4 << 20


def _lib():
    from torchmdnet import _native
    lib = _native.cabi()
    lib.tmdn_last_error.restype = ctypes.c_char_p
    lib.tmdn_csr_build_workspace_bytes.restype = ctypes.c_size_t
    lib.tmdn_csr_build_workspace_bytes.argtypes = [ctypes.c_int, ctypes.c_int64, ctypes.c_void_p, ctypes.c_double]
    common = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int,
              ctypes.c_double, ctypes.c_double, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
              ctypes.c_void_p]
    lib.tmdn_csr_build_f32.restype = ctypes.c_int
    lib.tmdn_csr_build_f32.argtypes = common + [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
    lib.tmdn_csr_build_static_f32.restype = ctypes.c_int
    lib.tmdn_csr_build_static_f32.argtypes = common + [ctypes.c_void_p]
    return lib


def _outputs(n, cap, dev):
    return (torch.full((n + 1,), -7, dtype=torch.int32, device=dev), torch.full((cap,), -7, dtype=torch.int32, device=dev),
            torch.full((cap, 4), -7.0, device=dev), torch.full((1,), -7, dtype=torch.int32, device=dev))


def _run(lib, static, strategy, pos, batch, box, periodic, hi, cap):
    n, dev = pos.shape[0], pos.device
    rp, col, geom, nf = _outputs(n, cap, dev)
    boxp = None if box is None else (ctypes.c_double * 9)(*[float(x) for x in np.asarray(box, dtype=np.float64).ravel()])
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    args = [strategy, pos.data_ptr(), batch.data_ptr(), n, boxp, int(periodic), 0.0, hi, cap, rp.data_ptr(), col.data_ptr(),
            geom.data_ptr(), nf.data_ptr()]
    if static:
        rc = lib.tmdn_csr_build_static_f32(*args, stream)
    else:
        nbytes = lib.tmdn_csr_build_workspace_bytes(strategy, n, boxp, hi)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        rc = lib.tmdn_csr_build_f32(*args, ws.data_ptr(), nbytes, stream)
    torch.cuda.synchronize()
    assert rc == 0, lib.tmdn_last_error()
    e = int(nf)
    return rp.cpu(), col[:e].cpu(), geom[:e].cpu(), e


def _cases():
    rng = np.random.RandomState(77)
    sizes = rng.randint(1, 40, 60)
    batch = np.repeat(np.arange(60, dtype=np.int64), sizes)
    yield "molecules", BRUTE, (rng.rand(batch.size, 3) * 6).astype(np.float32), batch, None, False, 5.0
    L = 40.0
    pos = (rng.rand(6000, 3) * L - L / 2).astype(np.float32)
    yield "periodic_cell", CELL, pos, np.zeros(6000, dtype=np.int64), np.diag([L, L, L]), True, 5.0
    pos2 = (rng.rand(900, 3) * 14).astype(np.float32)          # a different size next: stale scratch from the call before
    yield "single", BRUTE, pos2, np.zeros(900, dtype=np.int64), None, False, 4.0


def test_static_scratch_csr_build_matches_the_workspace_entry_point():
    lib = _lib()
    dev = torch.device("cuda:0")
    for name, strategy, pos, batch, box, periodic, hi in _cases():
        tp, tb = torch.from_numpy(pos).to(dev), torch.from_numpy(batch).to(dev)
        cap = 128 * pos.shape[0]
        ref = _run(lib, False, strategy, tp, tb, box, periodic, hi, cap)
        assert ref[3] > pos.shape[0], name
        for call in range(2):                                  # two calls in a row share the static array
            got = _run(lib, True, strategy, tp, tb, box, periodic, hi, cap)
            assert got[3] == ref[3], (name, call)
            assert torch.equal(got[0], ref[0]) and torch.equal(got[1], ref[1]), (name, call)
            assert torch.equal(got[2], ref[2]), (name, call)   # same kernels, same rounding: bit-identical geometry


def test_static_scratch_rejects_sizes_beyond_the_limit():
    lib = _lib()
    dev = torch.device("cuda:0")
    n = 200_000                                                # cell strategy: ~12 MB of scratch > 4 MiB
    L = 126.0
    box = np.diag([L, L, L])
    boxp = (ctypes.c_double * 9)(*box.ravel())
    assert lib.tmdn_csr_build_workspace_bytes(CELL, n, boxp, 5.0) > STATIC_BYTES
    pos = torch.zeros(n, 3, device=dev)
    batch = torch.zeros(n, dtype=torch.long, device=dev)
    rp, col, geom, nf = _outputs(n, 1024, dev)
    rc = lib.tmdn_csr_build_static_f32(CELL, pos.data_ptr(), batch.data_ptr(), n, boxp, 1, 0.0, 5.0, 1024, rp.data_ptr(),
                                       col.data_ptr(), geom.data_ptr(), nf.data_ptr(), None)
    assert rc != 0 and b"TMDN_CSR_STATIC_SCRATCH_BYTES" in lib.tmdn_last_error()
    torch.cuda.synchronize()
    assert int(nf) == -7                                       # nothing was launched
