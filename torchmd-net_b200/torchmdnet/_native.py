"""Loader of the native libraries (built in-tree by ``torchmd-net_b200/build.py``).

``load()`` registers the torch ops of ``libtmdn_torch.so`` (which links ``libtmdn.so``) exactly once and
raises if the libraries are missing -- nothing in this package computes without them.
"""
import ctypes
import os
import threading

import torch

LIBDIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib")
LIBTMDN = os.path.join(LIBDIR, "libtmdn.so")
LIBSHIM = os.path.join(LIBDIR, "libtmdn_torch.so")

_lock = threading.Lock()
_loaded = False


class NativeLibraryError(RuntimeError):
    pass


def load():
    """Load the C-ABI library and the torch op shim; idempotent."""
    global _loaded
    if _loaded:
        return
    with _lock:
        if _loaded:
            return
        for p in (LIBTMDN, LIBSHIM):
            if not os.path.isfile(p):
                raise NativeLibraryError(
                    "%s not found: build the native code first (python torchmd-net_b200/build.py). "
                    "This package has no CPU / pure-PyTorch fallback." % p)
        # libtmdn.so first and RTLD_GLOBAL so that the shim resolves against this very copy
        ctypes.CDLL(LIBTMDN, mode=ctypes.RTLD_GLOBAL)
        torch.ops.load_library(LIBSHIM)
        _loaded = True


def require_cuda(t, what):
    if not t.is_cuda:
        raise RuntimeError("%s: this build runs on CUDA (sm_100a) only; got a %s tensor and there is no CPU "
                           "fallback" % (what, t.device.type))


def cabi():
    """ctypes handle on libtmdn.so (used by tests to check the exported symbols)."""
    if not os.path.isfile(LIBTMDN):
        raise NativeLibraryError("%s not found: run python torchmd-net_b200/build.py" % LIBTMDN)
    return ctypes.CDLL(LIBTMDN, mode=ctypes.RTLD_GLOBAL)
