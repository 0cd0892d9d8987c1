"""Loader for liborbgpu.so (the C ABI of include/orbgpu.h).  Fails loudly when the library is
missing: there is no CPU fallback for any compute entry point."""
import atexit
import ctypes as C
import os
import subprocess
import weakref

from ._abi import OrbParams

_HERE = os.path.dirname(os.path.abspath(__file__))
# ORBGPU_LIB: another build of the same library (A/B timing of a kernel change); never a different implementation
LIB_PATH = os.environ.get("ORBGPU_LIB") or os.path.join(_HERE, "liborbgpu.so")
_LIB = None


class OrbGpuError(RuntimeError):
    def __init__(self, code, what, msg):
        super().__init__(f"liborbgpu: {what} failed with status {code}: {msg}")
        self.code = code


def build(verbose=False):
    """Compile liborbgpu.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    subprocess.check_call(["make", "-C", os.path.join(_HERE, "csrc")] + ([] if verbose else ["-s"]))


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise OrbGpuError(-2, "load", f"{LIB_PATH} not built; run `python -c 'import __graft_entry__ as g; g.build()'` "
                                         "(liborbgpu has no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        L.orb_last_error.restype = C.c_char_p
        L.orb_extractor_stream.restype = C.c_void_p
        L.orb_extractor_stream.argtypes = [C.c_void_p]
        L.orb_extractor_launch_count.restype = C.c_int64
        L.orb_extractor_launch_count.argtypes = [C.c_void_p]
        L.orb_extractor_graph_replays.restype = C.c_int64
        L.orb_extractor_graph_replays.argtypes = [C.c_void_p]
        L.orb_extractor_destroy.argtypes = [C.c_void_p]
        L.orb_extractor_destroy.restype = None
        for name in ("orb_frame_stream", "orb_get_call_stream"):
            getattr(L, name).restype = C.c_void_p
        L.orb_frame_stream.argtypes = [C.c_void_p]
        L.orb_frame_launch_count.restype = C.c_int64
        L.orb_frame_launch_count.argtypes = [C.c_void_p]
        L.orb_frame_size.argtypes = [C.c_void_p]
        L.orb_set_call_stream.argtypes = [C.c_void_p, C.c_int]
        _LIB = L
    return _LIB


def check(rc, what):
    if rc != 0:
        raise OrbGpuError(rc, what, lib().orb_last_error().decode(errors="replace"))


# Handles that are still alive when the interpreter exits are closed here, before the CUDA runtime's own exit
# handlers run: destroying them from __del__ during late shutdown would call into a torn-down runtime.
_live = weakref.WeakSet()


def register(obj):
    _live.add(obj)


@atexit.register
def _close_all():
    for o in list(_live):
        try:
            o.close()
        except Exception:
            pass
