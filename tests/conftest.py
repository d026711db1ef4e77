import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def golden(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


@pytest.fixture(scope="session")
def orc():
    from oracle import oracle_py as O
    return O.Oracle()


@pytest.fixture(scope="session")
def ref():
    """the unmodified reference (only where oracle/_ref was built); tests that need it skip otherwise"""
    from oracle import oracle_py as O
    if os.path.isdir("/root/reference/src"):
        O.build()
    if not O.ref_available():
        pytest.skip("oracle/_ref/libref_rt.so not present")
    return O.Ref()


@pytest.fixture(scope="session")
def rt():
    import rtp_b200
    rtp_b200.lib()
    return rtp_b200


@pytest.fixture(scope="session")
def emul():
    """host build of the product's device headers (tests/emul) -- test infrastructure, see emul_shim.h"""
    import ctypes as C
    d = os.path.join(ROOT, "tests", "emul")
    so = os.path.join(d, "libemul.so")
    csrc = os.path.join(ROOT, "ray-tracing-practice_b200", "csrc")
    srcs = [os.path.join(d, "emul.cpp"), os.path.join(csrc, "scene_pack.cpp"), os.path.join(csrc, "wide_bvh.cpp")]
    deps = srcs + [os.path.join(d, "emul_shim.h")] + [os.path.join(csrc, f) for f in os.listdir(csrc)]
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(f) for f in deps):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared", "-pthread",
                               "-I/usr/local/cuda/include", "-o", so] + srcs)
    E = C.CDLL(so)
    vp, ci, cf = C.c_void_p, C.c_int, C.c_float
    E.emu_scene_create.argtypes = [vp, ci, vp, ci, vp, ci, vp, ci, ci, ci, vp]
    E.emu_scene_create2.argtypes = [vp, ci, vp, ci, vp, ci, vp, ci, ci, ci, ci, vp]
    E.emu_lbvh_check.argtypes = [vp, ci, vp, ci, vp]
    E.emu_scene_destroy.argtypes = [vp]
    E.emu_trace.argtypes = [vp, vp, ci, cf, cf, ci, vp, vp, vp]
    E.emu_render.argtypes = [vp, vp, ci, ci, ci, C.c_ulonglong, vp, vp]
    E.emu_bvh_check.argtypes = [vp, ci, vp, ci, vp, vp]
    return E


class EmuScene:
    def __init__(self, E, sp, pl, mt, tex=None, tex_mat=-1, builder=0):
        import ctypes as C
        self.E = E
        self.h = C.c_void_p()
        self.keep = [np.ascontiguousarray(sp), np.ascontiguousarray(pl), np.ascontiguousarray(mt)]
        tw = th = 0
        tp = None
        if tex is not None:
            tex = np.ascontiguousarray(tex, np.float32)
            th, tw = tex.shape[:2]
            tp = tex.ctypes.data
        E.emu_scene_create2(self.keep[0].ctypes.data, len(sp), self.keep[1].ctypes.data, len(pl), self.keep[2].ctypes.data,
                            len(mt), tp, tw, th, tex_mat, builder, C.byref(self.h))

    def trace(self, rays, exact=1, tmin=0.001, tmax=1e30):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        n = len(rays)
        ty, ix, t = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n, np.float32)
        self.E.emu_trace(self.h, rays.ctypes.data, n, tmin, tmax, exact, ty.ctypes.data, ix.ctypes.data, t.ctypes.data)
        return ty, ix, t

    def render(self, cam, rng=1, exact=1, tex_mode=0, seed=0):
        import ctypes as C
        cam = np.ascontiguousarray(cam)
        W, H = int(cam["image_width"][0]), int(cam["image_height"][0])
        fb = np.zeros((H, W, 3), np.float32)
        cnt = (C.c_ulonglong * 6)()
        self.E.emu_render(self.h, cam.ctypes.data, rng, exact, tex_mode, seed, fb.ctypes.data, cnt)
        return fb, list(cnt)


@pytest.fixture(scope="session")
def EmuSceneCls():
    return EmuScene


def has_gpu():
    try:
        import rtp_b200
        return rtp_b200.device_count() > 0
    except Exception:
        return False
