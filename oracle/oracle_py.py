"""ctypes front-end for the two CPU checkers -- TEST INFRASTRUCTURE ONLY.

* ``Oracle()``  -> oracle/liboracle.so      (our restatement, oracle/oracle.cpp; always buildable)
* ``Ref()``     -> oracle/_ref/libref_rt.so (the UNMODIFIED reference, built by oracle/Makefile where
                   /root/reference exists; travels to the GPU box as a prebuilt file)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
The product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libref_rt.so")

# numpy images of the reference's PODs (sizes/offsets: SURVEY.md section 8, checked in tests against _ref)
SPHERE = np.dtype({"names": ["center", "radius", "material_idx"], "formats": [("<f4", 3), "<f4", "<i4"],
                   "offsets": [0, 12, 16], "itemsize": 32})
PLANE = np.dtype({"names": ["type", "D", "material_idx", "w", "u", "v", "base", "normal"],
                  "formats": ["<i4", "<f4", "<i4", ("<f4", 3), ("<f4", 3), ("<f4", 3), ("<f4", 3), ("<f4", 3)],
                  "offsets": [0, 4, 8, 12, 24, 36, 48, 60], "itemsize": 80})
MATERIAL = np.dtype({"names": ["type", "fuzz", "ir", "absorption", "albedo", "emit", "tex_obj", "cpu_tex"],
                     "formats": ["<i4", "<f4", "<f4", ("<f4", 3), ("<f4", 3), ("<f4", 3), "<u8", "<u8"],
                     "offsets": [0, 4, 8, 12, 24, 36, 48, 56], "itemsize": 64})
BVHNODE = np.dtype({"names": ["box", "left", "right", "type"], "formats": [("<f4", 6), "<i4", "<i4", "<i4"],
                    "offsets": [0, 24, 28, 32], "itemsize": 36})
CAMERA = np.dtype({"names": ["origin", "pixel00_loc", "pixel_delta_u", "pixel_delta_v", "background", "image_width",
                             "image_height", "samples_per_pixel", "max_depth"],
                   "formats": [("<f4", 3)] * 5 + ["<i4"] * 4, "offsets": [0, 12, 24, 36, 48, 60, 64, 68, 72],
                   "itemsize": 76})

LAMBERTIAN, METAL, DIELECTRIC, DIFFUSE_LIGHT = 0, 1, 2, 3
QUAD, ELLIPSE, TRIANGLE = 0, 1, 2


def build(force=False):
    """(Re)build liboracle.so and, where /root/reference exists, oracle/_ref/libref_rt.so."""
    if force or not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(
            os.path.join(HERE, "oracle.cpp")):
        subprocess.check_call(["make", "-s", "-C", HERE, ORACLE_SO])
    if os.path.isdir("/root/reference/src"):
        if force or not os.path.exists(REF_SO) or os.path.getmtime(REF_SO) < os.path.getmtime(
                os.path.join(HERE, "ref_harness.cu")):
            subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


def _fp(a):
    return a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


class _Checker:
    """Common surface of the two libraries; `p` is the symbol prefix ('orc_' or 'ref_')."""

    def __init__(self, path, p):
        self.lib = C.CDLL(path)
        self.p = p
        L = self.lib
        g = lambda n: getattr(L, p + n)
        g("wang_hash").restype = C.c_uint
        g("wang_hash").argtypes = [C.c_uint]
        g("random_float").restype = C.c_float
        g("reflectance").restype = C.c_float
        g("reflectance").argtypes = [C.c_float, C.c_float]
        g("hit_sphere").argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        g("hit_plane").argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        g("aabb_hit").argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float]
        g("material_scatter").argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_void_p]
        g("build_camera").argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_int, C.c_void_p]

    def fn(self, name):
        return getattr(self.lib, self.p + name)

    # ---- KATs
    def wang_hash(self, s):
        return self.fn("wang_hash")(C.c_uint(s & 0xFFFFFFFF))

    def random_floats(self, seed, n):
        s = C.c_uint(seed)
        return [self.fn("random_float")(C.byref(s)) for _ in range(n)], s.value

    def reflectance(self, c, r):
        return self.fn("reflectance")(c, r)

    def random_in_unit_sphere(self, seed):
        s = C.c_uint(seed)
        o = np.zeros(3, np.float32)
        self.fn("random_in_unit_sphere")(C.byref(s), _fp(o))
        return o, s.value

    def random_in_hemisphere(self, n, seed):
        s = C.c_uint(seed)
        o = np.zeros(3, np.float32)
        n = _f32(n)
        self.fn("random_in_hemisphere")(_fp(n), C.byref(s), _fp(o))
        return o, s.value

    def hit_sphere(self, ray6, tmin, tmax, cr4):
        rec = np.zeros(11, np.float32)
        ray6, cr4 = _f32(ray6), _f32(cr4)
        h = self.fn("hit_sphere")(_fp(ray6), tmin, tmax, _fp(cr4), _fp(rec))
        return bool(h), rec

    def hit_plane(self, ray6, tmin, tmax, buv9, ptype):
        rec = np.zeros(11, np.float32)
        der = np.zeros(7, np.float32)
        ray6, buv9 = _f32(ray6), _f32(buv9)
        h = self.fn("hit_plane")(_fp(ray6), tmin, tmax, _fp(buv9), int(ptype), _fp(rec), _fp(der))
        return bool(h), rec, der

    def aabb_hit(self, lo, hi, ray6, tmin, tmax):
        lo, hi, ray6 = _f32(lo), _f32(hi), _f32(ray6)
        return bool(self.fn("aabb_hit")(_fp(lo), _fp(hi), _fp(ray6), tmin, tmax))

    def material_scatter(self, ray6, point, normal, t, front, mat, seed):
        s = C.c_uint(seed)
        att = np.zeros(3, np.float32)
        out = np.zeros(6, np.float32)
        ray6, point, normal = _f32(ray6), _f32(point), _f32(normal)
        m = np.ascontiguousarray(mat)
        alive = self.fn("material_scatter")(_fp(ray6), _fp(point), _fp(normal), float(t), int(front), _fp(m), C.byref(s),
                                            _fp(att), _fp(out))
        return bool(alive), att, out, s.value

    def build_camera(self, H, W, lookfrom, lookat, vfov, spp, depth):
        cam = np.zeros(1, CAMERA)
        a, b = _f32(lookfrom), _f32(lookat)
        self.fn("build_camera")(H, W, _fp(a), _fp(b), float(vfov), spp, depth, _fp(cam))
        return cam

    # ---- scenes
    def scene(self, spheres, planes, mats, tex=None, tex_mat=-1):
        return _Scene(self, spheres, planes, mats, tex, tex_mat)


class _Scene:
    def __init__(self, lib, spheres, planes, mats, tex, tex_mat, handle=None):
        self.k = lib
        self.h = C.c_void_p()
        if handle is not None:
            self.h = handle
            return
        sp = np.ascontiguousarray(spheres, SPHERE)
        pl = np.ascontiguousarray(planes, PLANE)
        mt = np.ascontiguousarray(mats, MATERIAL)
        tw = th = 0
        tp = None
        if tex is not None:
            tex = _f32(tex)
            th, tw = tex.shape[0], tex.shape[1]
            tp = _fp(tex)
        name = "scene_create" if lib.p == "orc_" else "scene_from_arrays"
        f = lib.fn(name)
        f.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int,
                      C.c_void_p]
        rc = f(_fp(sp), len(sp), _fp(pl), len(pl), _fp(mt), len(mt), tp, tw, th, tex_mat, C.byref(self.h))
        assert rc == 0

    def counts(self):
        o = (C.c_int * 6)()
        f = self.k.fn("scene_counts")
        f.argtypes = [C.c_void_p, C.c_void_p]
        f(self.h, o)
        return dict(zip(["spheres", "planes", "materials", "nodes", "depth", "tex_mat"], list(o)))

    def nodes(self):
        n = np.zeros(self.counts()["nodes"], BVHNODE)
        if self.k.p == "orc_":
            f = self.k.fn("scene_nodes")
            f.argtypes = [C.c_void_p, C.c_void_p]
            f(self.h, _fp(n))
        else:
            f = self.k.fn("scene_copy")
            f.argtypes = [C.c_void_p] * 5
            f(self.h, None, None, None, _fp(n))
        return n

    def trace(self, rays6, mode=0, tmin=0.001, tmax=1e30, want_rec=True):
        rays6 = _f32(rays6).reshape(-1, 6)
        n = len(rays6)
        typ = np.zeros(n, np.int32)
        idx = np.zeros(n, np.int32)
        rec = np.zeros((n, 11), np.float32) if want_rec else None
        f = self.k.fn("trace")
        f.restype = C.c_int
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
        bad = f(self.h, _fp(rays6), n, mode, tmin, tmax, _fp(typ), _fp(idx), _fp(rec) if want_rec else None)
        assert bad == 0, "id-recording walk disagreed with the reference's hit_bvh on %d rays" % bad
        return typ, idx, rec

    def render(self, cam, nthreads=0, x0=0, xs=1, y0=0, ys=1, hit_mode=1, tex_mode=0, counted=True):
        """float32 [H,W,3] SUM image (unrendered pixels = 0) and the path-segment count."""
        cam = np.ascontiguousarray(cam, CAMERA)
        W, H = int(cam["image_width"][0]), int(cam["image_height"][0])
        fb = np.zeros((H, W, 3), np.float32)
        segs = C.c_ulonglong(0)
        if self.k.p == "orc_":
            f = self.k.fn("render")
            f.argtypes = [C.c_void_p, C.c_void_p] + [C.c_int] * 7 + [C.c_void_p, C.c_void_p]
            f(self.h, _fp(cam), nthreads, x0, xs, y0, ys, hit_mode, tex_mode, _fp(fb), C.byref(segs))
        else:
            if nthreads <= 0:
                nthreads = os.cpu_count() or 1
            f = self.k.fn("render_cpu")
            f.argtypes = [C.c_void_p, C.c_void_p] + [C.c_int] * 5 + [C.c_void_p, C.c_void_p, C.c_int]
            f(self.h, _fp(cam), nthreads, x0, xs, y0, ys, _fp(fb), C.byref(segs), 1 if counted else 0)
        return fb, segs.value

    def close(self):
        if self.h:
            f = self.k.fn("scene_destroy")
            f.argtypes = [C.c_void_p]
            f(self.h)
            self.h = C.c_void_p()


class Oracle(_Checker):
    def __init__(self):
        build()
        super().__init__(ORACLE_SO, "orc_")
        self.lib.orc_tex2d.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, C.c_void_p]
        self.lib.orc_plane_derive.argtypes = [C.c_void_p, C.c_int]

    def tex2d(self, tex, u, v, mode=0):
        tex = _f32(tex)
        o = np.zeros(3, np.float32)
        self.lib.orc_tex2d(_fp(tex), tex.shape[1], tex.shape[0], u, v, mode, _fp(o))
        return o

    def plane_derive(self, planes):
        planes = np.ascontiguousarray(planes, PLANE)
        self.lib.orc_plane_derive(_fp(planes), len(planes))
        return planes


def ref_available():
    return os.path.exists(REF_SO)


class Ref(_Checker):
    def __init__(self):
        super().__init__(REF_SO, "ref_")
        L = self.lib
        L.ref_tex2d_cpu.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p]
        L.ref_load_texels.restype = C.POINTER(C.c_float)
        L.ref_load_texels.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p]
        L.ref_free_texels.argtypes = [C.c_void_p]
        L.ref_scene_from_config.argtypes = [C.c_char_p, C.c_char_p, C.c_void_p]
        L.ref_scene_copy.argtypes = [C.c_void_p] * 5
        L.ref_scene_texture.restype = C.POINTER(C.c_float)
        L.ref_scene_texture.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_gpu_render.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.ref_selfcheck_color.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.ref_cpu_render_config.argtypes = [C.c_void_p]
        if hasattr(L, "ref_gpu_render_config"):
            L.ref_gpu_render_config.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        L.ref_render_cpu_unmodified.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_float, C.c_int,
                                                C.c_int, C.c_void_p]

    def tex2d(self, tex, u, v, mode=0):
        assert mode == 0
        tex = _f32(tex)
        o = np.zeros(3, np.float32)
        self.lib.ref_tex2d_cpu(_fp(tex), tex.shape[1], tex.shape[0], u, v, _fp(o))
        return o

    def load_texels(self, path):
        w, h = C.c_int(), C.c_int()
        p = self.lib.ref_load_texels(path.encode(), C.byref(w), C.byref(h))
        if not p:
            return None
        a = np.ctypeslib.as_array(p, shape=(h.value, w.value, 4)).copy()
        self.lib.ref_free_texels(p)
        return a

    def scene_from_config(self, text, texture_override=None):
        h = C.c_void_p()
        rc = self.lib.ref_scene_from_config(text.encode(), texture_override.encode() if texture_override else None,
                                            C.byref(h))
        assert rc == 0
        return _RefConfigScene(self, h)

    def struct_sizes(self):
        o = (C.c_int * 12)()
        self.lib.ref_struct_sizes(o)
        return list(o)


class _RefConfigScene(_Scene):
    def __init__(self, lib, h):
        super().__init__(lib, None, None, None, None, -1, handle=h)

    def arrays(self):
        c = self.counts()
        sp = np.zeros(c["spheres"], SPHERE)
        pl = np.zeros(c["planes"], PLANE)
        mt = np.zeros(c["materials"], MATERIAL)
        nd = np.zeros(c["nodes"], BVHNODE)
        self.k.lib.ref_scene_copy(self.h, _fp(sp), _fp(pl), _fp(mt), _fp(nd))
        return sp, pl, mt, nd

    def texture(self):
        w, h = C.c_int(), C.c_int()
        p = self.k.lib.ref_scene_texture(self.h, C.byref(w), C.byref(h))
        if not p:
            return None
        return np.ctypeslib.as_array(p, shape=(h.value, w.value, 4)).copy()

    def params(self):
        o = (C.c_int * 6)()
        f = self.k.lib.ref_scene_params_render
        f.argtypes = [C.c_void_p, C.c_void_p]
        f(self.h, o)
        fov = np.frombuffer(bytes(o), np.float32)[5]
        return dict(W=o[0], H=o[1], max_depth=o[2], sqrt_spp=o[3], num_frames=o[4], fov=float(fov))


def gpu_render_program_ref(cfg_scene):
    """The reference program's own gpu_render loop over the config's frames (writes params.output_path); returns the
    (frame, ms, total_rays) lines it printed."""
    buf = C.create_string_buffer(1 << 16)
    rc = cfg_scene.k.lib.ref_gpu_render_config(cfg_scene.h, buf, len(buf))
    if rc != 0:
        raise RuntimeError("ref_gpu_render_config failed: %d" % rc)
    return [(int(a), float(b), int(c)) for a, b, c in (l.split("\t") for l in buf.value.decode().strip().split("\n") if l.count("\t") == 2)]


def gpu_render_ref(scene, cam, reps=1):
    """Reference CUDA render_kernel on the current GPU (baseline). Returns (sum image, kernel ms)."""
    cam = np.ascontiguousarray(cam, CAMERA)
    W, H = int(cam["image_width"][0]), int(cam["image_height"][0])
    fb = np.zeros((H, W, 3), np.float32)
    ms = C.c_float(0)
    rc = scene.k.lib.ref_gpu_render(scene.h, _fp(cam), reps, _fp(fb), C.byref(ms))
    if rc != 0:
        raise RuntimeError("ref_gpu_render failed rc=%d" % rc)
    return fb, ms.value
