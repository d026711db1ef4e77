"""ctypes binding of librt_b200.so -- the C-ABI in include/rt_b200.h plus the host mirror (host/host_api.cpp).

This package holds no compute of its own: every render / trace call goes straight into the CUDA library.  If the
library is missing it is built in-tree with the package Makefile (nvcc, sm_100a); if that fails the import raises --
there is no Python or CPU fallback.

The directory name carries a hyphen (it is the reference's name + "_b200"), so import it through the repo-root shim:
``import rtp_b200 as rt``.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RTB_LIB") or os.path.join(HERE, "librt_b200.so")  # RTB_LIB: a tuning build (make variant)

# numpy images of the reference PODs that cross the boundary (include/rt_b200.h)
SPHERE = np.dtype({"names": ["center", "radius", "material_idx"], "formats": [("<f4", 3), "<f4", "<i4"],
                   "offsets": [0, 12, 16], "itemsize": 32})
PLANE = np.dtype({"names": ["type", "D", "material_idx", "w", "u", "v", "base", "normal"],
                  "formats": ["<i4", "<f4", "<i4", ("<f4", 3), ("<f4", 3), ("<f4", 3), ("<f4", 3), ("<f4", 3)],
                  "offsets": [0, 4, 8, 12, 24, 36, 48, 60], "itemsize": 80})
MATERIAL = np.dtype({"names": ["type", "fuzz", "ir", "absorption", "albedo", "emit", "tex_obj", "cpu_tex"],
                     "formats": ["<i4", "<f4", "<f4", ("<f4", 3), ("<f4", 3), ("<f4", 3), "<u8", "<u8"],
                     "offsets": [0, 4, 8, 12, 24, 36, 48, 56], "itemsize": 64})
CAMERA = np.dtype({"names": ["origin", "pixel00_loc", "pixel_delta_u", "pixel_delta_v", "background", "image_width",
                             "image_height", "samples_per_pixel", "max_depth"],
                   "formats": [("<f4", 3)] * 5 + ["<i4"] * 4, "offsets": [0, 12, 24, 36, 48, 60, 64, 68, 72],
                   "itemsize": 76})

LAMBERTIAN, METAL, DIELECTRIC, DIFFUSE_LIGHT = 0, 1, 2, 3
QUAD, ELLIPSE, TRIANGLE = 0, 1, 2
KERNEL_MEGA, KERNEL_WAVEFRONT, KERNEL_POOL = 0, 1, 2
RNG_PHILOX, RNG_REFERENCE = 0, 1
TEX_CPU, TEX_GPU = 0, 1
BVH_HOST_SAH, BVH_GPU_LBVH = 0, 1
TILE = 16

C_ABI_SYMBOLS = ["rt_scene_create", "rt_scene_create_ex", "rt_scene_rebuild", "rt_scene_destroy", "rt_scene_get_stats", "rt_render", "rt_render_device",
                 "rt_render_finish", "rt_render_u8", "rt_trace_primary", "rt_trace_primary_ex", "rt_trace_primary_noalloc",
                 "rt_noalloc_max_rays", "rt_tonemap_u8", "rt_tiles_pack",
                 "rt_tiles_unpack", "rt_tiles_packed_floats", "rt_tile_owner", "rt_comm_unique_id", "rt_comm_init_rank",
                 "rt_comm_init_all", "rt_comm_destroy", "rt_comm_rank", "rt_comm_size", "rt_render_tiles", "rt_last_error",
                 "rt_device_count", "rt_set_device", "rt_version"]
COMM_ID_BYTES = 128


class RtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("rt_b200 error %d: %s" % (code, msg))
        self.code = code


class RenderOpts(C.Structure):
    _fields_ = [("struct_size", C.c_int), ("kernel", C.c_int), ("rng", C.c_int), ("exact_math", C.c_int),
                ("tex_mode", C.c_int), ("seed", C.c_ulonglong), ("tile_rank", C.c_int), ("tile_nranks", C.c_int),
                ("count_work", C.c_int), ("top_nodes_in_smem", C.c_int), ("sample_chunk", C.c_int)]


class WorkCounters(C.Structure):
    _fields_ = [(n, C.c_ulonglong) for n in
                ("segments", "node_visits", "sphere_tests", "plane_tests", "shaded_hits", "samples")]


class SceneStats(C.Structure):
    _fields_ = [("num_spheres", C.c_int), ("num_planes", C.c_int), ("num_materials", C.c_int), ("wide_nodes", C.c_int),
                ("wide_depth", C.c_int), ("max_stack", C.c_int), ("leaves", C.c_int), ("sah_cost", C.c_float),
                ("build_ms", C.c_float), ("upload_ms", C.c_float), ("device_bytes", C.c_ulonglong),
                ("bvh_builder", C.c_int), ("input_upload_ms", C.c_float), ("num_hoisted", C.c_int),
                ("grid_spheres", C.c_int), ("grid_cells", C.c_int)]


class SceneOpts(C.Structure):
    _fields_ = [("struct_size", C.c_int), ("bvh_builder", C.c_int)]


class TraceOpts(C.Structure):
    _fields_ = [("struct_size", C.c_int), ("exact_math", C.c_int)]


class TilesInfo(C.Structure):
    _fields_ = [("struct_size", C.c_int), ("rank", C.c_int), ("nranks", C.c_int), ("segments_rank", C.c_ulonglong),
                ("segments_total", C.c_ulonglong), ("kernel_ms", C.c_float), ("device_ms", C.c_float),
                ("gathered_bytes", C.c_ulonglong), ("d2h_bytes", C.c_ulonglong)]


def build(force=False):
    """Compile librt_b200.so + rt_main in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    if force or not os.path.exists(LIB_PATH):
        subprocess.check_call(["make", "-s", "-C", HERE] + (["-B"] if force else []))
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        vp, ci, cf = C.c_void_p, C.c_int, C.c_float
        L.rt_last_error.restype = C.c_char_p
        L.rt_version.restype = C.c_char_p
        L.rt_scene_create.argtypes = [vp, ci, vp, ci, vp, ci, vp, ci, ci, ci, vp]
        L.rt_scene_create_ex.argtypes = [vp, ci, vp, ci, vp, ci, vp, ci, ci, ci, vp, vp]
        L.rt_scene_rebuild.argtypes = [vp, vp, ci, vp, ci]
        L.rt_scene_destroy.argtypes = [vp]
        L.rt_scene_get_stats.argtypes = [vp, vp]
        L.rt_render.argtypes = [vp, vp, vp, vp, vp, vp, vp]
        L.rt_render_device.argtypes = [vp, vp, vp, vp, vp]
        L.rt_render_finish.argtypes = [vp, vp, vp, vp]
        L.rt_render_u8.argtypes = [vp, vp, vp, ci, vp, vp, vp]
        L.rt_trace_primary.argtypes = [vp, vp, ci, cf, cf, vp, vp, vp]
        L.rt_trace_primary_ex.argtypes = [vp, vp, ci, cf, cf, vp, vp, vp, vp]
        L.rt_trace_primary_noalloc.argtypes = [vp, vp, ci, cf, cf, vp, vp, vp, vp]
        L.rt_tile_owner.restype = None
        L.rt_tile_owner.argtypes = [ci, ci, vp, vp]
        L.rt_comm_unique_id.argtypes = [vp]
        L.rt_comm_init_rank.argtypes = [vp, ci, ci, vp]
        L.rt_comm_init_all.argtypes = [ci, vp, vp]
        L.rt_comm_destroy.restype = None
        L.rt_comm_destroy.argtypes = [vp]
        L.rt_comm_rank.argtypes = [vp]
        L.rt_comm_size.argtypes = [vp]
        L.rt_render_tiles.argtypes = [vp, vp, vp, vp, vp, vp]
        L.rt_set_device.argtypes = [ci]
        L.rt_tonemap_u8.argtypes = [vp, vp, ci, ci, ci, vp]
        L.rt_tiles_pack.argtypes = [vp, vp, ci, ci, ci, ci, vp, vp]
        L.rt_tiles_unpack.argtypes = [vp, vp, ci, ci, ci, vp, vp]
        L.rt_tiles_packed_floats.restype = C.c_ulonglong
        L.rt_tiles_packed_floats.argtypes = [ci, ci, ci]
        # host mirror
        L.rth_scene_from_config.argtypes = [C.c_char_p, C.c_char_p, vp]
        L.rth_scene_empty.argtypes = [vp]
        L.rth_scene_destroy.argtypes = [vp]
        L.rth_scene_error.restype = C.c_char_p
        L.rth_scene_error.argtypes = [vp]
        L.rth_scene_counts.argtypes = [vp, vp]
        L.rth_scene_copy.argtypes = [vp, vp, vp, vp, vp]
        L.rth_scene_add_spheres.argtypes = [vp, ci, C.c_uint]
        L.rth_scene_params.argtypes = [vp, vp, vp]
        L.rth_camera_for_frame.argtypes = [vp, ci, vp]
        L.rth_build_camera.argtypes = [ci, ci, vp, vp, cf, ci, ci, vp]
        L.rth_make_planes.argtypes = [vp, vp, vp, ci, vp]
        L.rth_default_config.restype = C.c_char_p
        L.rth_tonemap.argtypes = [vp, C.c_ulonglong, ci, vp]
        L.rth_save.argtypes = [C.c_char_p, ci, ci, vp, ci]
        L.rth_load_texels.argtypes = [C.c_char_p, vp, vp, vp]
        L.rth_bvh_check.argtypes = [vp, ci, vp, ci, vp, vp, vp]
        L.rth_gpu_render.argtypes = [vp, ci, C.c_char_p, ci]
        L.rth_gpu_render_ex.argtypes = [vp, ci, ci, ci, C.c_char_p, ci]
        L.rth_set_output.argtypes = [vp, C.c_char_p, ci, ci, ci, ci, ci]
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise RtError(rc, lib().rt_last_error().decode())


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def device_count():
    return lib().rt_device_count()


# ------------------------------------------------------------------------------------------------ host mirror
class HostScene:
    """read_scene_params + create_scene (src/main.cu:346-550) on the host; arrays in the reference's layouts."""

    def __init__(self, config_text=None, texture=None):
        """texture: None keeps the config's path, "" skips the texture, a path overrides it."""
        self.h = C.c_void_p()
        L = lib()
        if config_text is None:
            L.rth_scene_empty(C.byref(self.h))
        else:
            rc = L.rth_scene_from_config(config_text.encode(), None if texture is None else texture.encode(),
                                         C.byref(self.h))
            if rc != 0:
                msg = L.rth_scene_error(self.h).decode()
                L.rth_scene_destroy(self.h)
                raise ValueError(msg)
        self.warning = L.rth_scene_error(self.h).decode()

    def add_spheres(self, n, seed=1234):
        lib().rth_scene_add_spheres(self.h, int(n), int(seed))
        return self

    def counts(self):
        o = (C.c_int * 5)()
        lib().rth_scene_counts(self.h, o)
        return dict(spheres=o[0], planes=o[1], materials=o[2], tex_w=o[3], tex_h=o[4])

    def arrays(self):
        c = self.counts()
        sp = np.zeros(c["spheres"], SPHERE)
        pl = np.zeros(c["planes"], PLANE)
        mt = np.zeros(c["materials"], MATERIAL)
        tex = np.zeros((c["tex_h"], c["tex_w"], 4), np.float32) if c["tex_w"] else None
        lib().rth_scene_copy(self.h, _ptr(sp), _ptr(pl), _ptr(mt), _ptr(tex))
        return sp, pl, mt, tex

    def params(self):
        o = (C.c_int * 5)()
        f = C.c_float()
        lib().rth_scene_params(self.h, o, C.byref(f))
        return dict(W=o[0], H=o[1], max_depth=o[2], sqrt_spp=o[3], num_frames=o[4], fov=f.value)

    def set_output(self, pattern=None, num_frames=0, W=0, H=0, max_depth=0, sqrt_spp=0):
        lib().rth_set_output(self.h, pattern.encode() if pattern else None, num_frames, W, H, max_depth, sqrt_spp)

    def camera(self, frame=0):
        cam = np.zeros(1, CAMERA)
        lib().rth_camera_for_frame(self.h, frame, _ptr(cam))
        return cam

    def gpu_render(self, fmt=0, parity=False, pipeline=True, split_tiles=False, gpus=0):
        """gpu_render(SceneParams&, SceneData&) of the reference (src/camera.cu:290) over the C-ABI. Returns the log.
        parity: the reference's random stream + arithmetic (rt_main --parity); pipeline=False: files written inside the
        frame loop like the reference; gpus / split_tiles: frames (or every frame's tiles) over several devices."""
        buf = C.create_string_buffer(1 << 16)
        flags = (1 if parity else 0) | (0 if pipeline else 2) | (4 if split_tiles else 0)
        rc = lib().rth_gpu_render_ex(self.h, fmt, flags, int(gpus), buf, len(buf))
        if rc != 0:
            raise RtError(rc, lib().rt_last_error().decode())
        return buf.value.decode()

    def close(self):
        if self.h:
            lib().rth_scene_destroy(self.h)
            self.h = C.c_void_p()


def build_camera(H, W, lookfrom, lookat, vfov, spp, max_depth):
    cam = np.zeros(1, CAMERA)
    a = np.ascontiguousarray(lookfrom, np.float32)
    b = np.ascontiguousarray(lookat, np.float32)
    lib().rth_build_camera(H, W, _ptr(a), _ptr(b), float(vfov), spp, max_depth, _ptr(cam))
    return cam


def make_planes(base, u, v, material_idx, ptype):
    """PlaneData host constructor (include/plane.h:19-28) for n planes."""
    buv = np.ascontiguousarray(np.concatenate([np.asarray(base, np.float32).reshape(-1, 3),
                                               np.asarray(u, np.float32).reshape(-1, 3),
                                               np.asarray(v, np.float32).reshape(-1, 3)], 1), np.float32)
    n = len(buv)
    mat = np.ascontiguousarray(np.broadcast_to(np.asarray(material_idx, np.int32), (n,)))
    typ = np.ascontiguousarray(np.broadcast_to(np.asarray(ptype, np.int32), (n,)))
    out = np.zeros(n, PLANE)
    lib().rth_make_planes(_ptr(buv), _ptr(mat), _ptr(typ), n, _ptr(out))
    return out


def default_config():
    return lib().rth_default_config().decode()


def tonemap_host(fb_sum, divisor):
    fb = np.ascontiguousarray(fb_sum, np.float32)
    out = np.zeros(fb.shape, np.uint8)
    lib().rth_tonemap(_ptr(fb), fb.size, int(divisor), _ptr(out))
    return out


def save_image(path, rgb_u8, fmt=0):
    rgb = np.ascontiguousarray(rgb_u8, np.uint8)
    H, W = rgb.shape[0], rgb.shape[1]
    return lib().rth_save(path.encode(), W, H, _ptr(rgb), fmt) == 0


def load_texels(path):
    w, h = C.c_int(), C.c_int()
    if lib().rth_load_texels(path.encode(), None, C.byref(w), C.byref(h)) != 0:
        return None
    out = np.zeros((h.value, w.value, 4), np.float32)
    lib().rth_load_texels(path.encode(), _ptr(out), C.byref(w), C.byref(h))
    return out


def bvh_check(spheres, planes):
    sp = np.ascontiguousarray(spheres, SPHERE)
    pl = np.ascontiguousarray(planes, PLANE)
    o = (C.c_int * 4)()
    sah, ms = C.c_float(), C.c_float()
    rc = lib().rth_bvh_check(_ptr(sp), len(sp), _ptr(pl), len(pl), o, C.byref(sah), C.byref(ms))
    return rc, dict(nodes=o[0], depth=o[1], max_stack=o[2], leaves=o[3], sah=sah.value, build_ms=ms.value)


# ------------------------------------------------------------------------------------------------ device scene
class Scene:
    """rt_scene handle: the arrays create_scene produces, uploaded as wide BVH + SoA buffers (needs a GPU)."""

    def __init__(self, spheres, planes, materials, tex=None, tex_material=-1, bvh_builder=None):
        self.h = C.c_void_p()
        sp = np.ascontiguousarray(spheres, SPHERE)
        pl = np.ascontiguousarray(planes, PLANE)
        mt = np.ascontiguousarray(materials, MATERIAL)
        tw = th = 0
        if tex is not None:
            tex = np.ascontiguousarray(tex, np.float32)
            th, tw = tex.shape[0], tex.shape[1]
        if bvh_builder is None:
            _check(lib().rt_scene_create(_ptr(sp), len(sp), _ptr(pl), len(pl), _ptr(mt), len(mt), _ptr(tex), tw, th,
                                         tex_material if tex is not None else -1, C.byref(self.h)))
        else:
            so = SceneOpts(C.sizeof(SceneOpts), int(bvh_builder))
            _check(lib().rt_scene_create_ex(_ptr(sp), len(sp), _ptr(pl), len(pl), _ptr(mt), len(mt), _ptr(tex), tw, th,
                                            tex_material if tex is not None else -1, C.byref(so), C.byref(self.h)))

    @classmethod
    def from_host(cls, host, bvh_builder=None):
        sp, pl, mt, tex = host.arrays()
        return cls(sp, pl, mt, tex, 0, bvh_builder=bvh_builder)

    def rebuild(self, spheres, planes):
        """rt_scene_rebuild: new geometry, device BVH build, same materials / texture"""
        sp = np.ascontiguousarray(spheres, SPHERE)
        pl = np.ascontiguousarray(planes, PLANE)
        _check(lib().rt_scene_rebuild(self.h, _ptr(sp), len(sp), _ptr(pl), len(pl)))

    def stats(self):
        s = SceneStats()
        _check(lib().rt_scene_get_stats(self.h, C.byref(s)))
        return {n: getattr(s, n) for n, _ in SceneStats._fields_}

    @staticmethod
    def _opts(kernel=KERNEL_MEGA, rng=RNG_PHILOX, exact=False, tex_mode=TEX_GPU, seed=0, tile_rank=0, tile_nranks=1,
              count_work=False, top_nodes=-1, sample_chunk=0):
        o = RenderOpts()
        o.struct_size = C.sizeof(RenderOpts)
        o.kernel, o.rng, o.exact_math, o.tex_mode, o.seed = kernel, rng, int(exact), tex_mode, seed
        o.tile_rank, o.tile_nranks, o.count_work, o.top_nodes_in_smem = tile_rank, tile_nranks, int(count_work), top_nodes
        o.sample_chunk = sample_chunk
        return o

    def render(self, cam, out=None, **kw):
        """One frame through rt_render (host framebuffer). Returns (sum image [H,W,3], info dict)."""
        cam = np.ascontiguousarray(cam, CAMERA)
        W, H = int(cam["image_width"][0]), int(cam["image_height"][0])
        fb = out if out is not None else np.zeros((H, W, 3), np.float32)
        o = self._opts(**kw)
        segs, ms, wc = C.c_ulonglong(), C.c_float(), WorkCounters()
        _check(lib().rt_render(self.h, _ptr(cam), C.byref(o), _ptr(fb), C.byref(segs), C.byref(ms), C.byref(wc)))
        info = {n: getattr(wc, n) for n, _ in WorkCounters._fields_}
        info.update(segments=segs.value, kernel_ms=ms.value)
        return fb, info

    def render_u8(self, cam, divisor, **kw):
        """One frame through rt_render_u8: render + device-side output stage; returns (uint8 [H,W,3], info)."""
        cam = np.ascontiguousarray(cam, CAMERA)
        W, H = int(cam["image_width"][0]), int(cam["image_height"][0])
        out = np.zeros((H, W, 3), np.uint8)
        o = self._opts(**kw)
        segs, ms = C.c_ulonglong(), C.c_float()
        _check(lib().rt_render_u8(self.h, _ptr(cam), C.byref(o), int(divisor), _ptr(out), C.byref(segs), C.byref(ms)))
        return out, dict(segments=segs.value, kernel_ms=ms.value)

    def render_device(self, cam, d_fb_ptr, stream=None, **kw):
        cam = np.ascontiguousarray(cam, CAMERA)
        o = self._opts(**kw)
        _check(lib().rt_render_device(self.h, _ptr(cam), C.byref(o), C.c_void_p(d_fb_ptr), C.c_void_p(stream or 0)))

    def finish(self):
        segs, ms, wc = C.c_ulonglong(), C.c_float(), WorkCounters()
        _check(lib().rt_render_finish(self.h, C.byref(segs), C.byref(ms), C.byref(wc)))
        info = {n: getattr(wc, n) for n, _ in WorkCounters._fields_}
        info.update(segments=segs.value, kernel_ms=ms.value)
        return info

    def trace_primary(self, rays6, tmin=0.001, tmax=1e30, exact=True):
        """closest hit ids / t.  exact=True: rt_trace_primary (reference leaf arithmetic); exact=False: rt_trace_primary_ex
        with the production arithmetic of the default render."""
        rays = np.ascontiguousarray(rays6, np.float32).reshape(-1, 6)
        n = len(rays)
        typ, idx, t = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n, np.float32)
        if exact:
            _check(lib().rt_trace_primary(self.h, _ptr(rays), n, tmin, tmax, _ptr(typ), _ptr(idx), _ptr(t)))
        else:
            to = TraceOpts(C.sizeof(TraceOpts), 0)
            _check(lib().rt_trace_primary_ex(self.h, _ptr(rays), n, tmin, tmax, C.byref(to), _ptr(typ), _ptr(idx), _ptr(t)))
        return typ, idx, t

    def trace_primary_noalloc(self, rays6, tmin=0.001, tmax=1e30, exact=True, out=None):
        """rt_trace_primary_noalloc: as trace_primary, staged through the library's static device array (no cudaMalloc in the
        call; at most rt_noalloc_max_rays() rays).  out = (type, index, t) arrays to fill, so a caller's loop allocates nothing"""
        rays = np.ascontiguousarray(rays6, np.float32).reshape(-1, 6)
        n = len(rays)
        typ, idx, t = out if out is not None else (np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n, np.float32))
        to = TraceOpts(C.sizeof(TraceOpts), 1 if exact else 0)
        _check(lib().rt_trace_primary_noalloc(self.h, _ptr(rays), n, tmin, tmax, C.byref(to), _ptr(typ), _ptr(idx), _ptr(t)))
        return typ, idx, t

    def render_tiles(self, comm, cam, out=None, want_host=True, **kw):
        """rt_render_tiles: this rank's share of one frame + the NCCL gather to rank 0.  Returns (frame or None, info)."""
        cam = np.ascontiguousarray(cam, CAMERA)
        W, H = int(cam["image_width"][0]), int(cam["image_height"][0])
        rank0 = comm is None or comm.rank == 0
        fb = None
        if rank0 and want_host:
            fb = out if out is not None else np.zeros((H, W, 3), np.float32)
        o = self._opts(**kw)
        ti = TilesInfo()
        ti.struct_size = C.sizeof(TilesInfo)
        _check(lib().rt_render_tiles(self.h, comm.h if comm is not None else None, _ptr(cam), C.byref(o), _ptr(fb), C.byref(ti)))
        return fb, {n: getattr(ti, n) for n, _ in TilesInfo._fields_ if n != "struct_size"}

    def tonemap_u8(self, W, H, divisor, d_fb_ptr=None):
        out = np.zeros((H, W, 3), np.uint8)
        _check(lib().rt_tonemap_u8(self.h, C.c_void_p(d_fb_ptr or 0), W, H, divisor, _ptr(out)))
        return out

    def tiles_pack(self, d_fb_ptr, W, H, rank, nranks, d_packed_ptr, stream=None):
        _check(lib().rt_tiles_pack(self.h, C.c_void_p(d_fb_ptr), W, H, rank, nranks, C.c_void_p(d_packed_ptr),
                                   C.c_void_p(stream or 0)))

    def tiles_unpack(self, d_all_ptr, W, H, nranks, d_fb_ptr, stream=None):
        _check(lib().rt_tiles_unpack(self.h, C.c_void_p(d_all_ptr), W, H, nranks, C.c_void_p(d_fb_ptr),
                                     C.c_void_p(stream or 0)))

    def close(self):
        if self.h:
            lib().rt_scene_destroy(self.h)
            self.h = C.c_void_p()


def tiles_packed_floats(W, H, nranks):
    return int(lib().rt_tiles_packed_floats(W, H, nranks))


def tile_owner(tile, nranks):
    """(rank, k) of a tile id under the interleaved partition (rt_tile_owner; host arithmetic, no GPU needed)"""
    r, k = C.c_int(), C.c_int()
    lib().rt_tile_owner(int(tile), int(nranks), C.byref(r), C.byref(k))
    return r.value, k.value


def set_device(dev):
    _check(lib().rt_set_device(int(dev)))


class Comm:
    """rt_comm: the NCCL communicator of a tile-partitioned frame, one rank per GPU."""

    def __init__(self, handle, rank, nranks):
        self.h, self.rank, self.nranks = handle, rank, nranks

    @staticmethod
    def unique_id():
        buf = (C.c_char * COMM_ID_BYTES)()
        _check(lib().rt_comm_unique_id(buf))
        return bytes(buf)

    @classmethod
    def init_rank(cls, uid, rank, nranks):
        h = C.c_void_p()
        _check(lib().rt_comm_init_rank(uid, rank, nranks, C.byref(h)))
        return cls(h, rank, nranks)

    @classmethod
    def init_all(cls, ndev, devices=None):
        hs = (C.c_void_p * ndev)()
        dv = (C.c_int * ndev)(*devices) if devices is not None else None
        _check(lib().rt_comm_init_all(ndev, dv, hs))
        return [cls(C.c_void_p(hs[i]), i, ndev) for i in range(ndev)]

    @classmethod
    def from_torch(cls, dist, device):
        """one rank per process under torch.distributed: rank 0's id is broadcast as a byte tensor"""
        import torch
        rank, world = dist.get_rank(), dist.get_world_size()
        t = torch.zeros(COMM_ID_BYTES, dtype=torch.uint8, device=device)
        if rank == 0:
            t.copy_(torch.frombuffer(bytearray(cls.unique_id()), dtype=torch.uint8))
        dist.broadcast(t, 0)
        return cls.init_rank(bytes(t.cpu().numpy().tobytes()), rank, world)

    def close(self):
        if self.h:
            lib().rt_comm_destroy(self.h)
            self.h = C.c_void_p()


def pixel_center_rays(cam):
    """Rays through the pixel centres of a CameraData (no jitter): N x 6 float32, row-major pixels."""
    c = np.ascontiguousarray(cam, CAMERA)[0]
    W, H = int(c["image_width"]), int(c["image_height"])
    ii, jj = np.meshgrid(np.arange(W, dtype=np.float32), np.arange(H, dtype=np.float32))
    p = (c["pixel00_loc"] + ii[..., None] * c["pixel_delta_u"]) + jj[..., None] * c["pixel_delta_v"]
    d = (p - c["origin"]).astype(np.float32)
    o = np.broadcast_to(c["origin"], d.shape)
    return np.ascontiguousarray(np.concatenate([o, d], -1).reshape(-1, 6), np.float32)


# ------------------------------------------------------------------------------------------------ configs
def test_config_text(W=200, H=100, fov=90, max_depth=5, sqrt_spp=2, texture="../floor2.jpg", output="test_output_%d.png",
                     num_frames=1):
    """The config create_test_config.py prints (create_test_config.py:4-79): the shipped 3-solid scene with a STATIC
    camera at (15 cos 3.14159, 15 sin 3.14159, 4.5) looking at (0,0,4.5).  Same token order as the reference format."""
    lines = [str(num_frames), output, "%d %d %g" % (W, H, fov),
             "15.0 4.5 3.14159", "0.0 0.0", "0.0 0.0 0.0", "0.0 0.0",
             "0.0 4.5 0.0", "0.0 0.0", "0.0 0.0 0.0", "0.0 0.0",
             "0.0 0.0 3.0", "0.3 0.0 0.0", "3.0", "1.5 0.1", "3",
             "4.0 0.0 6.0", "0.0 0.3 0.0", "3.0", "1.2 0.1", "2",
             "8.0 0.0 9.0", "0.0 0.0 0.3", "3.0", "1.0 0.1", "1",
             "-15.0 -15.0 -1.0", "-15.0 15.0 -1.0", "15.0 15.0 -1.0", "15.0 -15.0 -1.0",
             texture, "1.0 1.0 1.0", "0.3", "4",
             "-15.0 -15.0 10.0", "10.0 10.0 10.0", "-15.0 15.0 10.0", "10.0 10.0 10.0",
             "15.0 15.0 10.0", "10.0 10.0 10.0", "15.0 -15.0 10.0", "10.0 10.0 10.0",
             "%d %d" % (max_depth, sqrt_spp)]
    return "\n".join(lines) + "\n"


def shipped_config_text(W=1080, H=720, fov=50, max_depth=50, sqrt_spp=50, texture="../floor2.jpg", output="render_%d.png",
                        num_frames=100):
    """The token stream of the reference's shipped config.txt (config.txt:1-17; BASELINE.json configs[0]): three glass
    solids, floor quad, four lights, camera on the cylindrical path of src/camera.cu:301-315.  Output path and texture
    tokens are parameters because the shipped ones point at the author's machine / a missing file."""
    lines = [str(num_frames), output, "%d %d %g" % (W, H, fov),
             "15.0 4.5 3.14159    0.0 4.5    0.0 1.0 1.0    0.0 -1.57",
             "0.0 4.5 0.0    0.0 4.5    0.0 1.0 0.0    0.0 -1.57",
             "0.0 0.0 3.0     0.3 0.0 0.0     3.0     1.5     0.1     3",
             "4 0.0 6.0     0.0 0.3 0.0     3.0     1.2     0.1     2",
             "8 0.0 9.0     0.0 0.0 0.3     3.0     1     0.1     1",
             "-15.0 -15.0 -1.0      -15.0 15.0 -1.0       15.0 15.0 -1.0        15.0 -15.0 -1.0 " + texture,
             "1.0 1.0 1.0", "0.3", "4",
             "-15.0 -15.0 1  10.0 10.0 10.0", "-15.0 15.0 1   10.0 10.0 10.0",
             "15.0 15.0 1    10.0 10.0 10.0", "15.0 -15.0 1   10.0 10.0 10.0",
             "%d %d" % (max_depth, sqrt_spp)]
    return "\n".join(lines) + "\n"


def synthetic_texture(w=2000, h=1330, seed=7):
    """A float RGBA texture of floor.jpg's shape (2000x1330) for runs on boxes where the reference's asset is absent:
    linear-light checker + seeded noise, alpha 1 -- the layout stbi_loadf(..., 4) returns (src/main.cu:18)."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    chk = (((xx // 125) + (yy // 95)) & 1).astype(np.float32)
    tex = np.empty((h, w, 4), np.float32)
    tex[..., 0] = 0.15 + 0.6 * chk
    tex[..., 1] = 0.2 + 0.5 * (1 - chk)
    tex[..., 2] = 0.25 + 0.1 * chk
    tex[..., :3] *= (0.85 + 0.3 * rng.random((h, w, 1), dtype=np.float32))
    tex[..., 3] = 1.0
    return tex


# the synthetic configs of BASELINE.json: create_test_config.py scene + N seeded spheres
CONFIGS = {
    "test_config_200x100": dict(n=0, W=200, H=100, sqrt_spp=2, depth=5),
    "spheres10k_1080p_64spp_d16": dict(n=10000, W=1920, H=1080, sqrt_spp=8, depth=16),
    "spheres100k_1080p_256spp_d32": dict(n=100000, W=1920, H=1080, sqrt_spp=16, depth=32),
    "spheres1m_4k_64spp_d8": dict(n=1000000, W=3840, H=2160, sqrt_spp=8, depth=8),
    "spheres100k_8k_1024spp_d32": dict(n=100000, W=7680, H=4320, sqrt_spp=32, depth=32),
    # BASELINE.json configs[0]: the reference's shipped config.txt scene (94 spheres + 105 planes, camera path, depth 50,
    # 2500 spp) WITH a floor texture of floor.jpg's shape through the texture unit (floor.jpg itself cannot travel)
    "shipped_textured_1080x720_2500spp_d50": dict(kind="shipped", n=0, W=1080, H=720, sqrt_spp=50, depth=50),
}


def workload(name):
    """(cfg, HostScene, texture or None) of a named BASELINE workload"""
    cfg = CONFIGS[name]
    if cfg.get("kind") == "shipped":
        hs = HostScene(shipped_config_text(W=cfg["W"], H=cfg["H"], max_depth=cfg["depth"], sqrt_spp=cfg["sqrt_spp"]), texture="")
        return cfg, hs, synthetic_texture()
    hs = synthetic_scene(cfg["n"], W=cfg["W"], H=cfg["H"], sqrt_spp=cfg["sqrt_spp"], depth=cfg["depth"], seed=1234, texture="")
    return cfg, hs, None


def synthetic_scene(n_spheres, W=1920, H=1080, sqrt_spp=8, depth=16, seed=1234, texture=""):
    """HostScene of the test-config scene + n seeded spheres (mt19937(seed)); texture "" = untextured floor."""
    hs = HostScene(test_config_text(W, H, 90, depth, sqrt_spp), texture=texture)
    if n_spheres > 0:
        hs.add_spheres(n_spheres, seed)
    return hs
