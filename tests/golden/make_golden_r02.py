#!/usr/bin/env python
"""Round-2 fixtures, generated like make_golden.py from the UNMODIFIED reference (oracle/_ref/libref_rt.so):

    python tests/golden/make_golden_r02.py

textured_spheres.npz -- a scene whose TEXTURED material sits on spheres, so get_sphere_uv (include/sphere.h:16-22) feeds
tex2D_cpu (include/materials.h:20-51): the reference's host frame (Camera::render_cpu over ray_color_host) and the (u, v) of
its primary hits.  The shipped scene only ever textures the floor plane, so none of make_golden.py's fixtures reach that code.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle_py as O  # noqa: E402
from make_golden import procedural_texture, zero_pad  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def main():
    O.build()
    ref = O.Ref()
    import rtp_b200 as rt
    hs = rt.synthetic_scene(400, W=96, H=54, sqrt_spp=2, depth=8, seed=99, texture="")
    sp, pl, mt, _ = hs.arrays()
    # the Lambertian palette material most spheres use becomes the textured one
    lamb = [m for m in np.unique(sp["material_idx"]) if mt["type"][m] == O.LAMBERTIAN]
    tex_mat = int(max(lamb, key=lambda m: (sp["material_idx"] == m).sum()))
    mt["albedo"][tex_mat] = [0.9, 0.9, 0.9]
    big = np.argsort(-sp["radius"])[:40]           # and the 40 largest spheres wear it, so the camera sees plenty of it
    sp["material_idx"][big] = tex_mat
    tex = procedural_texture()
    cam = hs.camera(0)
    rs = ref.scene(sp, pl, mt, tex, tex_mat)
    fb, segs = rs.render(cam)
    trace_cam = rt.build_camera(90, 160, [-15, 0, 4.5], [0, 0, 4.5], 90.0, 1, 1)
    rays = rt.pixel_center_rays(trace_cam)
    ty, ix, rec = rs.trace(rays, 0)
    on_tex = (ty == 0) & (sp["material_idx"][np.clip(ix, 0, len(sp) - 1)] == tex_mat)
    assert on_tex.sum() > 500, on_tex.sum()
    np.savez_compressed(os.path.join(OUT, "textured_spheres.npz"), spheres=zero_pad(sp), planes=zero_pad(pl), materials=zero_pad(mt),
                        tex=tex, tex_mat=np.array([tex_mat]), camera=cam, fb_sum=fb, segments=np.array([segs]),
                        trace_camera=trace_cam, type_brute=ty.astype(np.int8), index_brute=ix, rec_brute=rec)
    print("textured_spheres.npz", os.path.getsize(os.path.join(OUT, "textured_spheres.npz")), "primary hits on textured spheres:", int(on_tex.sum()))


if __name__ == "__main__":
    main()
