#!/usr/bin/env python
"""Dumps what the texture-unit pin test compares (reference kernel's image of the head-on textured quad, the texture, the
camera) so the oracle's restatement of the hardware filter can be fitted offline.  Test infrastructure."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import rtp_b200 as rt
from oracle import oracle_py as O
from test_gpu_gates_r02 import _quad_scene

sp, pl, mt, tex, cam = _quad_scene(rt)
ref = O.Ref()
fr, _ = O.gpu_render_ref(ref.scene(sp, pl, mt, tex, 0), cam, reps=1)
sc = rt.Scene(sp, pl, mt, tex, 0)
fg, _ = sc.render(cam, rng=rt.RNG_REFERENCE, exact=True, tex_mode=rt.TEX_GPU)
orc = O.Oracle()
fo, _ = orc.scene(sp, pl, mt, tex, 0).render(cam, hit_mode=1, tex_mode=1)
np.savez_compressed(os.path.join(ROOT, "gpurun_out", "texpin.npz"), fr=fr, fg=fg, fo=fo, tex=tex, cam=cam, spheres=sp, planes=pl, materials=mt)
print("ref vs ours(TEX_GPU): max %.3g mean %.3g; ref vs oracle: max %.3g mean %.3g" % (np.abs(fr - fg).max(), np.abs(fr - fg).mean(), np.abs(fr - fo).max(), np.abs(fr - fo).mean()))
