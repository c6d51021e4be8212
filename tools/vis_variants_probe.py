"""Visibility pass with several builds of the library in ONE process (not part of the test-suite): the bench scene at a
quarter of its viewpoints, per-phase device times and a checksum of the result per build, so that a tuning build can be
timed and compared with the default build's result in a few seconds of GPU time.
usage: python tools/vis_variants_probe.py build_variants/lib_a.so build_variants/lib_b.so ..."""
import sys
import zlib

import numpy as np

sys.path.insert(0, ".")
from sspp_b200 import engine, scenes

nv = 16384
km, dm, occ, known = scenes.synthetic_map(n_vox=512, res=0.3, seed=42)
pts = scenes.surface_points(occ, known, km, dm, 0.3, 4_000_000, seed=45)
poses = scenes.viewpoints_near(pts, 65536, standoff=3.0, seed=46)[:nv]
ref = None
for path in sys.argv[1:]:
    engine._lib = None
    engine.LIB_GPU = path
    eng = engine.Engine(0)
    eng.map_create(0.3, km, dm, True)
    eng.map_upload_dense(occ, known)
    eng.set_model_points(pts)
    best = None
    for rep in range(4):
        st = eng.visibility(poses, 58.0, 45.0, 0.7, 6.0, fetch=False)
        eng.synchronize()
        t = eng.visibility_timings()
        if best is None or t[0] < best[0]:
            best = t.copy()
    st, vp, idx = eng.visibility(poses, 58.0, 45.0, 0.7, 6.0)
    crc = (zlib.crc32(np.ascontiguousarray(vp).tobytes()), zlib.crc32(np.ascontiguousarray(idx).tobytes()))
    if ref is None:
        ref = crc
    print("%s rays %d visible %d phases ms %s result %s" % (path, st.rays, st.visible_pairs, np.round(best[:5], 3), "same as first" if crc == ref else "DIFFERS"))
    eng.close()
