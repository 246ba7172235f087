"""Bisects the shading difference (prints, no asserts)."""
import os, sys, copy
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oracle
from prosper_b200 import rt_reference as rt, scene_data as sd, scene_zoo

w, h = 96, 64
base = scene_zoo.test_scene()

def run(scene, label, flags=rt.CFG_COUNTERS | rt.CFG_EXACT_FILTER):
    world = rt.World(scene, max_width=w, max_height=h, flags=flags)
    world.updateBuffers(); world.buildAccelerationStructures()
    orc = oracle.Oracle(scene)
    for dt in (5, 6, 7, 8, 9, 10, 0):
        f = sd.FLAG_SKIP_HISTORY | sd.FLAG_ACCUMULATE
        pc = sd.reference_pc(frame_index=1, flags=f, max_bounces=1, draw_type=dt)
        rt._check(world.lib.prt_record(world.handle, rt._p(np.ascontiguousarray(pc.reshape(1))), w, h, 1))
        out = rt.Output(w, h, world).illumination_f32()
        ref, _ = orc.render(pc, w, h, 1)
        d = np.abs(out - ref)
        print(f" [{label}] drawType {dt}: max |diff| per channel {d.reshape(-1,4).max(0)}  n px >1e-3: {(d[...,:3].max(-1)>1e-3).sum()}", flush=True)
        if dt == 0:
            ys, xs = np.nonzero(d[..., :3].max(-1) > 1e-3)
            for y, x in list(zip(ys, xs))[:3]:
                print("    px", x, y, "gpu", out[y, x], "orc", ref[y, x])
    world.close()

run(base, "full scene")
s2 = copy.copy(base); s2.point_lights = base.point_lights[:0]; s2.spot_lights = base.spot_lights[:0]
run(s2, "directional only")
s3 = copy.copy(base); s3.materials = base.materials.copy()
for f in ("baseColorTextureSampler", "metallicRoughnessTextureSampler", "normalTextureSampler"):
    s3.materials[f] = 0
run(s3, "no textures")
run(base, "hw filter", flags=rt.CFG_COUNTERS)
