"""How far the texture units' 8-bit-weight bilinear moves a 16-spp render from the oracle, by texture resolution
(context for the tolerance of tests/test_gpu_parity.py::test_radiance_atrium_with_lights_and_foliage)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oracle
from prosper_b200 import rt_reference as rt, scene_data as sd, scene_zoo

def rel_rmse(a, b, floor=1e-2):
    lum = np.array([0.2126, 0.7152, 0.0722], np.float32)
    la, lb = a[..., :3] @ lum, b[..., :3] @ lum
    return float(np.sqrt(np.mean(((la - lb) / (np.abs(lb) + floor)) ** 2)))

w, h, spp = 160, 90, 16
for tex in (64, 256, 1024):
    scene = scene_zoo.atrium(w, h, tex_size=tex, detail=0.08, bake=False, lights=32)
    orc = oracle.Oracle(scene)
    ref, _ = orc.render(sd.reference_pc(max_bounces=4), w, h, spp)
    other, _ = orc.render(sd.reference_pc(frame_index=1 + spp, max_bounces=4), w, h, spp)  # MC noise yardstick
    for flags, name in ((rt.CFG_EXACT_FILTER, "software filter"), (0, "texture units")):
        world = rt.World(scene, max_width=w, max_height=h, samples_per_launch=4, flags=flags)
        world.updateBuffers(); world.buildAccelerationStructures()
        r = rt.RtReference(); r.init(world); r.drawUi(maxBounces=4)
        out = r.record(world, (0, 0, w, h), rt.Options(ibl=True), frames=spp).illumination_f32()
        world.close()
        print(f"tex {tex:5d} {name:15s} relRMSE {rel_rmse(out, ref):.4f}  bias {abs(out[..., :3].mean() / ref[..., :3].mean() - 1):.5f}"
              f"   (two oracle seeds: relRMSE {rel_rmse(other, ref):.3f})", flush=True)
