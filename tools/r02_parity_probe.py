"""Diagnostic (GPU box): measures the quantities the r02 parity tests put bounds on. Prints instead of asserting."""
import os, sys, time, copy
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oracle
from prosper_b200 import rt_reference as rt, scene_data as sd, scene_zoo

LUM = np.array([0.2126, 0.7152, 0.0722], np.float32)
def rel_rmse(a, b):
    la, lb = a[..., :3] @ LUM, b[..., :3] @ LUM
    return float(np.sqrt(np.mean(((la - lb) / (np.abs(lb) + 1e-2)) ** 2)))
def mism(a, b):
    return (a["drawInstanceIndex"] != b["drawInstanceIndex"]) | (a["primitiveID"] != b["primitiveID"])
def world_for(scene, w, h, flags, spl=1):
    world = rt.World(scene, max_width=w, max_height=h, samples_per_launch=spl, flags=flags)
    world.updateBuffers(); world.buildAccelerationStructures()
    return world

which = sys.argv[1:] or ["dof", "moved", "lights", "bench"]
if "dof" in which:
    for name, scene in (("test_scene", scene_zoo.test_scene()), ("atrium_small", scene_zoo.atrium(160, 90, tex_size=64, detail=0.08))):
        w, h = 384, 256
        world = world_for(scene, w, h, rt.CFG_COUNTERS | rt.CFG_EXACT_FILTER)
        orc = oracle.Oracle(scene)
        flags = sd.FLAG_ACCUMULATE | sd.FLAG_CLAMP_INDIRECT | sd.FLAG_IBL | sd.FLAG_SKIP_HISTORY | sd.FLAG_DOF
        for f in (3, 17):
            pc = sd.reference_pc(frame_index=f, flags=flags, aperture=0.05, focus_distance=3.0, focal_len=0.035)
            ho, margin = orc.trace_primary(pc, w, h); hg = world.trace_primary(pc, w, h)
            bad = mism(ho, hg)
            print(f"dof {name} frame {f}: mismatches {int(bad.sum())} of {w*h} = {bad.mean():.2e}; excl alpha ties {int((bad & ~(margin < 1e-4)).sum())}", flush=True)
        world.close()
if "moved" in which:
    scene = copy.deepcopy(scene_zoo.atrium(160, 90, tex_size=64, detail=0.08))
    assert len(scene.tlas_instances) == 1
    m = sd.trs((0.3, -0.2, 0.5), (0.2, 1.0, 0.1), 0.35, (1.1, 0.9, 1.05))
    t = sd.instance_transforms(m)
    scene.transforms[0] = t; scene.tlas_instances[0]["transform"] = t["modelToWorld"]
    w, h = 160, 90
    world = world_for(scene, w, h, rt.CFG_COUNTERS | rt.CFG_EXACT_FILTER)
    orc = oracle.Oracle(scene)
    for f in (1, 2):
        pc = sd.reference_pc(frame_index=f)
        ho, margin = orc.trace_primary(pc, w, h); hg = world.trace_primary(pc, w, h)
        bad = mism(ho, hg); same = ~bad
        print(f"moved single instance frame {f}: mismatches {int(bad.sum())} excl ties {int((bad & ~(margin < 1e-4)).sum())} hits {(hg['drawInstanceIndex'] != sd.MISS).mean():.2f} bary exact {bool(np.array_equal(ho['bary'][same], hg['bary'][same]))}", flush=True)
    world.close()
if "lights" in which:
    scene = scene_zoo.atrium(160, 90, tex_size=64, detail=0.08, lights=1024)
    w, h, spp = 160, 90, 16
    orc = oracle.Oracle(scene)
    t0 = time.time(); ref, counters = orc.render(sd.reference_pc(max_bounces=4), w, h, spp); print(f"lights1024 oracle {time.time()-t0:.1f}s", flush=True)
    for flags, label in ((rt.CFG_COUNTERS | rt.CFG_EXACT_FILTER, "exact"), (rt.CFG_COUNTERS, "hw")):
        world = world_for(scene, w, h, flags, spl=4)
        r = rt.RtReference(); r.init(world); r.drawUi(maxBounces=4)
        out = r.record(world, (0, 0, w, h), rt.Options(ibl=True), frames=spp).illumination_f32()
        c = world.counters()
        print(f"lights1024 {label}: relRMSE {rel_rmse(out, ref):.4f} bias {(out[..., :3].mean()-ref[..., :3].mean())/ref[..., :3].mean():.2e} rays {c['closestRays']}/{int(counters[0])} {c['shadowRays']}/{int(counters[1])}", flush=True)
        world.close()
if "bench" in which:
    t0 = time.time(); scene = scene_zoo.atrium(480, 270, seed=1, tex_size=1024); print(f"bench scene gen {time.time()-t0:.1f}s tris {scene.triangle_count}", flush=True)
    w, h = 480, 270
    orc = oracle.Oracle(scene)
    for spp in (16, 64):
        t0 = time.time(); ref, counters = orc.render(sd.reference_pc(max_bounces=4), w, h, spp); print(f"bench oracle {spp} spp {time.time()-t0:.1f}s", flush=True)
        for flags, label in ((rt.CFG_COUNTERS, "hw"), (rt.CFG_COUNTERS | rt.CFG_EXACT_FILTER, "exact")):
            world = world_for(scene, w, h, flags, spl=4)
            r = rt.RtReference(); r.init(world); r.drawUi(maxBounces=4)
            out = r.record(world, (0, 0, w, h), rt.Options(ibl=True), frames=spp).illumination_f32()
            c = world.counters()
            print(f"bench scene {label} {spp} spp: relRMSE {rel_rmse(out, ref):.4f} bias {(out[..., :3].mean()-ref[..., :3].mean())/ref[..., :3].mean():.2e} rays {c['closestRays']}/{int(counters[0])} {c['shadowRays']}/{int(counters[1])}", flush=True)
            world.close()
    # oracle vs oracle with other frame indices: the Monte Carlo noise floor of the same metric
    ref2, _ = orc.render(sd.reference_pc(max_bounces=4, frame_index=1001), w, h, 64)
    print(f"bench scene oracle(frames 1..64) vs oracle(frames 1001..1064): relRMSE {rel_rmse(ref2, ref):.4f}", flush=True)
