"""First-contact diagnostic for a GPU box: prints instead of asserting. Usage: python tools/gpu_check.py"""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oracle
from prosper_b200 import rt_reference as rt, scene_data as sd, scene_zoo

def world_for(scene, w, h, flags=rt.CFG_COUNTERS | rt.CFG_EXACT_FILTER, spl=1):
    t = time.time()
    world = rt.World(scene, max_width=w, max_height=h, samples_per_launch=spl, flags=flags)
    world.updateBuffers(); world.buildAccelerationStructures(); world.synchronize()
    print(f"  world+build {time.time()-t:.2f}s", world.bvh_stats(), world.timings(), flush=True)
    return world

def mism(a, b):
    return (a["drawInstanceIndex"] != b["drawInstanceIndex"]) | (a["primitiveID"] != b["primitiveID"])

for name, scene, (w, h) in (("test_scene", scene_zoo.test_scene(), (96, 64)),
                            ("test_scene_bc7", scene_zoo.test_scene(fmt=sd.TEX_BC7), (96, 64)),
                            ("atrium_small_inst", scene_zoo.atrium(160, 90, tex_size=64, detail=0.08, bake=False, lights=32), (160, 90))):
    print(name, flush=True)
    try:
        world = world_for(scene, w, h)
        orc = oracle.Oracle(scene)
        pc = sd.reference_pc()
        ho, margin = orc.trace_primary(pc, w, h)
        hg = world.trace_primary(pc, w, h)
        bad = mism(ho, hg)
        print("  primary mismatches", int(bad.sum()), "of", w * h, "excl ties", int((bad & ~(margin < 1e-4)).sum()),
              "bary exact", bool(np.array_equal(ho["bary"][~bad], hg["bary"][~bad])), world.counters(), flush=True)
        if bad.sum():
            i = np.nonzero(bad)[0][:5]; print("   ", i, ho[i], hg[i], margin[i])
        if "bc7" in name:
            rng = np.random.default_rng(3)
            uv = rng.uniform(-1.5, 2.5, (256, 2)).astype(np.float32)
            for tex in range(1, len(scene.textures)):
                ts = tex
                ex = world.probe_texture(ts, uv, True); hwv = world.probe_texture(ts, uv, False)
                ref = np.array([orc.sample_texture(tex, 0, float(u), float(v)) for u, v in uv])
                print(f"  tex {tex}: exact maxerr {np.abs(ex-ref).max():.3g}  hw maxerr {np.abs(hwv-ref).max():.3g}", flush=True)
        r = rt.RtReference(); r.init(world); r.drawUi(maxBounces=4)
        out = r.record(world, (0, 0, w, h), rt.Options(ibl=True), frames=4).illumination_f32()
        ref, cnt = orc.render(sd.reference_pc(max_bounces=4), w, h, 4)
        d = np.abs(out - ref)
        print("  render maxdiff", float(d.max()), "px differing", int((d.max(-1) > 1e-4).sum()), "means", float(out[..., :3].mean()),
              float(ref[..., :3].mean()), "orc counters", cnt, world.counters(), world.timings(), flush=True)
        world.close()
    except Exception as e:
        print("  FAILED:", repr(e), flush=True)

# full-size atrium: build + throughput
print("atrium 1080p", flush=True)
t = time.time(); scene = scene_zoo.atrium(1920, 1080, tex_size=512); print(f"  gen {time.time()-t:.1f}s tris {scene.triangle_count}", flush=True)
for flags, label in ((0, "timed"), (rt.CFG_COUNTERS, "counters")):
    world = world_for(scene, 1920, 1080, flags=flags, spl=2)
    r = rt.RtReference(); r.init(world); r.drawUi(maxBounces=4)
    for it in range(3):
        world.reset_counters()
        t = time.time()
        r.record(world, (0, 0, 1920, 1080), rt.Options(ibl=True), frames=4)
        world.synchronize()
        dt = time.time() - t
        c = world.counters(); tm = world.timings()
        rays = c["closestRays"] + c["shadowRays"]
        print(f"  [{label}] 4 spp wall {dt*1e3:.1f} ms, record {tm['lastRecordMs']:.1f} ms, rays {rays/1e6:.1f} M -> {rays/tm['lastRecordMs']/1e3:.1f} Mrays/s;"
              f" closest {tm['lastTraceClosestMs']:.2f} ms", c, flush=True)
    if label == "timed":
        from prosper_b200 import imageio
        img = rt.Output(1920, 1080, world).illumination_f32()
        os.makedirs("gpurun_out", exist_ok=True)
        imageio.save_preview("gpurun_out/atrium_preview.png", img[::2, ::2])
    world.close()
