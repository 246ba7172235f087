"""Bisects GPU-vs-oracle differences (prints, no asserts)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oracle
from tests import hostsim
from prosper_b200 import rt_reference as rt, scene_data as sd, scene_zoo

def mism(a, b):
    return (a["drawInstanceIndex"] != b["drawInstanceIndex"]) | (a["primitiveID"] != b["primitiveID"])

rng = np.random.default_rng(5)
for name, scene in (("test_scene", scene_zoo.test_scene()), ("atrium_inst", scene_zoo.atrium(160, 90, tex_size=64, detail=0.08, bake=False, lights=32))):
    print("==", name, flush=True)
    world = rt.World(scene, max_width=160, max_height=90, flags=rt.CFG_COUNTERS | rt.CFG_EXACT_FILTER)
    world.updateBuffers(); world.buildAccelerationStructures()
    orc = oracle.Oracle(scene); hs = hostsim.HostSim(scene)
    n = 20000
    o = np.zeros((n, 4), np.float32); d = np.zeros((n, 4), np.float32)
    o[:, :3] = rng.uniform(-3, 3, (n, 3)) + np.array([0, 2, 0])
    v = rng.normal(size=(n, 3)); d[:, :3] = v / np.linalg.norm(v, axis=1, keepdims=True)
    o[:, 3] = rng.uniform(0, 0.2, n)
    d[:, 3] = np.where(rng.random(n) < 0.5, np.inf, rng.uniform(0.5, 10, n))
    d[:50, 0:3] = (0, -1, 0); d[50:100, 0:3] = (1, 0, 0); d[100:110, 0:3] = (0, 0, 0)
    seeds = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
    ho, margin = orc.trace_rays(o, d, seeds)
    hg = world.trace_rays(o, d, seeds)
    bad = mism(ho, hg)
    print(" closest: gpu vs oracle", int(bad.sum()), flush=True)
    for i in np.nonzero(bad)[0][:12]:
        print("  ray", i, "o", o[i], "d", d[i], "oracle", ho[i], "gpu", hg[i], "margin", margin[i])
    # same rays through the CPU copy of the traversal over the GPU-built BLASes
    for m in range(len(scene.model_offsets) - 1):
        nodes, tris = world.read_blas(m)
        if len(nodes):
            hs.replace_blas(m, nodes, tris)
    hh = hs.trace_rays(o, d, seeds)
    print(" closest: hostsim(GPU-built BLAS) vs oracle", int(mism(ho, hh).sum()), " vs gpu", int(mism(hg, hh).sum()), flush=True)
    # no tMin / tMax variation
    o2 = o.copy(); o2[:, 3] = 0; d2 = d.copy(); d2[:, 3] = np.inf
    print(" closest tmin=0,tmax=inf:", int(mism(orc.trace_rays(o2, d2, seeds)[0], world.trace_rays(o2, d2, seeds)).sum()), flush=True)
    so, sm = orc.trace_rays(o, d, seeds, shadow=True); sg = world.trace_rays(o, d, seeds, shadow=True)
    print(" shadow occlusion diffs", int(((so["drawInstanceIndex"] != sd.MISS) != (sg["drawInstanceIndex"] != sd.MISS)).sum()), flush=True)
    # render bisect
    w, h = 96, 64
    cam_scene = scene
    for mb, flags, label in ((1, sd.FLAG_SKIP_HISTORY | sd.FLAG_ACCUMULATE, "1 bounce, no ibl"),
                             (1, sd.FLAG_SKIP_HISTORY | sd.FLAG_ACCUMULATE | sd.FLAG_IBL, "1 bounce, ibl"),
                             (2, sd.FLAG_SKIP_HISTORY | sd.FLAG_ACCUMULATE | sd.FLAG_IBL | sd.FLAG_CLAMP_INDIRECT, "2 bounces"),
                             (4, sd.FLAG_SKIP_HISTORY | sd.FLAG_ACCUMULATE | sd.FLAG_IBL | sd.FLAG_CLAMP_INDIRECT, "4 bounces")):
        pc = sd.reference_pc(frame_index=1, flags=flags, max_bounces=mb)
        pcb = np.ascontiguousarray(pc.reshape(1))
        world.reset_counters()
        rt._check(world.lib.prt_record(world.handle, rt._p(pcb), w, h, 1))
        out = rt.Output(w, h, world).illumination_f32()
        ref, cnt = orc.render(pc, w, h, 1)
        hsr, hcnt = hs.render(pc, w, h, 1)
        dd = np.abs(out[..., :3] - ref[..., :3]).max(-1)
        c = world.counters()
        print(f" render [{label}] maxdiff {dd.max():.4g} n>1e-3 {int((dd > 1e-3).sum())} means gpu {out[...,:3].mean():.5f} orc {ref[...,:3].mean():.5f}"
              f" hostsim(gpu blas) {hsr[...,:3].mean():.5f}; rays gpu {c['closestRays']},{c['shadowRays']} orc {cnt[0]},{cnt[1]}", flush=True)
        ys, xs = np.nonzero(dd > 1e-3)
        for y, x in list(zip(ys, xs))[:4]:
            print("    px", x, y, "gpu", out[y, x], "orc", ref[y, x])
    world.close()
