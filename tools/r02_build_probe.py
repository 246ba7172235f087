"""Builder latency probe: warm BLAS rebuild and TLAS rebuild times on the bench scenes, plus the integer
trace counters of one fixed frame -- identical counters before and after a builder change mean an identical tree.
Usage: python tools/r02_build_probe.py [atrium] [district] [stress]   (GPU box)"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np

from prosper_b200 import rt_reference as rt, scene_data as sd, scene_zoo


def probe(name, scene, width, height):
    world = rt.World(scene, device=0, max_width=width, max_height=height, samples_per_launch=1, flags=rt.CFG_COUNTERS)
    world.buildAccelerationStructures()
    world.updateBuffers(scene)
    out = {"scene": name, "triangles": int(scene.triangle_count), "instances": int(len(scene.tlas_instances))}
    blas_ms, blas_wall = [], []
    for _ in range(5):
        world.synchronize()
        t0 = time.perf_counter()
        world.buildAccelerationStructures(rebuild_blas=True)
        world.synchronize()
        blas_wall.append((time.perf_counter() - t0) * 1e3)
        blas_ms.append(world.timings()["lastBlasMs"])
    out["blas_all_models_wall_ms"] = float(np.median(blas_wall))  # every model + one TLAS build
    out["blas_last_model_device_ms"] = float(np.median(blas_ms))
    tlas_ms, tlas_wall = [], []
    for _ in range(20):
        world.synchronize()
        t0 = time.perf_counter()
        world.buildTlas(scene.tlas_instances)
        world.synchronize()
        tlas_wall.append((time.perf_counter() - t0) * 1e3)
        tlas_ms.append(world.timings()["lastTlasMs"])
    out["tlas_device_ms"] = float(np.median(tlas_ms))
    out["tlas_wall_ms_from_host_records"] = float(np.median(tlas_wall))
    out["tlas_launches"] = int(world.timings().get("lastTlasLaunches", 0))
    out["bvh"] = world.bvh_stats()
    # one fixed frame: the counters are exact integers, so they fingerprint the tree
    ref = rt.RtReference()
    ref.init(world)
    ref.drawUi(accumulate=False, maxBounces=4)
    world.reset_counters()
    ref.record(world, (0, 0, width, height), nextFrame=1, frames=1)
    world.synchronize()
    cnt = world.counters()
    out["fingerprint"] = {k: int(cnt[k]) for k in ("closestRays", "shadowRays", "nodesVisited", "trisTested", "anyHitCalls", "stackOverflows")}
    world.close()
    return out


if __name__ == "__main__":
    which = sys.argv[1:] or ["atrium", "district", "stress"]
    for w in which:
        if w == "atrium":
            r = probe(w, scene_zoo.atrium(1920, 1080, seed=1), 1920, 1080)
        elif w == "district":
            r = probe(w, scene_zoo.district(1920, 1080), 1920, 1080)
        else:
            r = probe(w, scene_zoo.stress(1920, 1080, instances=10000), 1920, 1080)
        print(json.dumps(r), flush=True)
