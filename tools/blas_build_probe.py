"""One BLAS build of the bench scene (atrium, 257,884 triangles) and two warm rebuilds; run plainly, then under ncu
(tools/sah_profile.sh) for the per-kernel launch list of the binned-SAH builder."""
import sys

from prosper_b200 import rt_reference as rt, scene_zoo

scene = scene_zoo.atrium(1920, 1080, seed=1, lights=0, tex_size=256, detail=1.0)
world = rt.World(scene, device=0, max_width=64, max_height=64, samples_per_launch=1, flags=0)
world.updateBuffers()
world.buildAccelerationStructures()
ms = []
for _ in range(2):
    world.buildAccelerationStructures(rebuild_blas=True)
    world.synchronize()
    ms.append(world.timings()["lastBlasMs"])
print("triangles", scene.triangle_count, "warm BLAS rebuild ms", ms, file=sys.stderr)
world.close()
