"""TLAS rebuild launch list: python tools/tlas_profile.py [instances]  (run under ncu --metrics gpu__time_duration.sum)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from prosper_b200 import rt_reference as rt, scene_zoo
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
scene = scene_zoo.stress(1920, 1080, instances=n)
w = rt.World(scene, max_width=64, max_height=64)
w.updateBuffers(); w.buildAccelerationStructures(); w.synchronize()
for _ in range(3):
    w.buildTlas(scene.tlas_instances); w.synchronize()
    print(w.timings()["lastTlasMs"], w.timings()["lastTlasLaunches"])
w.close()
