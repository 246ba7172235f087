"""prosper_b200 -- B200-native (sm_100a) implementation of prosper's RtReference path tracer.

The product is libprt_b200.so (prosper_b200/csrc, C ABI in include/prt.h); this package holds the
host-side mirror of the reference's render interface (rt_reference.py), the data-contract
containers (scene_data.py) and the procedural scene generators (scenes.py, scene_zoo.py).
"""
