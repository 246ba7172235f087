"""Generates tests/golden/*.  Run from the repo root: python tests/golden/make_golden.py

kat.json          known-answer vectors for the integer RNG and the packers. The reference has no tests for
                  this path; the pcg / pcg3d values were derived from the reference
                  formulas (res/shader/common/random.glsl:7-28) with plain Python integers, the packing
                  values with the vendored glm. They pin the oracle.
test_scene_*.npz  oracle outputs (primary hit records, a 4-bounce image) on scene_zoo.test_scene(): they pin
                  the CUDA path on the GPU box, where neither /root/reference nor this generator's inputs
                  other than the repo itself are needed.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
HERE = os.path.dirname(os.path.abspath(__file__))

M = 0xFFFFFFFF


def pcg(v):
    state = (v * 747796405 + 2891336453) & M
    word = (((state >> ((state >> 28) + 4)) ^ state) * 277803737) & M
    return ((word >> 22) ^ word) & M


def pcg3d(v):
    x, y, z = [(c * 1664525 + 1013904223) & M for c in v]
    x = (x + y * z) & M; y = (y + z * x) & M; z = (z + x * y) & M
    x ^= x >> 16; y ^= y >> 16; z ^= z >> 16
    x = (x + y * z) & M; y = (y + z * x) & M; z = (z + x * y) & M
    return [x, y, z]


def main():
    kat = {
        "pcg": {str(v): pcg(v) for v in (0, 1, 2, 0xFFFFFFFF, 12345, 959, 0x80000000)},
        "pcg3d": [{"in": list(v), "out": pcg3d(v), "next": pcg3d(pcg3d(v))}
                  for v in ((0, 0, 1), (1, 2, 3), (959, 539, 64), (1919, 1079, 4095))],
        # the values the survey derived independently; make_golden must reproduce them
        "survey_pins": {"pcg(0)": 0x07bb2fe2, "pcg(1)": 0xa8beea3c, "pcg(2)": 0x7a7ecc88, "pcg(0xFFFFFFFF)": 0xe62a4902,
                        "pcg(12345)": 0xf45ead0e, "pcg3d(0,0,1)": [0x7f5db8d2, 0x68e0a5ac, 0x3b10c274],
                        "pcg3d(0,0,1).next": [0x350a78c5, 0x790a2562, 0xbbfc6aa8],
                        "pcg3d(1,2,3)": [0xfa9f79a6, 0x48f2f44c, 0x596f5ab1],
                        "pcg3d(1,2,3).next": [0xe8d7a3ea, 0xf713c12c, 0xc6bf8d86],
                        "pcg3d(959,539,64)": [0xd4ac7b16, 0x447e5c03, 0x16d39b9b],
                        "pcg3d(1919,1079,4095)": [0xa33a8a54, 0x4873dda6, 0x752968d3],
                        "packHalf4x16(1.5,-2,0.1,1)": 0x3c002e66c0003e00, "packSnorm3x10_1x2(0,0,1,0)": 0x1ff00000},
    }
    assert kat["pcg"]["0"] == 0x07bb2fe2 and kat["pcg3d"][1]["out"] == [0xfa9f79a6, 0x48f2f44c, 0x596f5ab1]
    with open(os.path.join(HERE, "kat.json"), "w") as f:
        json.dump(kat, f, indent=1)

    import oracle
    from prosper_b200 import scene_data as sd, scene_zoo
    for name, fmt in (("rgba8", sd.TEX_RGBA8), ("bc7", sd.TEX_BC7)):
        scene = scene_zoo.test_scene(fmt=fmt)
        orc = oracle.Oracle(scene)
        w, h = 96, 64
        out = {}
        for f in (1, 2, 4095):
            hits, margin = orc.trace_primary(sd.reference_pc(frame_index=f), w, h)
            out[f"hits_f{f}"] = hits.view(np.uint32).reshape(-1, 4)
            out[f"margin_f{f}"] = margin
        img, counters = orc.render(sd.reference_pc(max_bounces=4), w, h, 8)
        out["image_8spp"] = img.astype(np.float16)
        out["counters_8spp"] = counters
        np.savez_compressed(os.path.join(HERE, f"test_scene_{name}.npz"), **out)
    print("golden written")


if __name__ == "__main__":
    main()
