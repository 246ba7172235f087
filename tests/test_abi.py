"""CPU: the C-ABI library loads and exports every symbol include/prt.h declares; struct sizes match the
reference's shared CPU/GPU structs; no compute calls without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from prosper_b200 import rt_reference as rt
from prosper_b200 import scene_data as sd

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "prt.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(prt_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = rt.load_library()
    declared = _declared_symbols()
    assert len(declared) >= 30
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/prt.h but not exported"
    assert sorted(rt.ABI_SYMBOLS) == declared, "rt_reference.ABI_SYMBOLS must list exactly the header's entry points"


def test_struct_sizes_match_reference_layouts():
    # sizes measured with g++ on the reference's shader_structs
    assert sd.MATERIAL_DATA.itemsize == 48 and sd.GEOMETRY_METADATA.itemsize == 44
    assert sd.POINT_LIGHT.itemsize == 32 and sd.SPOT_LIGHT.itemsize == 48
    assert sd.CAMERA_UNIFORMS.itemsize == 532 and sd.DRAW_INSTANCE.itemsize == 12
    assert sd.MODEL_INSTANCE_TRANSFORMS.itemsize == 96 and sd.REFERENCE_PC.itemsize == 32
    assert sd.TLAS_INSTANCE.itemsize == 64 and sd.HIT.itemsize == 16
    assert C.sizeof(rt.PrtConfig) == 24 and C.sizeof(rt.PrtCounters) == 120 and C.sizeof(rt.PrtTimings) == 56


def test_header_compiles_as_c_and_cpp(tmp_path):
    import subprocess
    src = tmp_path / "t.c"
    src.write_text('#include "prt.h"\nint main(void){ return sizeof(prt_reference_pc) == 32 ? 0 : 1; }\n')
    for cc, std in (("gcc", "-std=c99"), ("g++", "-std=c++17")):
        exe = tmp_path / ("t_" + cc)
        subprocess.run([cc, std, "-x", "c" if cc == "gcc" else "c++", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
        assert subprocess.run([str(exe)]).returncode == 0


def test_no_cpu_fallback(test_scene):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError, match="no CUDA device|no CPU fallback"):
        rt.World(test_scene)


def test_product_does_not_touch_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "prosper_b200")):
        for f in files:
            if f.endswith((".py", ".h", ".cu", ".cuh")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert not re.search(r"^\s*(import|from)\s+(oracle|tests)|^\s*#\s*include.*(oracle|hostsim)|liboracle|libhostsim|dlopen", text, flags=re.M), f"{f} uses test infrastructure"


def test_reference_pc_defaults():
    pc = sd.reference_pc()
    assert pc["frameIndex"] == 1 and pc["maxBounces"] == 4 and pc["rouletteStartBounce"] == 3
    r = rt.RtReference()
    assert (r.m_accumulate, r.m_clampIndirect, r.m_rouletteStartBounce, r.m_maxBounces) == (True, True, 3, 6)  # RtRef.hpp:99-104
    r.drawUi(maxBounces=9)
    assert r.m_maxBounces == 6
