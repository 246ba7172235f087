"""ctypes binding of the kernel-logic simulator (tests/hostsim/libhostsim.so). TEST INFRASTRUCTURE.

The simulator runs the per-thread bodies of the CUDA kernels (prosper_b200/csrc/core_*.h) on the CPU so
that `pytest -m "not gpu"` covers the GPU builder / traversal / shading logic. Never imported by the product.
"""
from __future__ import annotations

import ctypes as C
import glob
import os
import subprocess

import numpy as np

import oracle
from prosper_b200.scene_data import HIT

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libhostsim.so")


def build(force: bool = False) -> str:
    root = os.path.dirname(os.path.dirname(_HERE))
    srcs = [os.path.join(_HERE, "hostsim.cpp"), os.path.join(_HERE, "Makefile"), os.path.join(root, "oracle", "oracle.h"),
            os.path.join(root, "include", "prt_types.h")]
    srcs += glob.glob(os.path.join(root, "prosper_b200", "csrc", "core_*.h")) + [os.path.join(root, "prosper_b200", "csrc", "hd.h")]
    newest = max(os.path.getmtime(s) for s in srcs)
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < newest:
        subprocess.run(["make", "-C", _HERE, "-B", "libhostsim.so"], check=True, capture_output=True)
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        l = C.CDLL(build())
        l.hs_create.restype = C.c_void_p
        l.hs_create.argtypes = [C.c_void_p, C.c_int]
        l.hs_destroy.argtypes = [C.c_void_p]
        l.hs_stats.argtypes = [C.c_void_p, C.c_void_p]
        l.hs_trace_counters.argtypes = [C.c_void_p, C.c_void_p]
        l.hs_trace_primary.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_int]
        l.hs_trace_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_void_p, C.c_int]
        l.hs_render.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_int, C.c_void_p]
        l.hs_bc7_decode_block.restype = C.c_int
        l.hs_bc7_encode.restype = None
        l.hs_bc7_encode.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]
        l.hs_build_check.restype = C.c_uint32
        l.hs_build_check.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_void_p]
        l.hs_replace_blas.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32]
        _lib = l
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and a.size else None


def bc7_encode(rgba8: np.ndarray, want_modes: bool = False):
    """The product's BC7 encoder core (core_bc7enc.h: modes 4 / 5 / 6) compiled for the host: uint8 [h, w, 4] -> uint8 blocks."""
    img = np.ascontiguousarray(rgba8, np.uint8)
    h, w, _ = img.shape
    blocks = np.zeros((w // 4) * (h // 4) * 16, np.uint8)
    modes = np.zeros((w // 4) * (h // 4), np.uint8)
    lib().hs_bc7_encode(_p(img), w, h, _p(blocks), _p(modes))
    return (blocks, modes) if want_modes else blocks


class HostSim:
    def __init__(self, scene, exact_material_filter: bool = True, ploc: bool = False, lbvh_blas: bool = False):
        self.scene = scene
        self._s, self._keep = oracle.pack_scene(scene)
        self._ctx = lib().hs_create(C.byref(self._s), (1 if exact_material_filter else 0) | (2 if ploc else 0) | (4 if lbvh_blas else 0))

    def close(self):
        if self._ctx:
            lib().hs_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def replace_blas(self, model, nodes, tris):
        nodes, tris = np.ascontiguousarray(nodes), np.ascontiguousarray(tris)
        lib().hs_replace_blas(self._ctx, model, _p(nodes), nodes.shape[0], _p(tris), tris.shape[0])

    def stats(self):
        out = np.zeros(8, np.uint64)
        lib().hs_stats(self._ctx, _p(out))
        keys = ("buildError", "blasNodes", "blasTris", "tlasNodes", "tlasInstances", "blasLevels", "tlasLevels", "blasRecords")
        return dict(zip(keys, (int(v) for v in out)))

    def trace_counters(self):
        out = np.zeros(5, np.uint64)
        lib().hs_trace_counters(self._ctx, _p(out))
        return dict(zip(("nodes", "tris", "instances", "anyHits", "overflow"), (int(v) for v in out)))

    def trace_primary(self, pc, width, height, threads=0):
        hits = np.zeros(width * height, HIT)
        pc = np.ascontiguousarray(pc.reshape(1))
        lib().hs_trace_primary(self._ctx, _p(pc), width, height, _p(hits), threads)
        return hits

    def trace_rays(self, origin_tmin, dir_tmax, seeds=None, shadow=False, threads=0):
        o = np.ascontiguousarray(origin_tmin, np.float32)
        d = np.ascontiguousarray(dir_tmax, np.float32)
        n = o.shape[0]
        sd = np.ascontiguousarray(seeds, np.uint32) if seeds is not None else np.zeros(n, np.uint32)
        hits = np.zeros(n, HIT)
        lib().hs_trace_rays(self._ctx, _p(o), _p(d), _p(sd), n, 1 if shadow else 0, _p(hits), threads)
        return hits

    def render(self, pc, width, height, samples, rgba=None, threads=0):
        if rgba is None:
            rgba = np.zeros((height, width, 4), np.float32)
        counters = np.zeros(4, np.uint64)
        pc = np.ascontiguousarray(pc.reshape(1))
        lib().hs_render(self._ctx, _p(pc), width, height, samples, _p(rgba), threads, _p(counters))
        return rgba, counters


def build_check(boxes, builder: str = "sah"):
    """Builds the BLAS hierarchy over bare boxes [n, 6] (lo xyz, hi xyz) and validates the wide tree.
    Returns dict(nodes, levels, depth, errors)."""
    import numpy as np
    b = np.ascontiguousarray(boxes, dtype=np.float32).reshape(-1, 6)
    out = (C.c_uint32 * 4)()
    lib().hs_build_check(b.ctypes.data_as(C.c_void_p), len(b), {"sah": 0, "lbvh": 1, "ploc": 2}[builder], out)
    return dict(nodes=out[0], levels=out[1], depth=out[2], errors=out[3])
