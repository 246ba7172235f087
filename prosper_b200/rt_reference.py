"""Host-side mirror of prosper's render interface for the RtReference path, over the C ABI of
libprt_b200.so (include/prt.h).

`World` stands for the acceleration-structure / scene-buffer half of scene::World the pass depends on
(src/scene/World.hpp:49, 64-65; World.cpp:468-961) and `RtReference` for render::RtReference
(src/render/RtReference.hpp:19-107): same method names (init / record / drawUi -> options /
releasePreserved), same Options / Output shapes, same tunables and defaults
(m_accumulate, m_clampIndirect, m_rouletteStartBounce = 3, m_maxBounces = 6), errors as RuntimeError
where the reference throws std::runtime_error (RtRef.cpp:112-113).

There is no CPU fallback: importing works anywhere (so CPU CI can check the ABI), but constructing a
World without a CUDA device raises.
"""
from __future__ import annotations

import ctypes as C
import dataclasses
import os
from typing import Optional

import numpy as np

from . import scene_data as sd

_HERE = os.path.dirname(os.path.abspath(__file__))
# PRT_LIB: an alternative build of the same library (kernel experiments); there is still no fallback
LIB_PATH = os.environ.get("PRT_LIB") or os.path.join(_HERE, "libprt_b200.so")

CFG_COUNTERS, CFG_EXACT_FILTER, CFG_DECODE_BC7, CFG_PLOC, CFG_LBVH_BLAS = 1, 2, 4, 8, 16


class PrtConfig(C.Structure):
    _fields_ = [("device", C.c_int32), ("maxWidth", C.c_uint32), ("maxHeight", C.c_uint32),
                ("samplesPerLaunch", C.c_uint32), ("maxInstances", C.c_uint32), ("flags", C.c_uint32)]


class PrtTextureDesc(C.Structure):
    _fields_ = [("format", C.c_uint32), ("width", C.c_uint32), ("height", C.c_uint32), ("mipLevels", C.c_uint32),
                ("data", C.c_void_p), ("dataBytes", C.c_uint64)]


class PrtCounters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("closestRays", "shadowRays", "nodesVisited", "trisTested", "instancesEntered",
                                          "anyHitCalls", "shadedHits", "textureTaps", "pixelSamples", "kernelLaunches",
                                          "stackOverflows", "shadowNodesVisited", "shadowTrisTested",
                                          "shadowInstancesEntered", "shadowAnyHitCalls")]


class PrtTimings(C.Structure):
    _fields_ = [("lastRecordMs", C.c_float), ("lastTlasMs", C.c_float), ("lastBlasMs", C.c_float),
                ("lastTraceClosestMs", C.c_float), ("lastTraceClosestLaunches", C.c_uint32),
                ("lastRecordLaunches", C.c_uint32), ("lastTlasLaunches", C.c_uint32), ("lastShadeMs", C.c_float),
                ("lastShadeLaunches", C.c_uint32), ("lastTraceShadowMs", C.c_float), ("lastTraceShadowLaunches", C.c_uint32),
                ("ringWraps", C.c_uint32), ("tlasGraphReplays", C.c_uint32), ("tlasGraphCaptures", C.c_uint32)]


class PrtBvhStats(C.Structure):
    _fields_ = [("tlasNodes", C.c_uint32), ("blasNodesTotal", C.c_uint32), ("blasTrianglesTotal", C.c_uint64),
                ("blasCount", C.c_uint32), ("tlasInstances", C.c_uint32), ("blasReferencesTotal", C.c_uint64)]


# every symbol include/prt.h declares (tests/test_abi.py checks the library exports exactly these)
ABI_SYMBOLS = (
    "prt_create", "prt_destroy", "prt_last_error", "prt_stream", "prt_synchronize", "prt_set_geometry_buffer",
    "prt_set_meshes", "prt_set_materials", "prt_set_samplers", "prt_set_texture", "prt_set_sky_cubemap", "prt_set_models",
    "prt_build_blas", "prt_build_tlas", "prt_set_draw_instances", "prt_set_instance_transforms", "prt_set_lights",
    "prt_set_camera", "prt_record", "prt_output_f32_device", "prt_read_illumination_f16", "prt_read_output_f32",
    "prt_release_preserved", "prt_trace_primary", "prt_trace_rays", "prt_export_sum_device", "prt_import_sum_device",
    "prt_export_sum_to_device", "prt_encode_bc7", "prt_compress_texture", "prt_set_tone_map_lut", "prt_tone_map", "prt_tone_map_device",
    "prt_get_counters", "prt_reset_counters", "prt_get_timings", "prt_get_bvh_stats", "prt_probe_texture",
    "prt_debug_read_blas", "prt_debug_read_tlas",
    "prt_rtdi_set_gbuffer", "prt_rtdi_gbuffer_from_primary", "prt_rtdi_read_gbuffer", "prt_rtdi_set_reservoirs", "prt_rtdi_initial_reservoirs",
    "prt_rtdi_read_reservoirs", "prt_rtdi_trace", "prt_rtdi_read_output_f32", "prt_rtdi_read_illumination_f16", "prt_rtdi_release_preserved",
    "prt_compress_texture_static",
)

_lib = None


def load_library():
    """Loads libprt_b200.so. Raises if it has not been built (no fallback of any kind)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(prosper_b200 has no CPU fallback)")
        l = C.CDLL(LIB_PATH)
        l.prt_last_error.restype = C.c_char_p
        l.prt_stream.restype = C.c_void_p
        l.prt_stream.argtypes = [C.c_void_p]
        l.prt_create.argtypes = [C.POINTER(PrtConfig), C.POINTER(C.c_void_p)]
        l.prt_destroy.argtypes = [C.c_void_p]
        l.prt_destroy.restype = None
        vp, u32, u64 = C.c_void_p, C.c_uint32, C.c_uint64
        sigs = {
            "prt_synchronize": [vp], "prt_set_geometry_buffer": [vp, u32, vp, u64], "prt_set_meshes": [vp, vp, vp, u32],
            "prt_set_materials": [vp, vp, u32], "prt_set_samplers": [vp, vp, u32],
            "prt_set_texture": [vp, u32, C.POINTER(PrtTextureDesc)], "prt_set_sky_cubemap": [vp, u32, vp],
            "prt_set_models": [vp, vp, vp, u32], "prt_build_blas": [vp, u32, C.POINTER(u64)], "prt_build_tlas": [vp, vp, u32],
            "prt_set_draw_instances": [vp, vp, u32], "prt_set_instance_transforms": [vp, vp, u32],
            "prt_set_lights": [vp, vp, vp, u32, vp, u32], "prt_set_camera": [vp, vp], "prt_record": [vp, vp, u32, u32, u32],
            "prt_output_f32_device": [vp, C.POINTER(vp)], "prt_read_illumination_f16": [vp, vp, u64],
            "prt_read_output_f32": [vp, vp, u64], "prt_release_preserved": [vp], "prt_trace_primary": [vp, vp, u32, u32, vp],
            "prt_trace_rays": [vp, vp, vp, vp, u32, C.c_int, vp], "prt_export_sum_device": [vp, C.POINTER(vp)],
            "prt_import_sum_device": [vp, vp], "prt_export_sum_to_device": [vp, vp],
            "prt_compress_texture": [vp, vp, u32, u32, u32, u32, vp, u64, C.POINTER(u32), C.POINTER(u32)],
            "prt_compress_texture_static": [vp, vp, u32, u32, u32, u32, vp, u64, C.POINTER(u32), C.POINTER(u32)],
            "prt_set_tone_map_lut": [vp, vp, u32], "prt_tone_map": [vp, vp, vp, u64], "prt_tone_map_device": [vp, vp, C.POINTER(vp)],
            "prt_encode_bc7": [vp, vp, u32, u32, vp, u64], "prt_get_counters": [vp, C.POINTER(PrtCounters)], "prt_reset_counters": [vp],
            "prt_get_timings": [vp, C.POINTER(PrtTimings)], "prt_get_bvh_stats": [vp, C.POINTER(PrtBvhStats)],
            "prt_probe_texture": [vp, u32, vp, vp, u32, C.c_int, vp],
            "prt_debug_read_blas": [vp, u32, vp, u64, vp, u64, C.POINTER(u32), C.POINTER(u32)],
            "prt_debug_read_tlas": [vp, vp, u64, vp, u64, C.POINTER(u32)],
            "prt_rtdi_set_gbuffer": [vp, u32, u32, vp, vp, vp], "prt_rtdi_gbuffer_from_primary": [vp, u32, u32, u32],
            "prt_rtdi_read_gbuffer": [vp, vp, vp, vp], "prt_rtdi_set_reservoirs": [vp, vp], "prt_rtdi_initial_reservoirs": [vp, u32],
            "prt_rtdi_read_reservoirs": [vp, vp], "prt_rtdi_trace": [vp, vp], "prt_rtdi_read_output_f32": [vp, vp, u64],
            "prt_rtdi_read_illumination_f16": [vp, vp, u64], "prt_rtdi_release_preserved": [vp],
        }
        for name, args in sigs.items():
            fn = getattr(l, name)
            fn.argtypes = args
            fn.restype = C.c_int
        _lib = l
    return _lib


def _check(status: int):
    if status != 0:
        raise RuntimeError(f"prt error {status}: {load_library().prt_last_error().decode()}")


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and a.size else None


class World:
    """scene::World as the RtReference pass sees it: scene buffers + BLAS/TLAS on one GPU."""

    def __init__(self, scene: sd.SceneData, device: int = 0, max_width: int = 1920, max_height: int = 1080,
                 samples_per_launch: int = 1, flags: int = 0):
        self.lib = load_library()
        self.scene = scene
        self.device = device
        cfg = PrtConfig(device, max_width, max_height, samples_per_launch, 0, flags)
        self.handle = C.c_void_p()
        st = self.lib.prt_create(C.byref(cfg), C.byref(self.handle))
        if st != 0:
            msg = self.lib.prt_last_error().decode()
            if self.handle:
                self.lib.prt_destroy(self.handle)
            self.handle = None
            raise RuntimeError(f"prt_create failed ({st}): {msg}")
        self.blas_references = np.zeros(len(scene.model_offsets) - 1, np.uint64)
        self._accum_pixels = 0
        self._upload_static()

    # WorldData::init + uploadMeshDatas / uploadMaterialDatas (WorldData.cpp:307-586)
    def _upload_static(self):
        s, l, h = self.scene, self.lib, self.handle
        for i, buf in enumerate(s.geometry_buffers):
            b = np.ascontiguousarray(buf)
            _check(l.prt_set_geometry_buffer(h, i, _p(b), b.nbytes))
        md, mi = np.ascontiguousarray(s.metadatas), np.ascontiguousarray(s.mesh_infos)
        _check(l.prt_set_meshes(h, _p(md), _p(mi), len(md)))
        mats = np.ascontiguousarray(s.materials)
        _check(l.prt_set_materials(h, _p(mats), len(mats)))
        smp = np.ascontiguousarray(s.samplers)
        _check(l.prt_set_samplers(h, _p(smp), len(smp)))
        for i, t in enumerate(s.textures):
            if t is None:
                continue
            data = np.ascontiguousarray(t.data)
            desc = PrtTextureDesc(t.format, t.width, t.height, t.mip_levels, data.ctypes.data, data.nbytes)
            _check(l.prt_set_texture(h, i, C.byref(desc)))
        if s.sky is not None:
            sky = np.ascontiguousarray(s.sky)
            _check(l.prt_set_sky_cubemap(h, s.sky_face_size, _p(sky)))
        off, sub = np.ascontiguousarray(s.model_offsets), np.ascontiguousarray(s.sub_models)
        _check(l.prt_set_models(h, _p(off), _p(sub), len(off) - 1))

    # World::updateBuffers (World.cpp:468-536) + Camera::updateBuffer (Camera.cpp:162-204)
    def updateBuffers(self, scene: Optional[sd.SceneData] = None):
        s = scene or self.scene
        l, h = self.lib, self.handle
        di, xf = np.ascontiguousarray(s.draw_instances), np.ascontiguousarray(s.transforms)
        _check(l.prt_set_draw_instances(h, _p(di), len(di)))
        _check(l.prt_set_instance_transforms(h, _p(xf), len(xf)))
        dl = np.ascontiguousarray(s.directional_light.reshape(1))
        pts, spots = np.ascontiguousarray(s.point_lights), np.ascontiguousarray(s.spot_lights)
        _check(l.prt_set_lights(h, _p(dl), _p(pts), len(pts), _p(spots), len(spots)))
        cam = np.ascontiguousarray(s.camera.reshape(1))
        _check(l.prt_set_camera(h, _p(cam)))

    # World::buildAccelerationStructures (World.cpp:538-575): pending BLAS builds, then the TLAS
    def buildAccelerationStructures(self, scene: Optional[sd.SceneData] = None, rebuild_blas: bool = False) -> bool:
        s = scene or self.scene
        l, h = self.lib, self.handle
        new_blas = False
        for m in range(len(self.blas_references)):
            if self.blas_references[m] == 0 or rebuild_blas:
                ref = C.c_uint64()
                _check(l.prt_build_blas(h, m, C.byref(ref)))
                self.blas_references[m] = ref.value
                new_blas = True
        self.buildTlas(s.tlas_instances)
        return new_blas

    def buildTlas(self, tlas_instances: np.ndarray):
        ti = np.ascontiguousarray(tlas_instances)
        _check(self.lib.prt_build_tlas(self.handle, _p(ti), len(ti)))

    def synchronize(self):
        _check(self.lib.prt_synchronize(self.handle))

    def stream_ptr(self) -> int:
        """The cudaStream_t every call of this context is ordered on (for torch.cuda.ExternalStream)."""
        return int(self.lib.prt_stream(self.handle))

    # ---- multi-GPU: the accumulation as (sum.rgb, count), reduced by the caller over NCCL ----
    def export_sum_ptr(self) -> int:
        p = C.c_void_p()
        _check(self.lib.prt_export_sum_device(self.handle, C.byref(p)))
        return int(p.value)

    def export_sum_to(self, device_ptr: int):
        """Writes the (sum.rgb, count) image into caller-owned device memory (complete on return)."""
        _check(self.lib.prt_export_sum_to_device(self.handle, C.c_void_p(device_ptr)))

    def import_sum(self, device_ptr: int):
        _check(self.lib.prt_import_sum_device(self.handle, C.c_void_p(device_ptr)))


    def counters(self) -> dict:
        c = PrtCounters()
        _check(self.lib.prt_get_counters(self.handle, C.byref(c)))
        return {n: int(getattr(c, n)) for n, _ in PrtCounters._fields_}

    def reset_counters(self):
        _check(self.lib.prt_reset_counters(self.handle))

    def timings(self) -> dict:
        t = PrtTimings()
        _check(self.lib.prt_get_timings(self.handle, C.byref(t)))
        return {n: getattr(t, n) for n, _ in PrtTimings._fields_ if n != "pad"}

    def bvh_stats(self) -> dict:
        b = PrtBvhStats()
        _check(self.lib.prt_get_bvh_stats(self.handle, C.byref(b)))
        return {n: int(getattr(b, n)) for n, _ in PrtBvhStats._fields_}

    def trace_primary(self, pc: np.ndarray, width: int, height: int) -> np.ndarray:
        hits = np.zeros(width * height, sd.HIT)
        pcb = np.ascontiguousarray(pc.reshape(1))
        _check(self.lib.prt_trace_primary(self.handle, _p(pcb), width, height, _p(hits)))
        return hits

    def trace_rays(self, origin_tmin, dir_tmax, seeds=None, shadow=False) -> np.ndarray:
        o = np.ascontiguousarray(origin_tmin, np.float32)
        d = np.ascontiguousarray(dir_tmax, np.float32)
        n = o.shape[0]
        sdv = np.ascontiguousarray(seeds, np.uint32) if seeds is not None else None
        hits = np.zeros(n, sd.HIT)
        _check(self.lib.prt_trace_rays(self.handle, _p(o), _p(d), _p(sdv) if sdv is not None else None, n,
                                       1 if shadow else 0, _p(hits)))
        return hits

    def probe_texture(self, texture_sampler: int, uv: Optional[np.ndarray], exact: bool, texel_xy: Optional[np.ndarray] = None) -> np.ndarray:
        """texture() at `uv` [n, 2] (hardware filter, or the software bilinear when `exact`), or -- with `texel_xy` int32
        [n, 2] -- the raw texels at integer coordinates as the texture units decode them. -> float32 [n, 4]."""
        if texel_xy is not None:
            xy = np.ascontiguousarray(texel_xy, np.int32)
            out = np.zeros((xy.shape[0], 4), np.float32)
            _check(self.lib.prt_probe_texture(self.handle, texture_sampler, None, _p(xy), xy.shape[0], 1 if exact else 0, _p(out)))
            return out
        uv = np.ascontiguousarray(uv, np.float32)
        out = np.zeros((uv.shape[0], 4), np.float32)
        _check(self.lib.prt_probe_texture(self.handle, texture_sampler, _p(uv), None, uv.shape[0], 1 if exact else 0, _p(out)))
        return out

    def encode_bc7(self, rgba: np.ndarray) -> np.ndarray:
        """Texture-cache build step (Texture.cpp:213-296): RGBA8 [h, w, 4] -> BC7 blocks (uint8), on the GPU."""
        img = np.ascontiguousarray(rgba, np.uint8)
        h, w, ch = img.shape
        if ch != 4:
            raise RuntimeError("encode_bc7 wants RGBA8")
        out = np.zeros((w // 4) * (h // 4) * 16, np.uint8)
        _check(self.lib.prt_encode_bc7(self.handle, _p(img), w, h, _p(out), out.nbytes))
        return out

    def set_tone_map_lut(self, lut_rgb9e5: np.ndarray):
        """ToneMap::init (ToneMap.cpp:31-42): the dim^3 R9G9B9E5 table of tone_map.comp (uint32, x fastest)."""
        lut = np.ascontiguousarray(lut_rgb9e5, np.uint32).reshape(-1)
        dim = round(lut.size ** (1.0 / 3.0))
        if dim ** 3 != lut.size:
            raise RuntimeError("the tone-map LUT must be a cube")
        _check(self.lib.prt_set_tone_map_lut(self.handle, _p(lut), dim))

    def tone_map(self, exposure: float, contrast: float) -> np.ndarray:
        """ToneMap::record (ToneMap.cpp:63-134) over the current illumination -> RGBA8 [h, w, 4]."""
        pc = np.array([exposure, contrast], np.float32)
        out = np.zeros(self._accum_pixels * 4, np.uint8)
        _check(self.lib.prt_tone_map(self.handle, _p(pc), _p(out), out.nbytes))
        return out

    def tone_map_device(self, exposure: float, contrast: float) -> int:
        """Same pass, result left on the GPU: device pointer of the RGBA8 image (stream-ordered, not synchronised)."""
        pc = np.array([exposure, contrast], np.float32)
        out = C.c_void_p()
        _check(self.lib.prt_tone_map_device(self.handle, _p(pc), C.byref(out)))
        return int(out.value)

    def read_blas(self, model: int):
        """(nodes uint8 [n, 80], tris uint8 [t, 48]) of the BLAS of `model` (test hook)."""
        nn, nt = C.c_uint32(), C.c_uint32()
        _check(self.lib.prt_debug_read_blas(self.handle, model, None, 0, None, 0, C.byref(nn), C.byref(nt)))
        nodes = np.zeros((nn.value, 80), np.uint8)
        tris = np.zeros((nt.value, 48), np.uint8)
        if nn.value:
            _check(self.lib.prt_debug_read_blas(self.handle, model, _p(nodes), nodes.nbytes, _p(tris), tris.nbytes, None, None))
        return nodes, tris

    def read_tlas(self):
        """(nodes uint8 [n, 80], instance records uint8 [n, 96]) of the current TLAS (test hook)."""
        n = C.c_uint32()
        _check(self.lib.prt_debug_read_tlas(self.handle, None, 0, None, 0, C.byref(n)))
        nodes = np.zeros((n.value, 80), np.uint8)
        inst = np.zeros((n.value, 96), np.uint8)
        if n.value:
            _check(self.lib.prt_debug_read_tlas(self.handle, _p(nodes), nodes.nbytes, _p(inst), inst.nbytes, None))
        return nodes, inst

    def close(self):
        if getattr(self, "handle", None):
            self.lib.prt_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


@dataclasses.dataclass
class Options:  # RtRef.hpp:44-50
    depthOfField: bool = False
    ibl: bool = False
    colorDirty: bool = False
    drawType: int = 0


@dataclasses.dataclass
class Output:  # RtRef.hpp:51-54: `illumination` is the fp16 RGBA image handle; here a host array getter
    width: int
    height: int
    _world: World

    def illumination(self) -> np.ndarray:
        """fp16 RGBA [h, w, 4]: the blit of RtRef.cpp:336-377."""
        out = np.zeros((self.height, self.width, 4), np.float16)
        _check(self._world.lib.prt_read_illumination_f16(self._world.handle, _p(out), out.nbytes))
        return out

    def illumination_f32(self) -> np.ndarray:
        """fp32 RGBA [h, w, 4] with the sample count in .a (m_previousIllumination)."""
        out = np.zeros((self.height, self.width, 4), np.float32)
        _check(self._world.lib.prt_read_output_f32(self._world.handle, _p(out), out.nbytes))
        return out


class RtReference:
    """render::RtReference (RtRef.hpp:19-107)."""

    sMaxBounces = 6  # RtRef.hpp:22
    sFramePeriod = 4096  # RtRef.cpp:31

    def __init__(self):
        self.m_initialized = False
        self.m_accumulationDirty = True
        self.m_accumulate = True  # RtRef.hpp:99-104
        self.m_clampIndirect = True
        self.m_frameIndex = 0
        self.m_rouletteStartBounce = 3
        self.m_maxBounces = self.sMaxBounces

    def init(self, world: World):
        if self.m_initialized:
            raise RuntimeError("RtReference::init called twice")  # WHEELS_ASSERT(!m_initialized), RtRef.cpp:108
        self.m_initialized = True

    def recompileShaders(self):
        """Nothing to recompile: the kernels are ahead-of-time sm_100a code (RtRef.cpp:122-146)."""
        return True

    def drawUi(self, accumulate=None, clampIndirect=None, rouletteStartBounce=None, maxBounces=None):
        """The four UI knobs of RtRef.cpp:148-159; changing one marks the accumulation dirty."""
        for name, v in (("m_accumulate", accumulate), ("m_clampIndirect", clampIndirect),
                        ("m_rouletteStartBounce", rouletteStartBounce), ("m_maxBounces", maxBounces)):
            if v is not None and getattr(self, name) != v:
                setattr(self, name, v)
                self.m_accumulationDirty = True
        self.m_maxBounces = min(self.m_maxBounces, self.sMaxBounces)
        self.m_rouletteStartBounce = min(self.m_rouletteStartBounce, self.m_maxBounces)

    def _pc(self, options: Options, camera_params: dict, skip_history: bool) -> np.ndarray:
        flags = 0  # pcFlags(), RtRef.cpp:68-88
        flags |= sd.FLAG_SKIP_HISTORY if skip_history else 0
        flags |= sd.FLAG_ACCUMULATE if self.m_accumulate else 0
        flags |= sd.FLAG_IBL if options.ibl else 0
        flags |= sd.FLAG_DOF if options.depthOfField else 0
        flags |= sd.FLAG_CLAMP_INDIRECT if self.m_clampIndirect else 0
        return sd.reference_pc(frame_index=self.m_frameIndex, flags=flags, max_bounces=self.m_maxBounces,
                               roulette_start=self.m_rouletteStartBounce, draw_type=options.drawType,
                               aperture=camera_params.get("apertureDiameter", 0.00001),
                               focus_distance=camera_params.get("focusDistance", 1.0),
                               focal_len=camera_params.get("focalLength", 0.035))

    def record(self, world: World, renderArea, options: Options = Options(), nextFrame: int = 0, frames: int = 1,
               camera_params: Optional[dict] = None, cameraChanged: bool = False) -> Output:
        """RtReference::record (RtRef.cpp:161-383). `frames` > 1 enqueues that many consecutive
        frames in one call (the reference records one per swapchain image)."""
        if not self.m_initialized:
            raise RuntimeError("RtReference::record before init")
        x, y, w, h = renderArea
        if x != 0 or y != 0:
            raise RuntimeError("renderArea.offset must be 0 (RtRef.cpp:327)")
        self.m_frameIndex = (self.m_frameIndex + 1) % self.sFramePeriod  # RtRef.cpp:170
        skip = cameraChanged or options.colorDirty or self.m_accumulationDirty  # RtRef.cpp:282-284
        pc = self._pc(options, camera_params or {}, skip)
        _check(world.lib.prt_record(world.handle, _p(np.ascontiguousarray(pc.reshape(1))), w, h, frames))
        world._accum_pixels = w * h
        self.m_frameIndex = (self.m_frameIndex + frames - 1) % self.sFramePeriod
        self.m_accumulationDirty = False
        return Output(w, h, world)

    def releasePreserved(self, world: World):
        _check(world.lib.prt_release_preserved(world.handle))  # RtRef.cpp:385-391
        self.m_accumulationDirty = True


class ToneMap:
    """render::ToneMap (src/render/ToneMap.hpp, ToneMap.cpp): the display transform after the path tracer."""

    def __init__(self):
        self.m_initialized = False
        self.m_exposure = 1.0  # ToneMap.hpp defaults, push_constants/tone_map.h
        self.m_contrast = 1.0

    def init(self, world: World, lut_rgb9e5: np.ndarray):
        """ToneMap.cpp:31-42; `lut_rgb9e5` is the payload of res/texture/tony_mc_mapface.dds (tonemap.read_lut_dds)."""
        if self.m_initialized:
            raise RuntimeError("ToneMap::init called twice")
        world.set_tone_map_lut(lut_rgb9e5)
        self.m_initialized = True

    def drawUi(self, exposure=None, contrast=None):
        """ToneMap.cpp:54-60: both sliders are clamped to [0.001, 10000]."""
        if exposure is not None:
            self.m_exposure = min(max(float(exposure), 0.001), 10000.0)
        if contrast is not None:
            self.m_contrast = min(max(float(contrast), 0.001), 10000.0)

    def record(self, world: World, inColor: Output) -> np.ndarray:
        """ToneMap.cpp:63-134: returns Output::toneMapped as RGBA8 [h, w, 4]."""
        if not self.m_initialized:
            raise RuntimeError("ToneMap::record before init")
        return world.tone_map(self.m_exposure, self.m_contrast).reshape(inColor.height, inColor.width, 4)
