"""ctypes binding of the CPU oracle (oracle/liboracle.so). TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module. The product (prosper_b200, libprt_b200.so) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")
_SOURCES = ["oracle.cpp", "oracle.h", "o_math.h", "o_shading.h", "o_scene.h", "o_bvh.h", "Makefile"]


def build(force: bool = False) -> str:
    """Compile liboracle.so with oracle/Makefile when it is missing or older than its sources."""
    newest = max(os.path.getmtime(os.path.join(_HERE, s)) for s in _SOURCES)
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < newest:
        subprocess.run(["make", "-C", _HERE, "-B", "liboracle.so"], check=True, capture_output=True)
    return _LIB_PATH


class _OrcTexture(C.Structure):
    _fields_ = [("format", C.c_uint32), ("width", C.c_uint32), ("height", C.c_uint32), ("mipLevels", C.c_uint32),
                ("data", C.c_void_p), ("dataBytes", C.c_uint64)]


class _OrcScene(C.Structure):
    _fields_ = [
        ("geometryBufferCount", C.c_uint32), ("geometryBuffers", C.POINTER(C.c_void_p)),
        ("geometryBufferBytes", C.POINTER(C.c_uint64)),
        ("meshCount", C.c_uint32), ("metadatas", C.c_void_p), ("meshInfos", C.c_void_p),
        ("materialCount", C.c_uint32), ("materials", C.c_void_p),
        ("samplerCount", C.c_uint32), ("samplers", C.c_void_p),
        ("textureCount", C.c_uint32), ("textures", C.POINTER(_OrcTexture)),
        ("modelCount", C.c_uint32), ("modelOffsets", C.c_void_p), ("subModels", C.c_void_p),
        ("instanceCount", C.c_uint32), ("tlasInstances", C.c_void_p),
        ("drawInstanceCount", C.c_uint32), ("drawInstances", C.c_void_p),
        ("transformCount", C.c_uint32), ("transforms", C.c_void_p),
        ("directionalLight", C.c_void_p),
        ("pointLightCount", C.c_uint32), ("pointLights", C.c_void_p),
        ("spotLightCount", C.c_uint32), ("spotLights", C.c_void_p),
        ("camera", C.c_void_p),
        ("skyFaceSize", C.c_uint32), ("skyRgba16f", C.c_void_p),
    ]


_lib = None


def lib():
    global _lib
    if _lib is None:
        l = C.CDLL(build())
        l.orc_create.restype = C.c_void_p
        l.orc_create.argtypes = [C.POINTER(_OrcScene), C.c_int]
        l.orc_destroy.argtypes = [C.c_void_p]
        l.orc_bvh_build_seconds.restype = C.c_double
        l.orc_bvh_build_seconds.argtypes = [C.c_void_p]
        l.orc_render.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_int,
                                 C.c_void_p]
        l.orc_trace_primary.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p,
                                        C.c_int]
        l.orc_trace_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int,
                                     C.c_void_p, C.c_void_p, C.c_int]
        l.orc_rtdi_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32] + [C.c_void_p] * 5 + [C.c_int, C.c_void_p]
        l.orc_rtdi_initial_reservoirs.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32] + [C.c_void_p] * 4 + [C.c_int]
        l.orc_rtdi_gbuffer.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32] + [C.c_void_p] * 3 + [C.c_int]
        l.orc_pcg.restype = C.c_uint32
        l.orc_pcg.argtypes = [C.c_uint32]
        l.orc_rng_to_01.restype = C.c_float
        l.orc_rng_to_01.argtypes = [C.c_uint32]
        l.orc_vndf_pdf.restype = C.c_float
        l.orc_vndf_pdf.argtypes = [C.c_void_p, C.c_void_p, C.c_float]
        l.orc_sample_vndf.argtypes = [C.c_void_p, C.c_float, C.c_void_p, C.c_void_p]
        l.orc_eval_brdf_times_nol.argtypes = [C.c_void_p] * 4 + [C.c_float, C.c_float, C.c_void_p]
        l.orc_thin_lens_ray.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float,
                                        C.c_void_p, C.c_void_p]
        l.orc_half_to_float.restype = C.c_float
        l.orc_half_to_float.argtypes = [C.c_uint16]
        l.orc_float_to_half.restype = C.c_uint16
        l.orc_float_to_half.argtypes = [C.c_float]
        l.orc_srgb_to_linear.restype = C.c_float
        l.orc_srgb_to_linear.argtypes = [C.c_float]
        l.orc_bc7_decode_block.restype = C.c_int
        l.orc_intersect_triangle.restype = C.c_int
        l.orc_intersect_triangle.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_void_p,
                                             C.c_void_p, C.c_void_p]
        l.orc_sample_texture.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_float, C.c_float, C.c_void_p]
        l.orc_sample_sky.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        l.orc_unpack_snorm10.argtypes = [C.c_uint32, C.c_void_p]
        l.orc_tone_map.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_float, C.c_float, C.c_void_p, C.c_uint32, C.c_void_p]
        l.orc_tone_map.restype = None
        l.orc_decode_rgb9e5.argtypes = [C.c_uint32, C.c_void_p]
        l.orc_decode_rgb9e5.restype = None
        _lib = l
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and a.size else None


def pack_scene(scene):
    """SceneData -> (orc_scene ctypes struct, list of objects that must outlive it)."""
    s = _OrcScene()
    keep = []

    def arr(a):
        a = np.ascontiguousarray(a)
        keep.append(a)
        return a

    bufs = [arr(b) for b in scene.geometry_buffers]
    nb = len(bufs)
    bufptrs = (C.c_void_p * nb)(*[b.ctypes.data for b in bufs])
    bufbytes = (C.c_uint64 * nb)(*[b.nbytes for b in bufs])
    keep += [bufptrs, bufbytes]
    s.geometryBufferCount = nb
    s.geometryBuffers = C.cast(bufptrs, C.POINTER(C.c_void_p))
    s.geometryBufferBytes = C.cast(bufbytes, C.POINTER(C.c_uint64))
    s.meshCount = len(scene.metadatas)
    s.metadatas = _p(arr(scene.metadatas))
    s.meshInfos = _p(arr(scene.mesh_infos))
    s.materialCount = len(scene.materials)
    s.materials = _p(arr(scene.materials))
    s.samplerCount = len(scene.samplers)
    s.samplers = _p(arr(scene.samplers))
    nt = len(scene.textures)
    tex = (_OrcTexture * nt)()
    keep.append(tex)
    for i, t in enumerate(scene.textures):
        if t is None:
            continue
        d = arr(t.data)
        tex[i] = _OrcTexture(t.format, t.width, t.height, t.mip_levels, d.ctypes.data, d.nbytes)
    s.textureCount = nt
    s.textures = C.cast(tex, C.POINTER(_OrcTexture))
    s.modelCount = len(scene.model_offsets) - 1
    s.modelOffsets = _p(arr(scene.model_offsets))
    s.subModels = _p(arr(scene.sub_models))
    s.instanceCount = len(scene.tlas_instances)
    s.tlasInstances = _p(arr(scene.tlas_instances))
    s.drawInstanceCount = len(scene.draw_instances)
    s.drawInstances = _p(arr(scene.draw_instances))
    s.transformCount = len(scene.transforms)
    s.transforms = _p(arr(scene.transforms))
    s.directionalLight = _p(arr(scene.directional_light.reshape(1)))
    s.pointLightCount = len(scene.point_lights)
    s.pointLights = _p(arr(scene.point_lights))
    s.spotLightCount = len(scene.spot_lights)
    s.spotLights = _p(arr(scene.spot_lights))
    s.camera = _p(arr(scene.camera.reshape(1)))
    s.skyFaceSize = scene.sky_face_size
    s.skyRgba16f = _p(arr(scene.sky)) if scene.sky is not None else None
    return s, keep


class Oracle:
    """One oracle context over a SceneData (prosper_b200.scene_data). Keeps the arrays alive."""

    def __init__(self, scene, brute_force: bool = False):
        self.scene = scene
        l = lib()
        s, self._keep = pack_scene(scene)
        self._s = s
        self._ctx = l.orc_create(C.byref(s), 1 if brute_force else 0)

    def close(self):
        if self._ctx:
            lib().orc_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def bvh_build_seconds(self) -> float:
        return lib().orc_bvh_build_seconds(self._ctx)

    def render(self, pc, width, height, samples, rgba=None, threads=0):
        """Returns (rgba float32 [h, w, 4], counters u64[4] = closest, shadow, shaded, pixel-samples)."""
        if rgba is None:
            rgba = np.zeros((height, width, 4), np.float32)
        counters = np.zeros(4, np.uint64)
        pc = np.ascontiguousarray(pc.reshape(1))
        lib().orc_render(self._ctx, _p(pc), width, height, samples, _p(rgba), threads, _p(counters))
        return rgba, counters

    def trace_primary(self, pc, width, height, threads=0):
        from prosper_b200.scene_data import HIT
        hits = np.zeros(width * height, HIT)
        margin = np.zeros(width * height, np.float32)
        pc = np.ascontiguousarray(pc.reshape(1))
        lib().orc_trace_primary(self._ctx, _p(pc), width, height, _p(hits), _p(margin), threads)
        return hits, margin

    def trace_rays(self, origin_tmin, dir_tmax, seeds=None, shadow=False, threads=0):
        from prosper_b200.scene_data import HIT
        o = np.ascontiguousarray(origin_tmin, np.float32)
        d = np.ascontiguousarray(dir_tmax, np.float32)
        n = o.shape[0]
        sd = np.ascontiguousarray(seeds, np.uint32) if seeds is not None else np.zeros(n, np.uint32)
        hits = np.zeros(n, HIT)
        margin = np.zeros(n, np.float32)
        lib().orc_trace_rays(self._ctx, _p(o), _p(d), _p(sd), n, 1 if shadow else 0, _p(hits), _p(margin), threads)
        return hits, margin

    # ---- ReSTIR DI trace pass ----
    def rtdi_gbuffer(self, frame_index, width, height, threads=0):
        """(albedoRoughness u32 [h, w], normalMetallic u32 [h, w], nonLinearDepth f32 [h, w]) from the primary hits."""
        ar, nm = np.zeros((height, width), np.uint32), np.zeros((height, width), np.uint32)
        z = np.zeros((height, width), np.float32)
        lib().orc_rtdi_gbuffer(self._ctx, C.c_uint32(frame_index), C.c_uint32(width), C.c_uint32(height), _p(ar), _p(nm), _p(z), C.c_int(threads))
        return ar, nm, z

    def rtdi_initial_reservoirs(self, frame_index, gbuffer, threads=0):
        ar, nm, z = (np.ascontiguousarray(a) for a in gbuffer)
        h, w = ar.shape
        out = np.zeros((h, w, 2), np.float32)
        lib().orc_rtdi_initial_reservoirs(self._ctx, C.c_uint32(frame_index), C.c_uint32(w), C.c_uint32(h), _p(ar), _p(nm), _p(z), _p(out), C.c_int(threads))
        return out

    def rtdi_trace(self, draw_type, frame_index, flags, gbuffer, reservoirs, rgba=None, threads=0):
        """One frame of rt/direct_illumination/main.rgen. Returns (rgba [h, w, 4], shadow rays traced)."""
        ar, nm, z = (np.ascontiguousarray(a) for a in gbuffer)
        h, w = ar.shape
        res = np.ascontiguousarray(reservoirs, np.float32)
        if rgba is None:
            rgba = np.zeros((h, w, 4), np.float32)
        pc = np.array([draw_type, frame_index, flags], np.uint32)
        rays = np.zeros(1, np.uint64)
        lib().orc_rtdi_trace(self._ctx, _p(pc), C.c_uint32(w), C.c_uint32(h), _p(ar), _p(nm), _p(z), _p(res), _p(rgba), C.c_int(threads), _p(rays))
        return rgba, int(rays[0])

    def sample_texture(self, tex, smp, u, v):
        out = np.zeros(4, np.float32)
        lib().orc_sample_texture(self._ctx, tex, smp, u, v, _p(out))
        return out

    def sample_sky(self, d):
        d = np.ascontiguousarray(d, np.float32)
        out = np.zeros(3, np.float32)
        lib().orc_sample_sky(self._ctx, _p(d), _p(out))
        return out


def tone_map(illumination_f16: np.ndarray, exposure: float, contrast: float, lut_rgb9e5: np.ndarray) -> np.ndarray:
    """tone_map.comp over an RGBA16F image [h, w, 4] -> RGBA8 [h, w, 4]."""
    img = np.ascontiguousarray(illumination_f16, np.float16)
    h, w, _ = img.shape
    lut = np.ascontiguousarray(lut_rgb9e5, np.uint32).reshape(-1)
    dim = round(lut.size ** (1.0 / 3.0))
    out = np.zeros((h, w, 4), np.uint8)
    lib().orc_tone_map(_p(img.view(np.uint16)), w, h, exposure, contrast, _p(lut), dim, _p(out))
    return out


def decode_rgb9e5(texel: int) -> np.ndarray:
    out = np.zeros(3, np.float32)
    lib().orc_decode_rgb9e5(int(texel), _p(out))
    return out


_REF_STBIR = os.path.join(_HERE, "_ref", "libstbir.so")


def build_ref() -> str:
    """oracle/_ref/libstbir.so from the reference's vendored stb header (only where /root/reference exists; the
    prebuilt file travels to the GPU box). Returns the path, or "" when neither the sources nor a prebuilt file exist."""
    if os.path.exists("/root/reference/ext/stb/stb_image_resize2.h"):
        src_time = os.path.getmtime(os.path.join(_HERE, "ref_stbir.c"))
        if not os.path.exists(_REF_STBIR) or os.path.getmtime(_REF_STBIR) < src_time:
            subprocess.run(["make", "-C", _HERE, "-B", "_ref/libstbir.so"], check=True, capture_output=True)
    return _REF_STBIR if os.path.exists(_REF_STBIR) else ""


def ref_stbir_resize(img: np.ndarray, out_w: int, out_h: int, srgb: bool) -> np.ndarray:
    """The reference's own mip filter (stbir_resize_uint8_srgb / _linear, Texture.cpp:193-208) on an RGBA8 image."""
    path = build_ref()
    if not path:
        raise FileNotFoundError("oracle/_ref/libstbir.so is not built")
    l = C.CDLL(path)
    l.ref_stbir_resize.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
    src = np.ascontiguousarray(img, np.uint8)
    h, w, ch = src.shape
    out = np.zeros((out_h, out_w, ch), np.uint8)
    if not l.ref_stbir_resize(_p(src), w, h, _p(out), out_w, out_h, ch, int(srgb)):
        raise RuntimeError("stbir_resize failed")
    return out
