"""glTF scenes through the scene host (libprt_host.so, include/prt_host.h): the Python face of WorldData's load path
and World's per-frame update for the RtReference path.

`load_gltf` runs the host (cgltf parse, material / model / animation / scene gather, MikkTSpace + meshoptimizer mesh
pipeline, `.prosper_mesh` cache write -- the reference's own third-party code, host/gltf_host.cpp), then does what
loadNextMesh / loadNextTexture do with the caches: reads every cache blob back into 64 MB geometry buffers
(mesh_cache.GeometryCache = uploadGeometryData) and turns every image into a mip chain in the reference's cache format
(texture_cache rules: full chain minus two levels, BC7 when every level is whole blocks). The result is the same
`SceneData` the generators emit, so World, the oracle and every test take it unchanged.

`GltfScene.update(time_s)` is updateAnimations + updateScene + updateBuffers + updateTlasInstances (World.cpp:349-536,
878-927): it rewrites transforms, draw instances, TLAS instance records, lights and the node camera in place.

Stated deviation: an image file that does not exist (the reference's own FlightHelmet copy ships 10 of its 15 images)
makes the reference throw (Texture.cpp:387-389); here the texture slot stays empty and every sampler that names it is
rewritten to texture 0 = "no texture" (materials.glsl:55-63), so the asset still loads.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional

import numpy as np

from . import bc7, mesh_cache, scene_data as sd, texture_cache

_HERE = os.path.dirname(os.path.abspath(__file__))
HOST_LIB_PATH = os.environ.get("PRT_HOST_LIB") or os.path.join(os.path.dirname(_HERE), "host", "_build", "libprt_host.so")
HOST_SYMBOLS = ("prth_last_error", "prth_load_gltf", "prth_destroy", "prth_get_counts", "prth_get_materials", "prth_get_samplers",
                "prth_get_mesh_infos", "prth_get_models", "prth_get_texture_source", "prth_decode_image", "prth_source_write_time",
                "prth_update", "prth_get_instances", "prth_get_lights", "prth_get_camera")


class PrthCounts(C.Structure):
    _fields_ = [(n, C.c_uint32) for n in ("meshes", "materials", "samplers", "textures", "models", "subModels", "nodes", "modelInstances",
                                          "drawInstances", "pointLights", "spotLights", "animatedNodes", "hasCamera")] + [("endTimeS", C.c_float)]


class PrthCamera(C.Structure):
    _fields_ = [("eye", C.c_float * 3), ("target", C.c_float * 3), ("up", C.c_float * 3), ("fov", C.c_float), ("zN", C.c_float), ("zF", C.c_float)]


_lib = None


def load_host_library():
    """Loads libprt_host.so. Raises when it has not been built (`make -C host`, needs /root/reference; no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(HOST_LIB_PATH):
            raise RuntimeError(f"{HOST_LIB_PATH} is missing: `make -C host` builds it against the reference's vendored cgltf / glm / "
                               "meshoptimizer / mikktspace / stb")
        l = C.CDLL(HOST_LIB_PATH)
        l.prth_last_error.restype = C.c_char_p
        l.prth_source_write_time.restype = C.c_int64
        l.prth_source_write_time.argtypes = [C.c_void_p]
        l.prth_load_gltf.argtypes = [C.c_char_p, C.c_char_p, C.c_int64, C.POINTER(C.c_void_p)]
        l.prth_destroy.argtypes = [C.c_void_p]
        l.prth_destroy.restype = None
        vp = C.c_void_p
        for name, args in (("prth_get_counts", [vp, C.POINTER(PrthCounts)]), ("prth_get_materials", [vp, vp]), ("prth_get_samplers", [vp, vp]),
                           ("prth_get_mesh_infos", [vp, vp]), ("prth_get_models", [vp, vp, vp]),
                           ("prth_get_texture_source", [vp, C.c_uint32, C.c_char_p, C.c_uint32, C.POINTER(C.c_uint32)]),
                           ("prth_decode_image", [C.c_char_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), vp, C.c_uint64]),
                           ("prth_update", [vp, C.c_float]), ("prth_get_instances", [vp, vp, vp, vp, vp]),
                           ("prth_get_lights", [vp, vp, vp, vp]), ("prth_get_camera", [vp, C.POINTER(PrthCamera)])):
            getattr(l, name).argtypes = args
            getattr(l, name).restype = C.c_int
        _lib = l
    return _lib


def _check(rc):
    if rc != 0:
        raise RuntimeError(load_host_library().prth_last_error().decode())


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def decode_image(path: str) -> np.ndarray:
    """stbi_load(path, 4 channels) (Texture.cpp:380-390) -> uint8 [h, w, 4]."""
    lib = load_host_library()
    w, h = C.c_uint32(), C.c_uint32()
    _check(lib.prth_decode_image(path.encode(), C.byref(w), C.byref(h), None, 0))
    out = np.zeros((h.value, w.value, 4), np.uint8)
    _check(lib.prth_decode_image(path.encode(), C.byref(w), C.byref(h), _p(out), out.nbytes))
    return out


def build_texture(rgba8: np.ndarray, linear: bool, max_size: Optional[int] = None) -> sd.Texture:
    """`compress` on the host (Texture.cpp:213-296): the level chain by the reference's rules, mips by the restatement of
    its stb filter (texture_cache.downsample_mitchell, pinned to the real stb within one code), blocks by the mode-6
    encoder. `max_size` first reduces level 0 with the same filter (fixtures only)."""
    level = np.ascontiguousarray(rgba8, np.uint8)
    while max_size is not None and max(level.shape[0], level.shape[1]) > max_size:
        level = texture_cache.downsample_mitchell(level, not linear)
    h, w, _ = level.shape
    levels = texture_cache.mip_level_count(w, h)
    is_bc7 = texture_cache.cache_format(w, h, levels) == texture_cache.dds.FORMAT_BC7_UNORM
    chunks: List[np.ndarray] = []
    for i in range(levels):
        chunks.append(bc7.encode_bc7_mode6(level) if is_bc7 else level.reshape(-1).copy())
        if i + 1 < levels:
            level = texture_cache.downsample_mitchell(level, not linear)
    return sd.Texture(sd.TEX_BC7 if is_bc7 else sd.TEX_RGBA8, w, h, levels, np.concatenate(chunks))


class GltfScene:
    """A loaded glTF: `.scene` is the SceneData, `.update(time_s)` advances animations and rewrites the per-frame arrays."""

    def __init__(self, gltf_path: str, width: int, height: int, cache_dir: Optional[str] = None, max_texture_size: Optional[int] = None,
                 source_write_time: int = 0, sky=None, load_textures: bool = True):
        self.lib = load_host_library()
        self.handle = C.c_void_p()
        self.width, self.height = width, height
        gltf_path = os.path.abspath(gltf_path)
        self.scene_dir = os.path.abspath(cache_dir) if cache_dir else os.path.dirname(gltf_path)
        _check(self.lib.prth_load_gltf(gltf_path.encode(), self.scene_dir.encode(), int(source_write_time), C.byref(self.handle)))
        c = PrthCounts()
        _check(self.lib.prth_get_counts(self.handle, C.byref(c)))
        self.counts = c
        materials = np.zeros(c.materials, sd.MATERIAL_DATA)
        _check(self.lib.prth_get_materials(self.handle, _p(materials)))
        samplers = np.zeros(c.samplers, sd.SAMPLER_DESC)
        _check(self.lib.prth_get_samplers(self.handle, _p(samplers)))
        infos = np.zeros(c.meshes, sd.MESH_INFO)
        _check(self.lib.prth_get_mesh_infos(self.handle, _p(infos)))
        offsets = np.zeros(c.models + 1, np.uint32)
        subs = np.zeros(c.subModels, sd.SUB_MODEL)
        _check(self.lib.prth_get_models(self.handle, _p(offsets), _p(subs)))
        # loadNextMesh: "Always read from the cache" + uploadGeometryData
        geo = mesh_cache.GeometryCache()
        for i in range(c.meshes):
            header, blob = mesh_cache.read_cache(mesh_cache.cache_path(self.scene_dir, i))
            if header is None:
                raise RuntimeError(f"mesh cache {i} missing after the host's cache pass")
            geo.upload(header, blob, int(infos[i]["materialIndex"]))
        buffers, metadatas, mesh_infos = geo.finish()
        mesh_infos["meshletCount"] = infos["meshletCount"]
        # loadNextTexture for every image
        textures: List[Optional[sd.Texture]] = [None]
        missing = set()
        self.texture_sources = [("", False)]
        gltf_dir = os.path.dirname(gltf_path)
        for t in range(1, c.textures):
            uri = C.create_string_buffer(1024)
            linear = C.c_uint32()
            _check(self.lib.prth_get_texture_source(self.handle, t, uri, 1024, C.byref(linear)))
            path = os.path.join(gltf_dir, uri.value.decode())
            self.texture_sources.append((path, bool(linear.value)))
            if load_textures and os.path.exists(path):
                textures.append(build_texture(decode_image(path), bool(linear.value), max_texture_size))
            else:
                textures.append(None)
                missing.add(t)
        for name in ("baseColorTextureSampler", "metallicRoughnessTextureSampler", "normalTextureSampler"):
            tex = materials[name] & 0xFFFFFF
            materials[name] = np.where(np.isin(tex, list(missing)), 0, materials[name]).astype(np.uint32)
        self.missing_textures = sorted(missing)
        face, sky_px = sky if sky is not None else (0, None)
        self.scene = sd.SceneData(
            geometry_buffers=buffers, metadatas=metadatas, mesh_infos=mesh_infos, materials=materials, samplers=samplers, textures=textures,
            model_offsets=offsets, sub_models=subs, tlas_instances=np.zeros(c.modelInstances, sd.TLAS_INSTANCE),
            draw_instances=np.zeros(c.drawInstances, sd.DRAW_INSTANCE), transforms=np.zeros(c.modelInstances, sd.MODEL_INSTANCE_TRANSFORMS),
            directional_light=np.zeros((), sd.DIRECTIONAL_LIGHT), point_lights=np.zeros(c.pointLights, sd.POINT_LIGHT),
            spot_lights=np.zeros(c.spotLights, sd.SPOT_LIGHT), camera=sd.make_camera((1, 0.5, 1), (0, 0, 0), width, height),
            sky_face_size=face, sky=sky_px, name=os.path.splitext(os.path.basename(gltf_path))[0])
        self.scales = np.zeros(c.modelInstances, np.float32)
        self._fetch()

    def _fetch(self):
        s = self.scene
        _check(self.lib.prth_get_instances(self.handle, _p(s.transforms), _p(self.scales), _p(s.draw_instances), _p(s.tlas_instances)))
        dl = np.zeros(1, sd.DIRECTIONAL_LIGHT)
        _check(self.lib.prth_get_lights(self.handle, _p(dl), _p(s.point_lights) if len(s.point_lights) else None,
                                        _p(s.spot_lights) if len(s.spot_lights) else None))
        s.directional_light = dl[0].copy()
        cam = PrthCamera()
        _check(self.lib.prth_get_camera(self.handle, C.byref(cam)))
        self.camera = cam
        if self.counts.hasCamera:
            s.camera = sd.make_camera(tuple(cam.eye), tuple(cam.target), self.width, self.height, fov_deg=float(np.degrees(cam.fov)),
                                      z_near=cam.zN, z_far=cam.zF, up=tuple(cam.up))

    def update(self, time_s: float):
        """World::updateAnimations(timeS) + updateScene + updateBuffers (+ updateTlasInstances)."""
        _check(self.lib.prth_update(self.handle, float(time_s)))
        self._fetch()

    def close(self):
        if self.handle:
            self.lib.prth_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---- cache directories ----------------------------------------------------------------------------------------------
# What a first run of the reference leaves next to a scene: prosper_cache/cache<i>.prosper_mesh (DeferredLoadingContext.cpp:
# 492-499) and one BC7 .dds per image (Texture.cpp:369-489). `save_cache` adds the texture files and the small arrays the glTF
# itself held (scene.npz) so that `load_cached` rebuilds the SceneData on a machine that has neither the glTF nor the host
# library's inputs (the GPU box).
_SMALL = ("materials", "samplers", "model_offsets", "sub_models", "tlas_instances", "draw_instances", "transforms", "directional_light",
          "point_lights", "spot_lights")


def save_cache(g: GltfScene, cache_dir: Optional[str] = None):
    from . import dds
    cache_dir = cache_dir or g.scene_dir
    s = g.scene
    for i, t in enumerate(s.textures):
        if t is None:
            continue
        fmt = dds.FORMAT_BC7_UNORM if t.format == sd.TEX_BC7 else dds.FORMAT_R8G8B8A8_UNORM
        dds.write_dds(dds.Dds(t.width, t.height, 1, fmt, t.mip_levels, t.data), os.path.join(cache_dir, "prosper_cache", f"texture{i}.dds"))
    np.savez_compressed(os.path.join(cache_dir, "scene.npz"), name=np.array(s.name), texture_count=np.array(len(s.textures)),
                        mesh_material=s.mesh_infos["materialIndex"], **{k: getattr(s, k) for k in _SMALL})


def load_cached(cache_dir: str, width: int, height: int, camera=None, sky=None) -> sd.SceneData:
    """SceneData from a cache directory written by `save_cache` (mesh caches read exactly as loadNextMesh reads them)."""
    from . import dds
    z = np.load(os.path.join(cache_dir, "scene.npz"))
    geo = mesh_cache.GeometryCache()
    for i, material in enumerate(z["mesh_material"]):
        header, blob = mesh_cache.read_cache(mesh_cache.cache_path(cache_dir, i))
        if header is None:
            raise RuntimeError(f"mesh cache {i} missing in {cache_dir}")
        geo.upload(header, blob, int(material))
    buffers, metadatas, mesh_infos = geo.finish()
    textures: List[Optional[sd.Texture]] = []
    for i in range(int(z["texture_count"])):
        path = os.path.join(cache_dir, "prosper_cache", f"texture{i}.dds")
        if not os.path.exists(path):
            textures.append(None)
            continue
        d = dds.read_dds(path)
        textures.append(sd.Texture(sd.TEX_BC7 if d.format == dds.FORMAT_BC7_UNORM else sd.TEX_RGBA8, d.width, d.height, d.mipLevelCount, d.data))
    face, sky_px = sky if sky is not None else (0, None)
    cam = camera if camera is not None else sd.make_camera((1, 0.5, 1), (0, 0, 0), width, height)  # Camera.hpp:24-26
    small = {k: z[k] for k in _SMALL}
    small["directional_light"] = small["directional_light"].reshape(())
    return sd.SceneData(geometry_buffers=buffers, metadatas=metadatas, mesh_infos=mesh_infos, textures=textures, camera=cam,
                        sky_face_size=face, sky=sky_px, name=str(z["name"]), **small)


FLIGHT_HELMET_DIR = os.path.join(os.path.dirname(_HERE), "tests", "golden", "flight_helmet")


def flight_helmet(width: int = 512, height: int = 512, cache_dir: Optional[str] = None) -> sd.SceneData:
    """BASELINE.json configs[0] on the one real asset the reference ships (res/glTF/FlightHelmet; DamagedHelmet is not in the
    image): the committed cache directory (tests/golden/make_flight_helmet.py; 256^2 textures) or a full-resolution one,
    a camera that frames the helmet, the procedural sky and the default directional light (lights.h:9,18-19)."""
    from . import scenes
    s = load_cached(cache_dir or FLIGHT_HELMET_DIR, width, height, camera=sd.make_camera((0.42, 0.12, 0.55), (0.0, -0.02, 0.0), width, height),
                    sky=scenes.procedural_sky(64))
    return s
