"""Host-side scene containers in the reference's data contract (numpy views of include/prt_types.h).

Everything here produces exactly the bytes prosper's host code hands to its shaders:
mesh blobs as packMeshData writes them (src/scene/DeferredLoadingContext.cpp:442-490),
GeometryMetadata / MaterialData / DrawInstance / ModelInstanceTransforms / light structs
(res/shader/shared/shader_structs/scene/*.h), CameraUniforms as Camera::updateBuffer fills it
(src/scene/Camera.cpp:105-204, 366-395) and the TLAS instance array of
World::Impl::updateTlasInstances (src/scene/World.cpp:878-928).
"""
from __future__ import annotations

import dataclasses
from typing import List, Optional

import numpy as np

GEOMETRY_METADATA = np.dtype([(n, "<u4") for n in (
    "bufferIndex", "indicesOffset", "positionsOffset", "normalsOffset", "tangentsOffset",
    "texCoord0sOffset", "meshletsOffset", "meshletBoundsOffset", "meshletVerticesOffset",
    "meshletTrianglesByteOffset", "usesShortIndices")])
MESH_INFO = np.dtype([(n, "<u4") for n in ("vertexCount", "indexCount", "meshletCount", "materialIndex")])
MATERIAL_DATA = np.dtype([
    ("baseColorFactor", "<f4", 4), ("metallicFactor", "<f4"), ("roughnessFactor", "<f4"),
    ("alphaCutoff", "<f4"), ("alphaMode", "<u4"), ("baseColorTextureSampler", "<u4"),
    ("metallicRoughnessTextureSampler", "<u4"), ("normalTextureSampler", "<u4"), ("pad", "<u4")])
DRAW_INSTANCE = np.dtype([(n, "<u4") for n in ("modelInstanceIndex", "meshIndex", "materialIndex")])
MODEL_INSTANCE_TRANSFORMS = np.dtype([("modelToWorld", "<f4", (3, 4)), ("normalToWorld", "<f4", (3, 4))])
TLAS_INSTANCE = np.dtype([("transform", "<f4", (3, 4)), ("customIndexAndMask", "<u4"),
                          ("sbtOffsetAndFlags", "<u4"), ("blasReference", "<u8")])
SUB_MODEL = np.dtype([("meshIndex", "<u4"), ("materialIndex", "<u4")])
SAMPLER_DESC = np.dtype([(n, "<u4") for n in ("magFilter", "minFilter", "addressU", "addressV")])
DIRECTIONAL_LIGHT = np.dtype([("irradiance", "<f4", 4), ("direction", "<f4", 4)])
POINT_LIGHT = np.dtype([("radianceAndRadius", "<f4", 4), ("position", "<f4", 4)])
SPOT_LIGHT = np.dtype([("radianceAndAngleScale", "<f4", 4), ("positionAndAngleOffset", "<f4", 4),
                       ("direction", "<f4", 4)])
CAMERA_UNIFORMS = np.dtype([
    ("worldToCamera", "<f4", (4, 4)), ("cameraToWorld", "<f4", (4, 4)), ("cameraToClip", "<f4", (4, 4)),
    ("clipToWorld", "<f4", (4, 4)), ("previousWorldToCamera", "<f4", (4, 4)),
    ("previousCameraToClip", "<f4", (4, 4)), ("eye", "<f4", 4), ("planes", "<f4", (6, 4)),
    ("resolution", "<u4", 2), ("currentJitter", "<f4", 2), ("previousJitter", "<f4", 2),
    ("near", "<f4"), ("far", "<f4"), ("maxViewScale", "<f4")])
REFERENCE_PC = np.dtype([("drawType", "<u4"), ("flags", "<u4"), ("frameIndex", "<u4"),
                         ("apertureDiameter", "<f4"), ("focusDistance", "<f4"), ("focalLength", "<f4"),
                         ("rouletteStartBounce", "<u4"), ("maxBounces", "<u4")])
HIT = np.dtype([("drawInstanceIndex", "<u4"), ("primitiveID", "<u4"), ("bary", "<f4", 2)])

assert GEOMETRY_METADATA.itemsize == 44 and MATERIAL_DATA.itemsize == 48 and DRAW_INSTANCE.itemsize == 12
assert MODEL_INSTANCE_TRANSFORMS.itemsize == 96 and TLAS_INSTANCE.itemsize == 64
assert POINT_LIGHT.itemsize == 32 and SPOT_LIGHT.itemsize == 48 and CAMERA_UNIFORMS.itemsize == 532
assert REFERENCE_PC.itemsize == 32 and HIT.itemsize == 16

ALPHA_OPAQUE, ALPHA_MASK, ALPHA_BLEND = 0, 1, 2
FLAG_SKIP_HISTORY, FLAG_ACCUMULATE, FLAG_IBL, FLAG_DOF, FLAG_CLAMP_INDIRECT = 1, 2, 4, 8, 16
TEX_RGBA8, TEX_BC7 = 0, 1
UNUSED = 0xFFFFFFFF
MISS = 0xFFFFFFFF


@dataclasses.dataclass
class Texture:
    format: int
    width: int
    height: int
    mip_levels: int
    data: np.ndarray  # uint8, level 0 first


@dataclasses.dataclass
class SceneData:
    geometry_buffers: List[np.ndarray]
    metadatas: np.ndarray
    mesh_infos: np.ndarray
    materials: np.ndarray
    samplers: np.ndarray
    textures: List[Optional[Texture]]  # index 0 is the reserved "no texture" slot
    model_offsets: np.ndarray
    sub_models: np.ndarray
    tlas_instances: np.ndarray
    draw_instances: np.ndarray
    transforms: np.ndarray
    directional_light: np.ndarray
    point_lights: np.ndarray
    spot_lights: np.ndarray
    camera: np.ndarray
    sky_face_size: int = 0
    sky: Optional[np.ndarray] = None  # float16 [6, n, n, 4]
    name: str = "scene"

    @property
    def triangle_count(self) -> int:
        """Unique (BLAS) triangles."""
        return int(self.mesh_infos["indexCount"].sum() // 3)

    @property
    def instanced_triangle_count(self) -> int:
        per_model = np.zeros(len(self.model_offsets) - 1, dtype=np.int64)
        for m in range(len(per_model)):
            sm = self.sub_models[self.model_offsets[m]:self.model_offsets[m + 1]]
            per_model[m] = int(self.mesh_infos["indexCount"][sm["meshIndex"]].sum() // 3)
        refs = self.tlas_instances["blasReference"]
        return int(per_model[(refs[refs > 0] - 1).astype(np.int64)].sum())


# ---- mesh packing (DeferredLoadingContext.cpp:442-490; glm packHalf4x16 / packSnorm3x10_1x2) ----
def pack_snorm3x10_1x2(v: np.ndarray) -> np.ndarray:
    v = np.asarray(v, dtype=np.float32)
    c = np.clip(v, -1.0, 1.0) * np.array([511.0, 511.0, 511.0, 1.0], dtype=np.float32)
    i = np.rint(c).astype(np.int32)
    return ((i[:, 0] & 0x3FF) | ((i[:, 1] & 0x3FF) << 10) | ((i[:, 2] & 0x3FF) << 20)
            | ((i[:, 3] & 0x3) << 30)).astype(np.uint32)


class GeometryBuilder:
    """Packs meshes tightly into one u32 geometry buffer the way uploadGeometryData does
    (DeferredLoadingContext.cpp:1192-1269): indices, positions, normals, tangents, uvs."""

    def __init__(self, buffer_index: int = 0):
        self.buffer_index = buffer_index
        self.words: List[np.ndarray] = []
        self.offset = 0  # in u32
        self.metadatas = []
        self.infos = []

    def _push(self, a: np.ndarray) -> int:
        start = self.offset
        self.words.append(a.astype(np.uint32, copy=False))
        self.offset += a.size
        return start

    def add_mesh(self, positions, normals, uvs, indices, material_index, tangents=None) -> int:
        positions = np.asarray(positions, dtype=np.float32)
        normals = np.asarray(normals, dtype=np.float32)
        uvs = np.asarray(uvs, dtype=np.float32)
        indices = np.asarray(indices, dtype=np.uint32).reshape(-1)
        nv = positions.shape[0]
        assert indices.size % 3 == 0 and indices.max() < nv
        short = nv <= 0xFFFF  # u16 indices when they fit
        if short:
            idx16 = indices.astype(np.uint16)
            if idx16.size % 2:
                idx16 = np.concatenate([idx16, np.zeros(1, np.uint16)])
            start = self._push(idx16.view(np.uint32))
            indices_offset = start * 2  # in u16 units
        else:
            indices_offset = self._push(indices)
        pos4 = np.concatenate([positions, np.ones((nv, 1), np.float32)], axis=1).astype(np.float16)
        positions_offset = self._push(pos4.reshape(-1).view(np.uint32))
        normals_offset = self._push(pack_snorm3x10_1x2(np.concatenate([normals, np.zeros((nv, 1), np.float32)], 1)))
        if tangents is not None:
            tangents_offset = self._push(pack_snorm3x10_1x2(np.asarray(tangents, np.float32)))
        else:
            tangents_offset = UNUSED
        uv_offset = self._push(uvs.astype(np.float16).reshape(-1).view(np.uint32))
        md = np.zeros((), GEOMETRY_METADATA)
        md["bufferIndex"] = self.buffer_index
        md["indicesOffset"] = indices_offset
        md["positionsOffset"] = positions_offset
        md["normalsOffset"] = normals_offset
        md["tangentsOffset"] = tangents_offset
        md["texCoord0sOffset"] = uv_offset
        for f in ("meshletsOffset", "meshletBoundsOffset", "meshletVerticesOffset", "meshletTrianglesByteOffset"):
            md[f] = UNUSED
        md["usesShortIndices"] = 1 if short else 0
        info = np.zeros((), MESH_INFO)
        info["vertexCount"] = nv
        info["indexCount"] = indices.size
        info["materialIndex"] = material_index
        self.metadatas.append(md)
        self.infos.append(info)
        return len(self.metadatas) - 1

    def finish(self):
        buf = np.concatenate(self.words) if self.words else np.zeros(0, np.uint32)
        return buf, np.array(self.metadatas, GEOMETRY_METADATA), np.array(self.infos, MESH_INFO)


def make_material(base_color=(1, 1, 1, 1), metallic=0.0, roughness=1.0, alpha_mode=ALPHA_OPAQUE, alpha_cutoff=0.5,
                  base_tex=0, mr_tex=0, normal_tex=0, sampler=0) -> np.ndarray:
    m = np.zeros((), MATERIAL_DATA)
    m["baseColorFactor"] = base_color
    m["metallicFactor"] = metallic
    m["roughnessFactor"] = roughness
    m["alphaCutoff"] = alpha_cutoff
    m["alphaMode"] = alpha_mode
    m["baseColorTextureSampler"] = (sampler << 24) | base_tex
    m["metallicRoughnessTextureSampler"] = (sampler << 24) | mr_tex
    m["normalTextureSampler"] = (sampler << 24) | normal_tex
    return m


# ---- transforms (World.cpp:359-466: modelToWorld = transpose(M), normalToWorld = mat3x4(inverse(M))) ----
def trs(translation=(0, 0, 0), rotation_axis=(0, 1, 0), angle=0.0, scale=(1, 1, 1)) -> np.ndarray:
    ax = np.asarray(rotation_axis, np.float64)
    ax = ax / np.linalg.norm(ax)
    c, s = np.cos(angle), np.sin(angle)
    x, y, z = ax
    r = np.array([[c + x * x * (1 - c), x * y * (1 - c) - z * s, x * z * (1 - c) + y * s],
                  [y * x * (1 - c) + z * s, c + y * y * (1 - c), y * z * (1 - c) - x * s],
                  [z * x * (1 - c) - y * s, z * y * (1 - c) + x * s, c + z * z * (1 - c)]])
    m = np.eye(4)
    m[:3, :3] = r @ np.diag(np.broadcast_to(np.asarray(scale, np.float64), 3))
    m[:3, 3] = translation
    return m


def instance_transforms(m: np.ndarray) -> np.ndarray:
    """ModelInstanceTransforms from an affine 4x4 M (math convention, column vectors).

    glm: modelToWorld = transpose(M) stored as mat3x4 => column c (4 floats) = row c of M.
    normalToWorld = mat3x4(inverse(M)) => column c = (inverse(M) column c).xyzw, i.e.
    [c][r] = inv[r][c], so `n * mat3(normalToWorld)` = transpose(inverse(M)) n."""
    t = np.zeros((), MODEL_INSTANCE_TRANSFORMS)
    m32 = m.astype(np.float32)
    t["modelToWorld"] = m32[:3, :]
    inv = np.linalg.inv(m.astype(np.float64)).astype(np.float32)
    t["normalToWorld"] = inv.T[:3, :]
    return t


# ---- camera (Camera.cpp:105-204, 366-395) ------------------------------------------------------
def make_camera(eye, target, width, height, fov_deg=59.0, z_near=0.1, z_far=100.0, up=(0, 1, 0)) -> np.ndarray:
    eye = np.asarray(eye, np.float32)
    target = np.asarray(target, np.float32)
    up = np.asarray(up, np.float32)

    def nrm(v):
        return (v / np.sqrt(np.dot(v, v))).astype(np.float32)

    fwd = nrm(target - eye)
    z = -fwd
    right = nrm(np.cross(up, z))
    new_up = nrm(np.cross(z, right))
    # glm mat4 constructor is column-major: w2c[col][row]
    w2c = np.array([[right[0], new_up[0], z[0], 0], [right[1], new_up[1], z[1], 0], [right[2], new_up[2], z[2], 0],
                    [-np.dot(right, eye), -np.dot(new_up, eye), -np.dot(z, eye), 1]], np.float32)
    c2w = np.linalg.inv(w2c.T.astype(np.float64)).T.astype(np.float32)
    ar = np.float32(width) / np.float32(height)
    zn, zf = np.float32(z_far), np.float32(z_near)  # swapped: reverse-Z
    tf = np.float32(1.0) / np.float32(np.tan(np.float32(np.radians(fov_deg)) * np.float32(0.5)))
    flip = np.array([[1, 0, 0, 0], [0, -1, 0, 0], [0, 0, 0.5, 0], [0, 0, 0.5, 1]], np.float32)  # columns
    proj = np.array([[tf / ar, 0, 0, 0], [0, tf, 0, 0], [0, 0, (zf + zn) / (zn - zf), -1],
                     [0, 0, 2 * zf * zn / (zn - zf), 0]], np.float32)  # columns
    c2c_math = (flip.T.astype(np.float64) @ proj.T.astype(np.float64))
    c2c = c2c_math.T.astype(np.float32)
    clip_to_world = np.linalg.inv(c2c_math @ w2c.T.astype(np.float64)).T.astype(np.float32)
    cam = np.zeros((), CAMERA_UNIFORMS)
    cam["worldToCamera"] = w2c
    cam["cameraToWorld"] = c2w
    cam["cameraToClip"] = c2c
    cam["clipToWorld"] = clip_to_world
    cam["previousWorldToCamera"] = w2c
    cam["previousCameraToClip"] = c2c
    cam["eye"] = (eye[0], eye[1], eye[2], 1.0)
    cam["resolution"] = (width, height)
    cam["near"] = z_near
    cam["far"] = z_far
    cam["maxViewScale"] = 1.0
    return cam


def focal_length(cam: np.ndarray, width: int, height: int, sensor_width: float = 0.036) -> float:
    """Camera::perspective: focalLength = sensorHeight * tf * 0.5 (Camera.cpp:150-152)."""
    tf = -float(cam["cameraToClip"][1][1])
    return sensor_width / (width / height) * tf * 0.5


def reference_pc(frame_index=1, flags=FLAG_ACCUMULATE | FLAG_CLAMP_INDIRECT | FLAG_IBL | FLAG_SKIP_HISTORY,
                 max_bounces=4, roulette_start=3, draw_type=0, aperture=0.00001, focus_distance=1.0,
                 focal_len=0.035) -> np.ndarray:
    """ReferencePC with the reference's defaults (RtRef.hpp:99-104, Camera.hpp:42-46)."""
    pc = np.zeros((), REFERENCE_PC)
    pc["drawType"] = draw_type
    pc["flags"] = flags
    pc["frameIndex"] = frame_index
    pc["apertureDiameter"] = aperture
    pc["focusDistance"] = focus_distance
    pc["focalLength"] = focal_len
    pc["rouletteStartBounce"] = roulette_start
    pc["maxBounces"] = max_bounces
    return pc
