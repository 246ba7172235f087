"""Deterministic procedural stand-ins for the scenes BASELINE.json names.

Sponza, DamagedHelmet and Bistro are not in the reference checkout and there is no network
(SURVEY.md section 7 "Assets"), so each config gets a seeded generator sized to it. All output is
in the reference's data contract (scene_data.py); nothing here is specific to the CUDA path.

  test_scene()        few hundred triangles exercising every feature (CPU-test sized)
  helmet(seed)        C1: helmet-class object (~95k tris, one BLEND visor), sky + 1 directional light
  atrium(seed)        C2: Sponza-like atrium, ~260k tris, >=20 % alpha-MASK foliage, BC7 textures
  atrium(lights=1024) C3: the same + 1024 point/spot lights (recipe: WorldData.cpp:1324-1353)
  district(seed)      C4: ~3.5M tris, >=1k animated instances (TLAS rebuilt per frame)
  instanced(seed)     C5: 10k instances of ~10k-tri meshes (~100M instanced tris)
"""
from __future__ import annotations

import numpy as np

from . import bc7
from .scene_data import (ALPHA_BLEND, ALPHA_MASK, ALPHA_OPAQUE, DIRECTIONAL_LIGHT, DRAW_INSTANCE, MATERIAL_DATA,
                         MODEL_INSTANCE_TRANSFORMS, POINT_LIGHT, SAMPLER_DESC, SPOT_LIGHT, SUB_MODEL, TEX_BC7,
                         TEX_RGBA8, TLAS_INSTANCE, GeometryBuilder, SceneData, Texture, instance_transforms,
                         make_camera, make_material, trs)


# ---- mesh primitives -------------------------------------------------------------------------
def param_surface(fn, nu, nv, uv_scale=(1.0, 1.0), flip=False):
    """Tessellate p = fn(u, v), u,v in [0,1], into nu x nv quads. Returns pos, normal, uv, tangent, idx."""
    u = np.linspace(0, 1, nu + 1)
    v = np.linspace(0, 1, nv + 1)
    uu, vv = np.meshgrid(u, v, indexing="xy")
    p = fn(uu, vv)  # [nv+1, nu+1, 3]
    e = 1e-4
    du = (fn(np.clip(uu + e, 0, 1), vv) - fn(np.clip(uu - e, 0, 1), vv))
    dv = (fn(uu, np.clip(vv + e, 0, 1)) - fn(uu, np.clip(vv - e, 0, 1)))
    n = np.cross(du, dv)
    if flip:
        n = -n
    nl = np.linalg.norm(n, axis=-1, keepdims=True)
    n = np.where(nl > 1e-20, n / np.maximum(nl, 1e-20), np.array([0.0, 1.0, 0.0]))
    tl = np.linalg.norm(du, axis=-1, keepdims=True)
    t = np.where(tl > 1e-20, du / np.maximum(tl, 1e-20), np.array([1.0, 0.0, 0.0]))
    t = t - n * (t * n).sum(-1, keepdims=True)
    t = t / np.maximum(np.linalg.norm(t, axis=-1, keepdims=True), 1e-20)
    sign = -1.0 if flip else 1.0
    tang = np.concatenate([t, np.full(t.shape[:-1] + (1,), sign)], axis=-1)
    uv = np.stack([uu * uv_scale[0], vv * uv_scale[1]], axis=-1)
    i = np.arange(nv)[:, None] * (nu + 1) + np.arange(nu)[None, :]
    a, b, c, d = i, i + 1, i + nu + 1, i + nu + 2
    tris = np.stack([a, b, d, a, d, c], axis=-1).reshape(-1, 3)
    if flip:
        tris = tris[:, ::-1]
    return (p.reshape(-1, 3).astype(np.float32), n.reshape(-1, 3).astype(np.float32),
            uv.reshape(-1, 2).astype(np.float32), tang.reshape(-1, 4).astype(np.float32),
            tris.reshape(-1).astype(np.uint32))


def merge(parts):
    pos, nrm, uv, tan, idx = [], [], [], [], []
    base = 0
    for p, n, t, g, i in parts:
        pos.append(p); nrm.append(n); uv.append(t); tan.append(g); idx.append(i + base)
        base += p.shape[0]
    return (np.concatenate(pos), np.concatenate(nrm), np.concatenate(uv), np.concatenate(tan), np.concatenate(idx))


def plane(sx, sz, nu, nv, uv_scale=(1, 1), y=0.0):
    return param_surface(lambda u, v: np.stack([(u - 0.5) * sx, np.full_like(u, y), (0.5 - v) * sz], -1), nu, nv,
                         uv_scale)


def box(sx, sy, sz, n=1):
    """Axis-aligned box centred at the origin (outward normals), n x n quads per face."""
    h = np.array([sx, sy, sz]) * 0.5
    faces = []
    axes = [(0, 1, 2), (1, 2, 0), (2, 0, 1)]
    for a, b, c in axes:
        for s in (1.0, -1.0):
            def fn(u, v, a=a, b=b, c=c, s=s):
                p = np.zeros(u.shape + (3,))
                p[..., a] = s * h[a]
                p[..., b] = (u - 0.5) * 2 * h[b] * s
                p[..., c] = (v - 0.5) * 2 * h[c]
                return p
            faces.append(param_surface(fn, n, n))
    return merge(faces)


def cylinder(radius, height, nu, nv, radius_fn=None):
    def fn(u, v):
        r = radius * (radius_fn(v) if radius_fn is not None else 1.0)
        ang = -u * 2 * np.pi
        return np.stack([r * np.cos(ang), v * height, r * np.sin(ang)], -1)
    return param_surface(fn, nu, nv, (4, 4))


def bumpy_sphere(radius, nu, nv, amp, freq, phase=0.0, squash=(1, 1, 1)):
    def fn(u, v):
        th = np.clip(v, 1e-4, 1 - 1e-4) * np.pi
        ph = -u * 2 * np.pi
        d = np.stack([np.sin(th) * np.cos(ph), np.cos(th), np.sin(th) * np.sin(ph)], -1)
        r = radius * (1 + amp * np.sin(freq * d[..., 0] + phase) * np.sin(freq * d[..., 1] * 1.3 + phase)
                      * np.cos(freq * d[..., 2] * 0.7))
        return d * r[..., None] * np.asarray(squash)
    return param_surface(fn, nu, nv, (2, 1), flip=True)


def cards(rng, count, size, spread, height):
    """Randomly oriented two-triangle foliage cards (alpha-masked leaves)."""
    parts = []
    for _ in range(count):
        c = np.array([rng.uniform(-spread, spread), rng.uniform(0, height), rng.uniform(-spread, spread)])
        a = rng.uniform(0, 2 * np.pi)
        tilt = rng.uniform(-0.6, 0.6)
        ux = np.array([np.cos(a), 0, np.sin(a)]) * size
        uy = np.array([-np.sin(a) * np.sin(tilt), np.cos(tilt), np.cos(a) * np.sin(tilt)]) * size
        parts.append(param_surface(lambda u, v: c + (u - 0.5)[..., None] * ux + (v - 0.5)[..., None] * uy, 1, 1))
    return merge(parts)


# ---- textures ----------------------------------------------------------------------------------
def _noise(rng, size, octaves=5):
    img = np.zeros((size, size))
    for o in range(octaves):
        n = min(2 ** (o + 2), size)
        g = rng.random((n, n))
        g = np.kron(g, np.ones((size // n, size // n)))
        k = size // n
        # cheap smoothing: box blur by rolling
        if k > 1:
            g = (g + np.roll(g, k // 2, 0) + np.roll(g, k // 2, 1) + np.roll(np.roll(g, k // 2, 0), k // 2, 1)) / 4
        img += g / 2 ** o
    img -= img.min()
    return img / img.max()


def albedo_texture(rng, size, base, contrast=0.35):
    n = _noise(rng, size)
    rgb = np.clip(np.asarray(base)[None, None, :] * (1 - contrast + 2 * contrast * n[..., None]), 0, 1)
    a = np.ones((size, size, 1))
    return (np.concatenate([rgb, a], -1) * 255 + 0.5).astype(np.uint8)


def mr_texture(rng, size, rough=0.6, metal=0.0):
    n = _noise(rng, size, 4)
    img = np.zeros((size, size, 4))
    img[..., 1] = np.clip(rough * (0.6 + 0.8 * n), 0.05, 1)
    img[..., 2] = metal
    img[..., 3] = 1
    return (img * 255 + 0.5).astype(np.uint8)


def normal_texture(rng, size, strength=2.0):
    h = _noise(rng, size, 5)
    dx = (np.roll(h, -1, 1) - np.roll(h, 1, 1)) * strength * size / 64
    dy = (np.roll(h, -1, 0) - np.roll(h, 1, 0)) * strength * size / 64
    n = np.stack([-dx, -dy, np.ones_like(h)], -1)
    n /= np.linalg.norm(n, axis=-1, keepdims=True)
    img = np.concatenate([n * 0.5 + 0.5, np.ones((size, size, 1))], -1)
    return (img * 255 + 0.5).astype(np.uint8)


def leaf_texture(rng, size):
    """Green leaves with an alpha mask: a rosette of leaf shapes, ~45 % coverage."""
    y, x = np.mgrid[0:size, 0:size] / size - 0.5
    r = np.sqrt(x * x + y * y)
    ang = np.arctan2(y, x)
    shape = 0.46 * np.abs(np.cos(2.5 * ang)) ** 0.7 + 0.05
    alpha = np.clip((shape - r) * 24 + 0.5, 0, 1)
    n = _noise(rng, size, 4)
    rgb = np.stack([0.15 + 0.2 * n, 0.45 + 0.4 * n, 0.1 + 0.1 * n], -1)
    return (np.concatenate([rgb, alpha[..., None]], -1) * 255 + 0.5).astype(np.uint8)


def glass_texture(size):
    y, x = np.mgrid[0:size, 0:size] / size
    a = 0.25 + 0.5 * (0.5 + 0.5 * np.sin(x * 12) * np.sin(y * 9))
    rgb = np.stack([0.6 + 0 * x, 0.8 + 0 * x, 0.9 + 0 * x], -1)
    return (np.concatenate([rgb, a[..., None]], -1) * 255 + 0.5).astype(np.uint8)


def as_texture(img, fmt=TEX_BC7) -> Texture:
    h, w, _ = img.shape
    if fmt == TEX_BC7:
        data, levels = bc7.build_mip_chain_bc7(img)
        return Texture(TEX_BC7, w, h, levels, data)
    return Texture(TEX_RGBA8, w, h, 1, np.ascontiguousarray(img).reshape(-1))


def procedural_sky(face=64, horizon=(0.9, 0.95, 1.0), zenith=(0.25, 0.45, 0.9), ground=(0.25, 0.22, 0.2), scale=1.0):
    """Smooth gradient sky as an RGBA16F cubemap (stands in for res/env/storm.ktx, a missing blob)."""
    out = np.zeros((6, face, face, 4), np.float32)
    s = (np.arange(face) + 0.5) / face * 2 - 1
    sc, tc = np.meshgrid(s, s, indexing="xy")
    one = np.ones_like(sc)
    dirs = [(one, -tc, -sc), (-one, -tc, sc), (sc, one, tc), (sc, -one, -tc), (sc, -tc, one), (-sc, -tc, -one)]
    for f, (x, y, z) in enumerate(dirs):
        d = np.stack([x, y, z], -1)
        d /= np.linalg.norm(d, axis=-1, keepdims=True)
        up = d[..., 1:2]
        col = np.where(up >= 0, np.asarray(horizon) + (np.asarray(zenith) - np.asarray(horizon)) * np.sqrt(np.abs(up)),
                       np.asarray(horizon) + (np.asarray(ground) - np.asarray(horizon)) * np.minimum(-up * 4, 1))
        out[f, ..., :3] = col * scale
        out[f, ..., 3] = 1
    return face, out.astype(np.float16)


# ---- scene assembly ------------------------------------------------------------------------------
class SceneBuilder:
    def __init__(self, name):
        self.name = name
        self.geo = GeometryBuilder(0)
        self.materials = [make_material()]  # material 0 = default (WorldData.cpp:756-828)
        self.textures = [None]
        self.samplers = [(1, 1, 0, 0)]  # linear / repeat default (WorldData.cpp:686-697)
        self.models = []  # list of [(mesh, material)]
        self.raw_models = []
        self.instances = []  # (model, 4x4, active)

    def add_texture(self, tex: Texture) -> int:
        self.textures.append(tex)
        return len(self.textures) - 1

    def add_material(self, m) -> int:
        self.materials.append(m)
        return len(self.materials) - 1

    def add_model(self, meshes) -> int:
        """meshes: list of (mesh tuple from param_surface/merge, material index)."""
        subs = []
        for (p, n, uv, t, idx), mat in meshes:
            subs.append((self.geo.add_mesh(p, n, uv, idx, mat, tangents=t), mat))
        self.models.append(subs)
        self.raw_models.append(meshes)
        return len(self.models) - 1

    def baked(self) -> "SceneBuilder":
        """Flatten every active instance into world space and merge per material: one model with one
        sub-model per material and a single identity instance, which is how a single-node glTF such as
        Sponza reaches the reference (one BLAS, fp16 world-space positions)."""
        out = SceneBuilder(self.name)
        out.materials, out.textures, out.samplers = self.materials, self.textures, self.samplers
        per_mat = {}
        for model, mat4, active in self.instances:
            if not active:
                continue
            lin = mat4[:3, :3]
            nmat = np.linalg.inv(lin).T
            for (p, n, uv, t, idx), mat in self.raw_models[model]:
                pw = p @ lin.T + mat4[:3, 3]
                nw = n @ nmat.T
                nw /= np.maximum(np.linalg.norm(nw, axis=-1, keepdims=True), 1e-20)
                tw = t[:, :3] @ lin.T
                tw /= np.maximum(np.linalg.norm(tw, axis=-1, keepdims=True), 1e-20)
                flip = np.linalg.det(lin) < 0
                idx2 = idx.reshape(-1, 3)[:, ::-1].reshape(-1) if flip else idx
                per_mat.setdefault(mat, []).append(
                    (pw.astype(np.float32), nw.astype(np.float32), uv, np.concatenate([tw, t[:, 3:4]], 1).astype(np.float32), idx2))
        model = out.add_model([(merge(parts), mat) for mat, parts in sorted(per_mat.items())])
        out.add_instance(model)
        return out

    def add_instance(self, model, matrix=None, active=True):
        self.instances.append((model, np.eye(4) if matrix is None else matrix, active))

    def finish(self, camera, directional=None, point_lights=None, spot_lights=None, sky=None) -> SceneData:
        buf, metadatas, infos = self.geo.finish()
        offsets = np.zeros(len(self.models) + 1, np.uint32)
        subs = []
        for m, sm in enumerate(self.models):
            offsets[m + 1] = offsets[m] + len(sm)
            subs += sm
        sub_models = np.array(subs, SUB_MODEL) if subs else np.zeros(0, SUB_MODEL)
        n = len(self.instances)
        tlas = np.zeros(n, TLAS_INSTANCE)
        transforms = np.zeros(n, MODEL_INSTANCE_TRANSFORMS)
        draws = []
        rti = 0
        for i, (model, mat, active) in enumerate(self.instances):
            transforms[i] = instance_transforms(mat)
            tlas[i]["transform"] = transforms[i]["modelToWorld"]
            tlas[i]["customIndexAndMask"] = (rti & 0xFFFFFF) | (0xFF << 24)
            tlas[i]["blasReference"] = model + 1 if active else 0
            for mesh, material in self.models[model]:
                draws.append((i, mesh, material))
            rti += len(self.models[model])
        dl = np.zeros((), DIRECTIONAL_LIGHT)
        dl["irradiance"] = (2, 2, 2, 0) if directional is None else directional[0]
        dl["direction"] = (-1, -1, -1, 0) if directional is None else directional[1]  # lights.h:9,18-19
        face, sky_px = sky if sky is not None else (0, None)
        return SceneData(
            geometry_buffers=[buf], metadatas=metadatas, mesh_infos=infos,
            materials=np.array(self.materials, MATERIAL_DATA), samplers=np.array(self.samplers, SAMPLER_DESC),
            textures=self.textures, model_offsets=offsets, sub_models=sub_models, tlas_instances=tlas,
            draw_instances=np.array(draws, DRAW_INSTANCE), transforms=transforms, directional_light=dl,
            point_lights=np.zeros(0, POINT_LIGHT) if point_lights is None else point_lights,
            spot_lights=np.zeros(0, SPOT_LIGHT) if spot_lights is None else spot_lights,
            camera=camera, sky_face_size=face, sky=sky_px, name=self.name)
