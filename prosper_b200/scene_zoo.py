"""The named scenes (see scenes.py for the building blocks and the config mapping)."""
from __future__ import annotations

import numpy as np

from .scene_data import (ALPHA_BLEND, ALPHA_MASK, POINT_LIGHT, SPOT_LIGHT, TEX_BC7, TEX_RGBA8, make_camera,
                         make_material, trs)
from .scenes import (SceneBuilder, albedo_texture, as_texture, box, bumpy_sphere, cards, cylinder, glass_texture,
                     leaf_texture, merge, mr_texture, normal_texture, param_surface, plane, procedural_sky)


def _pcg(v):
    """The repo's own pcg hash (res/shader/common/random.glsl:7-12), vectorised, for light placement."""
    v = np.asarray(v, dtype=np.uint64)
    state = (v * 747796405 + 2891336453) & 0xFFFFFFFF
    word = (((state >> ((state >> 28) + 4)) ^ state) * 277803737) & 0xFFFFFFFF
    return ((word >> 22) ^ word) & 0xFFFFFFFF


def make_lights(count, lo, hi, seed=3):
    """count point+spot lights (half each) uniform in the AABB [lo, hi]; radiance U(1,5)/(4 pi) per
    channel, radius = sqrt(luminance / 0.01) (WorldData.cpp:1324-1353 recipe, glTF conversion 1478-1527)."""
    lo = np.asarray(lo, np.float32)
    hi = np.asarray(hi, np.float32)
    n_point = count // 2
    n_spot = count - n_point
    idx = np.arange(count * 8, dtype=np.uint64) + np.uint64(seed * 0x9E3779B1 & 0xFFFFFFFF)
    u = (_pcg(idx).astype(np.float64) / 4294967296.0).reshape(count, 8)
    pos = lo + (hi - lo) * u[:, 0:3]
    radiance = (1 + 4 * u[:, 3:6]) / (4 * np.pi)
    lum = radiance @ np.array([0.2126, 0.7152, 0.0722])
    radius = np.sqrt(lum / 0.01)
    points = np.zeros(n_point, POINT_LIGHT)
    points["radianceAndRadius"][:, :3] = radiance[:n_point]
    points["radianceAndRadius"][:, 3] = radius[:n_point]
    points["position"][:, :3] = pos[:n_point]
    points["position"][:, 3] = 1
    spots = np.zeros(n_spot, SPOT_LIGHT)
    d = u[n_point:, 6:8]
    ang = d[:, 0] * 2 * np.pi
    direction = np.stack([0.5 * np.cos(ang), -np.ones(n_spot), 0.5 * np.sin(ang)], -1)
    direction /= np.linalg.norm(direction, axis=-1, keepdims=True)
    inner, outer = 0.3, 0.6 + 0.3 * d[:, 1]
    scale = 1.0 / np.maximum(0.001, np.cos(inner) - np.cos(outer))  # glTF KHR_lights_punctual
    offset = -np.cos(outer) * scale
    spots["radianceAndAngleScale"][:, :3] = radiance[n_point:] * 4
    spots["radianceAndAngleScale"][:, 3] = scale
    spots["positionAndAngleOffset"][:, :3] = pos[n_point:]
    spots["positionAndAngleOffset"][:, 3] = offset
    spots["direction"][:, :3] = direction
    return points, spots


def test_scene(width=96, height=64, fmt=TEX_RGBA8, seed=7):
    """Small scene touching every feature: textured opaque, MASK foliage, BLEND glass, normal maps,
    rotated / non-uniformly scaled / inactive instances, a multi-geometry model, u16 indices,
    point + spot lights and a sky."""
    rng = np.random.default_rng(seed)
    sb = SceneBuilder("test_scene")
    t_alb = sb.add_texture(as_texture(albedo_texture(rng, 32, (0.8, 0.5, 0.3)), fmt))
    t_mr = sb.add_texture(as_texture(mr_texture(rng, 32, 0.5, 0.0), fmt))
    t_nrm = sb.add_texture(as_texture(normal_texture(rng, 32), fmt))
    t_leaf = sb.add_texture(as_texture(leaf_texture(rng, 64), fmt))
    t_glass = sb.add_texture(as_texture(glass_texture(32), fmt))
    sb.samplers.append((0, 0, 2, 2))  # nearest / clamp
    sb.samplers.append((1, 1, 1, 1))  # linear / mirrored
    m_floor = sb.add_material(make_material(base_tex=t_alb, mr_tex=t_mr, normal_tex=t_nrm, roughness=1.0))
    m_metal = sb.add_material(make_material((0.9, 0.8, 0.6, 1), metallic=1.0, roughness=0.3))
    m_leaf = sb.add_material(make_material(base_tex=t_leaf, alpha_mode=ALPHA_MASK, alpha_cutoff=0.5, sampler=2))
    m_glass = sb.add_material(make_material((1, 1, 1, 0.8), base_tex=t_glass, alpha_mode=ALPHA_BLEND, roughness=0.2))
    m_plain = sb.add_material(make_material((0.3, 0.6, 0.9, 1), roughness=0.7, base_tex=t_alb, sampler=1))
    floor = sb.add_model([(plane(6, 6, 4, 4, (3, 3)), m_floor)])
    blob = sb.add_model([(bumpy_sphere(0.5, 16, 10, 0.15, 5.0), m_metal), (box(0.5, 0.2, 0.5), m_plain)])
    bush = sb.add_model([(cards(rng, 24, 0.5, 0.5, 1.0), m_leaf)])
    pane = sb.add_model([(param_surface(lambda u, v: np.stack([u - 0.5, v, 0 * u], -1), 2, 2), m_glass)])
    sb.add_instance(floor)
    sb.add_instance(blob, trs((0, 0.6, 0), (0, 1, 0), 0.4))
    sb.add_instance(blob, trs((1.4, 0.5, -0.6), (1, 1, 0), 1.1, (0.6, 1.3, 0.8)))
    sb.add_instance(bush, trs((-1.2, 0, 0.3), (0, 1, 0), 0.7))
    sb.add_instance(bush, trs((-1.2, 5, 0.3)), active=False)
    sb.add_instance(pane, trs((0.3, 0.0, 1.2), (0, 1, 0), -0.3, (1.5, 1.2, 1)))
    cam = make_camera((1.5, 1.2, 3.0), (0, 0.4, 0), width, height)
    points, spots = make_lights(6, (-2, 0.5, -2), (2, 2.5, 2), seed)
    return sb.finish(cam, point_lights=points, spot_lights=spots, sky=procedural_sky(16))


def helmet(width=512, height=512, seed=11, fmt=TEX_BC7, tex_size=512, detail=1.0):
    """C1 stand-in: helmet-class object, ~95k triangles in 5 meshes / 5 materials, one BLEND
    (cf. FlightHelmet: 94,722 tris, 4 opaque + 1 BLEND "LensesMat")."""
    rng = np.random.default_rng(seed)
    sb = SceneBuilder("helmet")
    d = lambda n: max(4, int(n * detail))
    mats = []
    for base, rough, metal in (((0.55, 0.35, 0.2), 0.7, 0.0), ((0.7, 0.7, 0.72), 0.35, 1.0),
                               ((0.2, 0.2, 0.22), 0.5, 0.0), ((0.8, 0.75, 0.6), 0.9, 0.0)):
        ta = sb.add_texture(as_texture(albedo_texture(rng, tex_size, base), fmt))
        tm = sb.add_texture(as_texture(mr_texture(rng, tex_size, rough, metal), fmt))
        tn = sb.add_texture(as_texture(normal_texture(rng, tex_size), fmt))
        mats.append(sb.add_material(make_material(base_tex=ta, mr_tex=tm, normal_tex=tn, metallic=1.0 if metal else 0.0)))
    tg = sb.add_texture(as_texture(glass_texture(64), fmt))
    m_lens = sb.add_material(make_material((1, 1, 1, 0.6), base_tex=tg, alpha_mode=ALPHA_BLEND, roughness=0.1))
    shell = bumpy_sphere(0.32, d(160), d(120), 0.05, 9.0, 0.3, (1, 1.1, 1.15))
    strap = merge([cylinder(0.05, 0.5, d(64), d(64), lambda v: 1 + 0.2 * np.sin(v * 30))])
    mask_ = bumpy_sphere(0.2, d(96), d(80), 0.12, 14.0, 1.0, (1, 0.7, 0.5))
    base = merge([cylinder(0.25, 0.08, d(96), d(8)), plane(0.6, 0.6, d(48), d(48), (2, 2), 0.08)])
    lens = bumpy_sphere(0.09, d(48), d(32), 0.0, 1.0)
    model = sb.add_model([(shell, mats[0]), (strap, mats[1]), (mask_, mats[2]), (base, mats[3]), (lens, m_lens)])
    sb.add_instance(model, trs((0, 0.05, 0)))
    lens_only = sb.add_model([(lens, m_lens)])
    sb.add_instance(lens_only, trs((0.12, 0.1, 0.34)))
    sb.add_instance(lens_only, trs((-0.12, 0.1, 0.34)))
    cam = make_camera((1.0, 0.5, 1.0), (0, 0, 0), width, height)  # Camera.hpp:24-26 default
    return sb.finish(cam, sky=procedural_sky(64))


def atrium(width=1920, height=1080, seed=1, lights=0, fmt=TEX_BC7, tex_size=1024, detail=1.0, bake=True):
    """C2/C3 stand-in: Sponza-like atrium (two-storey colonnade around a nave, arches, drapes, planters
    with alpha-masked foliage). ~260k triangles at detail=1, >= 20 % of them alpha-MASK."""
    rng = np.random.default_rng(seed)
    sb = SceneBuilder("atrium")
    d = lambda n: max(2, int(round(n * 0.745 * np.sqrt(detail))))

    def textured(base, rough, metal=0.0, nrm=True):
        ta = sb.add_texture(as_texture(albedo_texture(rng, tex_size, base), fmt))
        tm = sb.add_texture(as_texture(mr_texture(rng, tex_size, rough, metal), fmt))
        tn = sb.add_texture(as_texture(normal_texture(rng, tex_size), fmt)) if nrm else 0
        return sb.add_material(make_material(base_tex=ta, mr_tex=tm, normal_tex=tn, metallic=1.0 if metal else 0.0))

    m_floor = textured((0.55, 0.5, 0.45), 0.55)
    m_wall = textured((0.75, 0.68, 0.58), 0.85)
    m_column = textured((0.8, 0.78, 0.72), 0.6)
    m_drape_r = textured((0.7, 0.12, 0.1), 0.9)
    m_drape_g = textured((0.1, 0.45, 0.2), 0.9)
    m_drape_b = textured((0.12, 0.2, 0.6), 0.9)
    m_bronze = textured((0.8, 0.55, 0.3), 0.35, 1.0)
    t_leaf = sb.add_texture(as_texture(leaf_texture(rng, tex_size), fmt))
    m_leaf = sb.add_material(make_material(base_tex=t_leaf, alpha_mode=ALPHA_MASK, alpha_cutoff=0.5, roughness=0.8))

    L, W, H = 28.0, 12.0, 11.0  # nave length (x), width (z), height (y)
    shell = sb.add_model([
        (plane(L, W, d(56), d(24), (14, 6)), m_floor),
        (merge([param_surface(lambda u, v, s=s: np.stack([(u - 0.5) * L * s, v * H, np.full_like(u, s * W / 2)], -1),
                              d(56), d(22), (7, 3)) for s in (1.0, -1.0)]
               + [param_surface(lambda u, v, s=s: np.stack([np.full_like(u, s * L / 2), v * H, (0.5 - u) * W * s], -1),
                                d(24), d(22), (3, 3)) for s in (1.0, -1.0)]), m_wall),
        (merge([param_surface(lambda u, v, s=s: np.stack([(u - 0.5) * L, np.full_like(u, H * 0.5 + 0.2),
                                                          s * (W / 2 - 1.5) + (0.5 - v) * 3.0], -1),
                              d(56), d(6), (14, 1.5)) for s in (1.0, -1.0)]), m_floor),  # side galleries
    ])
    sb.add_instance(shell)
    column = sb.add_model([(cylinder(0.35, 4.6, d(48), d(40), lambda v: 1 + 0.12 * np.cos(v * np.pi * 2) ** 8
                                     + 0.04 * np.sin(v * 40)), m_column),
                           (box(0.95, 0.3, 0.95, 2), m_column)])
    arch = sb.add_model([(param_surface(
        lambda u, v: np.stack([1.75 * np.cos(np.pi * u) * (1 + 0.1 * v), 1.75 * np.sin(np.pi * u) * (1 + 0.1 * v),
                               (v - 0.5) * 0.8], -1), d(40), d(6), (4, 1)), m_wall)])
    n_cols = 8
    for storey, y in enumerate((0.0, H * 0.5 + 0.2)):
        for i in range(n_cols):
            x = -L / 2 + 2.0 + i * (L - 4.0) / (n_cols - 1)
            for s in (1.0, -1.0):
                sb.add_instance(column, trs((x, y, s * (W / 2 - 2.2)), (0, 1, 0), 0.3 * i))
                if i + 1 < n_cols:
                    sb.add_instance(arch, trs((x + (L - 4.0) / (n_cols - 1) / 2, y + 4.6, s * (W / 2 - 2.2)),
                                              (0, 1, 0), 0.0, (1.0, 0.55, 1.0)))
    drape = lambda ph: param_surface(
        lambda u, v: np.stack([(u - 0.5) * 2.6, -v * 3.5 * (1 - 0.25 * np.sin(np.pi * u)),
                               0.18 * np.sin(u * 19 + ph) * (0.3 + v) + 0.1 * np.sin(v * 7 + ph)], -1),
        d(90), d(60), (2, 2))
    drapes = [sb.add_model([(drape(ph), m)]) for ph, m in ((0.0, m_drape_r), (1.3, m_drape_g), (2.6, m_drape_b))]
    for i in range(n_cols - 1):
        x = -L / 2 + 2.0 + (i + 0.5) * (L - 4.0) / (n_cols - 1)
        for s in (1.0, -1.0):
            sb.add_instance(drapes[(i + (s > 0)) % 3], trs((x, H * 0.5 - 0.2, s * (W / 2 - 2.0))))
    statue = sb.add_model([(bumpy_sphere(0.9, d(140), d(100), 0.18, 11.0, 0.7, (1, 1.3, 1)), m_bronze),
                           (box(1.6, 0.8, 1.6, 4), m_column)])
    sb.add_instance(statue, trs((L / 2 - 3.0, 1.5, 0)))
    sb.add_instance(statue, trs((1.0, 1.2, 0.4), (0, 1, 0), 2.0, (0.8, 0.8, 0.8)))
    # planters: pot + alpha-masked foliage cards (the Sponza plants / chains)
    n_cards = int(2400 * detail)
    bushes = [sb.add_model([(cylinder(0.45, 0.7, d(24), d(4), lambda v: 0.7 + 0.3 * v), m_bronze),
                            (cards(np.random.default_rng(seed * 100 + k), n_cards, 0.55, 0.9, 2.4), m_leaf)])
              for k in range(3)]
    k = 0
    for x in np.linspace(-L / 2 + 4, L / 2 - 4, 6):
        for s in (1.0, -1.0):
            sb.add_instance(bushes[k % 3], trs((x, 0, s * 2.7), (0, 1, 0), 0.9 * k))
            k += 1
    cam = make_camera((-L / 2 + 1.5, 2.0, 0.5), (L / 2, 3.4, -0.3), width, height)
    points = spots = None
    if lights:
        points, spots = make_lights(lights, (-L / 2 + 1, 0.5, -W / 2 + 1), (L / 2 - 1, H - 1, W / 2 - 1), 3)
    if bake:
        sb = sb.baked()
    sc = sb.finish(cam, directional=((3, 2.8, 2.5, 0), (-0.3, -1, -0.2, 0)), point_lights=points, spot_lights=spots,
                   sky=procedural_sky(64))
    sc.name = "atrium" if not lights else f"atrium+{lights}lights"
    return sc


def stress(width=1920, height=1080, instances=10000, seed=5, mesh_tris=10000, models=4):
    """C5 stand-in: `instances` random-TRS instances of `models` ~`mesh_tris`-triangle meshes (about
    instances x mesh_tris instanced triangles, ~100 M at the defaults) for TLAS rebuild timing and the
    two-level traversal sweep. Untextured opaque materials, sun + sky."""
    rng = np.random.default_rng(seed)
    sb = SceneBuilder("stress")
    mats = [sb.add_material(make_material(base_color=(*c, 1.0), roughness=r, metallic=m))
            for c, r, m in (((0.8, 0.3, 0.25), 0.6, 0.0), ((0.3, 0.7, 0.35), 0.4, 0.0), ((0.3, 0.4, 0.8), 0.8, 0.0),
                            ((0.85, 0.8, 0.6), 0.3, 1.0))]
    nu = max(8, int(round(np.sqrt(mesh_tris / 2.0 * 1.03))))
    nv = max(6, int(round(mesh_tris / 2.0 / nu)))
    ids = [sb.add_model([(bumpy_sphere(0.5, nu, nv, 0.12 + 0.04 * k, 5.0 + 2 * k, 0.3 * k, (1, 1 + 0.2 * k, 1)), mats[k % len(mats)])])
           for k in range(models)]
    side = max(4.0, 1.6 * instances ** (1.0 / 3.0))  # keeps the instance density constant
    pos = rng.uniform((-side, 0.5, -side), (side, side * 0.6 + 1.0, side), (instances, 3))
    axis = rng.normal(size=(instances, 3))
    axis /= np.linalg.norm(axis, axis=1, keepdims=True)
    ang = rng.uniform(0, 2 * np.pi, instances)
    scl = rng.uniform(0.5, 1.5, (instances, 3))
    for i in range(instances):
        sb.add_instance(ids[i % models], trs(tuple(pos[i]), tuple(axis[i]), float(ang[i]), tuple(scl[i])))
    floor = sb.add_model([(plane(4 * side, 4 * side, 8, 8, (8, 8)), mats[3])])
    sb.add_instance(floor)
    cam = make_camera((-1.6 * side, side * 0.8, 1.9 * side), (0, side * 0.3, 0), width, height)
    sc = sb.finish(cam, directional=((3, 2.8, 2.5, 0), (-0.3, -1, -0.2, 0)), sky=procedural_sky(64))
    sc.name = f"stress({instances})"
    return sc


def district(width=3840, height=2160, seed=4, detail=1.0, animated=1024):
    """C4 stand-in ("Bistro-scale"): a street grid of facade / roof / tree models, ~3.5 M unique triangles at
    detail=1 over ~1300 instances, the first `animated` of which are meant to be moved every frame (see
    district_frame) so that the TLAS is rebuilt per frame. Untextured materials, sun + sky."""
    rng = np.random.default_rng(seed)
    sb = SceneBuilder("district")
    d = lambda n: max(3, int(round(n * np.sqrt(detail))))
    mats = [sb.add_material(make_material(base_color=(*c, 1.0), roughness=r, metallic=m))
            for c, r, m in (((0.7, 0.66, 0.6), 0.8, 0.0), ((0.5, 0.3, 0.25), 0.7, 0.0), ((0.35, 0.4, 0.45), 0.5, 0.0),
                            ((0.2, 0.45, 0.2), 0.9, 0.0), ((0.8, 0.8, 0.85), 0.25, 1.0))]
    models = []
    for k in range(16):  # facades: bumpy boxes, ~190k triangles each at detail=1
        hgt = 6.0 + 3.0 * (k % 4)
        wall = lambda u, v, s, k=k, hgt=hgt: np.stack(
            [(u - 0.5) * 8.0, v * hgt, np.full_like(u, 4.0) + 0.15 * np.sin(u * (20 + 3 * k)) * np.sin(v * (14 + k))], -1)
        parts = []
        for q in range(4):
            p, n, uv, t, idx = param_surface(lambda u, v: wall(u, v, 1.0), d(160), d(150), (4, 4))
            c, s_ = np.cos(q * np.pi / 2), np.sin(q * np.pi / 2)
            rot = np.array([[c, 0, s_], [0, 1, 0], [-s_, 0, c]], np.float32)
            parts.append((p @ rot.T, n @ rot.T, uv, np.concatenate([t[:, :3] @ rot.T, t[:, 3:4]], 1), idx))
        models.append(sb.add_model([(merge(parts), mats[k % 3]), (box(8.2, 0.4, 8.2, 2), mats[4])]))
    trees = [sb.add_model([(bumpy_sphere(1.6, d(160), d(120), 0.35, 9.0 + k, 0.4 * k, (1, 1.4, 1)), mats[3]),
                           (cylinder(0.25, 2.0, d(16), d(4)), mats[1])]) for k in range(8)]
    grid, pitch = 18, 14.0
    cells = [(i, j) for i in range(grid) for j in range(grid)]
    rng.shuffle(cells)
    for n, (i, j) in enumerate(cells):
        x, z = (i - grid / 2 + 0.5) * pitch, (j - grid / 2 + 0.5) * pitch
        sb.add_instance(models[n % len(models)], trs((x, 0, z), (0, 1, 0), 0.5 * np.pi * (n % 4)))
    for n in range(1000):
        i, j = rng.integers(0, grid, 2)
        x, z = (i - grid / 2) * pitch + rng.uniform(-1, 1), (j - grid / 2 + 0.5) * pitch + rng.uniform(-5, 5)
        sb.add_instance(trees[n % len(trees)], trs((x, 2.0, z), (0, 1, 0), float(rng.uniform(0, 6.28)), (1, float(rng.uniform(0.8, 1.5)), 1)))
    # the moving instances come first in the arrays: reorder so that district_frame can rotate [0, animated)
    sb.instances = sb.instances[::-1]
    ground = sb.add_model([(plane(grid * pitch, grid * pitch, 64, 64, (32, 32)), mats[0])])
    sb.add_instance(ground)
    half = grid * pitch / 2
    cam = make_camera((-half * 0.9, 22.0, half * 0.95), (0, 4.0, 0), width, height)
    sc = sb.finish(cam, directional=((3, 2.8, 2.5, 0), (-0.4, -1, -0.3, 0)), sky=procedural_sky(64))
    sc.name = "district"
    sc.animated = min(animated, len(sc.tlas_instances) - 1)
    return sc


def district_frame(scene, frame: int):
    """Per-frame update of an animated scene (World::updateAnimations + updateScene, World.cpp:349-466): rotates the first
    `scene.animated` instances about Y by 0.02 rad per frame around their own origin and rewrites their
    ModelInstanceTransforms and TLAS instance records in place (linear keyframes of a rotation track)."""
    from .scene_data import instance_transforms
    n = scene.animated
    base = getattr(scene, "_base_m2w", None)
    if base is None:
        base = scene.transforms["modelToWorld"][:n].copy()  # [n, 3, 4]
        scene._base_m2w = base
    a = 0.02 * frame
    c, s = np.cos(a), np.sin(a)
    ry = np.array([[c, 0, s, 0], [0, 1, 0, 0], [-s, 0, c, 0], [0, 0, 0, 1]], np.float64)
    for i in range(n):
        m = np.eye(4)
        m[:3, :4] = base[i]
        t = instance_transforms(m @ ry)
        scene.transforms[i] = t
        scene.tlas_instances[i]["transform"] = t["modelToWorld"]
