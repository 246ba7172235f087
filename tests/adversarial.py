"""Scene and rays for the box-test conservativeness checks (TEST INFRASTRUCTURE; round-1 advisor finding on
core_trace.h::intersectNode): flat axis-aligned triangles whose boxes have zero thickness and lie exactly on the
quantisation grid of large nodes, hit by rays that start ON other surfaces of the same large nodes (secondary-ray
style), far away, or almost inside the triangle's plane, and that aim at the triangles' edges and corners displaced by
0, a few ulps, 1e-5 or 3e-4. The oracle has its own BVH (other topology, other padding), so identical hit records on
these rays mean neither tree culled a triangle the exact test accepts."""
import numpy as np

from prosper_b200 import scene_data as sd
from prosper_b200.scene_data import make_camera, make_material
from prosper_b200.scenes import SceneBuilder, merge, param_surface, plane, procedural_sky


def _quad_xz(cx, cz, y, s):
    return param_surface(lambda u, v: np.stack([cx + (u - 0.5) * s, np.full_like(u, y), cz + (0.5 - v) * s], -1), 1, 1)


def _wall_x(x, s, n):
    return param_surface(lambda u, v: np.stack([np.full_like(u, x), v * s, (u - 0.5) * s], -1), n, n)


def _wall_z(z, s, n):
    return param_surface(lambda u, v: np.stack([(u - 0.5) * s, v * s, np.full_like(u, z)], -1), n, n)


def flat_grid_scene(stickers=300, seed=3):
    """Floor on a 2 m grid, four walls at +-8 m, `stickers` 1 m quads on integer coordinates at heights on a 0.25 m grid;
    instanced twice (the second one rotated, so the two-level path and the world->object transform take part)."""
    sb = SceneBuilder("flat_grid")
    m = sb.add_material(make_material((0.5, 0.5, 0.5, 1)))
    rng = np.random.default_rng(seed)
    parts = [plane(64.0, 64.0, 32, 32), _wall_x(-8.0, 16.0, 8), _wall_x(8.0, 16.0, 8), _wall_z(-8.0, 16.0, 8), _wall_z(8.0, 16.0, 8)]
    centres = []
    for _ in range(stickers):
        cx, cz, y = float(rng.integers(-7, 8)), float(rng.integers(-7, 8)), float(rng.integers(1, 40)) * 0.25
        parts.append(_quad_xz(cx, cz, y, 1.0))
        centres.append((cx, y, cz))
    model = sb.add_model([(merge(parts), m)])
    sb.add_instance(model)
    sb.add_instance(model, sd.trs((40.0, 0.0, 3.0), (0, 1, 0), 0.7))
    scene = sb.finish(make_camera((0.5, 3.0, 7.5), (0.0, 2.0, 0.0), 64, 48), sky=procedural_sky(8))
    return scene, np.array(centres, np.float32)


def edge_rays(centres, n, seed=17):
    """(origin.xyz, tMin), (direction.xyz, tMax) as two float32 [n, 4] arrays."""
    rng = np.random.default_rng(seed)
    c = centres[rng.integers(0, len(centres), n)]
    edge = rng.integers(0, 4, n)
    along = rng.uniform(-0.5, 0.5, n)
    off = rng.choice([0.0, 1e-7, -1e-7, 1e-5, -1e-5, 3e-4, -3e-4], n)
    tx = np.where(edge < 2, c[:, 0] + np.where(edge == 0, -0.5, 0.5) + off, c[:, 0] + along)
    tz = np.where(edge < 2, c[:, 2] + along, c[:, 2] + np.where(edge == 2, -0.5, 0.5) + off)
    target = np.stack([tx, c[:, 1], tz], 1)
    kind = rng.integers(0, 4, n)  # origin on a wall at x, on the floor, on a sticker plane, on a wall at z
    o = np.zeros((n, 3))
    o[:, 0] = np.where(kind == 0, rng.choice([-8.0, 8.0], n), rng.uniform(-7.9, 7.9, n))
    o[:, 1] = np.where(kind == 1, 0.0, np.where(kind == 2, rng.integers(1, 40, n) * 0.25, rng.uniform(0.01, 15, n)))
    o[:, 2] = np.where(kind == 3, rng.choice([-8.0, 8.0], n), rng.uniform(-7.9, 7.9, n))
    far = rng.random(n) < 0.15
    o[far] = target[far] + (o[far] - target[far]) * 4.0
    d = target - o
    d /= np.maximum(np.linalg.norm(d, axis=1, keepdims=True), 1e-12)
    graze = rng.random(n) < 0.25  # almost inside the sticker's plane
    d[graze, 1] = rng.choice([1e-4, -1e-4, 1e-6, -1e-6, 0.0], int(graze.sum()))
    d /= np.maximum(np.linalg.norm(d, axis=1, keepdims=True), 1e-12)
    origin = np.zeros((n, 4), np.float32)
    direction = np.zeros((n, 4), np.float32)
    origin[:, :3] = o
    direction[:, :3] = d
    origin[:, 3] = rng.choice([0.0, 1e-4, 1e-3], n)
    direction[:, 3] = np.where(rng.random(n) < 0.7, np.inf, rng.uniform(1.0, 20.0, n))
    return origin, direction
