"""Scene host (libprt_host.so, include/prt_host.h): glTF ingestion through the reference's own cgltf / MikkTSpace /
meshoptimizer / glm code, the `.prosper_mesh` caches it writes, and the per-frame scene update (node hierarchy +
animation samplers, World.cpp:349-466, Animations.hpp:71-125). CPU only; the GPU side is tests/test_gpu_gltf.py."""
import base64
import ctypes as C
import json
import os
import struct

import numpy as np
import pytest

import oracle
from prosper_b200 import gltf_scene, mesh_cache, scene_data as sd
from tests import hostsim

REFERENCE_GLTF = "/root/reference/res/glTF/FlightHelmet/glTF/FlightHelmet.gltf"


def test_host_library_exports_its_header():
    lib = gltf_scene.load_host_library()
    header = open(os.path.join(os.path.dirname(os.path.dirname(__file__)), "include", "prt_host.h")).read()
    for sym in gltf_scene.HOST_SYMBOLS:
        assert hasattr(lib, sym), sym
        assert sym + "(" in header, f"{sym} not declared in include/prt_host.h"


@pytest.mark.skipif(not os.path.exists(REFERENCE_GLTF), reason="the reference's FlightHelmet lives in this container only")
def test_flight_helmet_caches_are_reproducible(tmp_path):
    """The committed fixture IS what the host writes from the reference's asset: regenerate the mesh caches and compare bytes."""
    g = gltf_scene.GltfScene(REFERENCE_GLTF, 64, 64, cache_dir=str(tmp_path), source_write_time=1, load_textures=False)
    c = g.counts
    assert (c.meshes, c.models, c.modelInstances, c.drawInstances, c.materials, c.textures, c.nodes) == (5, 5, 5, 5, 6, 16, 6)
    assert g.scene.triangle_count == 94722
    assert [int(m) for m in g.scene.materials["alphaMode"]] == [0, 0, 0, sd.ALPHA_BLEND, 0, 0]  # "LensesMat" is BLEND
    for i in range(c.meshes):
        a = open(mesh_cache.cache_path(str(tmp_path), i), "rb").read()
        b = open(mesh_cache.cache_path(gltf_scene.FLIGHT_HELMET_DIR, i), "rb").read()
        assert a == b, f"cache{i}.prosper_mesh differs from the committed fixture"
        header, _ = mesh_cache.read_cache(mesh_cache.cache_path(str(tmp_path), i))
        assert header.meshletCount > 0 and header.sourceWriteTime == 1 and header.usesShortIndices == 1
    # a second load finds valid caches and must not rewrite them (cacheValid, DeferredLoadingContext.cpp:560-574)
    stamp = os.path.getmtime(mesh_cache.cache_path(str(tmp_path), 0))
    gltf_scene.GltfScene(REFERENCE_GLTF, 64, 64, cache_dir=str(tmp_path), source_write_time=1, load_textures=False).close()
    assert os.path.getmtime(mesh_cache.cache_path(str(tmp_path), 0)) == stamp
    # the root node's rotation reaches every instance: modelToWorld = Ry(180 deg) * T(child)
    m = g.scene.transforms["modelToWorld"]
    assert np.allclose(m[:, 0, 0], -1) and np.allclose(m[:, 2, 2], -1) and np.allclose(m[:, 1, 1], 1)
    g.close()


def test_flight_helmet_fixture_traces_identically_in_oracle_and_kernel_logic():
    scene = gltf_scene.flight_helmet(160, 160)
    assert scene.triangle_count == 94722 and len(scene.tlas_instances) == 5
    orc, hs = oracle.Oracle(scene), hostsim.HostSim(scene)
    for f in (1, 2):
        pc = sd.reference_pc(frame_index=f)
        ho, margin = orc.trace_primary(pc, 160, 160)
        hg = hs.trace_primary(pc, 160, 160)
        bad = ((ho["drawInstanceIndex"] != hg["drawInstanceIndex"]) | (ho["primitiveID"] != hg["primitiveID"])) & ~(margin < 1e-4)
        assert bad.sum() == 0
        assert (ho["drawInstanceIndex"] != sd.MISS).mean() > 0.2
        assert set(np.unique(ho["drawInstanceIndex"])) >= {0, 1, 2, 3, 4}  # every part of the helmet is hit, the BLEND lenses included


# ---- a synthetic animated glTF: hierarchy, all three interpolation modes, camera, punctual lights -------------------
def _animated_gltf(path):
    quad_pos = np.array([[-1, 0, -1], [1, 0, -1], [1, 0, 1], [-1, 0, 1]], np.float32)
    quad_nrm = np.tile(np.array([[0, 1, 0]], np.float32), (4, 1))
    quad_idx = np.array([0, 2, 1, 0, 3, 2], np.uint16)
    times = np.array([0.0, 1.0, 2.0], np.float32)
    trans = np.array([[0, 0, 0], [2, 0, 0], [2, 4, 0]], np.float32)
    rots = np.array([[0, 0, 0, 1], [0, np.sin(np.pi / 4), 0, np.cos(np.pi / 4)], [0, 1, 0, 0]], np.float32)  # about +Y: 0, 90, 180 deg
    scales = np.array([[1, 1, 1], [2, 2, 2], [3, 1, 1]], np.float32)
    # cubic spline: (in-tangent, value, out-tangent) per key
    cubic = np.array([[0, 0, 0], [0, 0, 0], [1, 0, 0], [0, 0, 0], [0, 1, 0], [0, 0, 1], [0, 0, 0], [0, 1, 2], [0, 0, 0]], np.float32)
    chunks, views, accessors = [], [], []

    def add(a, typ, comp, minmax=False):
        raw = a.tobytes()
        off = sum(len(c) for c in chunks)
        chunks.append(raw + b"\0" * ((4 - len(raw) % 4) % 4))
        views.append({"buffer": 0, "byteOffset": off, "byteLength": len(raw)})
        acc = {"bufferView": len(views) - 1, "componentType": comp, "count": int(a.shape[0]), "type": typ}
        if minmax:
            acc["min"] = [float(v) for v in np.atleast_1d(a.min(axis=0))]
            acc["max"] = [float(v) for v in np.atleast_1d(a.max(axis=0))]
        accessors.append(acc)
        return len(accessors) - 1

    a_pos = add(quad_pos, "VEC3", 5126, True)
    a_nrm = add(quad_nrm, "VEC3", 5126)
    a_idx = add(quad_idx, "SCALAR", 5123)
    a_time = add(times, "SCALAR", 5126, True)
    a_tr, a_rot, a_sc, a_cub = add(trans, "VEC3", 5126), add(rots, "VEC4", 5126), add(scales, "VEC3", 5126), add(cubic, "VEC3", 5126)
    blob = b"".join(chunks)
    gltf = {
        "asset": {"version": "2.0"}, "scene": 0, "scenes": [{"nodes": [0, 5]}],
        "extensionsUsed": ["KHR_lights_punctual"],
        "extensions": {"KHR_lights_punctual": {"lights": [
            {"type": "point", "color": [1, 0.5, 0.25], "intensity": 40.0},
            {"type": "spot", "color": [1, 1, 1], "intensity": 10.0, "spot": {"innerConeAngle": 0.2, "outerConeAngle": 0.6}},
            {"type": "directional", "color": [1, 0.9, 0.8], "intensity": 3.0}]}},
        "nodes": [
            {"name": "root", "translation": [0, 1, 0], "children": [1, 2, 3, 4]},
            {"name": "spinner", "mesh": 0, "rotation": [0, 0, 0, 1]},
            {"name": "scaler", "mesh": 0, "translation": [5, 0, 0], "children": [6]},
            {"name": "cam", "camera": 0, "translation": [0, 2, 8]},
            {"name": "lamp", "extensions": {"KHR_lights_punctual": {"light": 0}}, "translation": [1, 2, 3]},
            {"name": "sun", "extensions": {"KHR_lights_punctual": {"light": 2}}, "rotation": [float(-np.sin(np.pi / 4)), 0, 0, float(np.cos(np.pi / 4))],
             "children": [7]},
            {"name": "leaf", "mesh": 0, "translation": [0, 0, 1], "extensions": {"KHR_lights_punctual": {"light": 1}}},
            {"name": "static", "mesh": 0, "matrix": [2, 0, 0, 0, 0, 2, 0, 0, 0, 0, 2, 0, 3, 4, 5, 1]},
        ],
        "meshes": [{"primitives": [{"attributes": {"POSITION": a_pos, "NORMAL": a_nrm}, "indices": a_idx}]}],
        "cameras": [{"type": "perspective", "perspective": {"yfov": 0.8, "znear": 0.05, "zfar": 50.0}}],
        "animations": [{
            "samplers": [{"input": a_time, "output": a_tr, "interpolation": "LINEAR"}, {"input": a_time, "output": a_rot, "interpolation": "LINEAR"},
                         {"input": a_time, "output": a_sc, "interpolation": "STEP"}, {"input": a_time, "output": a_cub, "interpolation": "CUBICSPLINE"}],
            "channels": [{"sampler": 0, "target": {"node": 0, "path": "translation"}}, {"sampler": 1, "target": {"node": 1, "path": "rotation"}},
                         {"sampler": 2, "target": {"node": 2, "path": "scale"}}, {"sampler": 3, "target": {"node": 6, "path": "translation"}}]}],
        "buffers": [{"byteLength": len(blob), "uri": "data:application/octet-stream;base64," + base64.b64encode(blob).decode()}],
        "bufferViews": views, "accessors": accessors,
    }
    with open(path, "w") as f:
        json.dump(gltf, f)
    return times, trans, rots, scales, cubic


def _m2w(t):  # ModelInstanceTransforms.modelToWorld (rows of M) -> 4x4
    m = np.eye(4)
    m[:3, :] = t
    return m


def _rot_y(a):
    c, s = np.cos(a), np.sin(a)
    return np.array([[c, 0, s, 0], [0, 1, 0, 0], [-s, 0, c, 0], [0, 0, 0, 1]])


def _trs(t=(0, 0, 0), r=np.eye(4), s=(1, 1, 1)):
    m = np.eye(4)
    m[:3, 3] = t
    return m @ r @ np.diag([*s, 1.0])


def test_scene_update_hierarchy_and_animation_samplers(tmp_path):
    """updateAnimations + updateScene + updateBuffers against the formulas written out in numpy: STEP, LINEAR (lerp and
    slerp) and CUBICSPLINE samplers (Animations.hpp:71-125, keyframe search Accessors.cpp:25-74), parent * T * R * S down
    the node tree (World.cpp:392-398), camera / light poses (World.cpp:411-452), uniform-scale detection (World.cpp:484-498),
    draw instances and TLAS records (World.cpp:500-516, 878-927), light conversion (WorldData.cpp:1457-1527)."""
    path = str(tmp_path / "anim.gltf")
    times, trans, rots, scales, cubic = _animated_gltf(path)
    g = gltf_scene.GltfScene(path, 64, 64, cache_dir=str(tmp_path))
    c, s = g.counts, g.scene
    assert (c.meshes, c.modelInstances, c.drawInstances, c.nodes, c.pointLights, c.spotLights, c.hasCamera) == (1, 4, 4, 8, 1, 1, 1)
    assert c.animatedNodes == 6  # root and the five nodes under it (dynamic flags propagate); "sun" and "static" are not animated
    assert abs(c.endTimeS - 2.0) < 1e-6
    # lights (WorldData.cpp:1457-1527)
    rad = np.array([1, 0.5, 0.25]) * 40.0 / (4 * np.pi)
    assert np.allclose(s.point_lights[0]["radianceAndRadius"][:3], rad, rtol=1e-6)
    assert np.isclose(s.point_lights[0]["radianceAndRadius"][3], np.sqrt(rad @ [0.2126, 0.7152, 0.0722] / 0.01), rtol=1e-6)
    scale = 1.0 / max(0.001, np.cos(0.2) - np.cos(0.6))
    assert np.isclose(s.spot_lights[0]["radianceAndAngleScale"][3], scale, rtol=1e-5)
    assert np.isclose(s.spot_lights[0]["positionAndAngleOffset"][3], -np.cos(0.6) * scale, rtol=1e-5)
    assert np.allclose(s.directional_light["irradiance"][:3], np.array([1, 0.9, 0.8]) * 3.0, rtol=1e-6)
    assert np.allclose(s.directional_light["direction"][:3], [0, -1, 0], atol=1e-6)  # -Z rotated by -90 deg about X
    # model instances are numbered in gatherScene's visiting order (stack: last child first)
    order = {int(d["modelInstanceIndex"]): i for i, d in enumerate(s.draw_instances)}
    assert len(order) == 4 and np.array_equal(s.tlas_instances["customIndexAndMask"] & 0xFFFFFF, np.arange(4))
    assert np.all(s.tlas_instances["customIndexAndMask"] >> 24 == 0xFF) and np.all(s.tlas_instances["blasReference"] == 1)

    def expect(t):
        def key(t):
            if t <= times[0]:
                return 0, 0.0, 0.0
            if t >= times[-1]:
                return len(times) - 1, 0.0, 0.0
            k = int(np.searchsorted(times, t, side="right") - 1)
            return k, (t - times[k]) / (times[k + 1] - times[k]), t - times[k]
        k, u, td = key(t)
        tr = trans[k] if u == 0 else (1 - u) * trans[k] + u * trans[k + 1]
        if u == 0:
            q = rots[k]
        else:  # slerp
            a, b = rots[k].astype(np.float64), rots[k + 1].astype(np.float64)
            d = a @ b
            if d < 0:
                b, d = -b, -d
            th = np.arccos(min(d, 1.0))
            q = (np.sin((1 - u) * th) * a + np.sin(u * th) * b) / np.sin(th)
        ang = 2 * np.arctan2(q[1], q[3])
        sc = scales[k]
        if u == 0:
            cb = cubic[3 * k + 1]
        else:
            vk, bk, vk1, ak1 = cubic[3 * k + 1], cubic[3 * k + 2], cubic[3 * k + 4], cubic[3 * k + 3]
            t2, t3 = u * u, u * u * u
            cb = (2 * t3 - 3 * t2 + 1) * vk + td * (t3 - 2 * t2 + u) * bk + (-2 * t3 + 3 * t2) * vk1 + td * (t3 - t2) * ak1
        root = _trs(tr)
        out = {"spinner": root @ _trs(r=_rot_y(ang)), "scaler": root @ _trs((5, 0, 0), s=sc)}
        out["leaf"] = out["scaler"] @ _trs(cb)
        out["static"] = _trs(r=np.array([[1, 0, 0, 0], [0, 0, 1, 0], [0, -1, 0, 0], [0, 0, 0, 1.0]])) @ _trs((3, 4, 5), s=(2, 2, 2))
        out["cam"] = root @ _trs((0, 2, 8))
        out["lamp"] = root @ _trs((1, 2, 3))
        return out, sc

    # instance index by node: visiting order is root -> (children pushed 1,2,3,4, popped 4,3,2,1): scaler(2) -> leaf(6), spinner(1); then sun -> static
    names = ["scaler", "leaf", "spinner", "static"]
    for t in (0.0, 0.25, 1.0, 1.5, 1.999, 2.0, 7.0):
        g.update(t)
        want, sc = expect(t)
        for i, n in enumerate(names):
            got = _m2w(s.transforms[i]["modelToWorld"])
            assert np.allclose(got, want[n], atol=2e-5), f"t={t} {n}\n{got}\n{want[n]}"
            assert np.allclose(s.tlas_instances[i]["transform"], want[n][:3], atol=2e-5)
            inv = np.linalg.inv(want[n])
            assert np.allclose(s.transforms[i]["normalToWorld"], inv.T[:3, :], atol=2e-4)
        uniform = g.scales
        assert np.isclose(uniform[names.index("spinner")], 1.0, rtol=1e-4)
        assert np.isclose(uniform[names.index("static")], 2.0, rtol=1e-4)
        assert (uniform[names.index("scaler")] == 0.0) == (not np.allclose(sc, sc[0]))  # non-uniform scale is flagged with 0
        assert np.allclose(s.point_lights[0]["position"][:3], want["lamp"][:3, 3], atol=2e-5)
        assert np.allclose(s.spot_lights[0]["positionAndAngleOffset"][:3], want["leaf"][:3, 3], atol=2e-5)
        d = want["leaf"][:3, :3] @ [0, 0, -1]
        assert np.allclose(s.spot_lights[0]["direction"][:3], d, atol=2e-5)
        assert np.allclose(tuple(g.camera.eye), want["cam"][:3, 3], atol=2e-5)
        assert np.allclose(tuple(g.camera.target), (want["cam"] @ [0, 0, -1, 1])[:3], atol=2e-5)
        assert np.isclose(g.camera.fov, 0.8) and np.isclose(g.camera.zN, 0.05) and np.isclose(g.camera.zF, 50.0)
        assert np.allclose(s.camera["eye"][:3], want["cam"][:3, 3], atol=2e-5)
    g.close()


def test_host_errors_are_loud(tmp_path):
    lib = gltf_scene.load_host_library()
    with pytest.raises(RuntimeError, match="Unknown extension"):
        gltf_scene.GltfScene(str(tmp_path / "scene.obj"), 8, 8)
    with pytest.raises(RuntimeError, match="Failed to parse gltf"):
        gltf_scene.GltfScene(str(tmp_path / "missing.gltf"), 8, 8)
    with pytest.raises(RuntimeError, match="Failed to load texture"):
        gltf_scene.decode_image(str(tmp_path / "missing.png"))
    bad = tmp_path / "prosper_cache"
    bad.mkdir()
    (bad / "cache0.prosper_mesh").write_bytes(struct.pack("=QI", 0x1234, 4) + b"\0" * 60)
    path = str(tmp_path / "anim.gltf")
    _animated_gltf(path)
    with pytest.raises(RuntimeError, match="Expected a valid mesh cache"):  # readCache, DeferredLoadingContext.cpp:520-523
        gltf_scene.GltfScene(path, 8, 8, cache_dir=str(tmp_path))
    assert lib.prth_get_counts(None, None) != 0
