"""CPU: pins the oracle (the reference has no tests for this path): known-answer
vectors, analytic properties of the BRDF / sampling functions, BVH-vs-brute-force agreement."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import oracle
from prosper_b200 import bc7, scene_data as sd

HERE = os.path.dirname(os.path.abspath(__file__))
KAT = json.load(open(os.path.join(HERE, "golden", "kat.json")))


def _f3(*v):
    return np.array(v, np.float32)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def test_pcg_known_answers():
    l = oracle.lib()
    for k, v in KAT["pcg"].items():
        assert l.orc_pcg(int(k)) == v
    pins = KAT["survey_pins"]
    assert l.orc_pcg(0) == pins["pcg(0)"] and l.orc_pcg(0xFFFFFFFF) == pins["pcg(0xFFFFFFFF)"] and l.orc_pcg(12345) == pins["pcg(12345)"]


def test_pcg3d_known_answers():
    l = oracle.lib()
    for e in KAT["pcg3d"]:
        i = np.array(e["in"], np.uint32)
        o = np.zeros(3, np.uint32)
        l.orc_pcg3d(_ptr(i), _ptr(o))
        assert o.tolist() == e["out"]
        o2 = np.zeros(3, np.uint32)
        l.orc_pcg3d(_ptr(o), _ptr(o2))
        assert o2.tolist() == e["next"]
    pins = KAT["survey_pins"]
    assert KAT["pcg3d"][0]["out"] == pins["pcg3d(0,0,1)"] and KAT["pcg3d"][3]["out"] == pins["pcg3d(1919,1079,4095)"]


def test_rng_to_01_reaches_one():
    l = oracle.lib()
    assert l.orc_rng_to_01(0) == 0.0 and l.orc_rng_to_01(0xFFFFFFFF) == 1.0 and l.orc_rng_to_01(0x80000000) == 0.5


def test_packing_known_answers():
    l = oracle.lib()
    pins = KAT["survey_pins"]
    halves = [l.orc_float_to_half(C.c_float(v)) for v in (1.5, -2.0, 0.1, 1.0)]
    assert sum(h << (16 * i) for i, h in enumerate(halves)) == pins["packHalf4x16(1.5,-2,0.1,1)"]
    assert int(sd.pack_snorm3x10_1x2(np.array([[0, 0, 1, 0]], np.float32))[0]) == pins["packSnorm3x10_1x2(0,0,1,0)"]
    for h in (0, 1, 0x3FF, 0x400, 0x3C00, 0x7BFF, 0x8001, 0xC000):
        f = l.orc_half_to_float(h)
        assert np.float32(f) == np.array([h], np.uint16).view(np.float16)[0].astype(np.float32)
        assert l.orc_float_to_half(C.c_float(f)) == h
    out = np.zeros(3, np.float32)
    l.orc_unpack_snorm10(pins["packSnorm3x10_1x2(0,0,1,0)"], _ptr(out))
    assert np.allclose(out, (0, 0, 1))


def test_bc7_decoder_round_trip():
    rng = np.random.default_rng(0)
    img = rng.integers(0, 256, (16, 16, 4), dtype=np.uint8)
    img[:8] = np.repeat(np.repeat(rng.integers(0, 256, (2, 4, 4), dtype=np.uint8), 4, 0), 4, 1)  # flat blocks encode almost losslessly
    blocks = bc7.encode_bc7_mode6(img)
    ref = bc7.decode_bc7_mode6(blocks, 16, 16)
    l = oracle.lib()
    out = np.zeros((16, 4, 4, 4), np.uint8)
    for b in range(16):
        blk = np.ascontiguousarray(blocks[b * 16:(b + 1) * 16])
        assert l.orc_bc7_decode_block(_ptr(blk), _ptr(out[b])) == 0
    got = out.reshape(4, 4, 4, 4, 4).transpose(0, 2, 1, 3, 4).reshape(16, 16, 4)
    assert np.array_equal(got, ref)
    assert np.abs(got[:8].astype(int) - img[:8].astype(int)).max() <= 2


def test_vndf_pdf_is_the_references_unnormalised_form():
    l = oracle.lib()
    rng = np.random.default_rng(1)
    for alpha, ve in ((0.25, _f3(0.3, 0.1, 0.95)), (0.6, _f3(0.7, -0.2, 0.68))):
        ve = ve / np.linalg.norm(ve)
        # Monte-Carlo estimate of the integral of pdf over the hemisphere with uniform directions
        n = 200000
        z = rng.random(n)
        phi = rng.random(n) * 2 * np.pi
        r = np.sqrt(1 - z * z)
        dirs = np.stack([r * np.cos(phi), r * np.sin(phi), z], 1).astype(np.float32)
        vals = np.array([l.orc_vndf_pdf(_ptr(ve), _ptr(np.ascontiguousarray(d)), C.c_float(alpha)) for d in dirs[:20000]])
        est = vals.mean() * 2 * np.pi
        # reference quirk (sampling.glsl:81-93): the "pdf" carries the full two-direction Schlick-GGX G
        # instead of G1(V), so it integrates to LESS than one; the restatement must keep that
        assert 0.3 < est < 1.0 and np.all(np.isfinite(vals)) and np.all(vals >= 0)


def test_cosine_sample_is_in_hemisphere_and_unit():
    l = oracle.lib()
    rng = np.random.default_rng(2)
    n = _f3(0.0, 0.6, 0.8)
    for _ in range(200):
        u = rng.random(2).astype(np.float32)
        out = np.zeros(3, np.float32)
        l.orc_cosine_sample_hemisphere(_ptr(n), _ptr(u), _ptr(out))
        assert abs(np.linalg.norm(out) - 1) < 1e-5 and np.dot(out, n) >= -1e-6


def test_brdf_furnace_bound_and_reciprocity():
    l = oracle.lib()
    rng = np.random.default_rng(3)
    n = _f3(0, 0, 1)
    albedo = _f3(0.8, 0.8, 0.8)
    for rough in (0.3, 0.8):
        v = _f3(0.4, 0.0, 0.9165)
        acc = np.zeros(3)
        m = 4000
        for _ in range(m):
            z, phi = rng.random(), rng.random() * 2 * np.pi
            r = np.sqrt(1 - z * z)
            ldir = _f3(r * np.cos(phi), r * np.sin(phi), z)
            out = np.zeros(3, np.float32)
            l.orc_eval_brdf_times_nol(_ptr(ldir), _ptr(v), _ptr(n), _ptr(albedo), C.c_float(rough), C.c_float(0.0), _ptr(out))
            acc += out
        assert np.all(acc / m * 2 * np.pi < 1.15)  # energy roughly conserved (white furnace)


def test_offset_ray_moves_along_normal():
    l = oracle.lib()
    for p, n in ((_f3(1, 2, 3), _f3(0, 1, 0)), (_f3(-4, 0.001, 2), _f3(-0.6, 0.8, 0)), (_f3(0.01, 0.02, -0.01), _f3(0, 0, -1))):
        out = np.zeros(3, np.float32)
        l.orc_offset_ray(_ptr(p), _ptr(n), _ptr(out))
        assert np.dot(out - p, n) > 0 and np.abs(out - p).max() < 1e-2


def test_pinhole_centre_is_camera_forward(test_scene):
    l = oracle.lib()
    cam = np.ascontiguousarray(test_scene.camera.reshape(1))
    o, d = np.zeros(3, np.float32), np.zeros(3, np.float32)
    l.orc_pinhole_ray(_ptr(cam), _ptr(_f3(0.5, 0.5)[:2].copy()), _ptr(o), _ptr(d))
    w2c = cam["worldToCamera"][0]
    fwd = -np.array([w2c[0][2], w2c[1][2], w2c[2][2]])
    assert np.allclose(d, fwd, atol=1e-6) and np.allclose(o, cam["eye"][0][:3])


def test_intersect_triangle_contract():
    l = oracle.lib()
    v0, v1, v2 = _f3(0, 0, 0), _f3(1, 0, 0), _f3(0, 1, 0)
    tuv = np.zeros(3, np.float32)
    assert l.orc_intersect_triangle(_ptr(_f3(0.25, 0.25, 1)), _ptr(_f3(0, 0, -1)), C.c_float(0), C.c_float(np.inf), _ptr(v0), _ptr(v1), _ptr(v2), _ptr(tuv)) == 1
    assert np.allclose(tuv, (1, 0.25, 0.25))
    # two-sided, open interval, edge-inclusive, parallel rays miss
    assert l.orc_intersect_triangle(_ptr(_f3(0.25, 0.25, -1)), _ptr(_f3(0, 0, 1)), C.c_float(0), C.c_float(np.inf), _ptr(v0), _ptr(v1), _ptr(v2), _ptr(tuv)) == 1
    assert l.orc_intersect_triangle(_ptr(_f3(0.25, 0.25, 1)), _ptr(_f3(0, 0, -1)), C.c_float(0), C.c_float(1.0), _ptr(v0), _ptr(v1), _ptr(v2), _ptr(tuv)) == 0
    assert l.orc_intersect_triangle(_ptr(_f3(0.5, 0.0, 1)), _ptr(_f3(0, 0, -1)), C.c_float(0), C.c_float(np.inf), _ptr(v0), _ptr(v1), _ptr(v2), _ptr(tuv)) == 1
    assert l.orc_intersect_triangle(_ptr(_f3(0.25, 0.25, 1)), _ptr(_f3(1, 0, 0)), C.c_float(0), C.c_float(np.inf), _ptr(v0), _ptr(v1), _ptr(v2), _ptr(tuv)) == 0


@pytest.mark.parametrize("fixture", ["test_scene", "small_atrium_instanced"])
def test_bvh_agrees_with_brute_force(fixture, request):
    scene = request.getfixturevalue(fixture)
    w, h = (96, 64) if fixture == "test_scene" else (80, 45)
    a, b = oracle.Oracle(scene), oracle.Oracle(scene, brute_force=True)
    pc = sd.reference_pc(frame_index=5)
    ha, _ = a.trace_primary(pc, w, h)
    hb, _ = b.trace_primary(pc, w, h)
    assert np.array_equal(ha, hb)
    ia, ca = a.render(sd.reference_pc(max_bounces=3), w // 2, h // 2, 1)
    ib, cb = b.render(sd.reference_pc(max_bounces=3), w // 2, h // 2, 1)
    assert np.array_equal(ia, ib) and np.array_equal(ca, cb)


def test_oracle_reproduces_golden_fixtures():
    from prosper_b200 import scene_zoo
    for name, fmt in (("rgba8", sd.TEX_RGBA8), ("bc7", sd.TEX_BC7)):
        g = np.load(os.path.join(HERE, "golden", f"test_scene_{name}.npz"))
        orc = oracle.Oracle(scene_zoo.test_scene(fmt=fmt))
        hits, _ = orc.trace_primary(sd.reference_pc(frame_index=2), 96, 64)
        assert np.array_equal(hits.view(np.uint32).reshape(-1, 4), g["hits_f2"])
        img, counters = orc.render(sd.reference_pc(max_bounces=4), 96, 64, 8)
        assert np.array_equal(counters, g["counters_8spp"])
        assert np.allclose(img, g["image_8spp"].astype(np.float32), rtol=2e-3, atol=2e-3)


def test_empty_and_inactive_scenes(test_scene):
    import copy
    s = copy.copy(test_scene)
    s.tlas_instances = test_scene.tlas_instances.copy()
    s.tlas_instances["blasReference"] = 0  # every instance inactive (World.cpp:898-905)
    hits, _ = oracle.Oracle(s).trace_primary(sd.reference_pc(), 16, 8)
    assert np.all(hits["drawInstanceIndex"] == sd.MISS)
