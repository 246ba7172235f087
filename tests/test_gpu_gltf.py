"""-m gpu: real glTF content through the scene host, traced by the CUDA path and checked against the CPU oracle.
FlightHelmet is the one real asset the reference ships (res/glTF/FlightHelmet); its caches under tests/golden/flight_helmet
are what libprt_host.so wrote from it (tests/golden/make_flight_helmet.py), so `primitiveID` here is the triangle order the
reference's own cgltf + meshoptimizer code produces, not a generator's."""
import os

import numpy as np
import pytest

import oracle
from prosper_b200 import gltf_scene, rt_reference as rt, scene_data as sd

pytestmark = pytest.mark.gpu
LUM = np.array([0.2126, 0.7152, 0.0722], np.float32)


def _world(scene, w, h, flags=rt.CFG_COUNTERS | rt.CFG_EXACT_FILTER, spl=1):
    world = rt.World(scene, max_width=w, max_height=h, samples_per_launch=spl, flags=flags)
    world.updateBuffers()
    world.buildAccelerationStructures()
    return world


def _mismatch(a, b):
    return (a["drawInstanceIndex"] != b["drawInstanceIndex"]) | (a["primitiveID"] != b["primitiveID"])


def test_flight_helmet_primary_ids_bit_exact_512():
    """BASELINE.json configs[0] geometry at its resolution: 94,722 triangles in 5 BLASes (4 opaque + the BLEND lenses),
    512x512. IDs and barycentrics bit for bit; alpha-decision ties excluded and bounded (< 0.1 %)."""
    w = h = 512
    scene = gltf_scene.flight_helmet(w, h)
    world, orc = _world(scene, w, h), oracle.Oracle(scene)
    for f in (1, 2, 4095):
        pc = sd.reference_pc(frame_index=f)
        ho, margin = orc.trace_primary(pc, w, h)
        hg = world.trace_primary(pc, w, h)
        tie = margin < 1e-4
        assert (_mismatch(ho, hg) & ~tie).sum() == 0, f"frame {f}"
        assert tie.mean() < 1e-3
        same = ~_mismatch(ho, hg)
        assert np.array_equal(ho["bary"][same], hg["bary"][same])
        assert (hg["drawInstanceIndex"] != sd.MISS).mean() > 0.2
    c = world.counters()
    assert c["stackOverflows"] == 0 and c["anyHitCalls"] > 0  # the lenses went through scene.rahit
    world.close()


def test_flight_helmet_radiance_512_16spp():
    """configs[0]: 512x512, 16 spp, 4 bounces, sky + the default directional light. Luminance relRMSE (1e-2 floor) against
    the oracle < 2 % with the software filter, < 5 % through the texture units (256^2 fixture textures, all magnified);
    mean bias < 0.1 %; ray counts within 0.2 % / 0.5 %."""
    w = h = 512
    spp = 16
    scene = gltf_scene.flight_helmet(w, h)
    ref, counters = oracle.Oracle(scene).render(sd.reference_pc(max_bounces=4), w, h, spp)
    for flags, tol in ((rt.CFG_COUNTERS | rt.CFG_EXACT_FILTER, 0.02), (rt.CFG_COUNTERS, 0.05)):
        world = _world(scene, w, h, flags=flags, spl=4)
        r = rt.RtReference()
        r.init(world)
        r.drawUi(maxBounces=4)
        out = r.record(world, (0, 0, w, h), rt.Options(ibl=True), frames=spp).illumination_f32()
        c = world.counters()
        world.close()
        la, lb = out[..., :3] @ LUM, ref[..., :3] @ LUM
        rel = float(np.sqrt(np.mean(((la - lb) / (np.abs(lb) + 1e-2)) ** 2)))
        bias = float((out[..., :3].mean() - ref[..., :3].mean()) / ref[..., :3].mean())
        assert rel < tol, f"flags {flags}: relRMSE {rel:.4f}"
        assert abs(bias) < 1e-3, f"flags {flags}: bias {bias:.2e}"
        assert abs(c["closestRays"] - int(counters[0])) <= 0.002 * int(counters[0])
        assert abs(c["shadowRays"] - int(counters[1])) <= 0.005 * int(counters[1])
        assert c["stackOverflows"] == 0


def test_animated_gltf_tlas_rebuilt_from_host_update(tmp_path):
    """The frame loop of the reference on an animated glTF: updateAnimations + updateScene + updateBuffers on the host
    (libprt_host.so), ring uploads, TLAS rebuild from the host's instance records, trace -- IDs against the oracle on
    the same arrays at several animation times, camera taken from the glTF's animated camera node."""
    from tests.test_gltf_host import _animated_gltf
    path = str(tmp_path / "anim.gltf")
    _animated_gltf(path)
    w, h = 192, 128
    g = gltf_scene.GltfScene(path, w, h, cache_dir=str(tmp_path))
    world = _world(g.scene, w, h)
    seen = set()
    for t in (0.0, 0.4, 1.0, 1.7, 2.0):
        g.update(t)
        world.updateBuffers(g.scene)
        world.buildTlas(g.scene.tlas_instances)
        pc = sd.reference_pc(frame_index=1 + int(t * 10))
        ho, margin = oracle.Oracle(g.scene).trace_primary(pc, w, h)
        hg = world.trace_primary(pc, w, h)
        assert (_mismatch(ho, hg) & ~(margin < 1e-4)).sum() == 0, f"t = {t}"
        seen |= set(np.unique(hg["drawInstanceIndex"]).tolist())
    assert len(seen - {sd.MISS}) >= 2  # the camera node looks at the animated quads
    assert world.counters()["stackOverflows"] == 0
    world.close()
    g.close()
