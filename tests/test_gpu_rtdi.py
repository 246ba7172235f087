"""-m gpu: the ReSTIR DI trace pass through the C ABI against the oracle's restatement of
rt/direct_illumination/main.rgen and restir_di/initial_reservoirs.comp."""
import numpy as np
import pytest

import oracle
from prosper_b200 import rt_reference as rt, rtdi, scene_data as sd

pytestmark = pytest.mark.gpu
LUM = np.array([0.2126, 0.7152, 0.0722], np.float32)


def _world(scene, w, h, flags=rt.CFG_COUNTERS | rt.CFG_EXACT_FILTER):
    world = rt.World(scene, max_width=w, max_height=h, samples_per_launch=1, flags=flags)
    world.updateBuffers()
    world.buildAccelerationStructures()
    return world


def _codes(packed, shifts, mask):
    return np.stack([(packed >> s) & mask for s in shifts], -1).astype(np.int64)


@pytest.mark.parametrize("scene_name", ["test_scene", "small_atrium_instanced"])
def test_gbuffer_from_primary_hits_matches_oracle(scene_name, request):
    """The G-buffer a raster pass would write (gbuffer.frag:57-110), from the tracer's primary hits: coverage identical,
    albedo / roughness within 1 of 255 codes, encoded normal / metallic within 2 of 1023, depth to 1e-5 relative."""
    scene = request.getfixturevalue(scene_name)
    w, h = 160, 96
    world, orc = _world(scene, w, h), oracle.Oracle(scene)
    gb = rtdi.GBuffer.from_primary_hits(world, w, h, frame_index=3)
    ar, nm, z = orc.rtdi_gbuffer(3, w, h)
    covered = z > 0
    differ = (gb.depth > 0) != covered
    assert differ.mean() < 1e-3, "coverage differs only where an alpha decision ties"
    both = covered & (gb.depth > 0)
    assert both.mean() > 0.5
    d8 = np.abs(_codes(gb.albedoRoughness, (0, 8, 16, 24), 255) - _codes(ar, (0, 8, 16, 24), 255))[both]
    assert (d8 > 1).mean() < 2e-3
    d10 = np.abs(_codes(gb.normalMetalness, (0, 10, 20), 1023) - _codes(nm, (0, 10, 20), 1023))[both]
    assert (d10 > 2).mean() < 2e-3
    assert (((gb.normalMetalness >> 30) != (nm >> 30)) & both).mean() < 2e-3
    assert np.allclose(gb.depth[both], z[both], rtol=1e-5, atol=1e-7)
    world.close()


def test_initial_reservoirs_and_trace_match_oracle(small_atrium_instanced):
    """restir_di/initial_reservoirs.comp and rt/direct_illumination/main.rgen on identical inputs (the G-buffer the GPU
    produced, read back): the chosen light agrees on > 99.9 % of the pixels (a resampling decision within an ulp of its
    threshold may flip), its weight to 1e-4; three accumulated trace frames agree to relRMSE < 0.5 % on luminance with the
    shadow-ray count within 0.1 % (sun + 32 lights, alpha-masked foliage in front of them)."""
    scene = small_atrium_instanced
    w, h = 160, 96
    world, orc = _world(scene, w, h), oracle.Oracle(scene)
    gb = rtdi.GBuffer.from_primary_hits(world, w, h, frame_index=1)
    init, trace = rtdi.InitialReservoirs(), rtdi.Trace()
    trace.init(world)
    assert trace.recompileShaders()
    ref = None
    ref_rays = 0
    world.reset_counters()
    for frame in (1, 2, 3):
        res_gpu = init.record(world, w, h)
        res_cpu = orc.rtdi_initial_reservoirs(frame, gb.arrays())
        idx_gpu, idx_cpu = res_gpu[..., 0].view(np.int32), res_cpu[..., 0].view(np.int32)
        same = idx_gpu == idx_cpu
        assert same.mean() > 0.999, f"frame {frame}: {(~same).sum()} reservoirs picked another light"
        valid = same & (idx_cpu >= 0)
        assert valid.mean() > 0.3
        assert np.allclose(res_gpu[..., 1][valid], res_cpu[..., 1][valid], rtol=1e-4, atol=1e-6)
        out = trace.record(world, w, h, rtdi.TraceInput(gbuffer=None, reservoirs=None))
        # the oracle consumes the GPU's reservoirs so that the trace is compared on identical inputs
        ref, rays = orc.rtdi_trace(0, frame, sd.FLAG_ACCUMULATE | (sd.FLAG_SKIP_HISTORY if frame == 1 else 0), gb.arrays(), res_gpu, rgba=ref)
        ref_rays += rays
    img = out.illumination_f32()
    assert np.all(img[..., 3] == 3)
    la, lb = img[..., :3] @ LUM, ref[..., :3] @ LUM
    assert np.sqrt(np.mean(((la - lb) / (np.abs(lb) + 1e-2)) ** 2)) < 5e-3
    assert abs(float(la.mean() - lb.mean())) < 1e-3 * float(lb.mean())
    c = world.counters()
    assert abs(c["shadowRays"] - ref_rays) <= 1e-3 * ref_rays and c["stackOverflows"] == 0 and c["anyHitCalls"] > 0
    f16 = out.illumination()
    assert np.array_equal(f16, img.astype(np.float16))
    world.close()


def test_trace_interface_semantics(test_scene):
    """Trace::record's host behaviour: caller-supplied G-buffer and reservoirs (Trace::Input), resetAccumulation and camera
    moves restart the history (Trace.cpp:255-263), debug draw types bypass the reservoirs, a new extent recreates the history,
    and the calls fail loudly out of order."""
    w, h = 64, 48
    world, orc = _world(test_scene, w, h), oracle.Oracle(test_scene)
    fresh = rtdi.Trace()
    fresh.init(world)
    with pytest.raises(RuntimeError, match="no G-buffer yet"):
        fresh.record(world, w, h, rtdi.TraceInput(None, None))
    ar, nm, z = orc.rtdi_gbuffer(1, w, h)  # "raster" G-buffer from the caller
    res = orc.rtdi_initial_reservoirs(1, (ar, nm, z))
    t = rtdi.Trace()
    t.init(world)
    with pytest.raises(RuntimeError, match="already initialized"):
        t.init(world)
    a = t.record(world, w, h, rtdi.TraceInput(rtdi.GBuffer(ar, nm, z), res)).illumination_f32()
    ref, _ = orc.rtdi_trace(0, 1, sd.FLAG_SKIP_HISTORY | sd.FLAG_ACCUMULATE, (ar, nm, z), res)
    assert np.all(a[..., 3] == 1)
    assert np.allclose(a[..., :3], ref[..., :3], rtol=2e-3, atol=2e-4)
    b = t.record(world, w, h, rtdi.TraceInput(None, None)).illumination_f32()
    assert np.all(b[..., 3] == 2)
    c = t.record(world, w, h, rtdi.TraceInput(None, None), resetAccumulation=True).illumination_f32()
    assert np.all(c[..., 3] == 1)
    d = t.record(world, w, h, rtdi.TraceInput(None, None), cameraChanged=True).illumination_f32()
    assert np.all(d[..., 3] == 1)
    alb = t.record(world, w, h, rtdi.TraceInput(None, None), drawType=8).illumination_f32()  # DrawType_Albedo
    want = np.stack([(ar >> s) & 255 for s in (0, 8, 16)], -1) / 255.0
    assert np.allclose(alb[..., :3], want, atol=1e-6) and np.all(alb[..., 3] == 1)
    t.releasePreserved(world)
    e = t.record(world, w, h, rtdi.TraceInput(None, None)).illumination_f32()
    assert np.all(e[..., 3] == 1)
    with pytest.raises(RuntimeError, match="extent"):
        t.record(world, w, h, rtdi.TraceInput(rtdi.GBuffer(ar[:10], nm[:10], z[:10]), None))
    # a smaller extent recreates the pass's images: the reservoirs must be supplied again
    gb2 = orc.rtdi_gbuffer(1, 32, 24)
    with pytest.raises(RuntimeError, match="no reservoirs yet"):
        t.record(world, 32, 24, rtdi.TraceInput(rtdi.GBuffer(*gb2), None))
    world.close()
