"""ReSTIR DI trace pass, oracle side: properties that pin the restatement of rt/direct_illumination/main.rgen,
restir_di/initial_reservoirs.comp and the G-buffer helpers without a GPU."""
import numpy as np

import oracle
from prosper_b200 import scene_data as sd


def _unpack_gbuffer(ar, nm):
    albedo = np.stack([(ar >> s) & 255 for s in (0, 8, 16)], -1) / 255.0
    n = np.stack([nm & 1023, (nm >> 10) & 1023], -1) / 1023.0
    nz = (nm >> 30) / 3.0
    x, y = n[..., 0] - n[..., 1], n[..., 0] + n[..., 1] - 1.0
    z = (nz * 2 - 1) * (1 - np.abs(x) - np.abs(y))
    v = np.stack([x, y, z], -1)
    return albedo, v / np.maximum(np.linalg.norm(v, axis=-1, keepdims=True), 1e-20)


def test_gbuffer_depth_reconstructs_the_hit_point(test_scene):
    """depth = clip z / w of the hit; worldPos(uv, depth) (scene/camera.glsl:24-30) at the pixel centre must land on the primary
    ray, and the decoded octahedral normal must face the camera for front-facing hits."""
    w, h = 96, 64
    orc = oracle.Oracle(test_scene)
    ar, nm, z = orc.rtdi_gbuffer(1, w, h)
    hit = z > 0
    assert 0.5 < hit.mean() < 1.0 and np.all(ar[~hit] == 0) and np.all(nm[~hit] == 0)
    cam = test_scene.camera
    c2w = cam["clipToWorld"].astype(np.float64)  # [col][row]
    ys, xs = np.nonzero(hit)
    uv = np.stack([(xs + 0.5) / w, (ys + 0.5) / h], -1)
    clip = np.concatenate([uv * 2 - 1, z[hit][:, None].astype(np.float64), np.ones((len(xs), 1))], 1)
    world = clip @ c2w  # row vector times [col][row] == column-major matrix times column vector
    p = world[:, :3] / world[:, 3:4]
    eye = cam["eye"][:3].astype(np.float64)
    d = p - eye
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    # the same pixel's pinhole ray direction
    o = np.zeros(3, np.float32)
    dirs = np.zeros((len(xs), 3), np.float32)
    import ctypes as C
    for k in range(0, len(xs), 37):
        dd = np.zeros(3, np.float32)
        oracle.lib().orc_pinhole_ray(cam.reshape(1).ctypes.data_as(C.c_void_p), np.array(uv[k], np.float32).ctypes.data_as(C.c_void_p),
                                     o.ctypes.data_as(C.c_void_p), dd.ctypes.data_as(C.c_void_p))
        assert np.allclose(d[k], dd, atol=2e-4), (k, d[k], dd)
    _, normal = _unpack_gbuffer(ar, nm)
    assert np.allclose(np.linalg.norm(normal[hit], axis=-1), 1.0, atol=1e-6)


def test_restir_estimator_converges_to_the_sum_over_lights(test_scene):
    """Unbiasedness of initial_reservoirs.comp + main.rgen together: the mean over many frames of f(X) * W (fresh reservoirs
    every frame) converges to the direct illumination summed over ALL lights, which the same trace computes exactly when fed
    the reservoirs (i, W = 1) light by light."""
    w, h = 40, 28
    orc = oracle.Oracle(test_scene)
    gb = orc.rtdi_gbuffer(1, w, h)
    lights = 1 + len(test_scene.point_lights) + len(test_scene.spot_lights)
    truth = np.zeros((h, w, 3), np.float64)
    for i in range(lights):
        res = np.zeros((h, w, 2), np.float32)
        res[..., 0] = np.array([i], np.int32).view(np.float32)[0]
        res[..., 1] = 1.0
        img, rays = orc.rtdi_trace(0, 1, sd.FLAG_SKIP_HISTORY, gb, res)
        assert np.all(img[..., 3] == 1)
        truth += img[..., :3]
    frames = 400
    acc = None
    total_rays = 0
    for f in range(1, frames + 1):
        res = orc.rtdi_initial_reservoirs(f, gb)
        acc, rays = orc.rtdi_trace(0, f, sd.FLAG_ACCUMULATE | (sd.FLAG_SKIP_HISTORY if f == 1 else 0), gb, res, rgba=acc)
        total_rays += rays
    assert np.all(acc[..., 3] == frames)
    assert 0 < total_rays <= frames * w * h
    lum = np.array([0.2126, 0.7152, 0.0722])
    lt, la = truth @ lum, acc[..., :3].astype(np.float64) @ lum
    lit = lt > 1e-3
    assert lit.mean() > 0.3
    assert abs(la[lit].mean() - lt[lit].mean()) < 0.03 * lt[lit].mean()  # image mean within 3 % after 400 frames
    assert np.sqrt(np.mean(((la - lt)[lit] / (lt[lit] + 0.05)) ** 2)) < 0.35  # per pixel: Monte Carlo noise, no bias structure
    # an invalid reservoir (index < 0) contributes nothing and traces nothing
    res = np.zeros((h, w, 2), np.float32)
    res[..., 0] = np.array([-1], np.int32).view(np.float32)[0]
    img, rays = orc.rtdi_trace(0, 1, sd.FLAG_SKIP_HISTORY, gb, res)
    assert rays == 0 and not img[..., :3].any()


def test_rtdi_debug_views_and_history(test_scene):
    w, h = 32, 24
    orc = oracle.Oracle(test_scene)
    gb = orc.rtdi_gbuffer(1, w, h)
    albedo, _ = _unpack_gbuffer(gb[0], gb[1])
    res = orc.rtdi_initial_reservoirs(1, gb)
    img, _ = orc.rtdi_trace(8, 1, sd.FLAG_ACCUMULATE, gb, res)  # DrawType_Albedo: "GBuffer pass stores debug output in albedo"
    assert np.allclose(img[..., :3], albedo, atol=1e-6) and np.all(img[..., 3] == 1)
    pos, _ = orc.rtdi_trace(5, 1, sd.FLAG_ACCUMULATE, gb, res)  # DrawType_Position
    assert np.isfinite(pos[gb[2] > 0]).all()
    # !accumulate overwrites, accumulate averages
    a, _ = orc.rtdi_trace(0, 1, sd.FLAG_SKIP_HISTORY | sd.FLAG_ACCUMULATE, gb, res)
    b, _ = orc.rtdi_trace(0, 2, sd.FLAG_ACCUMULATE, gb, orc.rtdi_initial_reservoirs(2, gb), rgba=a.copy())
    assert np.all(b[..., 3] == 2)
    c, _ = orc.rtdi_trace(0, 2, 0, gb, orc.rtdi_initial_reservoirs(2, gb), rgba=a.copy())
    assert np.all(c[..., 3] == 1) and np.allclose(b[..., :3], (a[..., :3] + c[..., :3]) / 2, atol=1e-5)
