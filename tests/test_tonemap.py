"""Display transform (tone_map.comp, render::ToneMap). CPU: the oracle against a float64
evaluation on the reference's own Tony McMapface table (tests/golden/tonemap_tony.json, made by
make_tonemap_golden.py) and against analytic cases; the RGB9E5 codec. GPU: prt_tone_map against the oracle."""
import json
import os

import numpy as np
import pytest

import oracle
from prosper_b200 import tonemap

TONY_DDS_HEADER = (
    "444453207c0000000f1080003000000030000000c00000003000000001000000" + "00" * 44 +
    "2000000004000000445831300000000000000000000000000000000000000000" +
    "0810000000002000000000000000000000000000" + "4300000004000000000000000100000000000000")
GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "tonemap_tony.json")


def test_rgb9e5_codec_round_trip_and_known_words():
    # 1.0 = mantissa 256 at exponent 16 (256 / 512 * 2^1); the largest finite value is 511/512 * 2^16
    assert np.allclose(oracle.decode_rgb9e5((16 << 27) | 256), [1.0, 0.0, 0.0])
    assert np.allclose(oracle.decode_rgb9e5((31 << 27) | (511 << 18)), [0.0, 0.0, 65408.0])
    rng = np.random.default_rng(3)
    rgb = np.exp(rng.uniform(np.log(1e-4), np.log(100.0), (500, 3))).astype(np.float32)
    words = tonemap.pack_rgb9e5(rgb)
    back = tonemap.unpack_rgb9e5(words)
    for w, b in zip(words[:50], back[:50]):
        assert np.array_equal(oracle.decode_rgb9e5(int(w)), b)
    # 9-bit mantissas under a shared exponent: half a step of the largest channel's scale
    step = 2.0 ** (np.floor(np.log2(rgb.max(axis=1))) + 1 - 9)
    assert (np.abs(back - rgb).max(axis=1) <= 0.5 * step * 1.0001).all()
    # re-packing a decoded texel is the identity
    assert np.array_equal(tonemap.pack_rgb9e5(back), words)


def _golden():
    with open(GOLDEN) as f:
        g = json.load(f)
    lut = np.zeros(g["dim"] ** 3, np.uint32)
    for k, v in g["lut_texels"].items():
        lut[int(k)] = v
    colours = np.array(g["colours_f16_bits"], np.uint16).view(np.float16)
    return g, lut, colours


def test_oracle_matches_float64_on_the_reference_table():
    g, lut, colours = _golden()
    img = np.concatenate([colours, np.ones((len(colours), 1), np.float16)], axis=1).reshape(1, -1, 4)
    for case in g["cases"]:
        got = oracle.tone_map(img, case["exposure"], case["contrast"], lut)[0]
        want = np.array(case["display"]) * 255.0
        assert (got[:, 3] == 255).all()
        # fp32 shader arithmetic against float64: never more than one 8-bit step, nearly always the same code
        diff = np.abs(got[:, :3].astype(np.float64) - want)
        assert diff.max() <= 1.0, diff.max()
        assert (np.abs(got[:, :3].astype(np.int64) - np.rint(want).astype(np.int64)) > 0).mean() < 0.02


def test_oracle_analytic_cases():
    dim = 16
    # identity-in-encoded-space table: lut(x) = x, so display = (s / (s + 1)) ^ (1 / 2.2) for greys
    x = np.arange(dim) / (dim - 1)
    b, g, r = np.meshgrid(x, x, x, indexing="ij")
    lut = tonemap.pack_rgb9e5(np.stack([r, g, b], axis=-1)).reshape(-1)
    greys = np.array([0.0, 0.25, 1.0, 4.0, 1000.0], np.float32)
    img = np.stack([greys, greys, greys, np.ones_like(greys)], axis=1).astype(np.float16).reshape(1, -1, 4)
    out = oracle.tone_map(img, 1.0, 1.0, lut)[0]
    want = np.rint((greys / (greys + 1.0)) ** (1 / 2.2) * 255.0)
    assert np.abs(out[:, 0].astype(np.float64) - want).max() <= 1.0  # 9-bit table mantissas
    assert (out[:, 0] == out[:, 1]).all() and (out[:, 1] == out[:, 2]).all()
    # exposure scales the stimulus, contrast is a power on HSV value: grey 0.5 at exposure 2 == grey 1.0
    a = oracle.tone_map(np.array([[[0.5, 0.5, 0.5, 1]]], np.float16), 2.0, 1.0, lut)
    c = oracle.tone_map(np.array([[[1.0, 1.0, 1.0, 1]]], np.float16), 1.0, 1.0, lut)
    assert np.array_equal(a, c)
    d = oracle.tone_map(np.array([[[0.25, 0.25, 0.25, 1]]], np.float16), 1.0, 0.5, lut)  # 0.25 ^ 0.5 = 0.5
    e = oracle.tone_map(np.array([[[0.5, 0.5, 0.5, 1]]], np.float16), 1.0, 1.0, lut)
    assert np.array_equal(d, e)
    # hue is preserved by the HSV round trip: a pure red stays pure red through an identity table
    red = oracle.tone_map(np.array([[[2.0, 0.0, 0.0, 1]]], np.float16), 1.0, 1.7, lut)[0, 0]
    assert red[0] > 0 and red[1] == 0 and red[2] == 0
    # NaN / negative / inf inputs produce defined bytes
    odd = np.array([[[np.nan, 1, 1, 1], [-1, -2, -3, 1], [np.inf, 0, 0, 1]]], np.float16)
    assert oracle.tone_map(odd, 1.0, 1.0, lut).shape == (1, 3, 4)


def test_stand_in_lut_is_monotone_on_greys_and_dds_reader(tmp_path):
    lut = tonemap.stand_in_lut(48)
    rgb = tonemap.unpack_rgb9e5(lut).reshape(48, 48, 48, 3)
    diag = rgb[np.arange(48), np.arange(48), np.arange(48)]
    assert (np.diff(diag[:, 0]) >= 0).all() and diag[0].max() == 0 and diag[-1].min() > 0.95
    # the 148 header bytes of the reference's res/texture/tony_mc_mapface.dds (48^3 volume, DXGI format 67)
    hdr = bytes.fromhex(TONY_DDS_HEADER)
    path = tmp_path / "lut.dds"
    path.write_bytes(bytes(hdr) + lut.tobytes())
    assert np.array_equal(tonemap.read_lut_dds(str(path)), lut)
    tonemap.write_ppm(str(tmp_path / "x.ppm"), np.zeros((2, 3, 4), np.uint8))
    assert (tmp_path / "x.ppm").read_bytes().startswith(b"P6\n3 2\n255\n")


@pytest.mark.gpu
@pytest.mark.parametrize("exposure,contrast", [(1.0, 1.0), (0.4, 1.0), (3.0, 1.4)])
def test_gpu_tone_map_matches_oracle(exposure, contrast):
    from prosper_b200 import rt_reference as rt, scene_data as sd, scene_zoo
    scene = scene_zoo.test_scene(fmt=sd.TEX_BC7)
    w, h = 160, 96
    world = rt.World(scene, device=0, max_width=w, max_height=h, samples_per_launch=4)
    world.updateBuffers()
    world.buildAccelerationStructures()
    r = rt.RtReference()
    r.init(world)
    out = r.record(world, (0, 0, w, h), rt.Options(ibl=True), frames=8)
    tm = rt.ToneMap()
    with pytest.raises(RuntimeError):
        tm.record(world, out)  # before init, like WHEELS_ASSERT(m_initialized)
    lut = tonemap.stand_in_lut(48)
    tm.init(world, lut)
    tm.drawUi(exposure=exposure, contrast=contrast)
    ldr = tm.record(world, out)
    want = oracle.tone_map(out.illumination(), exposure, contrast, lut)
    world.close()
    assert ldr.shape == want.shape == (h, w, 4)
    d = np.abs(ldr.astype(np.int32) - want.astype(np.int32))
    # same arithmetic on both sides up to powf rounding and FMA contraction: a byte may tip over a rounding boundary
    assert d.max() <= 1, d.max()
    assert (d > 0).mean() < 0.005, (d > 0).mean()
    assert ldr[..., :3].std() > 10  # an actual picture, not a constant


@pytest.mark.gpu
def test_gpu_tone_map_error_paths_and_golden_colours():
    from prosper_b200 import rt_reference as rt, scene_data as sd, scene_zoo
    g, lut, colours = _golden()
    scene = scene_zoo.test_scene(fmt=sd.TEX_BC7)
    world = rt.World(scene, device=0, max_width=16, max_height=5, samples_per_launch=1)
    with pytest.raises(RuntimeError):
        world.set_tone_map_lut(np.zeros(10, np.uint32))  # not a cube
    world.set_tone_map_lut(lut)
    with pytest.raises(RuntimeError):
        world._accum_pixels = 80
        world.tone_map(1.0, 1.0)  # nothing recorded yet
    # feed the golden colours through the sum import (count 1 -> the running mean holds them exactly)
    import torch
    img = np.concatenate([colours.astype(np.float32), np.ones((80, 1), np.float32)], axis=1)
    world.updateBuffers()
    world.buildAccelerationStructures()
    r = rt.RtReference()
    r.init(world)
    r.record(world, (0, 0, 16, 5), rt.Options(), frames=1)
    t = torch.from_numpy(img).cuda()
    torch.cuda.synchronize()
    world.import_sum(t.data_ptr())
    world.synchronize()
    for case in g["cases"]:
        got = world.tone_map(case["exposure"], case["contrast"]).reshape(80, 4)
        want = np.array(case["display"]) * 255.0
        assert np.abs(got[:, :3].astype(np.float64) - want).max() <= 1.0
    world.close()
