"""BC7 (texture-cache build, Texture.cpp:257-287: ispc_texcomp's `alpha_ultrafast` profile = modes 4 / 5 / 6).

Decoders first, because everything else is judged through them: the oracle's, the product's (core_bc7.h, host build), the
numpy one and -- on the GPU -- the texture units are each pinned on HAND-ASSEMBLED blocks, one or two per mode, whose fields
are packed here from the format's layout and whose expected texels were worked out by hand from
((64 - w) * e0 + w * e1 + 32) >> 6 (the arithmetic is written out beside each table). Then the encoder
(core_bc7enc.h, modes 4 / 5 / 6 with least-squares refinement): every block decodes, quality is gated against the round-1
bounding-box mode-6 encoder image by image, and the CUDA kernel emits the bits of the host build of the same header."""
import ctypes as C

import numpy as np
import pytest

import oracle
from prosper_b200 import bc7
from tests import hostsim


def _pack(fields):
    """LSB-first bit stream from (value, bit count) pairs -> 16 bytes."""
    bits, pos = 0, 0
    for value, n in fields:
        assert 0 <= value < (1 << n)
        bits |= value << pos
        pos += n
    assert pos == 128, pos
    return np.frombuffer(bits.to_bytes(16, "little"), np.uint8).copy()


def _indices(idx, nbits):
    """16 indices of `nbits` bits; the anchor texel (0) is stored without its top bit, which must be zero."""
    assert idx[0] < (1 << (nbits - 1))
    return [(idx[0], nbits - 1)] + [(v, nbits) for v in idx[1:]]


# ---- mode 6: 0000001 | R0 R1 G0 G1 B0 B1 A0 A1 (7 bits each) | p0 p1 | 16 x 4-bit indices --------------------------------
# endpoints with their p-bit appended: R 0 -> 255, G 254 -> 1, B 64 -> 193, A 254 -> 255; texel i uses index i.
# weights 0 4 9 13 17 21 26 30 34 38 43 47 51 55 60 64, e.g. index 5: R = (43 * 0 + 21 * 255 + 32) >> 6 = 5387 >> 6 = 84,
# G = (43 * 254 + 21 * 1 + 32) >> 6 = 10975 >> 6 = 171, B = (43 * 64 + 21 * 193 + 32) >> 6 = 6837 >> 6 = 106,
# A = (43 * 254 + 21 * 255 + 32) >> 6 = 16309 >> 6 = 254 (255 from index 8 on: 16288 + w >= 16320 needs w >= 32).
MODE6 = _pack([(1 << 6, 7), (0, 7), (127, 7), (127, 7), (0, 7), (32, 7), (96, 7), (127, 7), (127, 7), (0, 1), (1, 1)]
              + _indices(list(range(16)), 4))
MODE6_TEXELS = list(zip([0, 16, 36, 52, 68, 84, 104, 120, 135, 151, 171, 187, 203, 219, 239, 255],
                        [254, 238, 218, 203, 187, 171, 151, 135, 120, 104, 84, 68, 52, 37, 17, 1],
                        [64, 72, 82, 90, 98, 106, 116, 124, 133, 141, 151, 159, 167, 175, 185, 193],
                        [254] * 8 + [255] * 8))

# ---- mode 5: 000001 | rotation 2 | R0 R1 G0 G1 B0 B1 (7 bits) | A0 A1 (8 bits) | 2-bit colour indices | 2-bit alpha indices
# 7 -> 8 bits: (v << 1) | (v >> 6): R 10 -> 20, 100 -> 201; G 127 -> 255 both; B 0 -> 0, 64 -> 129; alpha 200 -> 40.
# weights 0 21 43 64: colour R = 20, (43 * 20 + 21 * 201 + 32) >> 6 = 79, (21 * 20 + 43 * 201 + 32) >> 6 = 142, 201;
# B = 0, 2741 >> 6 = 42, 5579 >> 6 = 87, 129; alpha = 200, 9472 >> 6 = 148, 5952 >> 6 = 93, 40.
# rotation 1 swaps the decoded alpha and red: (r, g, b, a) = (alpha value, 255, B, R value).
_C5 = [0, 1, 2, 3] * 4
_A5 = [0, 0, 1, 1, 2, 2, 3, 3, 3, 3, 2, 2, 1, 1, 0, 0]
MODE5 = _pack([(1 << 5, 6), (1, 2), (10, 7), (100, 7), (127, 7), (127, 7), (0, 7), (64, 7), (200, 8), (40, 8)] + _indices(_C5, 2) + _indices(_A5, 2))
_R5, _B5, _ALPHA5 = [20, 79, 142, 201], [0, 42, 87, 129], [200, 148, 93, 40]
MODE5_TEXELS = [(_ALPHA5[_A5[i]], 255, _B5[_C5[i]], _R5[_C5[i]]) for i in range(16)]

# ---- mode 4: 00001 | rotation 2 | index selection 1 | R0 R1 G0 G1 B0 B1 (5 bits) | A0 A1 (6 bits) | 2-bit indices | 3-bit indices
# 5 -> 8 bits: (v << 3) | (v >> 2): 0 -> 0, 31 -> 255, 16 -> 132; 6 -> 8 bits: (v << 2) | (v >> 4): 0 -> 0, 63 -> 255.
# R 0 -> 255, G 255 -> 0, B 132 -> 132, alpha 0 -> 255. 3-bit weights 0 9 18 27 37 46 55 64 on 0 -> 255:
# 0, 2327 >> 6 = 36, 4622 >> 6 = 72, 6917 >> 6 = 108, 9467 >> 6 = 147, 11762 >> 6 = 183, 14057 >> 6 = 219, 255;
# 2-bit weights on 0 -> 255: 0, 5387 >> 6 = 84, 10997 >> 6 = 171, 255.
_RAMP3, _RAMP2 = [0, 36, 72, 108, 147, 183, 219, 255], [0, 84, 171, 255]
_HEAD4 = [(0, 5), (31, 5), (31, 5), (0, 5), (16, 5), (16, 5), (0, 6), (63, 6)]
# (a) index selection 1: the 3-bit stream drives the colour, the 2-bit stream the alpha; rotation 0
_TWO_A, _THREE_A = [1, 0, 3, 2] * 4, [3, 7, 0, 4, 1, 5, 2, 6] * 2
MODE4_SEL1 = _pack([(1 << 4, 5), (0, 2), (1, 1)] + _HEAD4 + _indices(_TWO_A, 2) + _indices(_THREE_A, 3))
MODE4_SEL1_TEXELS = [(_RAMP3[_THREE_A[i]], _RAMP3[7 - _THREE_A[i]], 132, _RAMP2[_TWO_A[i]]) for i in range(16)]
# (b) index selection 0: 2-bit colour, 3-bit alpha; rotation 2 swaps the decoded alpha and green
_TWO_B, _THREE_B = [0, 1, 2, 3] * 4, [0, 1, 2, 3, 4, 5, 6, 7, 7, 6, 5, 4, 3, 2, 1, 0]
MODE4_SEL0 = _pack([(1 << 4, 5), (2, 2), (0, 1)] + _HEAD4 + _indices(_TWO_B, 2) + _indices(_THREE_B, 3))
MODE4_SEL0_TEXELS = [(_RAMP2[_TWO_B[i]], _RAMP3[_THREE_B[i]], 132, _RAMP2[3 - _TWO_B[i]]) for i in range(16)]

KNOWN_ANSWERS = [("mode 6", MODE6, MODE6_TEXELS), ("mode 5, rotation 1", MODE5, MODE5_TEXELS),
                 ("mode 4, index selection 1", MODE4_SEL1, MODE4_SEL1_TEXELS), ("mode 4, index selection 0, rotation 2", MODE4_SEL0, MODE4_SEL0_TEXELS)]


def _c_decode(fn, block):
    out = np.zeros(64, np.uint8)
    assert fn(block.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p)) == 0
    return [tuple(int(v) for v in out[4 * i:4 * i + 4]) for i in range(16)]


@pytest.mark.parametrize("name,block,texels", KNOWN_ANSWERS, ids=[k[0] for k in KNOWN_ANSWERS])
def test_decoders_on_hand_assembled_blocks(name, block, texels):
    assert bc7._decode_block(block.tobytes()) == texels, "numpy-side decoder"
    assert _c_decode(oracle.lib().orc_bc7_decode_block, block) == texels, "oracle decoder"
    assert _c_decode(hostsim.lib().hs_bc7_decode_block, block) == texels, "product decoder (core_bc7.h, host build)"


def _test_images():
    rng = np.random.default_rng(21)
    n = 64
    yy, xx = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    leaf = np.zeros((n, n, 4), np.uint8)  # green leaf on a transparent ground: the alpha-MASK foliage case
    inside = ((xx - 32) ** 2 / 700.0 + (yy - 32) ** 2 / 260.0) < 1.0
    leaf[..., 0] = 40 + (xx % 7) * 3
    leaf[..., 1] = np.clip(120 + (yy - 32) * 2 + rng.integers(-6, 7, (n, n)), 0, 255)
    leaf[..., 2] = 30
    leaf[..., 3] = np.where(inside, 255, 0)
    albedo = np.stack([128 + 100 * np.sin(xx / 5.0), 128 + 100 * np.cos(yy / 7.0), (xx * yy) % 256, np.full((n, n), 255)], -1)
    albedo = np.clip(albedo + rng.normal(0, 4, albedo.shape), 0, 255).astype(np.uint8)
    noise = rng.integers(0, 256, (n, n, 4), dtype=np.uint8)
    smooth_alpha = np.stack([xx * 4, yy * 4, (xx + yy) * 2, (xx * 3 + yy * 5) % 256], -1).astype(np.uint8)
    flat = np.tile(np.array([17, 130, 251, 77], np.uint8), (n, n, 1))
    return {"leaf": leaf, "albedo": albedo, "noise": noise, "smooth_alpha": smooth_alpha, "flat": flat}


def _psnr(a, b):
    mse = ((a.astype(np.float64) - b.astype(np.float64)) ** 2).mean()
    return 99.0 if mse == 0 else 10 * np.log10(255.0 ** 2 / mse)


def test_encoder_quality_against_the_round_1_encoder_and_mode_use():
    """Image by image the modes-4/5/6 encoder is at least as good as the bounding-box mode-6 encoder (decoded by the numpy
    decoder pinned above), clearly better where alpha is uncorrelated with colour, and it does use all three modes."""
    used = np.zeros(8, np.int64)
    gains = {}
    for name, img in _test_images().items():
        blocks, modes = hostsim.bc7_encode(img, want_modes=True)
        used += np.bincount(modes, minlength=8)
        new = _psnr(bc7.decode_bc7(blocks, 64, 64), img)
        old = _psnr(bc7.decode_bc7(bc7.encode_bc7_mode6(img), 64, 64), img)
        gains[name] = (new, old)
        assert new >= old - 0.01, (name, new, old)
    assert set(np.nonzero(used)[0]) == {4, 5, 6}, used
    assert gains["leaf"][0] > gains["leaf"][1] + 3.0, gains["leaf"]     # separate alpha channel: cut-outs
    assert gains["noise"][0] > gains["noise"][1] + 2.0, gains["noise"]  # least-squares refinement
    assert gains["flat"][0] > 45.0, gains["flat"]                       # a constant block within one code per channel
    assert np.mean([g[0] - g[1] for g in gains.values()]) > 1.5, gains


def test_encoded_blocks_decode_identically_everywhere():
    img = _test_images()["smooth_alpha"]
    blocks = hostsim.bc7_encode(img)
    ref = bc7.decode_bc7(blocks, 64, 64)
    lo, lh = oracle.lib(), hostsim.lib()
    for b in range(0, 256, 5):
        blk = np.ascontiguousarray(blocks[16 * b:16 * b + 16])
        want = [tuple(int(v) for v in t) for t in ref[(b // 16) * 4:(b // 16) * 4 + 4, (b % 16) * 4:(b % 16) * 4 + 4].reshape(16, 4)]
        assert _c_decode(lo.orc_bc7_decode_block, blk) == want
        assert _c_decode(lh.hs_bc7_decode_block, blk) == want


@pytest.mark.gpu
def test_texture_units_decode_the_hand_assembled_blocks():
    """The hardware BC7 decode is the format's own definition: the four known-answer blocks as one 16 x 4 texture."""
    from prosper_b200 import rt_reference as rt, scene_data as sd, scene_zoo
    scene = scene_zoo.test_scene(fmt=sd.TEX_BC7)
    payload = np.concatenate([k[1] for k in KNOWN_ANSWERS])
    scene.textures[1] = sd.Texture(sd.TEX_BC7, 16, 4, 1, payload)
    world = rt.World(scene, device=0, max_width=16, max_height=16, samples_per_launch=1)
    try:
        xy = np.array([[x, y] for y in range(4) for x in range(16)], np.int32)
        got = np.rint(world.probe_texture(1, None, exact=False, texel_xy=xy) * 255.0).astype(int)
        for k, (name, _, texels) in enumerate(KNOWN_ANSWERS):
            for i in range(16):
                at = (i // 4) * 16 + 4 * k + i % 4
                assert tuple(got[at]) == texels[i], (name, i)
    finally:
        world.close()


@pytest.mark.gpu
def test_gpu_encoder_emits_the_bits_of_the_host_build_and_beats_the_old_encoder():
    from prosper_b200 import rt_reference as rt, scene_data as sd, scene_zoo
    world = rt.World(scene_zoo.test_scene(fmt=sd.TEX_BC7), device=0, max_width=16, max_height=16, samples_per_launch=1)
    try:
        for name, img in _test_images().items():
            gpu = world.encode_bc7(img)
            assert np.array_equal(gpu, hostsim.bc7_encode(img)), name
            assert _psnr(bc7.decode_bc7(gpu, 64, 64), img) >= _psnr(bc7.decode_bc7(bc7.encode_bc7_mode6(img), 64, 64), img) - 0.01
    finally:
        world.close()
