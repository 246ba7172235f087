"""Texture-cache build: the .dds container against the byte layout of writeDds / readDds (Dds.cpp),
the level-chain rules of `compress` (Texture.cpp:213-296) and, on the GPU, prt_compress_texture against the numpy mip
filter + the host build of the BC7 encoder core (core_bc7enc.h; quality and decoders: tests/test_bc7.py), bit for bit."""
import struct

import numpy as np
import pytest

from prosper_b200 import bc7, dds, texture_cache as tc


def _image(w, h, seed=0):
    rng = np.random.default_rng(seed)
    yy, xx = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
    img = np.stack([(xx * 255 // max(w - 1, 1)), (yy * 255 // max(h - 1, 1)), rng.integers(0, 256, (h, w)),
                    np.where((xx // 8 + yy // 8) % 2 == 0, 255, rng.integers(0, 256, (h, w)))], axis=-1)
    return img.astype(np.uint8)


def test_level_chain_rules():
    assert tc.mip_level_count(1024, 1024) == 9          # 11 levels down to 1x1, minus 2x2 and 1x1
    assert tc.mip_level_count(1024, 1024, False) == 1
    assert tc.mip_level_count(4, 4) == 1 and tc.mip_level_count(2, 1) == 1
    assert tc.mip_level_count(2048, 512) == 10
    assert tc.cache_format(1024, 1024, 9) == dds.FORMAT_BC7_UNORM     # last kept level is 4x4
    assert tc.cache_format(2048, 512, 10) == dds.FORMAT_R8G8B8A8_UNORM  # 512 >> 8 = 2 is not a whole block
    assert tc.cache_format(100, 60, 5) == dds.FORMAT_R8G8B8A8_UNORM
    d = dds.Dds(64, 32, 1, dds.FORMAT_BC7_UNORM, 4)                     # 64x32, 32x16, 16x8, 8x4
    assert d.levelByteOffsets == [0, 2048, 2560, 2688] and d.data.size == 2720
    with pytest.raises(RuntimeError):
        dds.Dds(64, 32, 1, dds.FORMAT_BC7_UNORM, 5)                      # 4x2 is not a whole block
    r = dds.Dds(6, 3, 1, dds.FORMAT_R8G8B8A8_UNORM, 2)
    assert r.levelByteOffsets == [0, 72] and r.data.size == 72 + 3 * 1 * 4


def test_dds_bytes_match_the_reference_writer(tmp_path):
    d = dds.Dds(16, 8, 1, dds.FORMAT_BC7_UNORM, 2, np.arange(128 + 32, dtype=np.uint8))
    path = str(tmp_path / "t.dds")
    dds.write_dds(d, path)
    blob = open(path, "rb").read()
    assert len(blob) == 148 + 160 and not (tmp_path / "t.dds_TMP").exists()
    words = struct.unpack_from("<37I", blob, 0)
    # magic, dwSize, flags (compressed), height, width, linear size of level 0, depth, mips (Dds.cpp:143-196)
    assert words[:8] == (0x20534444, 124, 0x000A1007, 8, 16, 128, 0, 2)
    assert words[8:19] == (0,) * 11
    assert words[19:27] == (32, 0x4, 0x30315844, 0, 0, 0, 0, 0) and words[27:32] == (0x00401000, 0, 0, 0, 0)
    assert words[32:37] == (98, 3, 0, 1, 0)  # BC7Unorm, Texture2d, arraySize 1 (Dds.cpp:198-203)
    back = dds.read_dds(path)
    assert (back.width, back.height, back.depth, back.format, back.mipLevelCount) == (16, 8, 0, 98, 2)
    assert np.array_equal(back.data, d.data) and np.array_equal(back.level(1), np.arange(128, 160, dtype=np.uint8))
    # uncompressed: pitch = w * h * 4, masks present
    u = dds.Dds(6, 3, 1, dds.FORMAT_R8G8B8A8_UNORM, 1, np.arange(72, dtype=np.uint8))
    dds.write_dds(u, path)
    words = struct.unpack_from("<37I", open(path, "rb").read(), 0)
    assert words[2] == 0x0002100F and words[5] == 72 and words[22:27] == (32, 0xFF, 0xFF00, 0xFF0000, 0xFF000000)
    assert np.array_equal(dds.read_dds(path).data, u.data)


def test_dds_reader_refusals(tmp_path):
    path = str(tmp_path / "t.dds")
    dds.write_dds(dds.Dds(8, 8, 1, dds.FORMAT_BC7_UNORM, 1), path)
    good = open(path, "rb").read()

    def reject(offset, value, match):
        b = bytearray(good)
        struct.pack_into("<I", b, offset, value)
        open(path, "wb").write(b)
        with pytest.raises(RuntimeError, match=match):
            dds.read_dds(path)

    reject(0, 0x20534445, "doesn't appear to be a dds")
    reject(4, 120, "DDS_HEADER size")
    reject(8, 0x1007, "DDS_FLAGS")
    reject(76, 24, "DDS_PIXEL_FORMAT size")
    reject(84, 0x31545844, "Dx10 header")          # 'DXT1'
    reject(108, 0x1000, "DDS_CAPS")
    reject(128, 71, "Only R8G8B8A8Unorm and BC7Unorm")  # BC1
    reject(132, 2, "resource dimension")
    reject(140, 6, "arrays are not supported")
    open(path, "wb").write(good[:-4])
    with pytest.raises(RuntimeError, match="truncated"):
        dds.read_dds(path)
    with pytest.raises(RuntimeError):
        dds.write_dds(dds.Dds(4, 4, 4, dds.FORMAT_R9G9B9E5_SHAREDEXP, 1), path)  # 3D writes are not implemented


def test_numpy_mip_filter_properties():
    flat = np.full((8, 8, 4), 137, np.uint8)
    assert np.array_equal(tc.downsample_mitchell(flat, True), np.full((4, 4, 4), 137, np.uint8))
    assert np.array_equal(tc.downsample_mitchell(flat, False), np.full((4, 4, 4), 137, np.uint8))
    taps = tc.mip_tables()["taps"]
    assert abs(float(taps.sum()) - 1.0) < 1e-6 and taps[0] < 0 and np.allclose(taps, taps[::-1])
    checker = np.zeros((16, 16, 4), np.uint8)
    checker[::2, ::2] = 255
    checker[1::2, 1::2] = 255
    checker[..., 3] = 255
    srgb, lin = tc.downsample_mitchell(checker, True), tc.downsample_mitchell(checker, False)
    # half black, half white: 50 % in stored values is code 128; 50 % in linear light is sRGB code 188
    assert (lin[2:-2, 2:-2, :3] == 128).all() and (srgb[2:-2, 2:-2, :3] == 188).all() and (srgb[..., 3] == 255).all()
    # alpha weighting: fully transparent texels do not bleed their colour into the average
    img = np.zeros((8, 8, 4), np.uint8)
    img[..., 0] = 255                       # red everywhere ...
    img[:, ::2] = [0, 255, 0, 0]            # ... except transparent green columns
    img[:, 1::2, 3] = 255
    out = tc.downsample_mitchell(img, False)
    assert (out[..., 0] == 255).all() and (out[..., 1] == 0).all()
    assert (np.abs(out[:, 1:-1, 3].astype(int) - 128) <= 2).all()  # interior columns; the edge ones see the clamped border
    assert tc.downsample_mitchell(_image(7, 5), False).shape == (2, 3, 4)


@pytest.mark.parametrize("srgb", [True, False])
def test_numpy_mip_filter_against_the_reference_stb(srgb):
    """The restatement against the reference's own vendored stb_image_resize2 (oracle/_ref/libstbir.so, built from
    /root/reference/ext/stb where it lies): same taps, same alpha weighting; stb's float pipeline sums in another order
    and re-encodes sRGB through a table with up to 0.55 code of error, so bytes may differ by one code, rarely."""
    import oracle
    if not oracle.build_ref():
        pytest.skip("oracle/_ref/libstbir.so not built (no /root/reference on this machine)")
    for w, h, seed in ((64, 64, 1), (128, 32, 2), (16, 48, 3)):
        img = _image(w, h, seed)
        img[..., 3] = np.where(img[..., 3] > 128, img[..., 3], img[..., 3] // 2 + 40)   # keep alpha away from 0
        want = oracle.ref_stbir_resize(img, w // 2, h // 2, srgb)
        got = tc.downsample_mitchell(img, srgb)
        d = np.abs(got.astype(int) - want.astype(int))
        assert d.max() <= 1, (w, h, d.max())
        assert (d > 0).mean() < 0.05, (d > 0).mean()
    # with transparent texels (alpha 0) the weighted colour takes over in both
    img = _image(32, 32, 9)
    img[::3, ::2, 3] = 0
    want = oracle.ref_stbir_resize(img, 16, 16, srgb)
    got = tc.downsample_mitchell(img, srgb)
    d = np.abs(got.astype(int) - want.astype(int))
    assert d.max() <= 1 and (d > 0).mean() < 0.05, (d.max(), (d > 0).mean())


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,space", [(256, 256, tc.COLOR_SPACE_SRGB), (256, 128, tc.COLOR_SPACE_SRGB), (64, 64, tc.COLOR_SPACE_LINEAR),
                                       (100, 60, tc.COLOR_SPACE_SRGB), (4, 4, tc.COLOR_SPACE_SRGB)])
def test_gpu_compress_matches_numpy_chain_and_host_build_of_the_encoder(tmp_path, w, h, space):
    from prosper_b200 import rt_reference as rt, scene_data as sd, scene_zoo
    from tests import hostsim
    world = rt.World(scene_zoo.test_scene(fmt=sd.TEX_BC7), device=0, max_width=16, max_height=16, samples_per_launch=1)
    img = _image(w, h, seed=w + h)
    path = str(tmp_path / "cache.dds")
    out = tc.compress_to_file(world, path, img, space)
    levels = tc.mip_level_count(w, h)
    assert out.mipLevelCount == levels and out.format == tc.cache_format(w, h, levels)
    level = img
    for i in range(levels):
        if i:
            level = tc.downsample_mitchell(level, space == tc.COLOR_SPACE_SRGB)
        if out.format == dds.FORMAT_BC7_UNORM:
            want = hostsim.bc7_encode(level)
            assert np.array_equal(out.level(i), np.asarray(want, np.uint8).reshape(-1)), f"level {i}"
        else:
            assert np.array_equal(out.level(i), level.reshape(-1)), f"level {i}"
    back = dds.read_dds(path)
    assert np.array_equal(back.data, out.data) and back.mipLevelCount == levels
    single = tc.compress(world, img, space, generate_mip_maps=False)
    assert single.mipLevelCount == 1 and single.format == tc.cache_format(w, h, 1)
    if single.format == dds.FORMAT_BC7_UNORM:
        assert np.array_equal(single.level(0), hostsim.bc7_encode(img).reshape(-1))
    else:
        assert np.array_equal(single.level(0), img.reshape(-1))
    with pytest.raises(RuntimeError):
        tc.compress(world, img[..., :3])
    world.close()
