"""prt_compress_texture_static (include/prt.h): the texture-cache level chain with no device allocation -- its
intermediate levels, block staging and filter tables live in one file-scope `static __device__` array.

No counterpart in the reference. The checker is the existing entry point,
prt_compress_texture: same kernels on the same input, so the bytes must be IDENTICAL (the tolerance of every byte-valued
test in this suite is zero). Two calls in a row with different images prove that the array keeping the last call's
contents does not leak into the next result. Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from prosper_b200 import dds, texture_cache as tc


def _image(w, h, seed):
    rng = np.random.default_rng(seed)
    yy, xx = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
    img = np.stack([(xx * 255 // max(w - 1, 1)), (yy * 255 // max(h - 1, 1)), rng.integers(0, 256, (h, w)),
                    np.where((xx // 8 + yy // 8) % 2 == 0, 255, rng.integers(0, 256, (h, w)))], axis=-1)
    return np.ascontiguousarray(img.astype(np.uint8))


def _compress_static(world, rt, img, space, mips=True):
    h, w, _ = img.shape
    levels = tc.mip_level_count(w, h, mips)
    out = dds.Dds(w, h, 1, tc.cache_format(w, h, levels), levels)
    fmt, count = C.c_uint32(), C.c_uint32()
    rt._check(world.lib.prt_compress_texture_static(world.handle, rt._p(img), w, h, space, int(mips), rt._p(out.data), out.data.nbytes,
                                                    C.byref(fmt), C.byref(count)))
    assert count.value == levels
    assert fmt.value == (rt.sd.TEX_BC7 if out.format == dds.FORMAT_BC7_UNORM else rt.sd.TEX_RGBA8)
    return out


@pytest.mark.gpu
def test_static_scratch_entry_point_matches_the_allocating_one_twice_in_a_row():
    from prosper_b200 import rt_reference as rt, scene_data as sd, scene_zoo
    world = rt.World(scene_zoo.test_scene(fmt=sd.TEX_BC7), device=0, max_width=16, max_height=16, samples_per_launch=1)
    try:
        # two calls in a row: a large BC7 chain first, then a smaller one that reuses (and must not see) the first one's levels;
        # then an RGBA8 chain (100 x 60 is not whole blocks), a single level, and the largest image the entry point takes
        cases = [(512, 256, tc.COLOR_SPACE_SRGB, True), (256, 256, tc.COLOR_SPACE_LINEAR, True), (100, 60, tc.COLOR_SPACE_SRGB, True),
                 (64, 64, tc.COLOR_SPACE_SRGB, False), (1024, 1024, tc.COLOR_SPACE_SRGB, True)]
        for n, (w, h, space, mips) in enumerate(cases):
            img = _image(w, h, seed=100 + n)
            got = _compress_static(world, rt, img, space, mips)
            want = tc.compress(world, img, space, generate_mip_maps=mips)
            assert got.format == want.format and got.mipLevelCount == want.mipLevelCount
            assert np.array_equal(got.data, want.data), f"case {n}: {w}x{h} differs from prt_compress_texture"
        # the same image again after other sizes went through the array: still identical
        img = _image(512, 256, seed=100)
        assert np.array_equal(_compress_static(world, rt, img, tc.COLOR_SPACE_SRGB).data, tc.compress(world, img, tc.COLOR_SPACE_SRGB).data)
        # beyond the stated maximum: refused with a message that names the limit, nothing launched
        big = np.zeros((1024, 1028, 4), np.uint8)
        out = np.zeros(big.size * 2, np.uint8)
        fmt, count = C.c_uint32(), C.c_uint32()
        rc = world.lib.prt_compress_texture_static(world.handle, rt._p(big), 1028, 1024, 0, 1, rt._p(out), out.nbytes, C.byref(fmt), C.byref(count))
        assert rc != 0 and b"PRT_STATIC_MAX_TEXELS" in world.lib.prt_last_error()
    finally:
        world.close()
