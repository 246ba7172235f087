"""KTX 1.1 environment-map reader (utils::readKtx, Ktx.cpp:49-166): layout, level indexing and the
refusals of the reference; on the GPU, a cubemap read from a file drives the miss shader's sky lookup."""
import struct

import numpy as np
import pytest

from prosper_b200 import ktx


def _cube(n, mips, seed=0):
    rng = np.random.default_rng(seed)
    return [rng.uniform(0, 4, (6, max(n >> m, 1), max(n >> m, 1), 4)).astype(np.float16) for m in range(mips)]


def test_round_trip_levels_faces_and_offsets(tmp_path):
    levels = _cube(16, 5)
    path = str(tmp_path / "env.ktx")
    ktx.write_ktx(path, levels, key_value=b"KTXorientation\0S=r,T=d\0")
    k = ktx.read_ktx(path)
    assert (k.width, k.height, k.depth, k.arrayLayerCount, k.faceCount, k.mipLevelCount) == (16, 16, 1, 1, 6, 5)
    assert len(k.levelByteOffsets) == 5 * 6
    # faces of a level are contiguous and tightly packed; levels follow each other
    assert k.levelByteOffsets[:7] == [0, 2048, 4096, 6144, 8192, 10240, 12288]
    for m in range(5):
        for f in range(6):
            assert np.array_equal(k.level(m, 0, f), levels[m][f])
    assert np.array_equal(k.cube_faces(), levels[0])
    assert k.data.size == sum(lv.nbytes for lv in levels)


def test_header_bytes_follow_the_ktx11_layout(tmp_path):
    path = str(tmp_path / "one.ktx")
    ktx.write_ktx(path, [np.ones((1, 2, 3, 4), np.float16)])
    blob = open(path, "rb").read()
    assert blob[:12] == b"\xabKTX 11\xbb\r\n\x1a\n"
    h = struct.unpack_from("<13I", blob, 12)
    assert h == (0x04030201, 0x140B, 2, 0x1908, 0x881A, 0x1908, 3, 2, 0, 0, 1, 1, 0)
    assert struct.unpack_from("<I", blob, 64)[0] == 2 * 3 * 8 and len(blob) == 68 + 48
    k = ktx.read_ktx(path)
    assert (k.width, k.height, k.faceCount) == (3, 2, 1) and k.level(0).shape == (2, 3, 4)
    with pytest.raises(RuntimeError):
        k.cube_faces()


def test_refusals(tmp_path):
    path = str(tmp_path / "bad.ktx")
    good = tmp_path / "good.ktx"
    ktx.write_ktx(str(good), _cube(4, 1))
    blob = bytearray(good.read_bytes())

    def reject(mutate, match):
        b = bytearray(blob)
        mutate(b)
        open(path, "wb").write(b)
        with pytest.raises(RuntimeError, match=match):
            ktx.read_ktx(path)

    reject(lambda b: b.__setitem__(slice(0, 12), ktx.IDENT_20), "KTX 2.0 is not supported")
    reject(lambda b: b.__setitem__(0, 0), "doesn't appear to be a KTX")
    reject(lambda b: struct.pack_into("<I", b, 12, 0x01020304), "endianness")
    reject(lambda b: struct.pack_into("<I", b, 16, 0x1406), "Only RGBA16F")       # GL_FLOAT
    reject(lambda b: struct.pack_into("<I", b, 28, 0x8814), "Only RGBA16F")       # GL_RGBA32F
    reject(lambda b: struct.pack_into("<I", b, 36, 0), "width")
    reject(lambda b: b.__delitem__(slice(len(b) - 8, len(b))), "truncated")
    k = ktx.read_ktx(str(good))
    with pytest.raises(RuntimeError, match="512"):
        ktx.check_skybox(k)
    big = ktx.Ktx(512, 512, 1, 1, 6, 4, np.zeros(0, np.uint8), [])
    with pytest.raises(RuntimeError, match="more than 4"):
        ktx.check_skybox(big)
    big.mipLevelCount = 10
    ktx.check_skybox(big)


@pytest.mark.gpu
def test_gpu_sky_from_ktx_file_matches_oracle(tmp_path):
    """A cubemap written to disk, read back with read_ktx and handed to prt_set_sky_cubemap: every primary ray of an
    empty-ish view misses into it, so the debug-free 0-bounce image is the sky lookup (scene.rmiss) of the oracle."""
    import oracle
    from prosper_b200 import rt_reference as rt, scene_data as sd, scene_zoo
    n = 32
    f = np.arange(6, dtype=np.float32)[:, None, None]
    yy, xx = np.meshgrid(np.arange(n, dtype=np.float32), np.arange(n, dtype=np.float32), indexing="ij")
    faces = np.stack([0.2 + 0.1 * f + 0 * xx, xx[None] / n + 0 * f, yy[None] / n + 0 * f, 1 + 0 * f + 0 * xx], axis=-1)
    path = str(tmp_path / "sky.ktx")
    ktx.write_ktx(path, [faces.astype(np.float16)])
    cube = ktx.read_ktx(path).cube_faces()
    scene = scene_zoo.test_scene(fmt=sd.TEX_BC7)
    scene.sky = cube
    scene.sky_face_size = n
    w, h = 96, 64
    world = rt.World(scene, device=0, max_width=w, max_height=h, samples_per_launch=2)
    world.updateBuffers()
    world.buildAccelerationStructures()
    r = rt.RtReference()
    r.init(world)
    r.drawUi(maxBounces=1)
    out = r.record(world, (0, 0, w, h), rt.Options(ibl=True), frames=4).illumination_f32()
    world.close()
    ref, _ = oracle.Oracle(scene).render(sd.reference_pc(max_bounces=1), w, h, 4)
    miss = (oracle.Oracle(scene).trace_primary(sd.reference_pc(), w, h)[0]["drawInstanceIndex"] == sd.MISS).reshape(h, w)
    inner = miss.copy()  # jittered samples of a pixel next to geometry may hit it: keep pixels surrounded by misses
    for dy in (-1, 0, 1):
        for dx in (-1, 0, 1):
            inner &= np.roll(np.roll(miss, dy, 0), dx, 1)
    miss = inner
    assert miss.sum() > 100
    # pixels whose primary ray misses are pure sky lookups: the same texels, the same filter
    assert np.abs(out[miss][:, :3] - ref[miss][:, :3]).max() < 2e-3
