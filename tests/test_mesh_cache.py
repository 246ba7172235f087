"""The .prosper_mesh container (prosper_b200/mesh_cache.py) against the data contract and the oracle: CPU only."""
import dataclasses
import struct

import numpy as np
import pytest

from oracle import Oracle
from prosper_b200 import mesh_cache as mc, scene_data as sd


def _quad_mesh(n=5, seed=0):
    rng = np.random.default_rng(seed)
    g = np.linspace(-1, 1, n, dtype=np.float32)
    x, y = np.meshgrid(g, g)
    pos = np.stack([x.ravel(), y.ravel(), rng.uniform(-0.1, 0.1, n * n).astype(np.float32)], 1)
    nrm = np.tile(np.array([[0, 0, 1]], np.float32), (n * n, 1))
    uv = np.stack([(x.ravel() + 1) / 2, (y.ravel() + 1) / 2], 1).astype(np.float32)
    tan = np.tile(np.array([[1, 0, 0, 1]], np.float32), (n * n, 1))
    idx = []
    for j in range(n - 1):
        for i in range(n - 1):
            a = j * n + i
            idx += [a, a + 1, a + n, a + 1, a + n + 1, a + n]
    return pos, nrm, uv, np.array(idx, np.uint32), tan


def test_header_layout_and_magic():
    assert struct.pack("<Q", mc.MESH_CACHE_MAGIC) == b"PRSPRMSH"  # DeferredLoadingContext.cpp:24 comment
    assert mc._HEADER.size == 8 + 13 * 4 and mc._PREFIX.size == 12
    assert [f.name for f in dataclasses.fields(mc.MeshCacheHeader)][:4] == ["sourceWriteTime", "indexCount", "vertexCount", "meshletCount"]
    assert mc.cache_path("/x", 7).endswith("prosper_cache/cache7.prosper_mesh")


@pytest.mark.parametrize("with_tangents,odd_indices", [(True, False), (False, True)])
def test_round_trip_matches_geometry_builder(tmp_path, with_tangents, odd_indices):
    pos, nrm, uv, idx, tan = _quad_mesh()
    if odd_indices:
        idx = idx[:-3]  # 45 indices: the u16 section needs its 2 bytes of padding
        assert idx.size % 2 == 1
    tangents = tan if with_tangents else None
    ref = sd.GeometryBuilder(0)
    ref.add_mesh(pos * 2, nrm, uv, idx, 3, tangents=tangents)  # something in front, so offsets are rebased
    ref.add_mesh(pos, nrm, uv, idx, 1, tangents=tangents)
    buf, mds, infos = ref.finish()

    cache = mc.GeometryCache()
    for m, (p, mat) in enumerate(((pos * 2, 3), (pos, 1))):
        h, blob = mc.pack_mesh(p, nrm, uv, idx, tangents=tangents)
        path = mc.write_cache(str(tmp_path), 1234567890123, m, h, blob)
        assert mc.cache_valid(path, 1234567890123) and not mc.cache_valid(path, 1234567890124)
        h2, blob2 = mc.read_cache(path)
        assert h2 == h and blob2 == blob and h2.meshletCount == 0
        assert h2.tangentsOffset == (mc.UNUSED if tangents is None else h2.normalsOffset + len(p))
        cache.upload(h2, blob2, mat)
    bufs, mds2, infos2 = cache.finish()
    assert len(bufs) == 1 and np.array_equal(bufs[0], buf)
    assert np.array_equal(infos2, infos)
    for f in ("bufferIndex", "indicesOffset", "positionsOffset", "normalsOffset", "tangentsOffset", "texCoord0sOffset", "usesShortIndices"):
        assert np.array_equal(mds2[f], mds[f]), f
    # and back out of the packed buffer
    h3, blob3 = mc.extract_mesh(buf, mds[1], infos[1])
    h_direct, blob_direct = mc.pack_mesh(pos, nrm, uv, idx, tangents=tangents)
    assert h3 == h_direct and blob3 == blob_direct


def test_u32_indices_and_refusals(tmp_path):
    n = 260  # 67600 vertices > 65535: u32 indices
    pos, nrm, uv, idx, _ = _quad_mesh(n)
    h, blob = mc.pack_mesh(pos, nrm, uv, idx)
    assert h.usesShortIndices == 0 and h.positionsOffset == idx.size and h.blobByteCount == len(blob)
    path = mc.write_cache(str(tmp_path), 5, 0, h, blob)
    assert mc.read_cache(path)[0].vertexCount == n * n
    assert mc.read_cache(str(tmp_path / "missing.prosper_mesh")) == (None, None)
    raw = bytearray(open(path, "rb").read())
    old = bytes(raw)
    raw[8:12] = struct.pack("=I", mc.MESH_CACHE_VERSION - 1)  # "Old cache data version": not an error, just a miss
    open(path, "wb").write(raw)
    assert mc.read_cache(path) == (None, None) and not mc.cache_valid(path, 5)
    open(path, "wb").write(b"NOTACACH" + old[8:])
    with pytest.raises(RuntimeError, match="Expected a valid mesh cache"):
        mc.read_cache(path)
    with pytest.raises(ValueError):
        mc.write_cache(str(tmp_path), 5, 1, h, blob[:-4])
    with pytest.raises(ValueError):
        mc.pack_mesh(pos, nrm[:-1], uv, idx)


def test_scene_rebuilt_from_cache_files_traces_identically(tmp_path, test_scene):
    """Every mesh of a scene written out as cache files, read back and uploaded in REVERSE order (all offsets and mesh
    indices change): the oracle's primary hit records must not."""
    scene = test_scene
    n = len(scene.mesh_infos)
    for m in range(n):
        h, blob = mc.extract_mesh(scene.geometry_buffers[int(scene.metadatas[m]["bufferIndex"])], scene.metadatas[m], scene.mesh_infos[m])
        mc.write_cache(str(tmp_path), 42, m, h, blob)
    cache = mc.GeometryCache()
    for m in reversed(range(n)):
        path = mc.cache_path(str(tmp_path), m)
        assert mc.cache_valid(path, 42)
        h, blob = mc.read_cache(path)
        cache.upload(h, blob, int(scene.mesh_infos[m]["materialIndex"]))
    bufs, mds, infos = cache.finish()
    sub = scene.sub_models.copy()
    sub["meshIndex"] = n - 1 - sub["meshIndex"]
    draws = scene.draw_instances.copy()
    draws["meshIndex"] = n - 1 - draws["meshIndex"]
    rebuilt = dataclasses.replace(scene, geometry_buffers=bufs, metadatas=mds, mesh_infos=infos, sub_models=sub, draw_instances=draws)
    pc = sd.reference_pc(frame_index=1)
    a, _ = Oracle(scene).trace_primary(pc, 96, 64)
    b, _ = Oracle(rebuilt).trace_primary(pc, 96, 64)
    assert np.array_equal(a, b)
    img_a, _ = Oracle(scene).render(sd.reference_pc(max_bounces=4), 48, 32, 2)
    img_b, _ = Oracle(rebuilt).render(sd.reference_pc(max_bounces=4), 48, 32, 2)
    assert np.array_equal(img_a, img_b)  # shading fetches (normals, tangents, uvs) go through the rebased offsets too
