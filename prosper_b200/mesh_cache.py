"""The `.prosper_mesh` cache container: host mirror of writeCache / readCache / cacheValid / uploadGeometryData
(src/scene/DeferredLoadingContext.cpp:492-499, 504-558, 560-574, 609-800, 1192-1269; header struct
DeferredLoadingContext.hpp:60-78).

A cache file is: u64 magic "PRSPRMSH", u32 version 4, the MeshCacheHeader fields in declaration order (the 8-byte
source write time first, then thirteen u32), then one blob: indices (u16 padded to 4 bytes when the mesh has at most
65535 vertices, else u32), positions fp16x4, normals snorm 3x10, tangents snorm 3x10 + sign, uvs fp16x2, meshlets,
meshlet bounds, meshlet vertices, meshlet triangles. Header offsets are in u32 units from the blob start (indices at
0; meshlet vertices in index units; meshlet triangles in bytes). The path tracer only reads the first five sections
(SURVEY.md 8 a26), so a cache written here carries zero meshlets (the meshoptimizer clustering feeds the raster
passes, out of scope) while a cache written by prosper, meshlets included, loads unchanged: `GeometryCache.upload`
copies the whole blob and rebases every offset exactly like uploadGeometryData.

Caches are not portable by the reference's own note (DeferredLoadingContext.cpp:755-757): native endianness, and the
write time is the raw 8-byte `std::filesystem::file_time_type` count, kept here as an opaque int64.
"""
from __future__ import annotations

import os
import struct
from dataclasses import dataclass
from typing import List, Optional, Tuple

import numpy as np

from . import scene_data as sd

MESH_CACHE_MAGIC = 0x4853_4D52_5053_5250  # "PRSPRMSH" (DeferredLoadingContext.cpp:24)
MESH_CACHE_VERSION = 4                     # DeferredLoadingContext.cpp:27
GEOMETRY_BUFFER_SIZE = 64 * 1024 * 1024    # sGeometryBufferSize (DeferredLoadingContext.cpp:22)
UNUSED = 0xFFFFFFFF
_HEADER = struct.Struct("=q13I")           # sourceWriteTime + 13 x u32, written field by field (no padding)
_PREFIX = struct.Struct("=QI")


@dataclass
class MeshCacheHeader:
    """DeferredLoadingContext.hpp:60-78 (same names, same defaults)."""
    sourceWriteTime: int = 0
    indexCount: int = 0
    vertexCount: int = 0
    meshletCount: int = 0
    positionsOffset: int = UNUSED
    normalsOffset: int = UNUSED
    tangentsOffset: int = UNUSED
    texCoord0sOffset: int = UNUSED
    meshletsOffset: int = UNUSED
    meshletBoundsOffset: int = UNUSED
    meshletVerticesOffset: int = UNUSED
    meshletTrianglesByteOffset: int = UNUSED
    usesShortIndices: int = 0
    blobByteCount: int = 0

    def pack(self) -> bytes:
        return _HEADER.pack(self.sourceWriteTime, self.indexCount, self.vertexCount, self.meshletCount, self.positionsOffset,
                            self.normalsOffset, self.tangentsOffset, self.texCoord0sOffset, self.meshletsOffset,
                            self.meshletBoundsOffset, self.meshletVerticesOffset, self.meshletTrianglesByteOffset,
                            self.usesShortIndices, self.blobByteCount)


def cache_path(scene_dir: str, mesh_index: int) -> str:
    """getCachePath (DeferredLoadingContext.cpp:492-499)."""
    return os.path.join(scene_dir, "prosper_cache", f"cache{mesh_index}.prosper_mesh")


def pack_mesh(positions, normals, uvs, indices, tangents=None, source_write_time: int = 0) -> Tuple[MeshCacheHeader, bytes]:
    """The packing half of writeCache (DeferredLoadingContext.cpp:609-738) on already de-duplicated vertex arrays:
    returns the header and the blob. Vertex formats are the ones packMeshData produces (fp16x4 positions with w = 1,
    snorm10 normals, snorm10 + sign tangents, fp16x2 uvs) -- the same bytes GeometryBuilder.add_mesh emits."""
    positions = np.asarray(positions, dtype=np.float32)
    normals = np.asarray(normals, dtype=np.float32)
    indices = np.asarray(indices, dtype=np.uint32).reshape(-1)
    nv = positions.shape[0]
    if normals.shape[0] != nv:
        raise ValueError("normals.size() == meshInfo.vertexCount")
    if tangents is not None and np.asarray(tangents).shape[0] != nv:
        raise ValueError("tangents.size() == meshInfo.vertexCount || tangents.empty()")
    if uvs is not None and np.asarray(uvs).shape[0] != nv:
        raise ValueError("texCoord0s.size() == meshInfo.vertexCount || texCoord0s.empty()")
    if indices.size and int(indices.max()) >= nv:
        raise ValueError("index out of range")
    short = nv <= 0xFFFF
    if short:
        idx = indices.astype(np.uint16)
        if idx.size % 2:  # "Let's pad to 4byte boundary to make things simpler later"
            idx = np.concatenate([idx, np.zeros(1, np.uint16)])
        sections: List[bytes] = [idx.tobytes()]
    else:
        sections = [indices.tobytes()]
    byte_count = len(sections[0])

    def push(raw: bytes) -> int:
        nonlocal byte_count
        offset = byte_count
        byte_count += len(raw)
        if byte_count % 4:
            raise ValueError("Mesh data is not aligned properly")
        sections.append(raw)
        return offset

    pos4 = np.concatenate([positions, np.ones((nv, 1), np.float32)], axis=1).astype(np.float16)
    h = MeshCacheHeader(sourceWriteTime=int(source_write_time), indexCount=int(indices.size), vertexCount=int(nv),
                        usesShortIndices=1 if short else 0)
    h.positionsOffset = push(pos4.tobytes()) // 4
    h.normalsOffset = push(sd.pack_snorm3x10_1x2(np.concatenate([normals, np.zeros((nv, 1), np.float32)], 1)).astype(np.uint32).tobytes()) // 4
    t_off = push(sd.pack_snorm3x10_1x2(np.asarray(tangents, np.float32)).astype(np.uint32).tobytes() if tangents is not None else b"") // 4
    uv_off = push(np.asarray(uvs, np.float32).astype(np.float16).tobytes() if uvs is not None else b"") // 4
    h.tangentsOffset = t_off if tangents is not None else UNUSED
    h.texCoord0sOffset = uv_off if uvs is not None else UNUSED
    # no meshlets: the four sections are empty, their offsets are where they would start (as computeOffset gives)
    h.meshletsOffset = push(b"") // 4
    h.meshletBoundsOffset = push(b"") // 4
    h.meshletVerticesOffset = push(b"") // (2 if short else 4)
    h.meshletTrianglesByteOffset = push(b"")
    h.blobByteCount = byte_count
    if byte_count >= GEOMETRY_BUFFER_SIZE:
        raise ValueError("The default size for geometry buffers doesn't fit the mesh")
    return h, b"".join(sections)


def write_cache(scene_dir: str, source_write_time: int, mesh_index: int, header: MeshCacheHeader, blob: bytes) -> str:
    """The file half of writeCache (DeferredLoadingContext.cpp:740-800): tmp file, then rename."""
    if header.blobByteCount != len(blob):
        raise ValueError("blobLen == header.blobByteCount")
    header.sourceWriteTime = int(source_write_time)
    path = cache_path(scene_dir, mesh_index)
    os.makedirs(os.path.dirname(path), exist_ok=True)
    if os.path.exists(path):
        os.remove(path)
    tmp = os.path.splitext(path)[0] + ".prosper_mesh_TMP"
    with open(tmp, "wb") as f:
        f.write(_PREFIX.pack(MESH_CACHE_MAGIC, MESH_CACHE_VERSION))
        f.write(header.pack())
        f.write(blob)
    os.replace(tmp, path)
    return path


def read_cache(path: str, with_blob: bool = True) -> Tuple[Optional[MeshCacheHeader], Optional[bytes]]:
    """readCache (DeferredLoadingContext.cpp:504-558): (None, None) when the file is missing or of another version,
    RuntimeError on a wrong magic."""
    if not os.path.exists(path):
        return None, None
    with open(path, "rb") as f:
        prefix = f.read(_PREFIX.size)
        magic, version = _PREFIX.unpack(prefix) if len(prefix) == _PREFIX.size else (0, 0)
        if magic != MESH_CACHE_MAGIC:
            raise RuntimeError(f"Expected a valid mesh cache in file '{path}'")
        if version != MESH_CACHE_VERSION:
            return None, None
        raw = f.read(_HEADER.size)
        if len(raw) != _HEADER.size:
            raise RuntimeError(f"Truncated mesh cache header in file '{path}'")
        header = MeshCacheHeader(*_HEADER.unpack(raw))
        blob = None
        if with_blob:
            blob = f.read(header.blobByteCount)
            if len(blob) != header.blobByteCount:
                raise RuntimeError(f"Truncated mesh cache blob in file '{path}'")
    return header, blob


def cache_valid(path: str, scene_write_time: int) -> bool:
    """cacheValid (DeferredLoadingContext.cpp:560-574)."""
    header, _ = read_cache(path, with_blob=False)
    return header is not None and header.sourceWriteTime == int(scene_write_time)


class GeometryCache:
    """uploadGeometryData (DeferredLoadingContext.cpp:1192-1269) over the data contract of scene_data: blobs are
    appended gap-free to 64 MB geometry buffers and every header offset is rebased to the blob's start."""

    def __init__(self):
        self.buffers: List[List[np.ndarray]] = [[]]
        self.used: List[int] = [0]  # bytes
        self.metadatas = []
        self.infos = []

    def _buffer_for(self, byte_count: int) -> int:
        # getGeometryBuffer: the first buffer with room, else a new one
        for i, used in enumerate(self.used):
            if GEOMETRY_BUFFER_SIZE - used >= byte_count:
                return i
        self.buffers.append([])
        self.used.append(0)
        return len(self.used) - 1

    def upload(self, header: MeshCacheHeader, blob: bytes, material_index: int) -> int:
        if header.blobByteCount == 0 or header.blobByteCount != len(blob):
            raise ValueError("cacheHeader.blobByteCount == dataBlob.size()")
        b = self._buffer_for(header.blobByteCount)
        start = self.used[b]
        if start % 4:
            raise ValueError("Mesh data should be aligned for u32")
        self.buffers[b].append(np.frombuffer(blob, dtype=np.uint32))
        self.used[b] += header.blobByteCount
        s32, s16 = start // 4, start // 2
        short = header.usesShortIndices == 1
        md = np.zeros((), sd.GEOMETRY_METADATA)
        md["bufferIndex"] = b
        md["indicesOffset"] = s16 if short else s32
        md["positionsOffset"] = s32 + header.positionsOffset
        md["normalsOffset"] = s32 + header.normalsOffset
        md["tangentsOffset"] = UNUSED if header.tangentsOffset == UNUSED else s32 + header.tangentsOffset
        md["texCoord0sOffset"] = UNUSED if header.texCoord0sOffset == UNUSED else s32 + header.texCoord0sOffset
        md["meshletsOffset"] = s32 + header.meshletsOffset
        md["meshletBoundsOffset"] = s32 + header.meshletBoundsOffset
        md["meshletVerticesOffset"] = (s16 if short else s32) + header.meshletVerticesOffset
        md["meshletTrianglesByteOffset"] = start + header.meshletTrianglesByteOffset
        md["usesShortIndices"] = header.usesShortIndices
        info = np.zeros((), sd.MESH_INFO)
        info["vertexCount"] = header.vertexCount
        info["indexCount"] = header.indexCount
        info["materialIndex"] = material_index
        self.metadatas.append(md)
        self.infos.append(info)
        return len(self.metadatas) - 1

    def finish(self):
        """(geometry buffers as u32 arrays, GeometryMetadata[], MeshInfo[])"""
        bufs = [np.concatenate(w) if w else np.zeros(0, np.uint32) for w in self.buffers]
        return bufs, np.array(self.metadatas, sd.GEOMETRY_METADATA), np.array(self.infos, sd.MESH_INFO)


def extract_mesh(geometry_buffer: np.ndarray, md, info, source_write_time: int = 0) -> Tuple[MeshCacheHeader, bytes]:
    """The inverse of GeometryCache.upload for one mesh of a loaded scene: slices its blob (indices .. uvs, no meshlets)
    out of a u32 geometry buffer and rebuilds the header, so a scene held in the data contract can be written out as
    prosper's cache files. Raises if the mesh is not packed gap-free in writeCache's section order."""
    nv, ni = int(info["vertexCount"]), int(info["indexCount"])
    short = int(md["usesShortIndices"]) == 1
    start = int(md["indicesOffset"]) * (2 if short else 4)
    if start % 4:
        raise ValueError("Mesh data should be aligned for u32")
    index_bytes = (ni * 2 + 3) // 4 * 4 if short else ni * 4
    h = MeshCacheHeader(sourceWriteTime=int(source_write_time), indexCount=ni, vertexCount=nv, usesShortIndices=1 if short else 0)
    cursor = index_bytes
    for name, size, optional in (("positionsOffset", 8 * nv, False), ("normalsOffset", 4 * nv, False),
                                 ("tangentsOffset", 4 * nv, True), ("texCoord0sOffset", 4 * nv, True)):
        off = int(md[name])
        if off == UNUSED:
            if not optional:
                raise ValueError(f"{name} missing")
            continue
        if off * 4 - start != cursor:
            raise ValueError(f"{name}: the mesh is not packed gap-free in cache order")
        setattr(h, name, cursor // 4)
        cursor += size
    h.meshletsOffset = h.meshletBoundsOffset = cursor // 4
    h.meshletVerticesOffset = cursor // (2 if short else 4)
    h.meshletTrianglesByteOffset = cursor
    h.blobByteCount = cursor
    words = np.ascontiguousarray(geometry_buffer, dtype=np.uint32)
    blob = words[start // 4:(start + cursor) // 4].tobytes()
    if len(blob) != cursor:
        raise ValueError("geometry buffer too short for the mesh")
    return h, blob
