"""The .dds container of prosper's texture cache: host-side mirror of utils::Dds, writeDds and
readDds (src/utils/Dds.hpp:24-46, Dds.cpp:84-328) -- same header words, same level layout, same refusals -- so cache
files written here load in prosper and the reference's own files (e.g. res/texture/tony_mc_mapface.dds) load here."""
import dataclasses
import os
import struct
from typing import List

import numpy as np

# utils::DxgiFormat (Dds.hpp:14-20)
FORMAT_R8G8B8A8_UNORM, FORMAT_R9G9B9E5_SHAREDEXP, FORMAT_BC7_UNORM = 28, 67, 98
DDS_MAGIC, DX10_MAGIC = 0x20534444, 0x30315844  # Dds.cpp:63-64
TEXTURE2D, TEXTURE3D = 3, 4  # D3d10ResourceDimension, Dds.cpp:47-52
_HEADER = struct.Struct("<7I11I8I5I")  # DdsHeader with its DdsPixelFormat, Dds.cpp:16-45: 124 bytes
_DXT10 = struct.Struct("<5I")  # DdsHeaderDxt10, Dds.cpp:54-61
assert _HEADER.size == 124


def _level_layout(width, height, depth, fmt, mip_level_count):
    """Dds::Dds (Dds.cpp:84-141): byte offset of every level and the total size."""
    offsets, total = [], 0
    if depth <= 1:
        for i in range(mip_level_count):
            w, h = max(width >> i, 1), max(height >> i, 1)
            if fmt == FORMAT_R8G8B8A8_UNORM:
                size = w * h * 4
            elif fmt == FORMAT_BC7_UNORM:
                if w % 4 or h % 4:
                    raise RuntimeError("BC7 mips should be divide evenly by 4x4")
                size = w // 4 * h // 4 * 16
            else:
                raise RuntimeError("Unknown DxgiFormat")
            offsets.append(total)
            total += size
    else:
        if mip_level_count != 1:
            raise RuntimeError("Volume textures with mips are not implemented")
        if fmt != FORMAT_R9G9B9E5_SHAREDEXP:
            raise RuntimeError("Unimplemented DxgiFormat")
        total = width * height * depth * 4
    return offsets, total


@dataclasses.dataclass
class Dds:  # Dds.hpp:24-46
    width: int
    height: int
    depth: int
    format: int
    mipLevelCount: int
    data: np.ndarray = None
    levelByteOffsets: List[int] = None

    def __post_init__(self):
        self.levelByteOffsets, total = _level_layout(self.width, self.height, self.depth, self.format, self.mipLevelCount)
        if self.data is None:
            self.data = np.zeros(total, np.uint8)
        self.data = np.ascontiguousarray(self.data, np.uint8).reshape(-1)
        if self.data.size != total:
            raise RuntimeError(f"Dds data is {self.data.size} bytes, the level chain needs {total}")

    def level(self, i: int) -> np.ndarray:
        end = self.levelByteOffsets[i + 1] if i + 1 < self.mipLevelCount else self.data.size
        return self.data[self.levelByteOffsets[i]:end]


def write_dds(dds: Dds, path: str):
    """writeDds (Dds.cpp:143-226): header words as the reference writes them, temp file then rename."""
    compressed = dds.format == FORMAT_BC7_UNORM
    if dds.format not in (FORMAT_BC7_UNORM, FORMAT_R8G8B8A8_UNORM):
        raise RuntimeError("Unkown DxgiFormat")  # isFormatCompressed, Dds.cpp:66-80 (spelling as there)
    if dds.depth != 1:
        raise RuntimeError("DDS writes for 3D textures are not implemented")
    bits = 0 if compressed else 32
    pitch = dds.width * dds.height * (0 if compressed else 4)
    if compressed:  # Dds.cpp:156-163, including its levelByteOffsets[0] (= 0) for single-level files
        pitch = dds.levelByteOffsets[0] if dds.mipLevelCount == 1 else dds.levelByteOffsets[1] - dds.levelByteOffsets[0]
    flags = 0x000A1007 if compressed else 0x0002100F
    masks = (0x000000FF, 0x0000FF00, 0x00FF0000, 0xFF000000) if bits == 32 else (0, 0, 0, 0)
    header = _HEADER.pack(124, flags, dds.height, dds.width, pitch, 0, dds.mipLevelCount, *([0] * 11),
                          32, 0x4, DX10_MAGIC, bits, *masks, 0x00401000, 0, 0, 0, 0)
    if os.path.exists(path):
        os.remove(path)
    tmp = os.path.splitext(path)[0] + ".dds_TMP"
    with open(tmp, "wb") as f:
        f.write(struct.pack("<I", DDS_MAGIC))
        f.write(header)
        f.write(_DXT10.pack(dds.format, TEXTURE2D, 0, 1, 0))
        f.write(dds.data.tobytes())
    os.replace(tmp, path)


def read_dds(path: str) -> Dds:
    """readDds (Dds.cpp:228-328) with its header checks."""
    with open(path, "rb") as f:
        blob = f.read()
    if len(blob) < 148 or struct.unpack_from("<I", blob, 0)[0] != DDS_MAGIC:
        raise RuntimeError("File doesn't appear to be a dds")
    h = _HEADER.unpack_from(blob, 4)
    size, flags, height, width, _pitch, depth, mips = h[:7]
    pf_size, pf_flags, fourcc, bits, rmask, gmask, bmask, amask = h[18:26]
    caps = h[26]
    if size != 124:
        raise RuntimeError("Unexpexted DDS_HEADER size")
    if flags not in (0x0002100F, 0x0080100F, 0x000A1007):
        raise RuntimeError("Unexpexted DDS_FLAGS")
    if pf_size != 32:
        raise RuntimeError("Unexpexted DDS_PIXEL_FORMAT size")
    if pf_flags != 0x4 or fourcc != DX10_MAGIC:
        raise RuntimeError("Expected a Dx10 header")
    if bits not in (32, 0) or rmask not in (0xFF, 0) or gmask not in (0xFF00, 0) or bmask not in (0xFF0000, 0) \
            or amask not in (0xFF000000, 0):
        raise RuntimeError("Expected a 32bit RGBA layout or 0")
    if caps not in (0x00401000, 0x00001008):
        raise RuntimeError("Unexpected DDS_CAPS")
    fmt, dimension, _misc, array_size, _misc2 = _DXT10.unpack_from(blob, 128)
    if fmt not in (FORMAT_R8G8B8A8_UNORM, FORMAT_R9G9B9E5_SHAREDEXP, FORMAT_BC7_UNORM):
        raise RuntimeError("Only R8G8B8A8Unorm and BC7Unorm DDS textures are supported")
    if dimension not in (TEXTURE2D, TEXTURE3D):
        raise RuntimeError("Only Texture2d and Texture3d DDS resource dimension is supported")
    if array_size != 1:
        raise RuntimeError("DDS texture arrays are not supported")
    _, total = _level_layout(width, height, depth, fmt, mips)
    if len(blob) < 148 + total:
        raise RuntimeError("truncated dds payload")
    return Dds(width, height, depth, fmt, mips, np.frombuffer(blob, np.uint8, total, 148).copy())
