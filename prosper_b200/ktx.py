"""KTX 1.1 reader for environment cubemaps: the host-side mirror of utils::readKtx
(src/utils/Ktx.cpp:49-166, struct Ktx in Ktx.hpp:15-28) and of the checks TextureCubemap's constructor makes
before upload (TextureCubemap::init, Texture.cpp:589-601). Same fields, same level indexing, same refusals. The level-0 faces go to
prt_set_sky_cubemap; `write_ktx` exists so tests and tools can produce files without the reference's assets."""
import dataclasses
import struct
from typing import List

import numpy as np

IDENT_10 = bytes([0xAB, 0x4B, 0x54, 0x58, 0x20, 0x31, 0x31, 0xBB, 0x0D, 0x0A, 0x1A, 0x0A])  # «KTX 11»\r\n\x1A\n
IDENT_20 = bytes([0xAB, 0x4B, 0x54, 0x58, 0x20, 0x32, 0x30, 0xBB, 0x0D, 0x0A, 0x1A, 0x0A])  # «KTX 20»\r\n\x1A\n
GL_HALF_FLOAT, GL_RGBA, GL_RGBA16F = 0x140B, 0x1908, 0x881A
_HEADER = struct.Struct("<13I")  # Ktx10Header, Ktx.cpp:32-47
BLOCK_BYTES = 8  # one RGBA16F texel


@dataclasses.dataclass
class Ktx:  # Ktx.hpp:15-28
    width: int
    height: int
    depth: int
    arrayLayerCount: int
    faceCount: int
    mipLevelCount: int
    data: np.ndarray  # uint8
    levelByteOffsets: List[int]  # (iMip * arrayLayerCount * faceCount) + (iLayer * faceCount) + iFace
    format: str = "R16G16B16A16_SFLOAT"

    def level(self, iMip: int, iLayer: int = 0, iFace: int = 0) -> np.ndarray:
        """float16 [h, w, 4] view of one face of one level."""
        w, h = max(self.width >> iMip, 1), max(self.height >> iMip, 1)
        off = self.levelByteOffsets[(iMip * self.arrayLayerCount * self.faceCount) + (iLayer * self.faceCount) + iFace]
        return self.data[off:off + w * h * BLOCK_BYTES].view(np.float16).reshape(h, w, 4)

    def cube_faces(self, iMip: int = 0) -> np.ndarray:
        """float16 [6, n, n, 4] in +X -X +Y -Y +Z -Z order (the file's face order), what prt_set_sky_cubemap takes."""
        if self.faceCount != 6:
            raise RuntimeError("not a cubemap (Texture.cpp:593 asserts faceCount == 6)")
        if self.width != self.height:
            raise RuntimeError("cubemap faces must be square")
        if self.arrayLayerCount != 1:
            raise RuntimeError("cubemap arrays are not supported (Texture.cpp:693)")
        return np.stack([self.level(iMip, 0, f) for f in range(6)])


def read_ktx(path: str) -> Ktx:
    with open(path, "rb") as f:
        blob = f.read()
    ident = blob[:12]
    if ident == IDENT_20:
        raise RuntimeError("KTX 2.0 is not supported")  # Ktx.cpp:60-61
    if ident != IDENT_10:
        raise RuntimeError("File doesn't appear to be a KTX")  # Ktx.cpp:62-66
    if len(blob) < 12 + _HEADER.size:
        raise RuntimeError("truncated KTX header")
    (endianness, glType, _glTypeSize, glFormat, glInternalFormat, glBaseInternalFormat, pixelWidth, pixelHeight,
     pixelDepth, numberOfArrayElements, numberOfFaces, numberOfMipmapLevels, bytesOfKeyValueData) = _HEADER.unpack_from(blob, 12)
    if endianness != 0x04030201:
        raise RuntimeError("KTX and program endianness don't match")  # Ktx.cpp:71-73
    if not (glType == GL_HALF_FLOAT and glFormat == GL_RGBA and glInternalFormat == GL_RGBA16F
            and glBaseInternalFormat == glFormat):
        raise RuntimeError("Only RGBA16F is supported")  # Ktx.cpp:75-80
    if pixelWidth == 0:
        raise RuntimeError("KTX width is zero")  # Ktx.cpp:103
    k = Ktx(width=pixelWidth, height=max(pixelHeight, 1), depth=max(pixelDepth, 1),
            arrayLayerCount=max(numberOfArrayElements, 1), faceCount=max(numberOfFaces, 1),
            mipLevelCount=max(numberOfMipmapLevels, 1), data=np.zeros(0, np.uint8), levelByteOffsets=[])
    is_cubemap = numberOfArrayElements == 1 or numberOfFaces == 6  # Ktx.cpp:105-106
    pos = 12 + _HEADER.size + bytesOfKeyValueData  # key/value data is skipped, Ktx.cpp:88
    chunks, total = [], 0
    for _ in range(k.mipLevelCount):  # mips[layers[faces[rows[texels]]]], Ktx.cpp:121-163
        if pos + 4 > len(blob):
            raise RuntimeError("truncated KTX level")
        (image_size,) = struct.unpack_from("<I", blob, pos)
        pos += 4
        if is_cubemap:
            if image_size % 4:
                raise RuntimeError("Parsing expects tightly packed faces")  # Ktx.cpp:134-135
            image_size *= 6  # cubemap imageSize is the size of one face
        if pos + image_size > len(blob):
            raise RuntimeError("truncated KTX level data")
        chunks.append(np.frombuffer(blob, np.uint8, image_size, pos))
        face_bytes = image_size // k.faceCount
        read_mip_bytes = 0
        for _layer in range(k.arrayLayerCount):
            for _face in range(k.faceCount):
                k.levelByteOffsets.append(total + read_mip_bytes)
                read_mip_bytes += face_bytes
        if read_mip_bytes != image_size:
            raise RuntimeError("level size does not match its faces")  # Ktx.cpp:156-159
        pos += image_size + (image_size % 4)  # mip padding as Ktx.cpp:161-162 computes it
        total += image_size
    k.data = np.concatenate(chunks) if chunks else np.zeros(0, np.uint8)
    return k


def write_ktx(path: str, levels: List[np.ndarray], key_value: bytes = b""):
    """levels[m] = float16 [faces, h, w, 4] (faces = 6 for a cubemap, 1 for a 2D image)."""
    faces = levels[0].shape[0]
    h, w = levels[0].shape[1:3]
    kv = key_value + b"\0" * (-len(key_value) % 4)
    with open(path, "wb") as f:
        f.write(IDENT_10)
        f.write(_HEADER.pack(0x04030201, GL_HALF_FLOAT, 2, GL_RGBA, GL_RGBA16F, GL_RGBA, w, h, 0, 0, faces, len(levels), len(kv)))
        f.write(kv)
        for lv in levels:
            lv = np.ascontiguousarray(lv, np.float16)
            face_bytes = lv[0].nbytes
            f.write(struct.pack("<I", face_bytes if faces == 6 else lv.nbytes))
            f.write(lv.tobytes())
            f.write(b"\0" * (-lv.nbytes % 4))


def check_skybox(k: Ktx):
    """The checks of TextureCubemap::init (Texture.cpp:593-601): 6 faces, 512 x 512, more than 4 levels."""
    if k.faceCount != 6:
        raise RuntimeError("skybox KTX must have 6 faces")
    if not (k.width == 512 and k.height == 512):
        raise RuntimeError("skybox KTX must be 512 x 512 (Texture.cpp:595-598)")
    if not k.mipLevelCount > 4:
        raise RuntimeError("skybox KTX must carry more than 4 mip levels (Texture.cpp:599-601)")
