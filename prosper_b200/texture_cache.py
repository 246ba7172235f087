"""Texture-cache build: mirror of `compress` (src/scene/Texture.cpp:213-296). The mip chain and the BC7
encode run on the GPU behind prt_compress_texture; this module sizes the level chain by the reference's rules, calls the
library and writes the .dds the reference's Texture2D::init reads back (Texture.cpp:369-489)."""
import ctypes as C

import numpy as np

from . import dds
from . import rt_reference as rt

COLOR_SPACE_SRGB, COLOR_SPACE_LINEAR = 0, 1  # TextureColorSpace (Texture.hpp)


def mip_level_count(width: int, height: int, generate_mip_maps: bool = True) -> int:
    """Texture.cpp:217-225: the full chain down to 1x1 (getMipCount, Utils.hpp:144-148) minus its last two levels."""
    full = max(width, height).bit_length() if generate_mip_maps else 1
    return max(full - 2, 1)


def cache_format(width: int, height: int, levels: int) -> int:
    """Texture.cpp:227-237: BC7 only if every kept level is whole 4x4 blocks, else RGBA8."""
    for i in range(levels):
        if max(width >> i, 1) % 4 or max(height >> i, 1) % 4:
            return dds.FORMAT_R8G8B8A8_UNORM
    return dds.FORMAT_BC7_UNORM


def compress(world: "rt.World", rgba8: np.ndarray, color_space: int = COLOR_SPACE_SRGB, generate_mip_maps: bool = True) -> dds.Dds:
    """RGBA8 [h, w, 4] -> utils::Dds with the level chain filled by the GPU (no file I/O)."""
    img = np.ascontiguousarray(rgba8, np.uint8)
    h, w, ch = img.shape
    if ch != 4:
        raise RuntimeError("compress wants RGBA8 (the loader expands to 4 channels, Texture.cpp:311-338)")
    levels = mip_level_count(w, h, generate_mip_maps)
    out = dds.Dds(w, h, 1, cache_format(w, h, levels), levels)
    fmt, count = C.c_uint32(), C.c_uint32()
    rt._check(world.lib.prt_compress_texture(world.handle, rt._p(img), w, h, color_space, int(generate_mip_maps),
                                             rt._p(out.data), out.data.nbytes, C.byref(fmt), C.byref(count)))
    want = rt.sd.TEX_BC7 if out.format == dds.FORMAT_BC7_UNORM else rt.sd.TEX_RGBA8
    if count.value != levels or fmt.value != want:
        raise RuntimeError("library and host disagree on the level chain")
    return out


def compress_to_file(world: "rt.World", target_path: str, rgba8: np.ndarray, color_space: int = COLOR_SPACE_SRGB,
                     generate_mip_maps: bool = True) -> dds.Dds:
    """`compress` including its writeDds (Texture.cpp:295)."""
    out = compress(world, rgba8, color_space, generate_mip_maps)
    dds.write_dds(out, target_path)
    return out


# ---- numpy restatement of the mip filter, for the tests (same tables, same operation order as k_downsample_box) ----
def _srgb_to_linear(s):
    s = np.asarray(s, np.float64)
    return np.where(s <= 0.04045, s / 12.92, ((s + 0.055) / 1.055) ** 2.4)


def downsample_box(level: np.ndarray, srgb: bool) -> np.ndarray:
    h, w, _ = level.shape
    dw, dh = max(w >> 1, 1), max(h >> 1, 1)
    xs0, xs1 = np.minimum(2 * np.arange(dw), w - 1), np.minimum(2 * np.arange(dw) + 1, w - 1)
    ys0, ys1 = np.minimum(2 * np.arange(dh), h - 1), np.minimum(2 * np.arange(dh) + 1, h - 1)
    to_linear = _srgb_to_linear(np.arange(256) / 255.0).astype(np.float32)
    mid_srgb = _srgb_to_linear((np.arange(255) + 0.5) / 255.0).astype(np.float32)
    mid_lin = (np.arange(255) + 0.5).astype(np.float32)
    out = np.zeros((dh, dw, 4), np.uint8)
    for ch in range(4):
        lin = (not srgb) or ch == 3
        c = level[..., ch]
        taps = [c[np.ix_(ys, xs)] for ys, xs in ((ys0, xs0), (ys0, xs1), (ys1, xs0), (ys1, xs1))]
        v = [t.astype(np.float32) if lin else to_linear[t] for t in taps]
        mean = ((v[0] + v[1]) + (v[2] + v[3])) * np.float32(0.25)
        out[..., ch] = np.searchsorted(mid_lin if lin else mid_srgb, mean, side="right").astype(np.uint8)
    return out
