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


# ---- numpy restatement of the mip filter: what k_downsample_mitchell computes, operation for operation --------------
# (stbir_resize_uint8_srgb / _linear for an exact 2:1 reduction of RGBA, default settings; ext/stb/stb_image_resize2.h:
# default downsampling filter Mitchell, :876-877; "fancy" alpha weighting, :6784-6797, 4008-4215; edge clamp.)
def _srgb_to_linear(s):
    s = np.asarray(s, np.float64)
    return np.where(s <= 0.04045, s / 12.92, ((s + 0.055) / 1.055) ** 2.4)


def _mitchell(x):
    return (16.0 + x * x * (21.0 * x - 36.0)) / 18.0 if x < 1.0 else (32.0 + x * (-60.0 + x * (36.0 - 7.0 * x))) / 18.0


def mip_tables():
    d = [1.75, 1.25, 0.75, 0.25, 0.25, 0.75, 1.25, 1.75]
    taps = np.array([_mitchell(x) for x in d], np.float64)
    codes = np.arange(256) / 255.0
    mids = (np.arange(255) + 0.5) / 255.0
    return {"colour": _srgb_to_linear(codes).astype(np.float32), "linear": codes.astype(np.float32),
            "colourMid": _srgb_to_linear(mids).astype(np.float32), "linearMid": mids.astype(np.float32),
            "taps": (taps / taps.sum()).astype(np.float32)}


def downsample_mitchell(level: np.ndarray, srgb: bool) -> np.ndarray:
    """RGBA8 [h, w, 4] -> [max(h/2,1), max(w/2,1), 4]; float32 throughout, one rounding per operation, the same order as
    the kernel (horizontal taps left to right, then vertical taps top to bottom)."""
    tb = mip_tables()
    h, w, _ = level.shape
    dw, dh = max(w >> 1, 1), max(h >> 1, 1)
    dec = tb["colour"] if srgb else tb["linear"]
    a = tb["linear"][level[..., 3]]
    rgb = dec[level[..., :3]]
    v = np.concatenate([rgb, a[..., None], rgb * a[..., None]], axis=-1).astype(np.float32)  # R G B A Rpm Gpm Bpm
    taps = tb["taps"]
    hacc = None
    for i in range(8):
        xs = np.clip(2 * np.arange(dw) - 3 + i, 0, w - 1)
        term = taps[i] * v[:, xs]
        hacc = term if hacc is None else hacc + term
    acc = None
    for j in range(8):
        ys = np.clip(2 * np.arange(dh) - 3 + j, 0, h - 1)
        term = taps[j] * hacc[ys]
        acc = term if acc is None else acc + term
    alpha = acc[..., 3]
    tiny = np.float32(2.0 ** -120)  # stbir__small_float
    with np.errstate(divide="ignore", invalid="ignore"):
        ia = (np.float32(1.0) / alpha).astype(np.float32)
        colour = np.where((alpha < tiny)[..., None], acc[..., :3], acc[..., 4:7] * ia[..., None]).astype(np.float32)
    out = np.zeros((dh, dw, 4), np.uint8)
    mid = tb["colourMid"] if srgb else tb["linearMid"]
    out[..., :3] = np.searchsorted(mid, colour, side="right").astype(np.uint8)
    out[..., 3] = np.searchsorted(tb["linearMid"], alpha, side="right").astype(np.uint8)
    return out
