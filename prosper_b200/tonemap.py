"""Host helpers of the display transform: the R9G9B9E5 3D LUT that render::ToneMap loads from
res/texture/tony_mc_mapface.dds (ToneMap.cpp:37-39) and a stand-in table for machines without the asset."""
import numpy as np

DXGI_FORMAT_R9G9B9E5_SHAREDEXP = 67


def pack_rgb9e5(rgb: np.ndarray) -> np.ndarray:
    """float rgb [..., 3] -> uint32 R9G9B9E5_SHAREDEXP texels (the packing rule of the format's definition:
    shared exponent from the largest channel, 9-bit mantissas rounded to nearest)."""
    c = np.clip(np.nan_to_num(np.asarray(rgb, np.float64), nan=0.0), 0.0, 65408.0)
    mx = np.maximum(c.max(axis=-1), 2.0 ** -16)
    e = np.maximum(np.floor(np.log2(mx)).astype(np.int64) + 1 + 15, 0)
    scale = 2.0 ** (e - 15 - 9)
    top = np.floor(mx / scale + 0.5)
    e = np.where(top == 512, e + 1, e)
    scale = 2.0 ** (e - 15 - 9)
    m = np.floor(c / scale[..., None] + 0.5).astype(np.uint32)
    return (m[..., 0] | (m[..., 1] << 9) | (m[..., 2] << 18) | (e.astype(np.uint32) << 27)).astype(np.uint32)


def unpack_rgb9e5(texels: np.ndarray) -> np.ndarray:
    t = np.asarray(texels, np.uint32)
    scale = np.ldexp(1.0, (t >> 27).astype(np.int64) - 24).astype(np.float32)
    return np.stack([((t >> s) & 511).astype(np.float32) * scale for s in (0, 9, 18)], axis=-1)


def stand_in_lut(dim: int = 48) -> np.ndarray:
    """A smooth display curve in the same parametrisation as the shipped table (input axis x = s / (s + 1) per
    channel, tone_map.comp:19-24): per-channel 1 - exp(-s) with a little cross-channel desaturation towards
    white. NOT Tony McMapface -- a deterministic table for tests and the benchmark where the asset is absent."""
    x = np.arange(dim, dtype=np.float64) / (dim - 1)
    s = x / np.maximum(1.0 - x, 1e-4)  # inverse of s / (s + 1)
    curve = 1.0 - np.exp(-s)
    b, g, r = np.meshgrid(curve, curve, curve, indexing="ij")  # z = blue, y = green, x = red (x fastest)
    rgb = np.stack([r, g, b], axis=-1)
    peak = rgb.max(axis=-1, keepdims=True)
    rgb = rgb + (peak - rgb) * (peak ** 4) * 0.5
    return pack_rgb9e5(rgb).reshape(-1)


def read_lut_dds(path: str) -> np.ndarray:
    """The payload of the volume .dds render::ToneMap loads (res/texture/tony_mc_mapface.dds): uint32 [d*h*w]."""
    from . import dds
    d = dds.read_dds(path)
    if d.format != dds.FORMAT_R9G9B9E5_SHAREDEXP or not (d.width == d.height == d.depth):
        raise RuntimeError("the LUT must be a cubic DXGI_FORMAT_R9G9B9E5_SHAREDEXP volume")
    return d.data.view(np.uint32).copy()


def write_ppm(path: str, rgba8: np.ndarray):
    """The image writer of 8(f)3: RGBA8 [h, w, 4] -> binary PPM (no image library in this environment)."""
    h, w, _ = rgba8.shape
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (w, h))
        f.write(np.ascontiguousarray(rgba8[..., :3]).tobytes())
