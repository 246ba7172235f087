"""BC7 block encode/decode on the host (numpy), mode 6 only.

The reference's texture cache stores BC7 UNORM blocks produced by ispc_texcomp's
`alpha_ultrafast` profile (src/scene/Texture.cpp:257-287; only modes 4/5/6 enabled,
ext/ispc_texcomp/ispc_texcomp/ispc_texcomp.cpp:198-229). That ISPC encoder cannot be built here,
so procedural textures are encoded with this bounding-box mode-6 encoder (7-bit RGBA endpoints +
p-bits, 4-bit indices); the payload is a conformant BC7 stream either way, which is all the
texture units and the oracle's decoder care about.
"""
from __future__ import annotations

import numpy as np

_W4 = np.array([0, 4, 9, 13, 17, 21, 26, 30, 34, 38, 43, 47, 51, 55, 60, 64], dtype=np.int32)


def _to_blocks(img: np.ndarray) -> np.ndarray:
    h, w, _ = img.shape
    assert h % 4 == 0 and w % 4 == 0
    b = img.reshape(h // 4, 4, w // 4, 4, 4).transpose(0, 2, 1, 3, 4)
    return b.reshape(-1, 16, 4)


def _quant_endpoint(v: np.ndarray):
    """v: [N,4] float 0..255 -> (e7 [N,4] int, p [N] int) minimising reconstruction error."""
    best = None
    for p in (0, 1):
        e = np.clip(np.rint((v - p) / 2.0), 0, 127).astype(np.int32)
        rec = e * 2 + p
        err = ((rec - v) ** 2).sum(axis=1)
        if best is None:
            best = (e, np.full(v.shape[0], p, np.int32), err)
        else:
            take = err < best[2]
            best = (np.where(take[:, None], e, best[0]), np.where(take, p, best[1]), np.minimum(err, best[2]))
    return best[0], best[1]


def encode_bc7_mode6(img: np.ndarray) -> np.ndarray:
    """img: uint8 [H, W, 4] (H, W multiples of 4) -> uint8 [H/4 * W/4 * 16] BC7 blocks, row-major."""
    px = _to_blocks(np.ascontiguousarray(img)).astype(np.float64)  # [N,16,4]
    n = px.shape[0]
    lo = px.min(axis=1)
    hi = px.max(axis=1)
    e0, p0 = _quant_endpoint(lo)
    e1, p1 = _quant_endpoint(hi)
    c0 = (e0 * 2 + p0[:, None]).astype(np.float64)
    c1 = (e1 * 2 + p1[:, None]).astype(np.float64)
    d = c1 - c0
    dd = (d * d).sum(axis=1)
    t = ((px - c0[:, None, :]) * d[:, None, :]).sum(axis=2) / np.maximum(dd, 1e-9)[:, None]
    t = np.clip(t, 0.0, 1.0) * 64.0
    idx = np.abs(t[:, :, None] - _W4[None, None, :]).argmin(axis=2).astype(np.int64)  # [N,16]
    # anchor (texel 0) must have its MSB clear: swap endpoints and invert indices otherwise
    swap = idx[:, 0] >= 8
    idx = np.where(swap[:, None], 15 - idx, idx)
    e0s = np.where(swap[:, None], e1, e0)
    e1s = np.where(swap[:, None], e0, e1)
    p0s = np.where(swap, p1, p0)
    p1s = np.where(swap, p0, p1)
    lo64 = np.zeros(n, dtype=np.uint64)
    hi64 = np.zeros(n, dtype=np.uint64)

    def put(val, bit, nbits):
        nonlocal lo64, hi64
        val = val.astype(np.uint64) & np.uint64((1 << nbits) - 1)
        if bit < 64:
            lo64 |= val << np.uint64(bit)
            if bit + nbits > 64:
                hi64 |= val >> np.uint64(64 - bit)
        else:
            hi64 |= val << np.uint64(bit - 64)

    put(np.full(n, 1 << 6, np.uint64), 0, 7)
    bit = 7
    for c in range(4):
        put(e0s[:, c], bit, 7)
        bit += 7
        put(e1s[:, c], bit, 7)
        bit += 7
    put(p0s, bit, 1)
    bit += 1
    put(p1s, bit, 1)
    bit += 1
    put(idx[:, 0], bit, 3)
    bit += 3
    for i in range(1, 16):
        put(idx[:, i], bit, 4)
        bit += 4
    assert bit == 128
    out = np.stack([lo64, hi64], axis=1).astype("<u8")
    return out.view(np.uint8).reshape(-1)


def decode_bc7_mode6(blocks: np.ndarray, width: int, height: int) -> np.ndarray:
    """Inverse of encode_bc7_mode6 (for host-side checks): -> uint8 [H, W, 4]."""
    b = np.ascontiguousarray(blocks).view("<u8").reshape(-1, 2)
    lo64, hi64 = b[:, 0], b[:, 1]

    def get(bit, nbits):
        if bit < 64:
            v = lo64 >> np.uint64(bit)
            if bit + nbits > 64:
                v = v | (hi64 << np.uint64(64 - bit))
        else:
            v = hi64 >> np.uint64(bit - 64)
        return (v & np.uint64((1 << nbits) - 1)).astype(np.int32)

    assert np.all(get(0, 7) == 64), "not mode 6"
    bit = 7
    e = np.zeros((b.shape[0], 4, 2), np.int32)
    for c in range(4):
        for k in range(2):
            e[:, c, k] = get(bit, 7)
            bit += 7
    p0 = get(bit, 1)
    p1 = get(bit + 1, 1)
    bit += 2
    e[:, :, 0] = e[:, :, 0] * 2 + p0[:, None]
    e[:, :, 1] = e[:, :, 1] * 2 + p1[:, None]
    idx = np.zeros((b.shape[0], 16), np.int32)
    idx[:, 0] = get(bit, 3)
    bit += 3
    for i in range(1, 16):
        idx[:, i] = get(bit, 4)
        bit += 4
    w = _W4[idx]  # [N,16]
    px = ((64 - w)[:, :, None] * e[:, None, :, 0] + w[:, :, None] * e[:, None, :, 1] + 32) >> 6
    px = px.astype(np.uint8).reshape(height // 4, width // 4, 4, 4, 4).transpose(0, 2, 1, 3, 4)
    return px.reshape(height, width, 4)


def build_mip_chain_bc7(img: np.ndarray):
    """RGBA8 image -> concatenated BC7 levels, chain truncated at 4x4 like Texture.cpp:217-225
    (mipLevelCount = max(fullMips - 2, 1)); box-filtered mips. Returns (bytes, levels)."""
    levels = []
    cur = img
    while True:
        levels.append(encode_bc7_mode6(cur))
        h, w, _ = cur.shape
        if h <= 4 or w <= 4 or (h // 2) % 4 or (w // 2) % 4:
            break
        cur = cur.reshape(h // 2, 2, w // 2, 2, 4).astype(np.uint16).sum(axis=(1, 3))
        cur = ((cur + 2) // 4).astype(np.uint8)
    return np.concatenate(levels), len(levels)
