"""BC7 on the host (numpy): a bounding-box mode-6 encoder for fixtures and a decoder for modes 4 / 5 / 6.

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


_WEIGHTS = {2: (0, 21, 43, 64), 3: (0, 9, 18, 27, 37, 46, 55, 64), 4: tuple(int(v) for v in _W4)}


def _decode_block(block: bytes):
    """One 16-byte block -> 16 (r, g, b, a) tuples, row-major. Modes 4, 5 and 6 (what the reference's texture cache holds),
    written from the format's field tables: LSB-first bit stream, mode = position of the first set bit.
      mode 4: rotation 2, index selection 1, RGB endpoints 5 bits (R0 R1 G0 G1 B0 B1), alpha endpoints 6 bits,
              16 x 2-bit indices (31 bits: the anchor texel drops its top bit), 16 x 3-bit indices (47 bits);
      mode 5: rotation 2, RGB endpoints 7 bits, alpha endpoints 8 bits, 2-bit colour indices (31), 2-bit alpha indices (31);
      mode 6: RGBA endpoints 7 bits (R0 R1 ... A0 A1), one p-bit per endpoint, 16 x 4-bit indices (63 bits).
    Endpoints are widened to 8 bits by replicating their top bits; a texel is ((64 - w) * e0 + w * e1 + 32) >> 6."""
    bits = int.from_bytes(block, "little")
    pos = 0

    def get(n):
        nonlocal pos
        v = (bits >> pos) & ((1 << n) - 1)
        pos += n
        return v

    mode = 0
    while mode < 8 and get(1) == 0:
        mode += 1
    lerp = lambda a, b, w: ((64 - w) * a + w * b + 32) >> 6
    if mode == 6:
        e = [[get(7) for _ in range(2)] for _ in range(4)]
        p = [get(1), get(1)]
        e = [[(e[c][k] << 1) | p[k] for k in range(2)] for c in range(4)]
        idx = [get(3 if i == 0 else 4) for i in range(16)]
        return [tuple(lerp(e[c][0], e[c][1], _WEIGHTS[4][idx[i]]) for c in range(4)) for i in range(16)]
    if mode in (4, 5):
        rot = get(2)
        sel = get(1) if mode == 4 else 0
        cb, ab = (5, 6) if mode == 4 else (7, 8)
        widen = lambda v, n: v if n == 8 else ((v << (8 - n)) | (v >> (2 * n - 8)))
        e = [[widen(get(cb), cb) for _ in range(2)] for _ in range(3)]
        e.append([widen(get(ab), ab) for _ in range(2)])
        first = [get(1 if i == 0 else 2) for i in range(16)]
        nb = 3 if mode == 4 else 2
        second = [get(nb - 1 if i == 0 else nb) for i in range(16)]
        out = []
        for i in range(16):
            if mode == 4 and sel == 1:
                cw, aw = _WEIGHTS[3][second[i]], _WEIGHTS[2][first[i]]
            else:
                cw, aw = _WEIGHTS[2][first[i]], _WEIGHTS[nb][second[i]]
            ch = [lerp(e[c][0], e[c][1], cw) for c in range(3)] + [lerp(e[3][0], e[3][1], aw)]
            if rot:
                ch[3], ch[rot - 1] = ch[rot - 1], ch[3]
            out.append(tuple(ch))
        return out
    raise ValueError(f"BC7 mode {mode}: the texture cache only holds modes 4, 5 and 6")


def decode_bc7(blocks: np.ndarray, width: int, height: int) -> np.ndarray:
    """BC7 blocks of modes 4 / 5 / 6 -> uint8 [H, W, 4]. Plain Python per block: for tests and small images."""
    raw = np.ascontiguousarray(blocks, np.uint8).reshape(-1, 16)
    out = np.zeros((height, width, 4), np.uint8)
    bx = width // 4
    for b in range(raw.shape[0]):
        px = np.array(_decode_block(raw[b].tobytes()), np.uint8).reshape(4, 4, 4)
        out[(b // bx) * 4:(b // bx) * 4 + 4, (b % bx) * 4:(b % bx) * 4 + 4] = px
    return out


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
