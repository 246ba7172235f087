// core_bc7.h -- BC7 block decode, modes 4/5/6: the modes the reference's texture cache holds
// (ispc_texcomp `alpha_ultrafast` profile, Texture.cpp:257-287; ispc_texcomp.cpp:198-229).
// Used by the PRT_CFG_DECODE_BC7 upload path and by tests that pin the texture units' own decode.
#pragma once
#include "hd.h"

namespace prt {

struct Bc7Bits
{
    uint64_t lo, hi;
    uint32_t pos;
    HD uint32_t get(uint32_t n)
    {
        uint64_t v;
        if (pos >= 64) v = hi >> (pos - 64);
        else
        {
            v = lo >> pos;
            if (pos + n > 64) v |= hi << (64 - pos);
        }
        pos += n;
        return (uint32_t)(v & ((1ull << n) - 1ull));
    }
};

HD uint32_t bc7Weight(uint32_t bits, uint32_t idx)
{
    if (bits == 2) return idx * 21u + (idx >= 2u ? 1u : 0u);          // 0 21 43 64
    if (bits == 3) return idx * 9u + (idx >= 4u ? 1u : 0u);           // 0 9 18 27 37 46 55 64
    // 4 bits: 0 4 9 13 17 21 26 30 34 38 43 47 51 55 60 64
    return (idx * 64u + 7u) / 15u;
}

HD uint32_t bc7Lerp(uint32_t a, uint32_t b, uint32_t w) { return ((64u - w) * a + w * b + 32u) >> 6; }

// out: 16 texels, RGBA8 packed little-endian (r | g << 8 | b << 16 | a << 24), row-major in the block.
// Returns false for a mode outside 4/5/6 (texels zeroed), like the oracle.
HD bool bc7DecodeBlock(uint64_t lo, uint64_t hi, uint32_t *out)
{
    Bc7Bits br{lo, hi, 0};
    uint32_t mode = 0;
    while (mode < 8 && br.get(1) == 0) mode++;
    if (mode == 6)
    {
        uint32_t e[4][2];
        for (int c = 0; c < 4; ++c)
            for (int k = 0; k < 2; ++k) e[c][k] = br.get(7);
        const uint32_t p0 = br.get(1), p1 = br.get(1);
        for (int c = 0; c < 4; ++c)
        {
            e[c][0] = (e[c][0] << 1) | p0;
            e[c][1] = (e[c][1] << 1) | p1;
        }
        for (int i = 0; i < 16; ++i)
        {
            const uint32_t w = bc7Weight(4, br.get(i == 0 ? 3 : 4));
            uint32_t px = 0;
            for (int c = 0; c < 4; ++c) px |= bc7Lerp(e[c][0], e[c][1], w) << (8 * c);
            out[i] = px;
        }
        return true;
    }
    if (mode == 4 || mode == 5)
    {
        const uint32_t rot = br.get(2);
        const uint32_t idxMode = mode == 4 ? br.get(1) : 0;
        const uint32_t cb = mode == 4 ? 5 : 7, ab = mode == 4 ? 6 : 8;
        uint32_t e[4][2];
        for (int c = 0; c < 3; ++c)
            for (int k = 0; k < 2; ++k)
            {
                const uint32_t v = br.get(cb);
                e[c][k] = (v << (8 - cb)) | (v >> (2 * cb - 8));
            }
        for (int k = 0; k < 2; ++k)
        {
            const uint32_t v = br.get(ab);
            e[3][k] = ab == 8 ? v : ((v << 2) | (v >> 4));
        }
        const uint32_t b0 = 2, b1 = mode == 4 ? 3 : 2;
        uint32_t i0 = 0, i1lo = 0, i1hi = 0; // packed index streams
        for (int i = 0; i < 16; ++i) i0 |= br.get(i == 0 ? b0 - 1 : b0) << (2 * i);
        for (int i = 0; i < 16; ++i)
        {
            const uint32_t v = br.get(i == 0 ? b1 - 1 : b1);
            if (i < 8) i1lo |= v << (3 * i);
            else i1hi |= v << (3 * (i - 8));
        }
        for (int i = 0; i < 16; ++i)
        {
            const uint32_t a0 = (i0 >> (2 * i)) & 3u;
            const uint32_t a1 = ((i < 8 ? i1lo >> (3 * i) : i1hi >> (3 * (i - 8)))) & 7u;
            uint32_t cw, aw;
            if (mode == 5) { cw = bc7Weight(2, a0); aw = bc7Weight(2, a1); }
            else if (idxMode == 0) { cw = bc7Weight(2, a0); aw = bc7Weight(3, a1); }
            else { cw = bc7Weight(3, a1); aw = bc7Weight(2, a0); }
            uint32_t ch[4];
            for (int c = 0; c < 3; ++c) ch[c] = bc7Lerp(e[c][0], e[c][1], cw);
            ch[3] = bc7Lerp(e[3][0], e[3][1], aw);
            if (rot >= 1 && rot <= 3)
            {
                const uint32_t t = ch[3];
                ch[3] = ch[rot - 1];
                ch[rot - 1] = t;
            }
            out[i] = ch[0] | (ch[1] << 8) | (ch[2] << 16) | (ch[3] << 24);
        }
        return true;
    }
    for (int i = 0; i < 16; ++i) out[i] = 0;
    return false;
}

// Random access to one texel (index 0..15, row-major in the block) without expanding the block: what the
// software bilinear of the any-hit stage uses, straight from the BC7 payload in HBM.
HD uint32_t bc7Field(uint64_t lo, uint64_t hi, uint32_t pos, uint32_t n)
{
    Bc7Bits br{lo, hi, pos};
    return br.get(n);
}

HD uint32_t bc7DecodeTexel(uint64_t lo, uint64_t hi, uint32_t i)
{
    const uint32_t low = (uint32_t)lo;
    if (low & 0x40u && !(low & 0x3Fu)) // mode 6: 0000001
    {
        const uint32_t p0 = bc7Field(lo, hi, 63, 1), p1 = bc7Field(lo, hi, 64, 1);
        const uint32_t idx = i == 0 ? bc7Field(lo, hi, 65, 3) : bc7Field(lo, hi, 64 + 4 * i, 4);
        const uint32_t w = bc7Weight(4, idx);
        uint32_t px = 0;
        for (uint32_t c = 0; c < 4; ++c)
        {
            const uint32_t e0 = (bc7Field(lo, hi, 7 + 14 * c, 7) << 1) | p0;
            const uint32_t e1 = (bc7Field(lo, hi, 14 + 14 * c, 7) << 1) | p1;
            px |= bc7Lerp(e0, e1, w) << (8 * c);
        }
        return px;
    }
    const bool mode4 = (low & 0x1Fu) == 0x10u, mode5 = (low & 0x3Fu) == 0x20u;
    if (mode4 || mode5)
    {
        const uint32_t head = mode4 ? 5u : 6u; // mode bits
        const uint32_t rot = bc7Field(lo, hi, head, 2);
        const uint32_t idxMode = mode4 ? bc7Field(lo, hi, head + 2, 1) : 0u;
        const uint32_t cb = mode4 ? 5u : 7u, ab = mode4 ? 6u : 8u;
        const uint32_t colorPos = head + 2 + (mode4 ? 1u : 0u);
        const uint32_t alphaPos = colorPos + 6 * cb;
        const uint32_t idx0Pos = alphaPos + 2 * ab; // 2-bit stream: 31 bits
        const uint32_t idx1Pos = idx0Pos + 31;
        const uint32_t b1 = mode4 ? 3u : 2u;
        const uint32_t a0 = i == 0 ? bc7Field(lo, hi, idx0Pos, 1) : bc7Field(lo, hi, idx0Pos + 2 * i - 1, 2);
        const uint32_t a1 = i == 0 ? bc7Field(lo, hi, idx1Pos, b1 - 1) : bc7Field(lo, hi, idx1Pos + b1 * i - 1, b1);
        uint32_t cw, aw;
        if (mode5) { cw = bc7Weight(2, a0); aw = bc7Weight(2, a1); }
        else if (idxMode == 0) { cw = bc7Weight(2, a0); aw = bc7Weight(3, a1); }
        else { cw = bc7Weight(3, a1); aw = bc7Weight(2, a0); }
        uint32_t ch[4];
        for (uint32_t c = 0; c < 3; ++c)
        {
            const uint32_t v0 = bc7Field(lo, hi, colorPos + 2 * c * cb, cb), v1 = bc7Field(lo, hi, colorPos + (2 * c + 1) * cb, cb);
            ch[c] = bc7Lerp((v0 << (8 - cb)) | (v0 >> (2 * cb - 8)), (v1 << (8 - cb)) | (v1 >> (2 * cb - 8)), cw);
        }
        const uint32_t v0 = bc7Field(lo, hi, alphaPos, ab), v1 = bc7Field(lo, hi, alphaPos + ab, ab);
        ch[3] = bc7Lerp(ab == 8 ? v0 : ((v0 << 2) | (v0 >> 4)), ab == 8 ? v1 : ((v1 << 2) | (v1 >> 4)), aw);
        if (rot >= 1 && rot <= 3)
        {
            const uint32_t t = ch[3];
            ch[3] = ch[rot - 1];
            ch[rot - 1] = t;
        }
        return ch[0] | (ch[1] << 8) | (ch[2] << 16) | (ch[3] << 24);
    }
    return 0;
}

} // namespace prt
