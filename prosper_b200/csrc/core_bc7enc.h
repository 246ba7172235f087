// core_bc7enc.h -- BC7 block encoder for the texture-cache build (`compress`, Texture.cpp:257-287), to the profile the
// reference asks of ispc_texcomp: GetProfile_alpha_ultrafast (ext/ispc_texcomp/ispc_texcomp/ispc_texcomp.cpp:198-229) =
// four channels, modes 4 and 5 with alpha as the separate channel (rotation 0; mode 4 with both index layouts) and mode 6,
// the cheapest of the candidates wins (kernel.ispc:1659-1786, 2064-2071).
//
// Per candidate, the reference's scheme: endpoints from the principal axis of the block (covariance, eight power
// iterations, extents of the projections; block_segment, kernel.ispc:928-998), endpoint quantisation to the mode's
// precision (ep_quant, 1077-1185), per-texel index choice between the two neighbours of the projection (block_quant,
// 1227-1287), then least-squares endpoint refinement against the chosen indices (opt_endpoints, 1292-1356) with one
// (modes 4/5, and the scalar alpha channel, opt_channel 1634-1657) or two (mode 6) rounds.
//
// A restatement, not a transcription -- ispc's fast rsqrt / rcp make its exact bits unreachable anyway. Differences, all
// on the side of lower error: texel errors are measured on what the decoder will reconstruct (integer endpoints, the
// spec's 6-bit weights) and summed in integers; the least-squares system uses those weights, not index / (levels - 1);
// the best of the start and every refinement round is kept, not the last. Floating point is single operations in a
// fixed order (hd.h x-ops), integers elsewhere, so the host build of this header (tests/hostsim) emits the same bits.
#pragma once
#include "core_bc7.h"

namespace prt {

struct Bc7Candidate
{
    uint32_t err;      // sum of squared channel errors over the block, as decoded
    uint32_t e[4][2];  // quantised endpoints (stored precision), channel x endpoint
    uint32_t p[2];     // mode 6 p-bits
    uint8_t ci[16];    // colour (or all-channel) indices
    uint8_t ai[16];    // alpha indices (modes 4 / 5)
};

// stored endpoint -> the 8-bit value the decoder interpolates
HD uint32_t bc7Expand(uint32_t v, uint32_t bits) { return bits >= 8u ? v : ((v << (8u - bits)) | (v >> (2u * bits - 8u))); }

HD uint32_t bc7QuantTo(float v, uint32_t bits)
{
    const int levels = (1 << bits) - 1;
    int q = (int)xadd(xmul(xdiv(v, 255.0f), (float)levels), 0.5f);
    q = q < 0 ? 0 : (q > levels ? levels : q);
    return (uint32_t)q;
}

// Principal axis through the block's `ch` channels and the segment that spans the projections (block_segment).
HD void bc7Segment(const int px[16][4], int first, int ch, float ep[2][4])
{
    float mean[4] = {0.f, 0.f, 0.f, 0.f};
    for (int c = 0; c < ch; ++c)
    {
        int s = 0;
        for (int i = 0; i < 16; ++i) s += px[i][first + c];
        mean[c] = xdiv((float)s, 16.0f);
    }
    float cov[4][4];
    for (int a = 0; a < ch; ++a)
        for (int b = a; b < ch; ++b)
        {
            float s = 0.f;
            for (int i = 0; i < 16; ++i) s = xadd(s, xmul(xsub((float)px[i][first + a], mean[a]), xsub((float)px[i][first + b], mean[b])));
            s = xmul(s, 1.0f / 65536.0f);
            if (a == b) s = xadd(s, 1.0e-6f); // keeps a flat block's axis at the grey diagonal
            cov[a][b] = cov[b][a] = s;
        }
    float v[4] = {1.f, 1.f, 1.f, 1.f};
    for (int it = 0; it < 8; ++it)
    {
        float w[4];
        for (int a = 0; a < ch; ++a)
        {
            float s = 0.f;
            for (int b = 0; b < ch; ++b) s = xadd(s, xmul(cov[a][b], v[b]));
            w[a] = s;
        }
        for (int a = 0; a < ch; ++a) v[a] = w[a];
        if (it & 1)
        {
            float n = 0.f;
            for (int a = 0; a < ch; ++a) n = xadd(n, xmul(v[a], v[a]));
            const float r = xdiv(1.0f, xsqrt(n));
            for (int a = 0; a < ch; ++a) v[a] = xmul(v[a], r);
        }
    }
    float lo = 3.0e38f, hi = -3.0e38f;
    for (int i = 0; i < 16; ++i)
    {
        float d = 0.f;
        for (int a = 0; a < ch; ++a) d = xadd(d, xmul(v[a], xsub((float)px[i][first + a], mean[a])));
        lo = d < lo ? d : lo;
        hi = d > hi ? d : hi;
    }
    if (xsub(hi, lo) < 1.0f) { lo = xsub(lo, 0.5f); hi = xadd(hi, 0.5f); } // keep collapsed endpoints apart
    for (int a = 0; a < ch; ++a)
    {
        const float e0 = xadd(xmul(lo, v[a]), mean[a]), e1 = xadd(xmul(hi, v[a]), mean[a]);
        ep[0][a] = e0 < 0.f ? 0.f : (e0 > 255.f ? 255.f : e0);
        ep[1][a] = e1 < 0.f ? 0.f : (e1 > 255.f ? 255.f : e1);
    }
}

// Index per texel for `ch` channels interpolated between the decoder-side endpoints d[c][0..1] with `bits`-bit indices:
// the two indices around the projection are decoded and the closer one kept. Returns the block's squared error.
HD uint32_t bc7PickIndices(const int px[16][4], int first, int ch, const uint32_t d[4][2], uint32_t bits, uint8_t idx[16])
{
    const int levels = 1 << bits;
    int dir[4];
    int len2 = 0;
    for (int c = 0; c < ch; ++c) { dir[c] = (int)d[c][1] - (int)d[c][0]; len2 += dir[c] * dir[c]; }
    uint32_t total = 0;
    for (int i = 0; i < 16; ++i)
    {
        int dotp = 0;
        for (int c = 0; c < ch; ++c) dotp += (px[i][first + c] - (int)d[c][0]) * dir[c];
        int q = 0;
        if (len2 > 0 && dotp > 0)
        {
            q = (int)(((long long)dotp * (levels - 1)) / len2); // floor of the position on the index scale
            q = q > levels - 2 ? levels - 2 : q;
        }
        uint32_t errs[2];
        for (int k = 0; k < 2; ++k)
        {
            const uint32_t w = bc7Weight(bits, (uint32_t)(q + k));
            uint32_t e = 0;
            for (int c = 0; c < ch; ++c)
            {
                const int diff = (int)bc7Lerp(d[c][0], d[c][1], w) - px[i][first + c];
                e += (uint32_t)(diff * diff);
            }
            errs[k] = e;
        }
        const int pick = errs[1] < errs[0] ? 1 : 0;
        idx[i] = (uint8_t)(q + pick);
        total += errs[pick];
    }
    return total;
}

// Least squares: the endpoints (A, B) that minimise sum |(64 - w_i) / 64 A + w_i / 64 B - x_i|^2 for the chosen indices.
// The normal equations are integers; only the final quotient is floating point. A singular system (all indices equal)
// collapses the endpoints to the mean.
HD void bc7Refit(const int px[16][4], int first, int ch, uint32_t bits, const uint8_t idx[16], float ep[2][4])
{
    long long suu = 0, suw = 0, sww = 0, tu[4] = {0, 0, 0, 0}, tw[4] = {0, 0, 0, 0}, sum[4] = {0, 0, 0, 0};
    for (int i = 0; i < 16; ++i)
    {
        const long long w = (long long)bc7Weight(bits, idx[i]), u = 64 - w;
        suu += u * u; suw += u * w; sww += w * w;
        for (int c = 0; c < ch; ++c)
        {
            const long long x = px[i][first + c];
            tu[c] += u * x; tw[c] += w * x; sum[c] += x;
        }
    }
    const long long det = suu * sww - suw * suw;
    for (int c = 0; c < ch; ++c)
    {
        float a, b;
        if (det == 0) a = b = xdiv((float)sum[c], 16.0f);
        else
        {
            a = (float)((double)(64 * (tu[c] * sww - tw[c] * suw)) / (double)det);
            b = (float)((double)(64 * (tw[c] * suu - tu[c] * suw)) / (double)det);
        }
        ep[0][c] = a < 0.f ? 0.f : (a > 255.f ? 255.f : a);
        ep[1][c] = b < 0.f ? 0.f : (b > 255.f ? 255.f : b);
    }
}

// One vector of a candidate: `ch` channels starting at `first`, endpoints stored with `epBits` bits (0: mode 6's seven bits
// plus a p-bit shared by the channels of an endpoint), `idxBits`-bit indices, `rounds` least-squares refinements.
HD void bc7FitVector(const int px[16][4], int first, int ch, uint32_t epBits, uint32_t idxBits, int rounds, uint32_t eOut[4][2], uint32_t pOut[2],
                     uint8_t idxOut[16], uint32_t &errOut)
{
    float ep[2][4];
    if (ch == 1)
    {
        int lo = 255, hi = 0; // a scalar channel starts from its range (opt_channel)
        for (int i = 0; i < 16; ++i) { lo = px[i][first] < lo ? px[i][first] : lo; hi = px[i][first] > hi ? px[i][first] : hi; }
        ep[0][0] = (float)lo; ep[1][0] = (float)hi;
    }
    else
        bc7Segment(px, first, ch, ep);
    errOut = 0xFFFFFFFFu;
    for (int round = 0; round <= rounds; ++round)
    {
        uint32_t e[4][2], d[4][2], p[2] = {0u, 0u};
        for (int k = 0; k < 2; ++k)
        {
            if (epBits)
            {
                for (int c = 0; c < ch; ++c) { e[c][k] = bc7QuantTo(ep[k][c], epBits); d[c][k] = bc7Expand(e[c][k], epBits); }
                continue;
            }
            // seven bits + p: the 8-bit value keeps the p-bit as its lowest bit; the p-bit with the smaller error wins
            uint32_t v[2][4];
            float err[2] = {0.f, 0.f};
            for (int b = 0; b < 2; ++b)
                for (int c = 0; c < ch; ++c)
                {
                    int q = (int)xadd(xmul(xsub(ep[k][c], (float)b), 0.5f), 0.5f) * 2 + b;
                    q = q < b ? b : (q > 254 + b ? 254 + b : q);
                    v[b][c] = (uint32_t)q;
                    const float diff = xsub(ep[k][c], (float)q);
                    err[b] = xadd(err[b], xmul(diff, diff));
                }
            const int b = err[1] < err[0] ? 1 : 0;
            p[k] = (uint32_t)b;
            for (int c = 0; c < ch; ++c) { d[c][k] = v[b][c]; e[c][k] = v[b][c] >> 1; }
        }
        uint8_t idx[16];
        const uint32_t err = bc7PickIndices(px, first, ch, d, idxBits, idx);
        if (err < errOut)
        {
            errOut = err;
            for (int c = 0; c < ch; ++c) { eOut[c][0] = e[c][0]; eOut[c][1] = e[c][1]; }
            pOut[0] = p[0]; pOut[1] = p[1];
            for (int i = 0; i < 16; ++i) idxOut[i] = idx[i];
        }
        if (round < rounds) bc7Refit(px, first, ch, idxBits, idx, ep);
    }
}

struct Bc7Writer
{
    uint64_t lo, hi;
    uint32_t pos;
    HD void put(uint32_t v, uint32_t n)
    {
        if (pos < 64u)
        {
            lo |= (uint64_t)v << pos;
            if (pos + n > 64u) hi |= (uint64_t)v >> (64u - pos);
        }
        else
            hi |= (uint64_t)v << (pos - 64u);
        pos += n;
    }
};

// The anchor texel's index is stored without its top bit: a vector whose first index has it set swaps its endpoints and
// complements its indices.
HD bool bc7AnchorFlip(uint8_t idx[16], uint32_t bits)
{
    if (!(idx[0] >> (bits - 1u))) return false;
    const uint8_t top = (uint8_t)((1u << bits) - 1u);
    for (int i = 0; i < 16; ++i) idx[i] = (uint8_t)(top - idx[i]);
    return true;
}
HD void bc7PutIndices(Bc7Writer &w, const uint8_t idx[16], uint32_t bits)
{
    for (int i = 0; i < 16; ++i) w.put(idx[i], i == 0 ? bits - 1u : bits);
}

// texels: 16 RGBA8 words, row-major in the block (r | g << 8 | b << 16 | a << 24). Returns the mode written.
HD uint32_t bc7EncodeBlock(const uint32_t texels[16], uint64_t &lo, uint64_t &hi)
{
    int px[16][4];
    for (int i = 0; i < 16; ++i)
        for (int c = 0; c < 4; ++c) px[i][c] = (int)((texels[i] >> (8 * c)) & 255u);
    uint32_t bestErr = 0xFFFFFFFFu, bestMode = 6u, bestIdxMode = 0u;
    Bc7Candidate best{};
    // modes 4 (both index layouts) and 5: RGB vector + alpha scalar, rotation 0
    for (int cand = 0; cand < 3; ++cand)
    {
        const uint32_t mode = cand < 2 ? 4u : 5u, idxMode = cand == 1 ? 1u : 0u;
        const uint32_t cEp = mode == 4u ? 5u : 7u, aEp = mode == 4u ? 6u : 8u;
        const uint32_t cIdx = mode == 4u ? (idxMode ? 3u : 2u) : 2u, aIdx = mode == 4u ? (idxMode ? 2u : 3u) : 2u;
        Bc7Candidate k{};
        uint32_t cErr, aErr, unused[2], ae[4][2] = {};
        bc7FitVector(px, 0, 3, cEp, cIdx, 1, k.e, unused, k.ci, cErr);
        bc7FitVector(px, 3, 1, aEp, aIdx, 1, ae, unused, k.ai, aErr);
        k.e[3][0] = ae[0][0]; k.e[3][1] = ae[0][1];
        k.err = cErr + aErr;
        if (k.err < bestErr) { bestErr = k.err; best = k; bestMode = mode; bestIdxMode = idxMode; }
    }
    {
        Bc7Candidate k{};
        bc7FitVector(px, 0, 4, 0u, 4u, 2, k.e, k.p, k.ci, k.err);
        if (k.err < bestErr) { bestErr = k.err; best = k; bestMode = 6u; bestIdxMode = 0u; }
    }
    Bc7Writer w{0ull, 0ull, 0u};
    if (bestMode == 6u)
    {
        if (bc7AnchorFlip(best.ci, 4u))
        {
            for (int c = 0; c < 4; ++c) { const uint32_t t = best.e[c][0]; best.e[c][0] = best.e[c][1]; best.e[c][1] = t; }
            const uint32_t t = best.p[0]; best.p[0] = best.p[1]; best.p[1] = t;
        }
        w.put(1u << 6, 7u);
        for (int c = 0; c < 4; ++c) { w.put(best.e[c][0], 7u); w.put(best.e[c][1], 7u); }
        w.put(best.p[0], 1u);
        w.put(best.p[1], 1u);
        bc7PutIndices(w, best.ci, 4u);
    }
    else
    {
        const uint32_t cEp = bestMode == 4u ? 5u : 7u, aEp = bestMode == 4u ? 6u : 8u;
        const uint32_t cIdx = bestMode == 4u ? (bestIdxMode ? 3u : 2u) : 2u, aIdx = bestMode == 4u ? (bestIdxMode ? 2u : 3u) : 2u;
        if (bc7AnchorFlip(best.ci, cIdx))
            for (int c = 0; c < 3; ++c) { const uint32_t t = best.e[c][0]; best.e[c][0] = best.e[c][1]; best.e[c][1] = t; }
        if (bc7AnchorFlip(best.ai, aIdx)) { const uint32_t t = best.e[3][0]; best.e[3][0] = best.e[3][1]; best.e[3][1] = t; }
        w.put(1u << bestMode, bestMode + 1u);
        w.put(0u, 2u); // rotation 0: alpha is the separate channel (mode45_channel0 = 3)
        if (bestMode == 4u) w.put(bestIdxMode, 1u);
        for (int c = 0; c < 3; ++c) { w.put(best.e[c][0], cEp); w.put(best.e[c][1], cEp); }
        w.put(best.e[3][0], aEp);
        w.put(best.e[3][1], aEp);
        // the first index stream is the 2-bit one, the second the 3-bit one (mode 4) or the alpha one (mode 5)
        if (bestMode == 4u && bestIdxMode) { bc7PutIndices(w, best.ai, 2u); bc7PutIndices(w, best.ci, 3u); }
        else { bc7PutIndices(w, best.ci, cIdx); bc7PutIndices(w, best.ai, aIdx); }
    }
    lo = w.lo;
    hi = w.hi;
    return bestMode;
}

} // namespace prt
