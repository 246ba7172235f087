"""Generates tests/golden/tonemap_tony.json (run in the build container, where /root/reference exists):
tone_map.comp evaluated in float64 numpy on sample colours against the reference's own Tony McMapface table
(res/texture/tony_mc_mapface.dds). Only the LUT texels those samples touch are stored (index -> R9G9B9E5 word), so
the fixture pins the oracle on real table values without carrying the asset."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from prosper_b200 import tonemap  # noqa: E402

LUT_PATH = "/root/reference/res/texture/tony_mc_mapface.dds"


def glsl_mod(a, b):
    return a - b * np.floor(a / b)


def tone_map_f64(rgb, exposure, contrast, lut_words, dim, touched):
    """One pixel, float64, straight from tone_map.comp:30-62 and math.glsl:18-77."""
    r, g, b = (float(x) * exposure for x in rgb)
    value, low = max(r, g, b), min(r, g, b)
    chroma = value - low
    if chroma == 0:
        hue = 0.0
    elif value == r:
        hue = glsl_mod((g - b) / chroma, 6.0)
    elif value == g:
        hue = (b - r) / chroma + 2.0
    else:
        hue = (r - g) / chroma + 4.0
    sat = 0.0 if value == 0 else chroma / value
    value = value ** contrast
    chroma = value * sat
    x = chroma * (1.0 - abs(glsl_mod(hue, 2.0) - 1.0))
    table = [(chroma, x, 0), (x, chroma, 0), (0, chroma, x), (0, x, chroma), (x, 0, chroma), (chroma, 0, x)]
    rgb2 = np.array(table[min(int(hue), 5)], np.float64) + (value - chroma)
    uv = rgb2 / (rgb2 + 1.0) * ((dim - 1.0) / dim) + 0.5 / dim
    c = uv * dim - 0.5
    i0 = np.floor(c).astype(int)
    fr = c - i0
    out = np.zeros(3)
    for k in range(8):
        idx = [min(max(i0[a] + ((k >> a) & 1), 0), dim - 1) for a in range(3)]
        w = np.prod([fr[a] if (k >> a) & 1 else 1.0 - fr[a] for a in range(3)])
        flat = (idx[2] * dim + idx[1]) * dim + idx[0]
        touched[flat] = int(lut_words[flat])
        out += w * tonemap.unpack_rgb9e5(lut_words[flat:flat + 1])[0].astype(np.float64)
    return out ** (1.0 / 2.2)


def main():
    lut = tonemap.read_lut_dds(LUT_PATH)
    dim = 48
    rng = np.random.default_rng(20240611)
    colours = np.exp(rng.uniform(np.log(1e-3), np.log(60.0), (80, 3)))
    colours[:8] = colours[:8, :1]                      # greys
    colours[8:16, 1:] *= 0.01                          # saturated reds
    colours[16] = 0.0
    colours[17] = [1.0, 1.0, 1.0]
    colours = colours.astype(np.float16).astype(np.float64)  # the pass reads an fp16 image
    cases, touched = [], {}
    for exposure, contrast in ((1.0, 1.0), (0.35, 1.0), (2.5, 1.3)):
        exact = np.array([tone_map_f64(c, exposure, contrast, lut, dim, touched) for c in colours])
        cases.append({"exposure": exposure, "contrast": contrast,
                      "display": np.clip(exact, 0, 1).round(6).tolist()})
    fixture = {"source": "res/texture/tony_mc_mapface.dds (48^3 R9G9B9E5), tone_map.comp in float64",
               "dim": dim, "colours_f16_bits": colours.astype(np.float16).view(np.uint16).tolist(),
               "lut_texels": {str(k): v for k, v in sorted(touched.items())}, "cases": cases}
    with open(os.path.join(ROOT, "tests", "golden", "tonemap_tony.json"), "w") as f:
        json.dump(fixture, f)
    print(len(touched), "texels,", len(colours), "colours,", len(cases), "cases")


if __name__ == "__main__":
    main()
