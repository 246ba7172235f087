#!/usr/bin/env python
"""CPU probe of the trace kernel's work per ray (EXPERIMENT TOOL, test infrastructure only: it drives tests/hostsim, the
g++ build of prosper_b200/csrc/core_*.h, never the product library).

Renders the bench scene (scene_zoo.atrium(seed=1), the bench camera) at 240x135 x 2 frames, 4 bounces, path by path through
the same traceRay the GPU kernels are built from and prints, per closest ray and per shadow ray, the node items, triangle
tests and any-hit calls -- plus the same counts when every closest ray is traced a second time with tMax = the hit it found
("ideal": what perfect front-to-back order and perfect culling would visit in this tree). The integer counters agree with
the GPU's per-step counters (bench.py `breakdown.counters_per_step`) to three digits, so builder and traversal-order ideas
can be priced here without GPU time.

usage: [PRT_PROBE_CSRC=/tmp/csrc_copy] python tools/cpu_item_probe.py [-DPRT_SAH_GRANULE=2 ...]     (extra flags go to g++)
"""
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def patched_source():
    src = open(os.path.join(ROOT, "tests", "hostsim", "hostsim.cpp")).read()
    # PRT_PROBE_CSRC: an edited copy of prosper_b200/csrc to price (its core_scene.h includes ../../include/prt_types.h)
    csrc = os.environ.get("PRT_PROBE_CSRC", os.path.join(ROOT, "prosper_b200", "csrc"))
    src = src.replace('"../../prosper_b200/csrc/', '"%s/' % csrc).replace('"../../oracle/', '"%s/oracle/' % ROOT)

    def sub(old, new):
        nonlocal src
        assert old in src, "hostsim.cpp changed: update tools/cpu_item_probe.py (%r)" % old[:60]
        src = src.replace(old, new, 1)

    sub('extern "C" {', '#include <cstdio>\nstatic std::atomic<uint64_t> g_c[3], g_s[3], g_i[3], g_mismatch;\nextern "C" {')
    sub("        uint64_t nc = 0, nsr = 0, nsh = 0;",
        "        uint64_t nc = 0, nsr = 0, nsh = 0; TraceCounters tcC = {0,0,0,0,0}, tcS = {0,0,0,0,0}, tcI = {0,0,0,0,0};")
    sub("traceRay<false, true>(c->as, c->sv, ray.o, ray.d, ray.tMin, ray.tMax, seed, h, &tc, &ovf);\n                    nc++;",
        "traceRay<false, true>(c->as, c->sv, ray.o, ray.d, ray.tMin, ray.tMax, seed, h, &tcC, &ovf);\n                    nc++;\n"
        "                    { HitRec h2; const float tm = h.drawInstanceIndex != MISS_INDEX ? h.t * 1.000001f : ray.tMax;\n"
        "                      traceRay<false, true>(c->as, c->sv, ray.o, ray.d, ray.tMin, tm, seed, h2, &tcI, &ovf);\n"
        "                      if (h2.drawInstanceIndex != h.drawInstanceIndex || h2.primitiveID != h.primitiveID) g_mismatch++; }")
    sub("r.shadowTMax, r.shadowSeed, sh, &tc, &ovf)", "r.shadowTMax, r.shadowSeed, sh, &tcS, &ovf)")
    sub("        closest += nc; shadowRays += nsr; shaded += nsh;",
        "        closest += nc; shadowRays += nsr; shaded += nsh; flush(c, tcC); flush(c, tcS);\n"
        "        g_c[0] += tcC.nodes; g_c[1] += tcC.tris; g_c[2] += tcC.anyHits; g_s[0] += tcS.nodes; g_s[1] += tcS.tris;\n"
        "        g_s[2] += tcS.anyHits; g_i[0] += tcI.nodes; g_i[1] += tcI.tris; g_i[2] += tcI.anyHits;")
    sub("    if (counters)\n    {\n        counters[0] += closest;",
        '    fprintf(stderr, "closest rays %llu: nodes %.3f tris %.3f any-hit %.3f per ray | ideal (tMax = the hit): nodes %.3f tris %.3f | "\n'
        '            "shadow rays %llu: nodes %.3f tris %.3f any-hit %.3f per ray | ideal re-trace found another hit: %llu\\n",\n'
        "            (unsigned long long)closest.load(), g_c[0] / (double)closest, g_c[1] / (double)closest, g_c[2] / (double)closest,\n"
        "            g_i[0] / (double)closest, g_i[1] / (double)closest, (unsigned long long)shadowRays.load(), g_s[0] / (double)shadowRays,\n"
        "            g_s[1] / (double)shadowRays, g_s[2] / (double)shadowRays, (unsigned long long)g_mismatch.load());\n"
        "    if (counters)\n    {\n        counters[0] += closest;")
    return src


def main():
    tmp = tempfile.mkdtemp(prefix="prt_probe_")
    cpp, so = os.path.join(tmp, "probe.cpp"), os.path.join(tmp, "libprobe.so")
    open(cpp, "w").write(patched_source())
    subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", "-fno-fast-math", "-pthread", "-shared", "-o", so, cpp]
                   + sys.argv[1:], check=True)
    from tests import hostsim
    hostsim.build = lambda force=False: so
    from prosper_b200 import scene_data as sd, scene_zoo
    scene = scene_zoo.atrium(1920, 1080, seed=1, lights=0, tex_size=64)
    hs = hostsim.HostSim(scene)
    hs.render(sd.reference_pc(max_bounces=4), 240, 135, 2)
    print("wide nodes in the BLAS:", hs.stats()["blasNodes"])


if __name__ == "__main__":
    main()
