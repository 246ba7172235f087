"""Generates tests/golden/flight_helmet/ from the reference's res/glTF/FlightHelmet (run here, where /root/reference is):

    python tests/golden/make_flight_helmet.py [--full DIR]

prosper_cache/cache<i>.prosper_mesh  the mesh caches libprt_host.so writes: cgltf parse + (MikkTSpace) + meshoptimizer +
                                     meshlets + glm packing in the reference's call order (DeferredLoadingContext.cpp:
                                     804-865), i.e. the files a first run of prosper leaves in its own prosper_cache/.
prosper_cache/texture<i>.dds         the ten images the asset ships (of fifteen it names), reduced to 256^2 with the
                                     texture-cache mip filter so that the fixture stays small, BC7 mode 6 + mips.
scene.npz                            the small arrays of the glTF (materials, models, node transforms after updateScene, ...).

The source write time in the cache headers is pinned to 1 so that the files are reproducible byte for byte;
tests/test_gltf_host.py regenerates the mesh caches where the reference exists and compares.
--full DIR writes the same at the images' own 2048^2 (53 MB; git-ignored, used by bench.py --workload helmet when present)."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from prosper_b200 import gltf_scene  # noqa: E402

GLTF = "/root/reference/res/glTF/FlightHelmet/glTF/FlightHelmet.gltf"

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--full", default=None)
    a = ap.parse_args()
    out = a.full or gltf_scene.FLIGHT_HELMET_DIR
    os.makedirs(out, exist_ok=True)
    g = gltf_scene.GltfScene(GLTF, 512, 512, cache_dir=out, max_texture_size=None if a.full else 256, source_write_time=1)
    gltf_scene.save_cache(g, out)
    print("written", out, "triangles", g.scene.triangle_count, "missing textures", g.missing_textures)
