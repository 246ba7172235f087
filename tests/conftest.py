import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")


@pytest.fixture(scope="session")
def test_scene():
    from prosper_b200 import scene_zoo
    return scene_zoo.test_scene()


@pytest.fixture(scope="session")
def test_scene_bc7():
    from prosper_b200 import scene_zoo
    from prosper_b200.scene_data import TEX_BC7
    return scene_zoo.test_scene(fmt=TEX_BC7)


@pytest.fixture(scope="session")
def small_helmet():
    from prosper_b200 import scene_zoo
    return scene_zoo.helmet(128, 128, tex_size=64, detail=0.35)


@pytest.fixture(scope="session")
def small_atrium():
    from prosper_b200 import scene_zoo
    return scene_zoo.atrium(160, 90, tex_size=64, detail=0.08)


@pytest.fixture(scope="session")
def small_atrium_instanced():
    from prosper_b200 import scene_zoo
    return scene_zoo.atrium(160, 90, tex_size=64, detail=0.08, bake=False, lights=32)


@pytest.fixture(scope="session")
def small_stress():
    from prosper_b200 import scene_zoo
    return scene_zoo.stress(160, 90, instances=150, mesh_tris=600)


@pytest.fixture(scope="session")
def small_district():
    from prosper_b200 import scene_zoo
    return scene_zoo.district(160, 90, detail=0.002, animated=1000)
