import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (run on the B200 box)")


def _ensure_built():
    spec = importlib.util.spec_from_file_location("ngpb_build", os.path.join(ROOT, "instant-ngp-rendering_b200", "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    # On the GPU box the prebuilt .so files travel with the snapshot; building is a no-op then.
    if not (os.path.exists(mod.LIB_PATH) and os.path.exists(mod.ORACLE_LIB)) or os.environ.get("NGPB_REBUILD"):
        mod.build_all()
    return mod


_build = _ensure_built()


@pytest.fixture(scope="session")
def ngp():
    mod = importlib.import_module("instant-ngp-rendering_b200")
    sys.modules.setdefault("ngp_b200", mod)
    return mod


@pytest.fixture(scope="session")
def oracle():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding
    return oracle_binding
