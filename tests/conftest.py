import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


def _ensure_built():
    """The C restatement and the CUDA library must exist before test modules import them."""
    from oracle import flat

    flat.build()
    # load the build script by path: importing the package itself needs the library to exist already
    import importlib.util

    spec = importlib.util.spec_from_file_location("_b2_build", os.path.join(ROOT, "nv_genai_b200", "build.py"))
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    if b.needs_build():
        b.build()


_ensure_built()


def pytest_collection_modifyitems(config, items):
    # keep files in name order; test_zz_* (if any) therefore runs last
    items.sort(key=lambda it: it.fspath.basename)
