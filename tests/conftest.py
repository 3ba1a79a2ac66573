import importlib.util
import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


def load_pkg():
    """The package directory is named cuda-mapreduce_b200 (a hyphen), so import it by path."""
    name = "cuda_mapreduce_b200"
    if name in sys.modules:
        return sys.modules[name]
    path = os.path.join(ROOT, "cuda-mapreduce_b200", "__init__.py")
    spec = importlib.util.spec_from_file_location(name, path, submodule_search_locations=[os.path.dirname(path)])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="session")
def golden():
    with open(os.path.join(ROOT, "tests", "golden", "ref_vectors.json")) as f:
        d = json.load(f)
    out = []
    for v in d["vectors"]:
        data = None if v["input_hex"] is None else bytes.fromhex(v["input_hex"])
        out.append((v["name"], data, bytes.fromhex(v["stdout_hex"])))
    return out


@pytest.fixture(scope="session")
def orc():
    from oracle import orc as o
    o.lib()
    return o


@pytest.fixture(scope="session")
def pkg():
    return load_pkg()
