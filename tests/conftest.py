import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "cuda-bundle-adjustment_b200"), os.path.join(ROOT, "oracle"), ROOT,
          os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def graphs():
    """Cache of generated graphs / packed problems keyed by (name, kwargs)."""
    from cugo_b200 import pack, synth
    cache = {}

    def get(name, **kw):
        key = (name, tuple(sorted(kw.items())))
        if key not in cache:
            pk = {k: kw.pop(k) for k in list(kw) if k in ("per_edge_information", "per_edge_camera")}
            g = synth.make_graph(name, **kw)
            cache[key] = (g, pack.pack(g, **pk))
        return cache[key]

    return get
