import gzip
import hashlib
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


def _gz(name):
    with gzip.open(os.path.join(GOLDEN, name + ".gz"), "rb") as f:
        return np.frombuffer(f.read(), dtype=np.uint8)


@pytest.fixture(scope="session")
def manifest():
    with open(os.path.join(GOLDEN, "MANIFEST.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def oracle():
    from oracle.oracle import Oracle, build
    build()  # compiles the restatement, and oracle/_ref when /root/reference is present
    return Oracle()


@pytest.fixture(scope="session")
def golden(manifest, oracle):
    """name -> (txt, snappy) for the reference's six fixture pairs (xml.txt rebuilt with the oracle)."""
    out = {}
    for name, m in manifest.items():
        sn = _gz(name + ".snappy")
        assert hashlib.md5(sn.tobytes()).hexdigest() == m["snappy_md5"]
        if m["txt_stored"]:
            txt = _gz(name + ".txt")
        else:
            rc, txt = oracle.decompress(sn)
            assert rc == 0
        assert txt.size == m["txt_bytes"] and hashlib.md5(txt.tobytes()).hexdigest() == m["txt_md5"], name
        out[name] = (txt, sn)
    return out


@pytest.fixture(scope="session")
def P():
    import pim_compression_b200 as pkg
    pkg.build()
    pkg.lib()
    return pkg


def edge_inputs():
    """Small inputs around the compressor's boundaries."""
    rng = np.random.default_rng(12345)
    text = (b"the quick brown fox jumps over the lazy dog and then the dog jumps over the quick fox again " * 40)
    cases = {"empty": b""}
    for n in (1, 4, 14, 15, 16, 17, 59, 60, 61, 62, 63, 64, 65, 100, 255, 256, 257, 1000):
        cases[f"text{n}"] = text[:n]
    cases["zeros100k"] = bytes(100000)
    cases["rand70k"] = rng.integers(0, 256, 70000, dtype=np.uint8).tobytes()
    cases["lowent"] = rng.integers(0, 3, 50000, dtype=np.uint8).tobytes()
    cases["period7"] = (b"abcdefg" * 9000)
    cases["period33"] = (bytes(range(33)) * 2500)
    cases["period300"] = (rng.integers(0, 256, 300, dtype=np.uint8).tobytes() * 200)
    cases["runs"] = b"".join(bytes([i & 255]) * (1 + (i * 7) % 90) for i in range(1500))
    return {k: np.frombuffer(v, dtype=np.uint8) for k, v in cases.items()}
