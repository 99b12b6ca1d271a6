"""CPU: the per-warp compressor (pim-compression_b200/csrc/compress_warp.cuh, the source the CUDA kernel
includes) run on the lock-step warp emulator (tests/emu) and compared byte for byte with the oracle.
Covers the edge inputs of SURVEY.md Appendix A, a golden fixture, and a seeded fuzz over low-entropy,
periodic, run-heavy and text-like data with odd block sizes -- the cases where lanes of one pass share
hash slots and forwarded candidates fall inside matches.  The three lane orders model the unspecified
winner of same-slot shared-memory stores on the GPU."""
import numpy as np
import pytest

from conftest import edge_inputs
from emu.emu import Emu


@pytest.fixture(scope="module")
def emu():
    return Emu()


def _same(a, b):
    return a.size == b.size and np.array_equal(a, b)


def test_edge_inputs(emu, oracle):
    for order in (0, 1, 2):
        emu.set_order(order)
        for name, data in edge_inputs().items():
            if data.size > 20000:
                data = data[:20000]
            for bs in (32768, 1000, 4096):
                assert _same(emu.compress(data, bs), oracle.compress(data, bs)), (name, bs, order)


def test_long_inputs_full_blocks(emu, oracle):
    emu.set_order(0)
    for name, data in edge_inputs().items():
        if data.size >= 50000:
            for bs in (32768, 65536):
                assert _same(emu.compress(data, bs), oracle.compress(data, bs)), (name, bs)


def test_golden_fixture(emu, oracle, golden):
    txt, sn = golden["terror2"]
    emu.set_order(2)
    assert _same(emu.compress(txt, 32768), sn)
    st = emu.stats()
    assert st[2] > 10000 and st[0] + st[1] < st[2]  # several pairs per pass: the multi-pair walk is exercised


def _fuzz_case(rng, text):
    kind = int(rng.integers(0, 7))
    n = int(rng.integers(1, 5000))
    if kind == 0:
        return rng.integers(0, int(rng.integers(2, 6)), n, dtype=np.uint8)
    if kind == 1:
        per = int(rng.integers(1, 40))
        return np.tile(rng.integers(0, 256, per, dtype=np.uint8), n // per + 1)[:n].copy()
    if kind == 2:
        s = int(rng.integers(0, text.size - n))
        return text[s:s + n].copy()
    if kind == 3:
        parts, total = [], 0
        while total < n:
            k = int(rng.integers(1, 30))
            parts.append(np.full(k, int(rng.integers(97, 101)), dtype=np.uint8))
            total += k
        return np.concatenate(parts)[:n]
    if kind == 4:
        s = int(rng.integers(0, text.size - n))
        d = text[s:s + n].copy()
        m = rng.random(d.size) < 0.05
        d[m] = rng.integers(0, 256, int(m.sum()), dtype=np.uint8)
        return d
    if kind == 5:
        words = [rng.integers(97, 100, int(rng.integers(1, 6)), dtype=np.uint8) for _ in range(6)]
        return np.concatenate([words[int(i)] for i in rng.integers(0, 6, n)])[:n]
    return rng.integers(0, 256, n, dtype=np.uint8)


def test_fuzz(emu, oracle, golden):
    text = golden["plrabn12"][0]
    rng = np.random.default_rng(20261104)
    for i in range(250):
        d = _fuzz_case(rng, text)
        bs = int(rng.choice([int(rng.integers(1, 70)), int(rng.integers(70, 3000)), 4096, 32768, 65536]))
        emu.set_order(i % 3)
        assert _same(emu.compress(d, bs), oracle.compress(d, bs)), (i, d.size, bs)
    st = emu.stats()
    assert st[5] > 0 and st[6] > st[5]  # shared-slot passes and serial re-resolves both occurred


def test_global_table_variant(emu, oracle, golden):
    """parse_block_warp<true> (pairs whose hash table lives in global memory: slot sharing inside a pass is found
    with match.any instead of a store + read-back) gives the same bytes."""
    text = golden["plrabn12"][0]
    emu.set_global_table(True)
    try:
        for order in (0, 1, 2):
            emu.set_order(order)
            for name, data in edge_inputs().items():
                d = data[:20000]
                for bs in (32768, 1000):
                    assert _same(emu.compress(d, bs), oracle.compress(d, bs)), (name, bs, order)
        rng = np.random.default_rng(7771)
        for i in range(150):
            d = _fuzz_case(rng, text)
            bs = int(rng.choice([int(rng.integers(1, 70)), int(rng.integers(70, 3000)), 4096, 32768]))
            emu.set_order(i % 3)
            assert _same(emu.compress(d, bs), oracle.compress(d, bs)), (i, d.size, bs)
        txt, sn = golden["terror2"]
        assert _same(emu.compress(txt, 32768), sn)
        st = emu.stats()
        assert st[6] > 0  # shared-slot passes occurred
    finally:
        emu.set_global_table(False)


def test_extended_blocks_reproduce_the_references_truncations(emu, oracle, golden):
    """above 65,536 bytes per block the reference's 16-bit table entries and copy offsets
    (snappy_compress.c:139,241-243,347) make it emit streams that do not decode; the warp compressor must emit the same
    bytes (the oracle restatement is pinned to the reference itself for these sizes in tests/test_oracle.py)."""
    txt = golden["world192"][0]
    rng = np.random.default_rng(5)
    low = rng.integers(0, 4, 300000, dtype=np.uint8)
    for gt, order in ((False, 0), (True, 2)):
        emu.set_global_table(gt)
        emu.set_order(order)
        for data, bs in ((txt[:500000], 131072), (txt[:700000], 1048576), (low, 262144), (txt[:200000], 65537)):
            assert _same(emu.compress(data, bs), oracle.compress(data, bs)), (bs, gt, order)
    emu.set_global_table(False)
    emu.set_order(0)
