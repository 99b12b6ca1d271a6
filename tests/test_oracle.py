"""CPU: pin the oracle (restatement) against the reference's golden vectors and, when it was
compiled here, against the reference's own host codec (oracle/_ref)."""
import numpy as np
import pytest

from conftest import edge_inputs

BLOCK_SIZES = [1000, 4096, 8192, 16384, 20000, 32768, 65535, 65536]


def test_golden_decompress(golden, oracle):
    # the reference's own test: snappy/Makefile:54-56 (decompress every .snappy, cmp with .txt)
    for name, (txt, sn) in golden.items():
        rc, out = oracle.decompress(sn)
        assert rc == 0 and np.array_equal(out, txt), name
        rc, out = oracle.decompress_strict(sn)
        assert rc == 0 and np.array_equal(out, txt), name


def test_golden_compress(golden, oracle):
    # every fixture is snappy_compress_host(.txt, 32768) byte for byte
    for name, (txt, sn) in golden.items():
        assert np.array_equal(oracle.compress(txt, 32768), sn), name


def test_header_and_empty(oracle):
    c = oracle.compress(np.zeros(0, np.uint8), 32768)
    assert c.tobytes() == b"\x00\x80\x80\x02"  # Appendix A: empty file -> 4 bytes
    rc, out = oracle.decompress(c)
    assert rc == 0 and out.size == 0


def test_known_sizes(oracle):
    # Appendix A: 59/60/61/62 bytes of text -> 68/69/71/72 byte files; 100,000 zeros @32K -> 4,719 B
    for n, want in ((59, 68), (60, 69), (61, 71), (62, 72)):  # header 4 B + size 4 B + literal header + n
        assert oracle.compress(np.arange(n, dtype=np.uint8), 32768).size == want
    assert oracle.compress(np.zeros(100000, np.uint8), 32768).size == 4719
    rnd = np.random.default_rng(7).integers(0, 256, 32768, dtype=np.uint8)
    c = oracle.compress(rnd, 32768)
    assert bytes(c[10:13]) == b"\xf4\xff\x7f"  # literal header of a 32,768-byte literal


def test_roundtrip_edges(oracle):
    for name, data in edge_inputs().items():
        for bs in (1, 7, 64, 1000, 32768, 65535):
            if data.size // bs > 20000:
                continue
            c = oracle.compress(data, bs)
            rc, out = oracle.decompress_strict(c)
            assert rc == 0 and np.array_equal(out, data), (name, bs)
            rc, out = oracle.decompress(c)
            assert rc == 0 and np.array_equal(out, data), (name, bs)


def test_reference_quirk_64k_literal(oracle):
    # row Q(1): an all-literal 65,536-byte block breaks the REFERENCE decoder (uint16 length wraps);
    # the strict decoder (what the CUDA path implements) is correct.
    rnd = np.random.default_rng(3).integers(0, 256, 65536, dtype=np.uint8)
    c = oracle.compress(rnd, 65536)
    rc, out = oracle.decompress_strict(c)
    assert rc == 0 and np.array_equal(out, rnd)
    rc, out = oracle.decompress(c)
    assert rc != 0 or not np.array_equal(out, rnd)


def test_negative(oracle):
    good = oracle.compress(edge_inputs()["text1000"], 256)
    bad = good.copy()
    bad[0:5] = 0xFF  # five continuation bytes
    assert oracle.decompress(bad, cap=1 << 20)[0] == 1
    assert oracle.decompress(good, cap=10)[0] == 2  # dlength > max -> BUFFER_TOO_SMALL
    # copy reaching before the start of the output
    evil = np.frombuffer(b"\x08\x80\x80\x02" + b"\x03\x00\x00\x00" + bytes([0x01 | (4 << 2), 0x09, 0x00])[:3], dtype=np.uint8)
    assert oracle.decompress_strict(evil)[0] == 1
    assert oracle.decompress(evil)[0] == 1


def _ref():
    from oracle.oracle import Reference
    if not Reference.available():
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")
    return Reference()


def test_against_reference_golden(golden):
    ref = _ref()
    for name, (txt, sn) in golden.items():
        assert np.array_equal(ref.compress(txt, 32768), sn), name
        rc, out = ref.decompress(sn)
        assert rc == 0 and np.array_equal(out, txt), name


def test_against_reference_sweep(golden, oracle, P):
    ref = _ref()
    inputs = {"terror2": golden["terror2"][0], "world192_head": golden["world192"][0][:300000]}
    for kind in range(4):
        inputs[f"synth{kind}"] = P.synth_host(kind, 0xABC0 + kind, 700001)
    inputs.update({k: v for k, v in edge_inputs().items() if v.size})
    for name, data in inputs.items():
        for bs in BLOCK_SIZES + [64, 100]:
            if data.size // bs > 5000:
                continue
            a, b = oracle.compress(data, bs), ref.compress(data, bs)
            assert np.array_equal(a, b), (name, bs)
            if bs <= 65535:
                rc, out = ref.decompress(a)
                assert rc == 0 and np.array_equal(out, data), (name, bs)


def test_against_reference_above_64k(golden, oracle, P):
    """Extended domain: the restatement reproduces the reference's bytes -- 16-bit truncations
    included -- for block sizes up to 1 MiB, so it can be the checker for the GPU compressor's extended mode."""
    ref = _ref()
    inputs = {"world192": golden["world192"][0], "bin": P.synth_host(P.SYNTH_BIN, 6, 700000), "text": P.synth_host(P.SYNTH_TEXT, 5, 600000)}
    for name, data in inputs.items():
        for bs in (65537, 131072, 262144, 1048576):
            assert np.array_equal(oracle.compress(data, bs), ref.compress(data, bs)), (name, bs)
