"""GPU: mutation fuzz of the decoder.  Valid streams (text, binary mix, runs; small blocks so that a mutation hits block
headers, tags, length bytes and offsets alike) get a few bytes flipped, bytes removed or the tail cut; the GPU decoder
must then agree with the oracle's strict decoder: both reject, or both accept with the same bytes -- and it must come
back (no hang, no fault) every time.  The reference has no negative tests; its decoder clamps silently where this one
rejects (snappy_decompress.c:145-147,174-175)."""
import numpy as np
import pytest


def _mutate(rng, good, hdr_len):
    s = good.copy()
    kind = int(rng.integers(0, 5))
    if kind == 0:      # flip 1-4 bytes of the body
        for _ in range(int(rng.integers(1, 5))):
            s[int(rng.integers(hdr_len, s.size))] ^= np.uint8(1 << int(rng.integers(0, 8)))
    elif kind == 1:    # overwrite a byte with a wild tag / length byte
        s[int(rng.integers(hdr_len, s.size))] = np.uint8(rng.choice([0xFF, 0xFC, 0xF8, 0xF4, 0xF0, 0x03, 0x02, 0x01, 0x00]))
    elif kind == 2:    # cut the tail
        s = s[: int(rng.integers(hdr_len, s.size))].copy()
    elif kind == 3:    # drop a byte (everything behind it shifts)
        at = int(rng.integers(hdr_len, s.size))
        s = np.concatenate([s[:at], s[at + 1:]])
    else:              # damage a block's 4-byte size field
        at = hdr_len + int(rng.integers(0, 4))
        s[at] = np.uint8(rng.integers(0, 256))
    return s


@pytest.mark.gpu
def test_mutated_streams_agree_with_the_strict_oracle(P, oracle):
    rng = np.random.default_rng(20261105)
    inputs = [(P.synth_host(P.SYNTH_TEXT, 0xF0221, 20000), 4096), (P.synth_host(P.SYNTH_BIN, 0xF0222, 150000), 32768),
              (np.tile(np.frombuffer(b"abcabcabd", dtype=np.uint8), 3000), 1000), (P.synth_host(P.SYNTH_MIX, 0xF0223, 70000), 65536)]
    accepted = rejected = 0
    for data, bs in inputs:
        good = oracle.compress(data, bs)
        hdr_len, _, _, _ = P.block_offsets(good)
        assert np.array_equal(P.decompress(good), data)
        for _ in range(150):
            s = _mutate(rng, good, hdr_len)
            rc, want = oracle.decompress_strict(s)
            try:
                got = P.decompress(s, max_len=1 << 24)
                assert rc == 0, "the GPU decoder accepted a stream the strict oracle rejects"
                assert np.array_equal(got, want)
                accepted += 1
            except P.SnappyError as e:
                assert rc != 0, "the GPU decoder rejected a stream the strict oracle accepts"
                assert e.status in (P.SNAPPY_INVALID_INPUT, P.SNAPPY_BUFFER_TOO_SMALL)
                rejected += 1
        assert np.array_equal(P.decompress(good), data)  # the device is still healthy
    assert rejected > 250 and accepted + rejected == 600


@pytest.mark.gpu
def test_random_inputs_and_block_sizes_compress_like_the_oracle(P, oracle, golden):
    """Seeded fuzz of the compressor on the GPU (the same families the CPU warp emulator is fuzzed with, tests/test_compress_emu.py):
    low-entropy, periodic, run-heavy, text and random data, ragged lengths, block sizes from 1 to 65536 -- byte for byte the
    oracle's stream, and the round trip gives the input back."""
    rng = np.random.default_rng(0xB200F)
    text = golden["world192"][0]
    for case in range(120):
        kind = case % 6
        n = int(rng.integers(1, 200000))
        if kind == 0:
            data = rng.integers(0, int(rng.integers(2, 6)), n, dtype=np.uint8)
        elif kind == 1:
            per = int(rng.integers(1, 70))
            data = np.tile(rng.integers(0, 256, per, dtype=np.uint8), n // per + 1)[:n].copy()
        elif kind == 2:
            s = int(rng.integers(0, text.size - n))
            data = text[s:s + n].copy()
        elif kind == 3:
            data = np.repeat(rng.integers(0, 256, n // 16 + 1, dtype=np.uint8), rng.integers(1, 40, n // 16 + 1))[:n].copy()
        elif kind == 4:
            data = rng.integers(0, 256, n, dtype=np.uint8)
        else:
            data = np.concatenate([text[:n // 2], rng.integers(0, 256, n - n // 2, dtype=np.uint8)])
        bs = int(rng.choice([1, 7, 15, 16, 64, 100, 1000, 4096, 5000, 16384, 20000, 32768, 50000, 65535, 65536]))
        if data.size // bs > 30000:
            bs = 4096
        want = oracle.compress(data, bs)
        got = P.compress(data, bs)
        assert got.size == want.size and np.array_equal(got, want), (case, kind, data.size, bs)
        assert np.array_equal(P.decompress(got), data), (case, kind, data.size, bs)
