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
