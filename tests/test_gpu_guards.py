"""GPU: out-of-bounds writes and nondeterminism, checked by hand (compute-sanitizer is closed on this pool --
profiles/r2_sanitizer_note.txt).  Every device buffer the two device entry points are given sits between two guard bands
filled with a pattern; after compressing and decompressing, the bands must be untouched, the results must equal the
oracle's, and five repetitions must give identical bytes (the parser -> emitter hand-off and the shared hash-table
slots are the places a race would show).  The same test runs a second time in a subprocess with PIM_SNAPPY_NSM=1, which
makes the compressor use global-memory hash tables (the pair layout is fixed per process)."""
import os
import subprocess
import sys

import numpy as np
import pytest

GUARD = 4096
PAT = 0xC3


def _cases(P):
    rng = np.random.default_rng(11)
    return [("mix", P.synth_host(P.SYNTH_MIX, 0x6A7D, (2 << 20) + 4321), 32768),
            ("mix64k", P.synth_host(P.SYNTH_MIX, 0x6A7E, (1 << 20) + 77), 65536),   # largest block size, ragged last block
            ("text1000", P.synth_host(P.SYNTH_TEXT, 0x6A7F, 70001), 1000),           # odd block size: unaligned blocks
            ("zeros", np.zeros(200000, dtype=np.uint8), 32768),                        # every copy overlaps itself
            ("period300", np.tile(rng.integers(0, 256, 300, dtype=np.uint8), 400), 4096),
            ("bin", P.synth_host(P.SYNTH_BIN, 0x6A80, 700001), 16384),
            ("tiny", P.synth_host(P.SYNTH_TEXT, 0x6A81, 17), 32768)]


def _guarded(torch, nbytes, dev):
    t = torch.full((nbytes + 2 * GUARD,), PAT, dtype=torch.uint8, device=dev)
    return t, t.data_ptr() + GUARD


def _bands_ok(t, nbytes):
    return bool((t[:GUARD] == PAT).all().item()) and bool((t[GUARD + nbytes:] == PAT).all().item())


@pytest.mark.gpu
def test_guard_bands_and_determinism(P, oracle):
    import torch
    L = P.lib()
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    for name, data, bs in _cases(P):
        n = data.size
        nb = int(L.pim_snappy_num_blocks(n, bs))
        want = oracle.compress(data, bs)
        sizes = {"in": n, "scratch": int(L.pim_snappy_compress_scratch_bytes(n, bs)), "comp": int(L.pim_snappy_compress_bound(n, bs)),
                 "off": 8 * (nb + 1), "back": n, "status": 4}
        buf = {k: _guarded(torch, v, dev) for k, v in sizes.items()}
        buf["in"][0][GUARD:GUARD + n] = torch.from_numpy(data.copy()).to(dev)
        first = None
        for rep in range(5):
            buf["comp"][0][GUARD:GUARD + sizes["comp"]] = 0x5A
            buf["back"][0][GUARD:GUARD + n] = 0xA5
            P.compress_device(buf["in"][1], n, bs, buf["scratch"][1], sizes["scratch"], buf["comp"][1], buf["off"][1], st)
            P.decompress_device(buf["comp"][1], buf["off"][1], nb, bs, n, buf["back"][1], buf["status"][1], st)
            torch.cuda.synchronize()
            for k, v in sizes.items():
                assert _bands_ok(buf[k][0], v), (name, k, rep)
            off = buf["off"][0][GUARD:GUARD + 8 * (nb + 1)].view(torch.int64).cpu().numpy()
            total = int(off[-1])
            comp = buf["comp"][0][GUARD:GUARD + total].cpu().numpy()
            assert 0 < total < want.size and np.array_equal(comp, want[want.size - total:]), (name, rep)  # the oracle's stream body
            assert (buf["comp"][0][GUARD + total:GUARD + sizes["comp"]] == 0x5A).all().item(), (name, "bytes past the compacted stream were written")
            assert int(buf["status"][0][GUARD:GUARD + 4].view(torch.int32).item()) == 0
            assert np.array_equal(buf["back"][0][GUARD:GUARD + n].cpu().numpy(), data), (name, rep)
            assert np.array_equal(buf["in"][0][GUARD:GUARD + n].cpu().numpy(), data), (name, "input modified")
            if first is None:
                first = (comp.copy(), off.copy())
            else:
                assert np.array_equal(first[0], comp) and np.array_equal(first[1], off), (name, rep, "repetitions differ")


@pytest.mark.gpu
def test_guard_bands_with_global_memory_tables(P):
    if os.environ.get("PIM_SNAPPY_NSM"):
        pytest.skip("already the inner run")
    env = dict(os.environ, PIM_SNAPPY_NSM="1")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-q", "-m", "gpu", "-k", "determinism", "-p", "no:cacheprovider"],
                       capture_output=True, text=True, env=env, cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert "1 passed" in r.stdout
