"""GPU parity tests (run on the B200 box with -m gpu).  Every call goes through the C ABI
(ctypes -> libpim_snappy.so); the oracle is only the checker."""
import os
import subprocess

import numpy as np
import pytest

from conftest import edge_inputs

pytestmark = pytest.mark.gpu

IN_DOMAIN = [4096, 8192, 16384, 32768, 65535]


def _roundtrip_check(P, oracle, name, data, bs, ref_decoder_ok=True):
    want = oracle.compress(data, bs)
    got = P.compress(data, bs)
    assert got.size == want.size and np.array_equal(got, want), f"compress mismatch {name} bs={bs}: {got.size} vs {want.size}, first diff {np.flatnonzero(got[:min(got.size, want.size)] != want[:min(got.size, want.size)])[:4]}"
    back = P.decompress(want)
    assert back.size == data.size and np.array_equal(back, data), f"decompress mismatch {name} bs={bs}, first diff {np.flatnonzero(back[:data.size] != data[:back.size])[:4]}"
    if ref_decoder_ok:
        rc, o = oracle.decompress(got)
        assert rc == 0 and np.array_equal(o, data)


def test_golden_fixtures(P, oracle, golden):
    # the reference's `make test_dpu` (snappy/Makefile:58-60) + the implicit compression KATs
    for name, (txt, sn) in golden.items():
        out = P.decompress(sn)
        assert np.array_equal(out, txt), name
        comp = P.compress(txt, 32768)
        assert np.array_equal(comp, sn), name


def test_edge_cases(P, oracle):
    for name, data in edge_inputs().items():
        for bs in (1000, 32768, 65535):
            _roundtrip_check(P, oracle, name, data, bs)


def test_tiny_blocks(P, oracle):
    data = edge_inputs()["text1000"]
    for bs in (1, 2, 3, 14, 15, 16, 17, 60, 61, 64, 100, 255, 257):
        _roundtrip_check(P, oracle, "text1000", data, bs)


def test_block_size_sweep_corpus(P, oracle, golden):
    for name in ("terror2", "plrabn12", "world192", "xml"):
        txt = golden[name][0]
        for bs in IN_DOMAIN + [1000, 20000]:
            _roundtrip_check(P, oracle, name, txt, bs)


def test_block_65536(P, oracle, golden):
    # row Q: compress-bytes parity + GPU round trip vs the original (the REFERENCE decoder breaks on
    # all-literal 64 KiB blocks, so it is not consulted here)
    for name, data in (("world192", golden["world192"][0]), ("rand", np.random.default_rng(5).integers(0, 256, 300000, dtype=np.uint8)),
                       ("mix", None)):
        if data is None:
            data = P.synth_host(P.SYNTH_MIX, 0x64, 3 << 20)
        _roundtrip_check(P, oracle, name, data, 65536, ref_decoder_ok=False)


def test_synthetic_kinds(P, oracle):
    for kind in (P.SYNTH_TEXT, P.SYNTH_RANDOM, P.SYNTH_MIX, P.SYNTH_BIN):
        data = P.synth_host(kind, 0x5BA0 + kind, (5 << 20) + 12345)
        for bs in (4096, 32768, 65535):
            _roundtrip_check(P, oracle, f"synth{kind}", data, bs)


def test_config2_dickens_standin(P, oracle, golden):
    # BASELINE.json config 2 (dickens.txt is absent upstream): SYN-TEXT-10M and CORPUS-10M, 32 KB blocks
    n = 10192446
    _roundtrip_check(P, oracle, "SYN-TEXT-10M", P.synth_host(P.SYNTH_TEXT, 0xD1C0, n), 32768)
    parts = np.concatenate([golden[k][0] for k in ("world192", "plrabn12", "terror2", "xml")])
    corpus = np.tile(parts, n // parts.size + 1)[:n]
    _roundtrip_check(P, oracle, "CORPUS-10M", corpus, 32768)


def test_config3_block_sweep_standins(P, oracle):
    # BASELINE.json config 3 stand-ins (mozilla / spamfile absent): SYN-BIN-51M, SYN-TEXT-84M
    bin51 = P.synth_host(P.SYNTH_BIN, 0x0B1A, 51220480)
    txt84 = P.synth_host(P.SYNTH_TEXT, 0x5BA0, 84217482)
    for bs in IN_DOMAIN:
        _roundtrip_check(P, oracle, "SYN-BIN-51M", bin51, bs)
    for bs in (4096, 32768):
        _roundtrip_check(P, oracle, "SYN-TEXT-84M", txt84, bs)
    for bs in (131072, 1 << 20):  # outside the reference's valid domain (row Q): rejected, not mis-compressed
        with pytest.raises(P.SnappyError) as e:
            P.compress(bin51[: 1 << 21], bs)
        assert e.value.status == P.SNAPPY_BUFFER_TOO_SMALL


def test_copy4_and_long_literal_elements(P, oracle):
    # the compressor never emits COPY_4 or 4-byte literal lengths, the decoder must accept them
    # (snappy_decompress.c:118-133,278-283; :64-74)
    lit = np.arange(200, dtype=np.uint8)
    body = bytes([60 << 2, 199]) + lit.tobytes()                 # literal, 1 extra length byte
    body += bytes([0x03 | (9 << 2)]) + (150).to_bytes(4, "little")  # COPY_4 len 10 offset 150
    body += bytes([63 << 2]) + (49).to_bytes(4, "little") + bytes(range(50))  # literal, 4 extra length bytes
    body += bytes([0x02 | (63 << 2)]) + (1).to_bytes(2, "little")  # COPY_2 len 64 offset 1 (overlap)
    dlen = 200 + 10 + 50 + 64
    stream = np.frombuffer(bytes([dlen & 0x7F | 0x80, dlen >> 7]) + b"\x80\x80\x02" + len(body).to_bytes(4, "little") + body, dtype=np.uint8)
    rc, want = oracle.decompress_strict(stream)
    assert rc == 0 and want.size == dlen
    assert np.array_equal(P.decompress(stream), want)


def test_malformed_streams(P, oracle):
    good = oracle.compress(edge_inputs()["text1000"], 256)
    with pytest.raises(P.SnappyError) as e:  # five continuation bytes in the length varint
        bad = good.copy(); bad[0:5] = 0xFF
        P.decompress(bad)
    assert e.value.status == P.SNAPPY_INVALID_INPUT
    with pytest.raises(P.SnappyError) as e:  # dlength > output->max (snappy_decompress.c:201-204)
        P.decompress(good, max_len=10)
    assert e.value.status == P.SNAPPY_BUFFER_TOO_SMALL
    hdr = b"\x08\x80\x80\x02"
    cases = {
        "copy before start": hdr + b"\x03\x00\x00\x00" + bytes([0x01 | (4 << 2), 0x09, 0x00]),
        "zero offset": hdr + b"\x06\x00\x00\x00" + bytes([3 << 2]) + b"abcd" + bytes([0x01, 0x00]),
        "literal overruns input": hdr + b"\x03\x00\x00\x00" + bytes([7 << 2]) + b"ab",
        "output too long": hdr + b"\x0c\x00\x00\x00" + bytes([10 << 2]) + b"abcdefghijk",
        "output too short": hdr + b"\x04\x00\x00\x00" + bytes([2 << 2]) + b"abc",
        "truncated block": hdr + b"\x40\x00\x00\x00" + bytes([2 << 2]) + b"abc",
        "truncated copy2": hdr + b"\x07\x00\x00\x00" + bytes([4 << 2]) + b"abcde" + bytes([0x02 | (2 << 2)]),
    }
    for name, raw in cases.items():
        s = np.frombuffer(raw, dtype=np.uint8)
        assert oracle.decompress_strict(s)[0] == 1, name
        with pytest.raises(P.SnappyError) as e:
            P.decompress(s)
        assert e.value.status == P.SNAPPY_INVALID_INPUT, name
    # random garbage bodies must be rejected or decoded without faulting, never hang
    rng = np.random.default_rng(99)
    for i in range(20):
        body = rng.integers(0, 256, 5000, dtype=np.uint8).tobytes()
        s = np.frombuffer(b"\xa0\x9c\x01" + b"\x80\x80\x02" + len(body).to_bytes(4, "little") + body, dtype=np.uint8)
        try:
            out = P.decompress(s)
            rc, want = oracle.decompress_strict(s)
            assert rc == 0 and np.array_equal(out, want)
        except P.SnappyError as e:
            assert e.status == P.SNAPPY_INVALID_INPUT and oracle.decompress_strict(s)[0] == 1
    assert np.array_equal(P.decompress(good), edge_inputs()["text1000"])  # the device is still healthy


def test_device_resident_api_and_generators(P, oracle):
    import torch
    dev = torch.device("cuda:0")
    n, bs = (8 << 20) + 333, 32768
    host = P.synth_host(P.SYNTH_MIX, 0xFEED, n)
    d_in = torch.empty(n, dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    P.synth_device(P.SYNTH_MIX, 0xFEED, d_in.data_ptr(), n, st)
    assert np.array_equal(d_in.cpu().numpy(), host)  # host and device generators agree bit for bit
    L = P.lib()
    nb = L.pim_snappy_num_blocks(n, bs)
    scratch = torch.empty(L.pim_snappy_compress_scratch_bytes(n, bs), dtype=torch.uint8, device=dev)
    d_out = torch.empty(L.pim_snappy_compress_bound(n, bs), dtype=torch.uint8, device=dev)
    d_off = torch.empty(nb + 1, dtype=torch.int64, device=dev)
    P.compress_device(d_in.data_ptr(), n, bs, scratch.data_ptr(), scratch.numel(), d_out.data_ptr(), d_off.data_ptr(), st)
    torch.cuda.synchronize()
    off = d_off.cpu().numpy().astype(np.uint64)
    want = oracle.compress(host, bs)
    hlen, dlen, bs2, woff = P.block_offsets(want)
    assert np.array_equal(off, woff)
    assert np.array_equal(d_out[: int(off[-1])].cpu().numpy(), want[hlen:])
    d_back = torch.zeros(n, dtype=torch.uint8, device=dev)
    d_status = torch.ones(1, dtype=torch.int32, device=dev)
    P.decompress_device(d_out.data_ptr(), d_off.data_ptr(), nb, bs, n, d_back.data_ptr(), d_status.data_ptr(), st)
    torch.cuda.synchronize()
    assert int(d_status.item()) == 0 and torch.equal(d_back, d_in)


def test_batch_api_matches_oracle_per_stream(P, oracle):
    """pim_snappy_compress_batch_device / pim_snappy_decompress_batch_device: ragged streams -- empty, one
    byte, shorter than a block, text / random / binary -- through ONE launch give, per stream, the oracle's bytes and
    offsets, and decode back; a malformed stream is reported in its own status word only.  40 streams also cover the
    split into launches of 32."""
    import torch
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    L = P.lib()
    for bs in (32768, 4096):
        lens = [0, 1, 14, 15, bs - 1, bs, bs + 1, 3 * bs + 17, (2 << 20) + 5, 700001] + [50000 + 977 * i for i in range(30)]
        hosts = [P.synth_host((P.SYNTH_TEXT, P.SYNTH_RANDOM, P.SYNTH_MIX, P.SYNTH_BIN)[i % 4], 0xBA7C0 + i, n) for i, n in enumerate(lens)]
        d_in = [torch.from_numpy(h.copy()).to(dev) if h.size else torch.empty(0, dtype=torch.uint8, device=dev) for h in hosts]
        nbs = [int(L.pim_snappy_num_blocks(n, bs)) for n in lens]
        scr = [torch.empty(int(L.pim_snappy_compress_scratch_bytes(n, bs)), dtype=torch.uint8, device=dev) for n in lens]
        outs = [torch.empty(int(L.pim_snappy_compress_bound(n, bs)), dtype=torch.uint8, device=dev) for n in lens]
        offs = [torch.empty(nb + 1, dtype=torch.int64, device=dev) for nb in nbs]
        P.compress_batch_device([t.data_ptr() for t in d_in], lens, bs, [t.data_ptr() for t in scr], [t.numel() for t in scr],
                                [t.data_ptr() for t in outs], [t.data_ptr() for t in offs], st)
        torch.cuda.synchronize()
        for i, h in enumerate(hosts):
            want = oracle.compress(h, bs)
            hlen, dlen, bs2, woff = P.block_offsets(want)
            off = offs[i].cpu().numpy().astype(np.uint64)
            assert np.array_equal(off, woff), (bs, i)
            assert np.array_equal(outs[i][: int(off[-1])].cpu().numpy(), want[hlen:]), (bs, i)
        backs = [torch.zeros(n, dtype=torch.uint8, device=dev) for n in lens]
        status = torch.ones(len(lens), dtype=torch.int32, device=dev)
        args = ([t.data_ptr() for t in outs], [t.data_ptr() for t in offs], nbs, bs, lens, [t.data_ptr() for t in backs], status.data_ptr(), st)
        P.decompress_batch_device(*args)
        torch.cuda.synchronize()
        assert int(status.abs().sum().item()) == 0
        for i in range(len(lens)):
            assert torch.equal(backs[i], d_in[i]), (bs, i)
        # corrupt one element byte of stream 8 (a copy offset becomes 0xFFFF) -> only its status word is set
        victim = 8
        bad = outs[victim].clone()
        raw = bad[: int(offs[victim][-1].item())].cpu().numpy()
        raw[4:40] = 0xFE  # copy-2 tags with offset 0xFEFE at the block start: reaches before the block
        bad[: raw.size] = torch.from_numpy(raw).to(dev)
        ptrs = list(args[0])
        ptrs[victim] = bad.data_ptr()
        P.decompress_batch_device(ptrs, *args[1:])
        torch.cuda.synchronize()
        stv = status.cpu().numpy()
        assert stv[victim] == 1 and int(np.abs(stv).sum()) == 1, stv


def test_large_roundtrip_properties(P):
    """BASELINE-size property test (no oracle): 1 GiB mixed stream, compress -> decompress == input,
    block offsets consistent, result independent of how the work is repeated (determinism)."""
    import torch
    dev = torch.device("cuda:0")
    n, bs = 1 << 30, 32768
    L = P.lib()
    st = torch.cuda.current_stream().cuda_stream
    d_in = torch.empty(n, dtype=torch.uint8, device=dev)
    P.synth_device(P.SYNTH_MIX, 0x64000000, d_in.data_ptr(), n, st)
    nb = L.pim_snappy_num_blocks(n, bs)
    scratch = torch.empty(L.pim_snappy_compress_scratch_bytes(n, bs), dtype=torch.uint8, device=dev)
    d_out = torch.empty(L.pim_snappy_compress_bound(n, bs), dtype=torch.uint8, device=dev)
    d_off = torch.empty(nb + 1, dtype=torch.int64, device=dev)
    P.compress_device(d_in.data_ptr(), n, bs, scratch.data_ptr(), scratch.numel(), d_out.data_ptr(), d_off.data_ptr(), st)
    torch.cuda.synchronize()
    total = int(d_off[-1].item())
    assert 0.70 * n < total < 0.85 * n  # SYN-MIX compresses to ~0.77
    sizes = (d_off[1:] - d_off[:-1])
    assert int(sizes.min().item()) > 4 and int(sizes.max().item()) <= bs + bs // 6 + 36
    chk1 = int(d_out[:total].to(torch.int64).sum().item())
    d_back = torch.zeros(n, dtype=torch.uint8, device=dev)
    d_status = torch.ones(1, dtype=torch.int32, device=dev)
    P.decompress_device(d_out.data_ptr(), d_off.data_ptr(), nb, bs, n, d_back.data_ptr(), d_status.data_ptr(), st)
    torch.cuda.synchronize()
    assert int(d_status.item()) == 0 and torch.equal(d_back, d_in)
    d_out.zero_()
    P.compress_device(d_in.data_ptr(), n, bs, scratch.data_ptr(), scratch.numel(), d_out.data_ptr(), d_off.data_ptr(), st)
    torch.cuda.synchronize()
    assert int(d_off[-1].item()) == total and int(d_out[:total].to(torch.int64).sum().item()) == chk1


def test_streamed_host_path_matches_single_shot(P, oracle):
    """The chunked pipeline of the host entry points (H2D | kernels | D2H overlapped) must give the
    same bytes as one-piece processing and as the oracle, for ragged chunk/block combinations."""
    data = P.synth_host(P.SYNTH_MIX, 0xC0FFEE, (5 << 20) + 12345)
    try:
        for bs, chunk in ((32768, 1 << 20), (4096, 1 << 20), (65536, 3 << 19), (1000, 1 << 20)):
            want = oracle.compress(data, bs)
            P.set_chunk_bytes(0)
            one = P.compress(data, bs)
            P.set_chunk_bytes(chunk)
            got = P.compress(data, bs)
            assert got.size == want.size and np.array_equal(got, want), (bs, chunk)
            assert np.array_equal(one, want)
            assert np.array_equal(P.decompress(want), data), (bs, chunk)
        # a malformed block in a late chunk is still reported
        bad = oracle.compress(data, 32768).copy()
        _, _, _, off = P.block_offsets(bad)
        hdr = bad.size - int(off[-1])
        bad[hdr + int(off[-3]) + 4] = 0xFF  # copy4 tag with a wild offset at the start of a late block
        with pytest.raises(P.SnappyError):
            P.decompress(bad)
        # a header that promises far more output than the (truncated) file holds must come back as an error, not hang
        # (streamed path: declared length > 2 chunks; round-1 advisor finding)
        for body in (b"", b"\x05\x00\x00", b"\x05\x00\x00\x00" + bytes([4 << 2]) + b"abcde"):
            s = np.frombuffer(b"\x80\x80\x80\x20" + b"\x80\x80\x02" + body, dtype=np.uint8)  # dlen = 64 MiB, 32 KiB blocks
            with pytest.raises(P.SnappyError) as e:
                P.decompress(s)
            assert e.value.status == P.SNAPPY_INVALID_INPUT
        assert np.array_equal(P.decompress(oracle.compress(data, 32768)), data)  # still healthy
    finally:
        P.set_chunk_bytes(32 << 20)  # the library's default


def test_multi_gpu_shards_match_single_gpu(P, oracle):
    """with N GPUs the shim splits the blocks into contiguous ranges (snappy_compress.c:495),
    one host thread per GPU, and stitches on the host -- the bytes must not depend on N."""
    import torch
    ndev = torch.cuda.device_count()
    if ndev < 2:
        pytest.skip("needs at least two GPUs")
    data = P.synth_host(P.SYNTH_MIX, 0x2B200, (9 << 20) + 4097)
    try:
        for bs in (32768, 4096, 65536):
            want = oracle.compress(data, bs)
            for n in sorted({2, ndev}):
                assert P.set_devices(list(range(n))) == n
                for chunk in (0, 1 << 20):  # every shard in one piece / every GPU streams its shard in groups
                    P.set_chunk_bytes(chunk)
                    got = P.compress(data, bs)
                    assert got.size == want.size and np.array_equal(got, want), (bs, n, chunk)
                    assert np.array_equal(P.decompress(want), data), (bs, n, chunk)
        P.set_chunk_bytes(1 << 20)
        bad = oracle.compress(data, 32768).copy()  # a malformed block in the last GPU's shard is still reported
        _, _, _, off = P.block_offsets(bad)
        bad[bad.size - int(off[-1]) + int(off[-3]) + 4] = 0xFF
        with pytest.raises(P.SnappyError):
            P.decompress(bad)
        # fewer blocks than GPUs
        tiny = data[:40000]
        assert np.array_equal(P.compress(tiny, 32768), oracle.compress(tiny, 32768))
        assert np.array_equal(P.decompress(oracle.compress(tiny, 32768)), tiny)
    finally:
        P.set_chunk_bytes(32 << 20)
        P.set_devices([0])


def test_gather_shards_over_peer_copies(P, oracle):
    """shards compressed on different GPUs, gathered peer to peer into one block stream on GPU 0,
    which must be the oracle's stream body and decode back on GPU 0."""
    import torch
    ndev = torch.cuda.device_count()
    if ndev < 2:
        pytest.skip("needs at least two GPUs")
    L = P.lib()
    bs = 32768
    data = P.synth_host(P.SYNTH_MIX, 0x6A7, (6 << 20) + 1234)
    nb = int(L.pim_snappy_num_blocks(data.size, bs))
    devs, shards, offs, nbs = list(range(min(ndev, 4))), [], [], []
    for g in devs:
        lo, hi = P.shard_range(nb, len(devs), g)
        part = data[lo * bs:min(data.size, hi * bs)]
        with torch.cuda.device(g):
            d_in = torch.from_numpy(part.copy()).to(f"cuda:{g}")
            scratch = torch.empty(int(L.pim_snappy_compress_scratch_bytes(part.size, bs)), dtype=torch.uint8, device=f"cuda:{g}")
            d_out = torch.empty(int(L.pim_snappy_compress_bound(part.size, bs)), dtype=torch.uint8, device=f"cuda:{g}")
            d_off = torch.empty(hi - lo + 1, dtype=torch.int64, device=f"cuda:{g}")
            P.compress_device(d_in.data_ptr(), part.size, bs, scratch.data_ptr(), scratch.numel(), d_out.data_ptr(), d_off.data_ptr(),
                              torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize()
        shards.append(d_out); offs.append(d_off); nbs.append(hi - lo)
    sizes = [int(o[-1].item()) for o in offs]
    total = sum(sizes)
    with torch.cuda.device(0):
        d_all = torch.zeros(total, dtype=torch.uint8, device="cuda:0")
        d_all_off = torch.zeros(nb + 1, dtype=torch.int64, device="cuda:0")
        st = torch.cuda.current_stream().cuda_stream
        P.gather_shards(devs, [t.data_ptr() for t in shards], sizes, [t.data_ptr() for t in offs], nbs, 0, d_all.data_ptr(), d_all_off.data_ptr(), st)
        d_back = torch.zeros(data.size, dtype=torch.uint8, device="cuda:0")
        d_status = torch.ones(1, dtype=torch.int32, device="cuda:0")
        P.decompress_device(d_all.data_ptr(), d_all_off.data_ptr(), nb, bs, data.size, d_back.data_ptr(), d_status.data_ptr(), st)
        torch.cuda.synchronize()
    want = oracle.compress(data, bs)
    hdr = want.size - total
    assert hdr > 0 and np.array_equal(d_all.cpu().numpy(), want[hdr:])
    assert int(d_all_off[-1].item()) == total and int(d_status.item()) == 0
    assert np.array_equal(d_back.cpu().numpy(), data)


def test_cli_matches_reference_contract(P, golden, tmp_path):
    # same lines the reference CLI prints (dpu_snappy.c:152,159,212-227), same round trip as `make test`
    cli = os.path.join(os.path.dirname(P.LIB_PATH), "dpu_snappy")
    txt, sn = golden["terror2"]
    src, comp, back = tmp_path / "t.txt", tmp_path / "t.snappy", tmp_path / "t.out"
    src.write_bytes(txt.tobytes())
    r = subprocess.run([cli, "-d", "-c", "-b", "32768", "-i", str(src), "-o", str(comp)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert comp.read_bytes() == sn.tobytes()
    lines = r.stdout.splitlines()
    assert lines[0] == f"Using input file {src}" and lines[1] == f"Using output file {comp}"
    assert lines[2] == f"Compressed {sn.size} bytes to: {comp}" and lines[3].startswith("Compression ratio: 0.50")
    assert [l.split(":")[0] for l in lines[4:]] == ["Pre-processing time", "Alloc time", "Load time", "Copy in time", "Host time", "Copy out time", "Free time"]
    r = subprocess.run([cli, "-d", "-i", str(comp), "-o", str(back)], capture_output=True, text=True)
    assert r.returncode == 0 and back.read_bytes() == txt.tobytes()
    assert f"Decompressed {txt.size} bytes to: {back}" in r.stdout
    assert subprocess.run([cli], capture_output=True).returncode == 255      # missing -i  -> -1
    assert subprocess.run([cli, "-x"], capture_output=True).returncode == 254  # bad option -> -2


def test_sweep_tool_writes_reference_csv_layout(P, golden, tmp_path):
    # the reference's sweep driver, on GPUs; CSV columns as run_tests.py:134 writes them
    import csv
    import sys
    txt, _ = golden["terror2"]
    src = tmp_path / "terror2.txt"
    src.write_bytes(txt.tobytes())
    tool = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools", "run_tests.py")
    r = subprocess.run([sys.executable, tool, "-f", str(src), "-g", "1", "-b", "32768", "4096", "-o", str(tmp_path / "results"), "--reps", "1"],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    rows = list(csv.reader(open(tmp_path / "results" / "terror2_compression_breakdown.csv")))
    assert rows[0] == ["prepare", "alloc", "load", "copy_in", "run", "copy_out", "free", "dpus"] and rows[1][-1] == "1"
    blk = list(csv.reader(open(tmp_path / "results" / "compression_vs_blksize.csv")))
    assert [b[1] for b in blk[1:]] == ["32768", "4096"] and int(blk[1][2]) == golden["terror2"][1].size


def test_periodic_runs(P, oracle):
    """Runs of back-to-back copies with one distance (zero pages, repeated records) are decoded from the pattern
    before the run instead of element by element: every period around the element and ring sizes, runs that cross
    refills, blocks and the 4 KB ring, interleaved with literals."""
    rng = np.random.default_rng(0xBEEF)
    parts = []
    for period in (1, 2, 3, 4, 5, 7, 15, 16, 17, 31, 32, 33, 37, 59, 60, 61, 63, 64, 65, 66, 100, 127, 128, 129, 1000, 2047, 2048, 2049, 4095, 4097, 9000):
        pat = rng.integers(0, 256, period, dtype=np.uint8)
        for reps_bytes in (period + 3, 200, 4000, 70000):
            parts.append(np.tile(pat, reps_bytes // period + 2)[:max(reps_bytes, period + 1)])
            parts.append(rng.integers(0, 256, int(rng.integers(1, 90)), dtype=np.uint8))
    data = np.concatenate(parts)
    for bs in (32768, 65536, 4096, 1000):
        _roundtrip_check(P, oracle, f"periodic/{bs}", data, bs, ref_decoder_ok=bs < 65536)


def test_extended_domain_compressor(P, oracle, golden):
    """SURVEY 8(f) N4 / BASELINE config 3 "sweep to 1 MB": off by default; when switched on the compressor gives the
    reference's bytes above 65,536 bytes per block (streams that do not decode in the reference either), the decoder
    still refuses them."""
    data = np.concatenate([golden["world192"][0], P.synth_host(P.SYNTH_BIN, 0x0B1A, 3 << 20), P.synth_host(P.SYNTH_TEXT, 0x5BA0, 2 << 20)])
    with pytest.raises(P.SnappyError) as e:
        P.compress(data, 131072)
    assert e.value.status == P.SNAPPY_BUFFER_TOO_SMALL
    P.set_extended_blocks(True)
    try:
        for bs in (65537, 131072, 262144, 1048576, 1000000):
            want = oracle.compress(data, bs)
            for chunk in (0, 2 << 20):
                P.set_chunk_bytes(chunk)
                got = P.compress(data, bs)
                assert got.size == want.size and np.array_equal(got, want), (bs, chunk)
        with pytest.raises(P.SnappyError) as e:
            P.decompress(want)
        assert e.value.status == P.SNAPPY_BUFFER_TOO_SMALL
        with pytest.raises(P.SnappyError):
            P.compress(data, 1048577)
        assert np.array_equal(P.compress(data, 65536), oracle.compress(data, 65536))  # the valid domain is unchanged
    finally:
        P.set_extended_blocks(False)
        P.set_chunk_bytes(32 << 20)
