"""CPU: the C-ABI library builds, loads and exports every symbol include/pim_snappy.h declares;
struct layouts match the reference's; the product has no CPU fallback and never touches oracle/."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "pim-compression_b200")


def test_exports_match_header(P):
    hdr = open(os.path.join(ROOT, "include", "pim_snappy.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b([a-z_0-9]+)\s*\([^;{]*\)\s*;", hdr))
    declared -= {"defined"}
    assert {"snappy_compress_dpu", "snappy_decompress_dpu", "setup_compression", "setup_decompression"} <= declared
    from pim_compression_b200.api import EXPORTED_SYMBOLS
    assert declared == set(EXPORTED_SYMBOLS), declared ^ set(EXPORTED_SYMBOLS)
    L = P.lib()
    for name in declared:
        assert hasattr(L, name), name
    out = subprocess.run(["nm", "-D", "--defined-only", P.LIB_PATH], capture_output=True, text=True, check=True).stdout
    for name in declared:
        assert re.search(rf"\bT {name}\b", out), name


def test_synth_library_is_separate(P):
    """include/pim_synth.h is served by libpim_synth.so, not by the codec library (bench.py's reference arm loads only it)."""
    hdr = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "pim_synth.h")).read(), flags=re.S)
    declared = set(re.findall(r"\b([a-z_0-9]+)\s*\([^;{]*\)\s*;", hdr))
    assert declared == {"pim_snappy_synth_host", "pim_snappy_synth_device"}
    nm = lambda path: subprocess.run(["nm", "-D", "--defined-only", path], capture_output=True, text=True, check=True).stdout
    synth, codec = nm(P.api.SYNTH_LIB_PATH), nm(P.LIB_PATH)
    for name in declared:
        assert re.search(rf"\bT {name}\b", synth), name
        assert name not in codec
    assert "snappy_compress" not in synth and "snappy_decompress" not in synth


def test_out_of_domain_lengths_rejected_before_any_device_work(P):
    """Streams of 4 GiB and more do not exist in the format (varint32 lengths, snappy_compress.c:461):
    rejected with SNAPPY_BUFFER_TOO_SMALL, and so is a block size above 65536, before a buffer or a device is touched."""
    L = P.lib()
    inp, out, rt = P.HostBufferContext(), P.HostBufferContext(), P.ProgramRuntime()
    dummy = (C.c_uint8 * 16)()
    inp.buffer = inp.curr = C.addressof(dummy)
    out.buffer = out.curr = C.addressof(dummy)
    inp.max = out.max = 2**63
    inp.length = 1 << 32
    assert L.snappy_compress_dpu(C.byref(inp), C.byref(out), 32768, C.byref(rt)) == P.SNAPPY_BUFFER_TOO_SMALL
    inp.length = 10
    assert L.snappy_compress_dpu(C.byref(inp), C.byref(out), 65537, C.byref(rt)) == P.SNAPPY_BUFFER_TOO_SMALL
    assert L.snappy_compress_dpu(C.byref(inp), C.byref(out), 0, C.byref(rt)) == P.SNAPPY_BUFFER_TOO_SMALL
    assert bytes(dummy) == bytes(16)


def test_struct_layout(P):
    # dpu_snappy.h:37-55 on LP64: 5 x 8 bytes and 7 doubles
    assert C.sizeof(P.HostBufferContext) == 40
    assert [f[0] for f in P.HostBufferContext._fields_] == ["file_name", "buffer", "curr", "length", "max"]
    assert C.sizeof(P.ProgramRuntime) == 56
    assert (P.SNAPPY_OK, P.SNAPPY_INVALID_INPUT, P.SNAPPY_BUFFER_TOO_SMALL) == (0, 1, 2)


def test_pure_helpers(P):
    L = P.lib()
    assert L.pim_snappy_num_blocks(0, 32768) == 0
    assert L.pim_snappy_num_blocks(105438, 32768) == 4
    assert L.pim_snappy_num_blocks(1 << 30, 32768) == 32768
    for n, bs in ((1, 1), (1000, 7), (1 << 20, 4096), (1 << 20, 65536)):
        nb = L.pim_snappy_num_blocks(n, bs)
        assert L.pim_snappy_compress_bound(n, bs) >= n + n // 6 + 36 * nb
        assert L.pim_snappy_compress_scratch_bytes(n, bs) >= nb * (bs + bs // 6 + 36)


def test_no_cpu_fallback_sources():
    """The product never references the oracle or a host codec."""
    for dirpath, _, files in os.walk(PKG):
        for f in files:
            if f.endswith((".cu", ".cuh", ".c", ".h", ".py")) or f == "Makefile":
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle_" not in text and "liboracle" not in text and "_ref" not in text.replace("__restrict__", ""), f
    # the library exports no host codec symbol
    import pim_compression_b200 as P
    out = subprocess.run(["nm", "-D", "--defined-only", P.LIB_PATH], capture_output=True, text=True, check=True).stdout
    assert "snappy_compress_host" not in out and "snappy_decompress_host" not in out


def test_generators_deterministic(P):
    a = P.synth_host(P.SYNTH_MIX, 42, 3 * (1 << 20) + 5)
    b = P.synth_host(P.SYNTH_MIX, 42, 3 * (1 << 20) + 5)
    assert np.array_equal(a, b)
    assert not np.array_equal(a[:65536], P.synth_host(P.SYNTH_MIX, 43, 65536))
    # MIX = 1 MiB text then 1 MiB random
    assert set(np.unique(a[: 1 << 20]).tolist()) <= set(b" \nabcdefghijklmnopqrstuvwxyz")
    assert len(np.unique(a[1 << 20: 2 << 20])) == 256


@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="only meaningful on a machine without a GPU")
def test_fails_loudly_without_gpu(P):
    data = np.arange(1000, dtype=np.uint8)
    with pytest.raises((P.SnappyError, MemoryError)):
        P.compress(data)
    cli = os.path.join(PKG, "dpu_snappy")
    r = subprocess.run([cli, "-c", "-i", __file__, "-o", "/tmp/_pim_x"], capture_output=True, text=True)
    assert r.returncode != 0 and "GPU" in r.stderr


def test_block_offsets_walk(P, golden):
    txt, sn = golden["terror2"]
    hlen, dlen, bs, off = P.block_offsets(sn)
    assert (dlen, bs, len(off)) == (105438, 32768, 5) and hlen + int(off[-1]) == sn.size


def test_entry_points_fail_loudly_without_a_gpu(P, oracle):
    """No CPU fallback: on a machine without a usable sm_100 GPU the codec entry points return SNAPPY_INVALID_INPUT and
    produce nothing, for compression and for decompression (the buffers setup_* hand out for small calls are plain malloc
    and exist either way; pim_snappy_host_free releases them)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("this is the no-GPU behaviour")
    L = P.lib()
    data = np.frombuffer(b"no gpu, no codec " * 100, dtype=np.uint8).copy()
    inp, out, rt = P.HostBufferContext(), P.HostBufferContext(), P.ProgramRuntime()
    inp.buffer = inp.curr = data.ctypes.data
    inp.length, inp.max, out.max = data.size, 2**63, 2**63
    L.setup_compression(C.byref(inp), C.byref(out), C.byref(rt))
    assert out.buffer
    assert L.snappy_compress_dpu(C.byref(inp), C.byref(out), 32768, C.byref(rt)) == P.SNAPPY_INVALID_INPUT
    L.pim_snappy_host_free(out.buffer)
    with pytest.raises(P.SnappyError) as e:
        P.decompress(oracle.compress(data, 32768))
    assert e.value.status == P.SNAPPY_INVALID_INPUT
    assert P.device_count() <= 0
