"""The reference's OWN main() (/root/reference/snappy/dpu_snappy.c, unmodified, compiled with the reference's own
headers by oracle/Makefile target `refmain`) linked against libpim_snappy.so: the drop-in boundary exercised by the
caller it was written for.  CPU part: the binary binds the four entry points from the product library and fails
cleanly without a GPU.  GPU part: the reference's `make test_dpu` (snappy/Makefile:46,58-60) over the six fixtures,
and compression with the unmodified main, which relies on the library writing output->file_name itself
(snappy_compress.c:626-627,697-703; dpu_snappy.c:209) -- $PIM_SNAPPY_WRITE_FILE=1."""
import os
import re
import subprocess

import numpy as np
import pytest

from oracle.oracle import REFMAIN_BIN


def _need_binary():
    if not os.path.exists(REFMAIN_BIN):
        pytest.skip("oracle/_ref/dpu_snappy_refmain is built only where /root/reference exists (it travels to the GPU box)")


def test_reference_main_binds_the_product_library(P, oracle):
    _need_binary()
    und = subprocess.run(["nm", "-D", "--undefined-only", REFMAIN_BIN], capture_output=True, text=True, check=True).stdout
    for sym in ("snappy_compress_dpu", "snappy_decompress_dpu", "setup_compression", "setup_decompression"):
        assert re.search(rf"\bU {sym}\b", und), sym
    needed = subprocess.run(["readelf", "-d", REFMAIN_BIN], capture_output=True, text=True, check=True).stdout
    assert "libpim_snappy.so" in needed


def test_reference_main_fails_cleanly_without_a_gpu(P, tmp_path, golden):
    _need_binary()
    import torch
    if torch.cuda.is_available():
        pytest.skip("this is the no-GPU behaviour")
    f = tmp_path / "terror2.snappy"
    golden["terror2"][1].tofile(f)
    r = subprocess.run([REFMAIN_BIN, "-d", "-i", str(f), "-o", str(tmp_path / "out")], capture_output=True, text=True)
    assert r.returncode != 0 and not (tmp_path / "out").exists()  # no CPU fallback: nothing is decoded


@pytest.mark.gpu
def test_reference_make_test_dpu_over_the_fixtures(P, tmp_path, golden):
    _need_binary()
    for name, (txt, sn) in golden.items():
        f, out = tmp_path / (name + ".snappy"), tmp_path / (name + ".dpu_uncompressed")
        sn.tofile(f)
        r = subprocess.run([REFMAIN_BIN, "-d", "-i", str(f), "-o", str(out)], capture_output=True, text=True)
        assert r.returncode == 0, (name, r.stderr)
        assert f"Decompressed {txt.size} bytes to: {out}" in r.stdout
        assert np.array_equal(np.fromfile(out, dtype=np.uint8), txt), name  # `cmp` of snappy/Makefile:59


@pytest.mark.gpu
def test_reference_main_compresses_with_the_library_writing_the_file(P, tmp_path, golden):
    _need_binary()
    env = dict(os.environ, PIM_SNAPPY_WRITE_FILE="1")
    for name, (txt, sn) in golden.items():
        f, out = tmp_path / (name + ".txt"), tmp_path / (name + ".snappy")
        txt.tofile(f)
        r = subprocess.run([REFMAIN_BIN, "-d", "-c", "-i", str(f), "-o", str(out)], capture_output=True, text=True, env=env)
        assert r.returncode == 0, (name, r.stderr)
        assert np.array_equal(np.fromfile(out, dtype=np.uint8), sn), name  # byte-identical to the reference's own .snappy
    # without the option the unmodified main writes nothing (dpu_snappy.c:209 skips write_output_host): documented behaviour
    out = tmp_path / "nofile.snappy"
    r = subprocess.run([REFMAIN_BIN, "-d", "-c", "-i", str(tmp_path / "alice.txt"), "-o", str(out)], capture_output=True, text=True,
                       env={k: v for k, v in os.environ.items() if k != "PIM_SNAPPY_WRITE_FILE"})
    assert r.returncode == 0 and not out.exists()
