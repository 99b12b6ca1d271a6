"""Host logic of tools/run_tests.py (the GPU counterpart of scripts/asplos21/run_tests.py): the CLI's timing
lines are parsed as the reference's parse_output_file.py does and land in the reference's CSV layout."""
import csv
import importlib.util
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("run_tests_tool", os.path.join(ROOT, "tools", "run_tests.py"))
T = importlib.util.module_from_spec(spec)
spec.loader.exec_module(T)

SAMPLE = """Using input file a.txt
Using output file a.snappy
Compressed 52525 bytes to: a.snappy
Compression ratio: 0.501840
Pre-processing time: 0.000100
Alloc time: 0.002000
Load time: 0.030000
Copy in time: 0.000400
Host time: 0.000500
Copy out time: 0.000060
Free time: 0.000000
"""


def test_parse_and_breakdown_layout(tmp_path):
    r = T.parse_cli_output(SAMPLE)
    assert r["prepare"] == 0.0001 and r["run"] == 0.0005 and r["copy_out"] == 0.00006 and r["bytes"] == 52525
    assert abs(r["ratio"] - 0.50184) < 1e-9
    rows = T.breakdown_rows({2: r, 1: r})
    assert [row[-1] for row in rows] == [1, 2] and rows[0][:7] == [0.0001, 0.002, 0.03, 0.0004, 0.0005, 0.00006, 0.0]
    p = tmp_path / "x_compression_breakdown.csv"
    T.write_csv(p, ["prepare", "alloc", "load", "copy_in", "run", "copy_out", "free", "dpus"], rows)
    got = list(csv.reader(open(p)))
    assert got[0] == ["prepare", "alloc", "load", "copy_in", "run", "copy_out", "free", "dpus"] and len(got) == 3


def test_missing_line_is_an_error():
    import pytest
    with pytest.raises(ValueError):
        T.parse_cli_output(SAMPLE.replace("Copy in time", "Copy-in time"))


def test_host_time_lookup(tmp_path):
    d = tmp_path / "compression"
    d.mkdir()
    (d / "alice_host.txt").write_text("Compressed 1 bytes to: x\nHost time: 0.125000\n")
    assert T.host_time(str(tmp_path), "compression", "alice") == 0.125
    assert T.host_time(str(tmp_path), "decompression", "alice") is None and T.host_time(None, "compression", "alice") is None
