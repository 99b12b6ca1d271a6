#!/usr/bin/env python3
"""GPU counterpart of the reference's sweep driver (snappy/scripts/asplos21/run_tests.py).

Runs the dpu_snappy CLI of this repository over files x GPU counts (x block sizes) and writes the CSV
files the reference's chart scripts read, with the same column layout:

  results/<file>_{compression,decompression}_breakdown.csv   prepare,alloc,load,copy_in,run,copy_out,free,dpus
                                                             (run_tests.py:134 / chart_breakdown.py; `dpus` = GPUs here)
  results/{compression,decompression}_speedup_dpu.csv        version,time,dpus   (run_tests.py:83-97; `time` = host / GPU total)
  results/compression_vs_blksize.csv                         file,block_size,compressed_bytes,ratio,seconds

The per-phase times are parsed from the seven timing lines the CLI prints (dpu_snappy.c:212-227), exactly as
parse_output_file.py does.  `run` is the CLI's "Host time" line, which for the device path holds the kernels'
span (the reference derives it from DPU cycle counts instead).  The host baseline of the speed-up files comes
from --host-results DIR: output files of the reference's CLI run WITHOUT -d, named <file>_host.txt in
DIR/compression and DIR/decompression (this repository ships no CPU codec); without it `time` is seconds.
"""
import argparse
import csv
import os
import pathlib
import re
import subprocess
import sys

ROOT = pathlib.Path(__file__).resolve().parent.parent
CLI = ROOT / "pim-compression_b200" / "dpu_snappy"
PHASES = (("prepare", "Pre-processing time"), ("alloc", "Alloc time"), ("load", "Load time"), ("copy_in", "Copy in time"),
          ("run", "Host time"), ("copy_out", "Copy out time"), ("free", "Free time"))


def parse_cli_output(text):
    """The seven timing lines (seconds) plus the byte count / ratio lines of one CLI run."""
    out = {}
    for key, label in PHASES:
        m = re.search(rf"^{re.escape(label)}: ([0-9.eE+-]+)\s*$", text, re.M)
        if not m:
            raise ValueError(f"no '{label}' line in the CLI output")
        out[key] = float(m.group(1))
    m = re.search(r"^(Compressed|Decompressed) (\d+) bytes to: ", text, re.M)
    if m:
        out["bytes"] = int(m.group(2))
    m = re.search(r"^Compression ratio: ([0-9.eE+-]+)", text, re.M)
    if m:
        out["ratio"] = float(m.group(1))
    return out


def breakdown_rows(runs):
    """runs: {gpus: parsed output} -> rows in the reference's breakdown layout."""
    return [[r["prepare"], r["alloc"], r["load"], r["copy_in"], r["run"], r["copy_out"], r["free"], g] for g, r in sorted(runs.items())]


def write_csv(path, header, rows):
    path.parent.mkdir(parents=True, exist_ok=True)
    with open(path, "w", newline="") as f:
        w = csv.writer(f, delimiter=",")
        w.writerow(header)
        w.writerows(rows)


def host_time(results_dir, mode, name):
    """'Host time' of a reference host run (parse_output_file.py:get_host_runtime), or None."""
    if not results_dir:
        return None
    p = pathlib.Path(results_dir) / mode / f"{name}_host.txt"
    if not p.exists():
        return None
    m = re.search(r"^Host time: ([0-9.eE+-]+)", p.read_text(), re.M)
    return float(m.group(1)) if m else None


def run_cli(args, log_path):
    r = subprocess.run([str(CLI)] + args, capture_output=True, text=True)
    log_path.parent.mkdir(parents=True, exist_ok=True)
    log_path.write_text(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError(f"{' '.join(args)} failed: {r.stderr.strip()}")
    return parse_cli_output(r.stdout)


def main():
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("-f", "--files", nargs="+", required=True, help="inputs (uncompressed); name = file stem")
    ap.add_argument("-g", "--gpus", nargs="+", type=int, default=[1], help="GPU counts to sweep (the reference sweeps DPUs)")
    ap.add_argument("-b", "--block-sizes", nargs="+", type=int, default=[32768])
    ap.add_argument("-o", "--results", default="results")
    ap.add_argument("--host-results", default=None)
    ap.add_argument("--reps", type=int, default=3, help="runs per point; the fastest total is kept")
    a = ap.parse_args()
    res = pathlib.Path(a.results)
    for mode in ("compression", "decompression"):  # the reference's scripts expect these to exist, too
        (res / mode).mkdir(parents=True, exist_ok=True)
    speed = {"compression": [["host", "1", "0"]], "decompression": [["host", "1", "0"]]}
    blk_rows = []
    for f in a.files:
        name = pathlib.Path(f).stem
        runs = {"compression": {}, "decompression": {}}
        for g in a.gpus:
            comp = res / "compression" / f"{name}_gpus={g}.snappy"
            back = res / "decompression" / f"{name}_gpus={g}.out"
            for mode, args in (("compression", ["-d", "-c", "-g", str(g), "-b", str(a.block_sizes[0]), "-i", f, "-o", str(comp)]),
                               ("decompression", ["-d", "-g", str(g), "-i", str(comp), "-o", str(back)])):
                best = None
                for _ in range(a.reps):
                    r = run_cli(args, res / mode / f"{name}_gpus={g}.txt")
                    tot = sum(r[k] for k, _ in PHASES)
                    if best is None or tot < best[0]:
                        best = (tot, r)
                runs[mode][g] = best[1]
                h = host_time(a.host_results, mode, name)
                speed[mode].append([name, h / best[0] if h else best[0], g])
            if open(f, "rb").read() != open(back, "rb").read():
                raise RuntimeError(f"round trip of {f} on {g} GPU(s) differs from the input")
        for mode in runs:
            write_csv(res / f"{name}_{mode}_breakdown.csv", ["prepare", "alloc", "load", "copy_in", "run", "copy_out", "free", "dpus"],
                      breakdown_rows(runs[mode]))
        for bs in a.block_sizes:
            out = res / "compression" / f"{name}_bs={bs}.snappy"
            r = run_cli(["-d", "-c", "-g", str(a.gpus[0]), "-b", str(bs), "-i", f, "-o", str(out)], res / "compression" / f"{name}_bs={bs}.txt")
            blk_rows.append([name, bs, r.get("bytes"), r.get("ratio"), sum(r[k] for k, _ in PHASES)])
    for mode in speed:
        write_csv(res / f"{mode}_speedup_dpu.csv", ["version", "time", "dpus"], speed[mode])
    write_csv(res / "compression_vs_blksize.csv", ["file", "block_size", "compressed_bytes", "ratio", "seconds"], blk_rows)
    print(f"wrote {res}/*.csv")
    return 0


if __name__ == "__main__":
    sys.exit(main())
