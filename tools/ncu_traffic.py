#!/usr/bin/env python
"""Summarise DRAM traffic per kernel from an `ncu --set full` capture.

usage: tools/ncu_traffic.py raw.csv NCW SOURCE_LABEL > profiles/ncu_traffic.json
  raw.csv = `ncu -i capture.ncu-rep --page raw --csv`, NCW = codewords per launch in the captured run.
Output: {bench kernel name: {"bytes_per_codeword": dram read + write per codeword (mean over launches), ...}}
which bench.py scales to its own batch size for `roofline.traffic`.
"""
import csv
import json
import sys

NAMES = [("rs255_fft_kernel<1>", "rs255_fft_kernel<ENCODE>"), ("rs255_fft_kernel<(int)1>", "rs255_fft_kernel<ENCODE>"),
         ("rs255_fft_kernel<0>", "rs255_fft_kernel<SCREEN>"), ("rs255_fft_kernel<(int)0>", "rs255_fft_kernel<SCREEN>"),
         ("rs255_parity_kernel", "rs255_parity_kernel"), ("rs255_rows_kernel", "rs255_rows_kernel"),
         ("rs255_fft_compact_kernel", "rs255_compact_kernel"), ("rs255_solve_kernel", "rs255_solve_kernel"),
         ("rs255_correct_kernel", "rs255_correct_kernel")]
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    ncw = int(sys.argv[2])
    label = sys.argv[3] if len(sys.argv) > 3 else sys.argv[1]
    h, units = rows[0], rows[1]
    ix = {k: i for i, k in enumerate(h)}
    acc = {}
    for r in rows[2:]:
        kn = r[ix["Kernel Name"]]
        for pat, name in NAMES:
            if pat in kn:
                rd = float(r[ix["dram__bytes_read.sum"]]) * UNIT[units[ix["dram__bytes_read.sum"]]]
                wr = float(r[ix["dram__bytes_write.sum"]]) * UNIT[units[ix["dram__bytes_write.sum"]]]
                t = float(r[ix["gpu__time_duration.sum"]])
                if name == "rs255_solve_kernel" and rd + wr < 1e6:
                    break  # a launch on an empty queue (clean-decode pass)
                alu = float(r[ix["sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"]])
                a = acc.setdefault(name, {"rd": 0.0, "wr": 0.0, "n": 0, "t": 0.0, "alu": 0.0})
                a["rd"] += rd; a["wr"] += wr; a["n"] += 1; a["t"] += t; a["alu"] += alu
                break
    out = {}
    for name, a in acc.items():
        n = a["n"]
        out[name] = {"bytes_per_codeword": (a["rd"] + a["wr"]) / n / ncw, "read_per_codeword": a["rd"] / n / ncw,
                     "write_per_codeword": a["wr"] / n / ncw, "launches_captured": n, "codewords_per_launch": ncw,
                     "alu_pipe_pct": a["alu"] / n,
                     "ncu_duration_" + units[ix["gpu__time_duration.sum"]]: a["t"] / n, "source": label}
    json.dump(out, sys.stdout, indent=1)
    print()


if __name__ == "__main__":
    main()
