#!/usr/bin/env python
"""Executed fp32 flop and DRAM bytes of ONE captured kernel launch, from an ncu report taken with --set full --import-source on:
    ncu -i X.ncu-rep --page source --csv > src.csv ; ncu -i X.ncu-rep --page raw --csv > raw.csv
    python tools/ncu_flops.py src.csv raw.csv [--write profiles/traffic.json "<source note>"]
flop = sum over SASS instructions of predicated-on thread executions x {FFMA2 4, FADD2/FMUL2 2, FFMA 2, FADD/FMUL 1}
(ncu's sass_thread_inst_executed_op_ffma counter does not see FFMA2)."""
import csv
import json
import sys

W = {"FFMA2": 4, "FADD2": 2, "FMUL2": 2, "FFMA": 2, "FADD": 1, "FMUL": 1}


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    col = {n: i for i, n in enumerate(rows[h])}
    flop, per = 0, {}
    for r in rows[h + 1:]:
        if len(r) != len(rows[h]):
            continue
        t = r[col["Source"]].split()
        op = t[1] if t and t[0].startswith("@") and len(t) > 1 else (t[0] if t else "")
        op = op.split(".")[0]
        if op in W:
            n = int(r[col["Predicated-On Thread Instructions Executed"]] or 0)
            flop += W[op] * n
            per[op] = per.get(op, 0) + n
    raw = list(csv.reader(open(sys.argv[2])))
    hr = raw[0]
    vals = raw[2] if len(raw) > 2 else raw[1]
    g = lambda name: float(vals[hr.index(name)].replace(",", ""))
    unit = lambda name: raw[1][hr.index(name)]
    def to_bytes(name):
        v, u = g(name), unit(name)
        return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
    dram = to_bytes("dram__bytes_read.sum") + to_bytes("dram__bytes_write.sum")
    out = {"fused_kernel_dram_bytes_per_launch": dram, "fused_kernel_fp32_flop_executed_per_launch": flop,
           "thread_instructions": per, "gpu_time_ns": g("gpu__time_duration.sum") * {"ns": 1, "us": 1e3, "ms": 1e6}.get(unit("gpu__time_duration.sum"), 1)}
    print(json.dumps(out))
    if "--write" in sys.argv:
        i = sys.argv.index("--write")
        out["source"] = sys.argv[i + 2]
        out["algorithmic_def_b_bytes"] = 881606656
        json.dump(out, open(sys.argv[i + 1], "w"), indent=1)


if __name__ == "__main__":
    main()
