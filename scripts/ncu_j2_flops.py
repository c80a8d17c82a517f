#!/usr/bin/env python
"""Executed-flop model of element_fd_kernel<3,27,64,J2> from one `ncu --set full --import-source on` capture.

The J2 return mapping (bracketed Newton, mimi::solvers::ScalarSolve) runs a data-dependent number of hardening-law
evaluations, so a static flop count does not exist. This script reads

  raw.csv     `ncu -i rep --page raw --csv`: smsp__sass_thread_inst_executed_op_{dadd,dmul,dfma}_pred_on of the launch
  source.csv  `ncu -i rep --page source --csv --print-source cuda,sass`: thread instructions per SASS instruction,
              attributed to source lines (libdevice pow / log bodies are attributed to the lines that call them)
  bench.json  the bench.py --workload c5 line of the same configuration (device counters: return mappings and
              hardening evaluations of one assembly)

and writes profiles/j2_flops.json:

  flop_per_qp = flop_fixed_per_qp + flop_per_hardening_eval * hardening_evals_per_qp

flop = DADD + DMUL + 2 DFMA thread instructions (predicated-on). The split is by source line: everything inside the
hardening laws and the scalar solve (mb2_math.cuh, hardening_evaluate .. j2_scalar_solve) is the variable part.
bench.py recomputes flop_per_qp of a run from that run's own counters.

usage: ncu_j2_flops.py raw.csv source.csv bench.json out.json [capture-name]"""
import csv
import json
import re
import sys


def num(x):
    try:
        return float(x.replace(",", ""))
    except ValueError:
        return 0.0


def main():
    raw, src, bench, out = sys.argv[1:5]
    capture = sys.argv[5] if len(sys.argv) > 5 else raw
    rows = list(csv.reader(open(raw, newline="")))
    hdr, vals = rows[0], rows[2]
    g = lambda k: num(vals[hdr.index(k)])
    cyc = g("smsp__cycles_elapsed.avg")
    ops = {k: g("smsp__sass_thread_inst_executed_op_%s_pred_on.sum.per_cycle_elapsed" % k) * cyc for k in ("dadd", "dmul", "dfma")}
    flop_total = ops["dadd"] + ops["dmul"] + 2.0 * ops["dfma"]
    # ---- source page: FP64 thread instructions by (file, line)
    rows = list(csv.reader(open(src, newline="")))
    math_lo = math_hi = None
    # line range of the variable part: from `struct AD1` to the end of j2_scalar_solve in mb2_math.cuh
    text = open(__file__.rsplit("/scripts/", 1)[0] + "/mimi_b200/csrc/mb2_math.cuh").read().split("\n")
    for i, l in enumerate(text, 1):
        if math_lo is None and l.startswith("struct AD1"):
            math_lo = i
        if l.startswith("// J2::PlasticStress<ACC>"):
            math_hi = i
            break
    cur_file, cur_line, hdr, col = None, None, None, {}
    var = {"DADD": 0.0, "DMUL": 0.0, "DFMA": 0.0}
    tot = {"DADD": 0.0, "DMUL": 0.0, "DFMA": 0.0}
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur_file = r[1]
            continue
        if r[0] == "Line No":
            hdr = r
            col = {}
            for i, h in enumerate(hdr):
                col.setdefault(h, i)
            continue
        if hdr is None or len(r) < len(hdr):
            continue
        if r[0] != "":
            cur_line = int(r[0]) if r[0].isdigit() else None
            continue
        txt = r[3].strip().split()
        if not txt:
            continue
        op = (txt[1] if txt[0].startswith("@") and len(txt) > 1 else txt[0]).split(".")[0]
        if op not in tot:
            continue
        n = num(r[col["Predicated-On Thread Instructions Executed"]])
        tot[op] += n
        if cur_file and cur_file.endswith("mb2_math.cuh") and cur_line is not None and math_lo <= cur_line < math_hi:
            var[op] += n
    flop_src = tot["DADD"] + tot["DMUL"] + 2.0 * tot["DFMA"]
    flop_var = var["DADD"] + var["DMUL"] + 2.0 * var["DFMA"]
    B = json.load(open(bench))
    roof = B["roofline"]
    n_eval = roof["stress_evaluations_per_assembly"]
    yf, hpe = roof["yielded_fraction"], roof["hardening_evals_per_return_mapping"]
    qp = B["config"]["qp_per_gpu"]
    hard_evals = n_eval * yf * hpe
    # the source page's totals may differ from the raw page's by replay effects: scale the split to the raw total
    share_var = flop_var / flop_src
    J = {
        "kernel": "element_fd_kernel<3,27,64,J2>",
        "capture": capture,
        "thread_inst": {k: ops[k] for k in ops},
        "flop_per_launch": flop_total,
        "qp_per_launch": qp,
        "flop_per_qp": flop_total / qp,
        "fused_flop_share": 2.0 * ops["dfma"] / flop_total,
        "yielded_fraction": yf,
        "hardening_evals_per_return_mapping": hpe,
        "hardening_evals_per_qp": hard_evals / qp,
        "variable_share": share_var,
        "flop_per_hardening_eval": share_var * flop_total / hard_evals,
        "flop_fixed_per_qp": (1.0 - share_var) * flop_total / qp,
        "fused_share_variable": 2.0 * var["DFMA"] / flop_var if flop_var else 0.0,
        "fused_share_fixed": 2.0 * (tot["DFMA"] - var["DFMA"]) / (flop_src - flop_var),
        "source_vs_raw_flop_ratio": flop_src / flop_total,
        "variable_part": "mb2_math.cuh lines %d-%d (hardening laws, return-mapping residual, scalar solve; libdevice pow / log bodies included)" % (math_lo, math_hi - 1),
    }
    json.dump(J, open(out, "w"), indent=1)
    print(json.dumps(J, indent=1))


if __name__ == "__main__":
    main()
