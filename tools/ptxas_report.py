#!/usr/bin/env python
"""Registers / spills / stack of every solve-kernel instantiation, from the ptxas -v logs `build.py -v` leaves in
benchmarking-mpc-solvers_b200/build/*.log.  usage: ptxas_report.py [substring]"""
import glob
import os
import re
import subprocess
import sys

root = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "benchmarking-mpc-solvers_b200", "build")
pat = sys.argv[1] if len(sys.argv) > 1 else "mppi_solve_kernel"
for log in sorted(glob.glob(os.path.join(root, "*.log"))):
    txt = open(log).read().split("\n")
    rows = []
    for i, l in enumerate(txt):
        m = re.search(r"Compiling entry function '(\S+)'", l)
        if not m or pat not in m.group(1):
            continue
        blob = " ".join(txt[i:i + 4])
        regs = re.search(r"Used (\d+) registers", blob)
        spill = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", blob)
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = re.sub(r"\(mpc::MppiArgs.*", "", name).replace("mpc::", "").replace("(bool)", "").replace("(int)", "")
        rows.append((name, regs.group(1) if regs else "?", spill.groups() if spill else ("?",) * 3))
    if rows:
        print("==", os.path.basename(log))
        for name, r, (st, ss, sl) in rows:
            print("  %-70s regs %3s  stack %3s  spill st/ld %s/%s" % (name[:70], r, st, ss, sl))
