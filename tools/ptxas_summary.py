#!/usr/bin/env python3
"""Summarise `-Xptxas -v` output of the per-object build logs: kernel, registers, stack, spills."""
import re
import subprocess
import sys
import glob
import os

def main():
    pat = sys.argv[1] if len(sys.argv) > 1 else ""
    logs = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..",
                                         "singularity-eos_b200", "lib", "*.o.log")))
    for log in logs:
        txt = open(log).read()
        ents = re.split(r"ptxas info\s+: Compiling entry function '", txt)[1:]
        for e in ents:
            name = e.split("'")[0]
            try:
                dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
            except OSError:
                dem = name
            dem = re.sub(r"\(.*", "", dem).replace("void sgx::", "")
            if pat and pat not in dem:
                continue
            regs = re.search(r"Used (\d+) registers", e)
            stack = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", e)
            smem = re.search(r"(\d+) bytes smem", e)
            print(f"{os.path.basename(log)[:-6]:18s} {dem[:70]:70s} regs={regs.group(1) if regs else '?':>3} "
                  f"stack={stack.group(1) if stack else '?':>5} spill={stack.group(2) if stack else '?'}/{stack.group(3) if stack else '?'}"
                  f" smem={smem.group(1) if smem else 0}")

if __name__ == "__main__":
    main()
