#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` export: totals, stall mix and per-address-bucket
executed instructions / active lanes.  usage: tools/sass_profile.py file.csv [bucket]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
B = int(sys.argv[2]) if len(sys.argv) > 2 else 400
hdr, data = rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}


def f(r, k):
    try:
        return float(r[ix[k]])
    except (ValueError, KeyError):
        return 0.0


n = len(data)
ie = sum(f(r, "Instructions Executed") for r in data)
te = sum(f(r, "Thread Instructions Executed") for r in data)
ts = sum(f(r, "# Samples") for r in data)
print(f"{n} SASS instructions; {ie:.3e} warp-instructions executed; {te / ie:.1f} lanes active; "
      f"{ts:.0f} samples")
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = {h: sum(f(r, h) for r in data) for h in stalls}
print({k: round(100 * v / ts, 1) for k, v in sorted(tot.items(), key=lambda x: -x[1])[:8]})
print("  addr   Minstr  lanes  samples%  no_inst%  barrier%")
for b in range(0, n, B):
    seg = data[b:b + B]
    e = sum(f(r, "Instructions Executed") for r in seg)
    t = sum(f(r, "Thread Instructions Executed") for r in seg)
    s = sum(f(r, "# Samples") for r in seg)
    ni = sum(f(r, "stall_no_inst") for r in seg)
    ba = sum(f(r, "stall_barrier") for r in seg)
    print(f"{b:6d} {e / 1e6:8.2f} {t / max(e, 1):6.1f} {100 * s / ts:8.1f} "
          f"{100 * ni / max(s, 1):8.1f} {100 * ba / max(s, 1):8.1f}")
