"""Opcode mix of one kernel from an ncu report's source page: executed warp instructions and stall samples per SASS opcode.
usage: python tools/sass_mix.py report.ncu-rep kernel-regex [top]"""
import collections, csv, subprocess, sys

rep, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 18
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{pat}"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = None
ex, st = collections.Counter(), collections.Counter()
for r in rows:
    if r and r[0] == "Address":
        if hdr is not None:
            break  # first matching launch only
        hdr = {h: i for i, h in enumerate(r)}
        continue
    if hdr is None or len(r) < len(hdr):
        continue
    parts = r[hdr["Source"]].split()
    op = parts[1] if parts and parts[0].startswith("@") else (parts[0] if parts else "?")
    op = op.split(".")[0] + ("." + op.split(".")[1] if op.startswith(("LDG", "STG", "LDS", "STS", "BAR", "MUFU", "F2FP")) and "." in op else "")
    ex[op] += int(r[hdr["Instructions Executed"]] or 0)
    st[op] += int(r[hdr["# Samples"]] or 0)
tot, tots = sum(ex.values()), sum(st.values())
print(f"total warp instructions {tot}, stall samples {tots}")
for op, n in ex.most_common(top):
    print(f"  {op:14s} {n:10d} {100 * n / tot:5.1f}%   samples {100 * st[op] / max(tots, 1):5.1f}%")
