"""Aggregate an ncu launch list (--metrics gpu__time_duration.sum --csv) per kernel: launches, total time, share.
usage: python tools/launch_shares.py launches.csv [top]"""
import collections, csv, re, sys

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = [r for r in csv.reader(open(path, errors="replace")) if r]
hdr_i = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
hdr = {h: i for i, h in enumerate(rows[hdr_i])}
tot = collections.Counter(); cnt = collections.Counter()
for r in rows[hdr_i + 1:]:
    if len(r) <= hdr["Metric Value"] or r[hdr["Metric Name"]] != "gpu__time_duration.sum":
        continue
    name = r[hdr["Kernel Name"]].replace("void ", "").replace("at::native::", "").replace("(anonymous namespace)::", "")[:100]
    val = float(r[hdr["Metric Value"]].replace(",", ""))
    unit = r[hdr["Metric Unit"]]
    us = val / 1e3 if unit in ("ns", "nsecond") else (val if unit in ("us", "usecond") else val * 1e3)
    tot[name] += us; cnt[name] += 1
total = sum(tot.values())
print(f"{sum(cnt.values())} launches, {total / 1e3:.2f} ms of kernel time (serialised, cold cache)")
ours = sum(v for k, v in tot.items() if any(t in k for t in ("b200k", "rmsnorm", "swiglu", "rope_", "ce_", "colsum", "pw_kernel", "layernorm", "softmax_", "bias_act", "pointwise", "uniform_philox")))
print(f"b200 kernels: {ours / 1e3:.2f} ms = {100 * ours / total:.1f} %")
for name, us in tot.most_common(top):
    print(f"  {us / 1e3:9.3f} ms {100 * us / total:5.1f}%  x{cnt[name]:<5d} {name}")
