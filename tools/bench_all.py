"""Run bench.py for every workload and print a per-kernel roofline table (used to fill profiles/ and DESIGN.md)."""
import json, os, subprocess, sys

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root)
import bench

extra = sys.argv[1:]
for name in bench.WORKLOADS:
    steps = "200" if name == "cross_entropy" else "1000"
    p = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--workload", name, "--steps", steps, "--warmup", "20", "--no-cpu-baseline"] + extra,
                       capture_output=True, text=True)
    try:
        d = json.loads(p.stdout.strip().splitlines()[-1])
        with open(os.path.join(root, "gpurun_out", f"bench_{name}.json"), "w") as f:
            f.write(json.dumps(d) + "\n")
        for k in d["kernels"]:
            print("%-22s %9.1f us  %7.0f GB/s  %5.1f%% of measured %5.1f%% of 8 TB/s   (%.1f MB)" % (
                k["kernel"], k["us_per_launch"], k["achieved_gbps"], 100 * k["achieved_gbps"] / d["roofline"]["peak"], 100 * k["achieved_gbps"] / 8000, k["algorithmic_bytes"] / 1e6), flush=True)
        print("   step %.1f us, value %.0f GB/s, e2e %s, clocks %s" % (d["ms_per_step"] * 1e3, d["value"], d["e2e"] and round(d["e2e"]["value"], 1), d["clocks"]), flush=True)
    except Exception as e:
        print(name, "FAILED", e, p.stdout[-300:], p.stderr[-800:], flush=True)
