"""Run bench.py under several env-var settings (A/B tuning of kernel knobs) and print one line each.
usage: python tools/sweep_env.py "K1=V1,K2=V2" "K1=V3" ...   (each arg is one configuration)"""
import json, os, subprocess, sys

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for cfg in sys.argv[1:]:
    env = dict(os.environ)
    for kv in filter(None, cfg.split(",")):
        k, v = kv.split("=")
        env[k] = v
    p = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--no-cpu-baseline", "--no-e2e", "--no-train-step", "--steps", "500", "--warmup", "20"],
                       env=env, capture_output=True, text=True)
    try:
        d = json.loads(p.stdout.strip().splitlines()[-1])
        print(cfg or "default", "| step %.1f us |" % (d["ms_per_step"] * 1e3), " ".join("%s %.1f us %.0f GB/s;" % (k["kernel"], k["us_per_launch"], k["achieved_gbps"]) for k in d["kernels"]), flush=True)
    except Exception as e:
        print(cfg, "FAILED", e, p.stderr[-500:], flush=True)
