"""A/B helper: runs bench.py (headline only) under several environment settings and prints one line each.
usage: python profiles/ab.py NAME=VALUE[,NAME=VALUE] ...   (use '-' for the default environment)"""
import json, os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for spec in sys.argv[1:]:
    env = dict(os.environ)
    if spec != "-":
        for kv in spec.split(","):
            k, v = kv.split("=")
            env[k] = v
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--no-extras", "--no-dream", "--no-cpu-baseline"],
                         capture_output=True, text=True, env=env)
    try:
        d = json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][-1])
    except Exception:
        print(spec, "FAILED", out.stderr[-2000:])
        continue
    kc = {k: "%.3f/%d" % (v["ms_per_step"], v["launches_per_step"]) for k, v in d["kernel_classes"].items()}
    print("%-34s value %7.0f  e2e %7.0f  ms/step %.3f  %s" % (spec, d["value"], d["e2e"]["value"], d["ms_per_step"], kc), flush=True)
