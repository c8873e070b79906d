"""A/B of an upload knob on the headline e2e leg: runs bench.py (e2e only) once per value of the environment
variable and prints value / ms per step / phases.   python tools/e2e_ab.py SPMB_EVENT_RAW_PCT 10 100 10 100"""
import json, os, subprocess, sys
var, vals = sys.argv[1], sys.argv[2:]
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for v in vals:
    env = dict(os.environ); env[var] = v
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--no-cpu", "--no-myers", "--no-c5", "--no-small",
                        "--e2e-steps", "6"], env=env, capture_output=True, text=True)
    try:
        e = json.loads(r.stdout.strip().splitlines()[-1])["e2e"]
        print(var, v, "->", round(e["value"], 2), "Gbp/s", round(e["ms_per_step"], 2), "ms",
              {k: round(x, 2) for k, x in e["phase_ms"].items()},
              "event chunks raw/packed", e["upload"]["event_raw_chunks"], e["upload"]["event_host_packed_chunks"], flush=True)
    except Exception as ex:
        print(var, v, "failed", ex, r.stderr[-500:], flush=True)
