"""Developer helper: run bench.py and print the headline numbers + per-kernel milliseconds."""
import json, subprocess, sys
out = subprocess.run([sys.executable, "bench.py"] + sys.argv[1:], capture_output=True, text=True)
line = [l for l in out.stdout.splitlines() if l.startswith("{")]
if not line:
    print(out.stdout[-2000:], out.stderr[-2000:]); sys.exit(1)
d = json.loads(line[-1])
print("fps %.0f  e2e %.0f  ms/step %.3f  kernels %s  roof %s %.4f" % (d["value"], d["e2e"]["value"], d["ms_per_step"],
      {k: round(v, 3) for k, v in d["roofline"]["kernel_ms_per_launch"].items()}, d["roofline"]["kernel"], d["roofline"]["frac"]))
