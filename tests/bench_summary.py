"""Print a compact per-kernel summary of bench.py JSON lines (helper for GPU runs)."""
import json
import sys

for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception:
        print(f, "ERR", open(f).read()[-400:])
        continue
    print(f, "ms/step %.3f" % d["ms_per_step"], "G el-steps/s %.3f" % (d["value"] / 1e9),
          "step frac %.3f" % d["roofline"]["step"]["frac"])
    print("    " + "  ".join("%s %.3f" % (k, v["avg_ms"]) for k, v in d["roofline"]["kernels"].items()))
