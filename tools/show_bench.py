"""Print the essentials of bench.py JSON lines (files given on the command line)."""
import json
import sys

for f in sys.argv[1:]:
    try:
        d = json.loads([l for l in open(f) if l.startswith('{')][-1])
    except Exception as e:
        print(f, e)
        continue
    print(f, "value %.3g ms %.2f e2e_ms %.2f launches %d F %.2f act %.2f" % (
        d["value"], d["ms_per_step"], d["e2e"]["ms_per_step"], d["gpu_launches"], d["free_energy"],
        d.get("mean_active_latents_per_state", 0)))
    if d.get("roofline"):
        print(" roofline", {k: (round(v, 4) if isinstance(v, float) else v) for k, v in d["roofline"].items() if k not in ("peak_source", "simt")})
    print(" comm", d.get("comm"))
    print(" kernels", {k: (round(v["mean_ms"], 3), v["launches"]) for k, v in d["kernels"].items()})
    ad = d.get("at_density")
    if ad:
        print(" at_density ms %.2f act %.2f" % (ad["ms_per_step"], ad["mean_active_latents_per_state"]),
              {k: round(v["mean_ms"], 3) for k, v in ad["kernels"].items()})
    if d.get("graph"):
        print(" graph ms %.3f it/s %.1f" % (d["graph"]["ms_per_step"], d["graph"]["em_iters_per_sec"]))
    for k, v in (d.get("also") or {}).items():
        if v.get("graph"):
            print("  also", k, "graph ms %.3f it/s %.1f" % (v["graph"]["ms_per_step"], v["graph"]["em_iters_per_sec"]))
        print("  also", k, "ms %.2f value %.3g" % (v["ms_per_step"], v["value"]), {kk: round(vv["mean_ms"], 2) for kk, vv in v["kernels"].items()},
              "roof", round(v["roofline"]["frac"], 4) if v["roofline"] else None)
    if "cpu_baseline" in d:
        print(" cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["kind"], d["cpu_baseline"]["cores"])
