"""Pretty-print bench.py JSON lines: python tools/show_bench.py file.json [...]"""
import json
import sys

for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as ex:
        print(f, "unreadable:", ex)
        continue
    print(f, "impl", d.get("impl"), "value", round(d["value"], 2), "ms/step", round(d["ms_per_step"], 3), "e2e",
          round(d.get("e2e", {}).get("value", 0), 2), "n_gpus", d.get("n_gpus"), "launches", d.get("gpu_launches"))
    c = d.get("clocks") or {}
    print("  clocks", c.get("sm_mhz"), c.get("reasons"), " cfg", {k: d["config"].get(k) for k in ("num_rendered_R", "visible", "sum_n_contrib")})
    if "roofline" in d:
        r = d["roofline"]
        print("  roofline", r.get("kernel"), r["bound"], r["achieved"] and round(r["achieved"], 1), "/", round(r["peak"], 1), r["unit"],
              "frac", r["frac"] and round(r["frac"], 3), "share", round(r.get("share_of_step") or 0, 3))
    tot = 0
    for s in d.get("stages", []):
        tot += s["ms_per_step"]
        print("   %-18s n/step=%-3s ms/launch=%.4f ms/step=%.4f %s ach=%s frac=%s" % (
            s["stage"], s.get("launches_per_step", s.get("launches")), s["ms_per_launch"], s["ms_per_step"], s["bound"],
            s["achieved"] and round(s["achieved"], 1), s["frac"] and round(s["frac"], 3)))
    if tot:
        print("   stage sum ms/step %.3f" % tot)
    if "cpu_baseline" in d:
        print("  cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["kind"], d["cpu_baseline"]["cores"])
