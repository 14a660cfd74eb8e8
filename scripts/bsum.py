import json, sys
for line in open(sys.argv[1]):
    if not line.startswith("{"): continue
    j = json.loads(line)
    r = j.get("roofline") or {}
    print("qps", round(j["value"]), "ms", round(j["ms_per_step"], 3), "ef", j["config"].get("ef_search"), "recall", j["config"].get("recall_at_10"),
          "| parity", j.get("parity_point"), "| build", {k: round(v, 2) if isinstance(v, float) else v for k, v in (j.get("build") or {}).items()},
          "| e2e", round(j["e2e"]["value"]), "| roof", round(r.get("frac", 0), 3), "kms", round(r.get("kernel_ms", 0), 3),
          "| cpu", (j.get("cpu_baseline") or {}).get("value"), "| clocks", j.get("clocks"))
