import json, sys
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    print(d["config"], "%.3e ev/s" % d["events_per_s"], "ok" if d["rows_ok"] == d["n"] else "ROWS NOT OK", {k: round(v, 3) for k, v in d.items() if k.startswith("frac") or k == "hbm_frac"}, d.get("cpu_oracle", ""))
