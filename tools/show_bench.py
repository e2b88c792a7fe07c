import json,sys
for f in sys.argv[1:]:
    try:
        d=json.load(open(f))
    except Exception as e:
        print(f, "unreadable", e); continue
    print("==", f)
    print(" value %.1f e2e %.1f ms/step %.4f" % (d["value"], d["e2e"]["value"], d["ms_per_step"]))
    if "roofline" in d:
        r=d["roofline"]; print(" kernel_ms %.4f frac %.3f exch_us %s" % (r["kernel_ms"], r["frac"], r.get("exchange_us")))
    pc=d.get("parity_check"); 
    if pc: print(" parity ok=%s queries=%d shards_hit=%d paths=%s" % (pc["ok"], pc["queries"], pc["ranks_with_planted_hits"], [(p["path"][:28],p["ok"]) for p in pc["paths"]]))
    for k in ("batch256","c5","c2","c1","alt_exchange","batch1_shadow_bf16"):
        v=d.get(k)
        if v: print(" %s: %.1f %s" % (k, v["value"], {kk:(round(vv,3) if isinstance(vv,float) else vv) for kk,vv in v.items() if kk in ("ms_per_batch","ms_per_step","tensor_frac_tf32","hbm_frac","p50_us","p99_us","by_vector_p50_us")}))
    if d.get("cpu_baseline"):
        c=d["cpu_baseline"]; print(" cpu: %.2f qps cores=%s rows=%s extrap=%s faithful=%.3f" % (c["value"], c["cores"], c["sample_rows"], c["extrapolated"], c["faithful_1thread_value"]))
    print(" clocks", d.get("clocks"))
