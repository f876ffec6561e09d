import json,sys
for f in sys.argv[1:]:
    try:
        d=json.load(open(f)); r=d["roofline"]
        print(f, "ms/step %.3f"%d["ms_per_step"], "K1b %.3f"%r["kernel_ms"], "K1a %.3f"%r["k_minimizer_partition_ms"], {k:round(v,3) for k,v in r["stage_ms"].items()}, "e2e %.1f"%(d["e2e"]["ms_per_step"] if d.get("e2e") else -1))
    except Exception as e: print(f, "ERR", e)
