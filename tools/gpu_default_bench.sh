#!/bin/bash
# the driver's end-of-round commands: smoke, default bench (timed), reference arm
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
T0=$(date +%s)
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
echo "bench.py rc=$? wall $(( $(date +%s) - T0 )) s"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_default.json").read().strip().splitlines()[-1])
def walk(o,path=""):
    if isinstance(o,dict):
        for k,v in o.items():
            if k=="error": print("ERROR at",path,v)
            walk(v,path+"/"+k)
    elif isinstance(o,list):
        for i,v in enumerate(o): walk(v,path+f"[{i}]")
walk(d)
print({k:(v if not isinstance(v,(dict,list)) else "...") for k,v in d.items()})
print("e2e",d["e2e"]["value"],"u8",d["e2e_u8"]["value"],"plugin",d["e2e_plugin"]["value"])
print("roofline",{k:d["roofline"].get(k) for k in ("bound","achieved","peak","frac","traffic","avg_launch_ms","launches_per_step","sum_launch_ms","tensor_frac_of_probe")})
print("nets",{k:(v.get("img_s"),v.get("ms_per_step")) for k,v in d["nets"].items()})
print("gemm",json.dumps(d["gemm"])[:900])
print("conv1",json.dumps(d["conv1"])[:400])
print("cpu",d["cpu_baseline"])
PY
T0=$(date +%s)
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 2>/dev/null | tail -1 | cut -c1-300
echo "reference arm wall $(( $(date +%s) - T0 )) s"
