V=${1:-vX}; O=gpurun_out; mkdir -p $O
( time timeout 900 python bench.py ) > $O/r2_${V}_bench.json 2> $O/r2_${V}_bench.err; echo "bench rc=$?"; grep -v "LOG(INFO)" $O/r2_${V}_bench.err | tail -6
python - $O/r2_${V}_bench.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("config2", round(d["value"]/1e6,1), "M/s", round(d["ms_per_step"],3), "ms; e2e", round(d["e2e"]["ms_per_step"],2), d["parity"]["sentences_mismatching"], "frac", round(d["roofline"]["frac"],4), "traffic", d["roofline"]["traffic"])
for k,v in d.get("configs",{}).items():
    if "error" in v: print(k, v); continue
    if k=="pieces": print(k, round(v["value"]/1e6,1), "M/s", round(v["ms_per_step"],3), "ms", v.get("stage_ms_per_step"), v.get("parity"), "wall", v.get("wall_s"))
    elif k=="estep":
        for n in ("final","seed"): print(k,n, round(v[n]["value"]/1e6,1),"Mwords/s", round(v[n]["ms_per_step"],3), v[n]["parity"]["ok"])
    else: print(k, round(v["value"]/1e6,1), "M/s", round(v["ms_per_step"],3), "ms; e2e", round(v["e2e"]["ms_per_step"],2), v["parity"]["sentences_mismatching"], "wall", v.get("wall_s"))
PY
