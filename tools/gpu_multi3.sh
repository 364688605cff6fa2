# Multi-GPU pass: bench.py (headline + configs incl. the E-step with its all-reduce) on N ranks, no config 5.
V=${1:-vX}; N=${2:-2}; O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv,noheader | wc -l; nproc
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N ) > $O/r2_${V}_bench_${N}gpu.json 2> $O/r2_${V}_bench_${N}gpu.err; echo "bench rc=$?"; grep -v "LOG(INFO)" $O/r2_${V}_bench_${N}gpu.err | grep -v "^$" | tail -4
python - $O/r2_${V}_bench_${N}gpu.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("config2", d["n_gpus"], "GPUs", round(d["value"]/1e6,1), "M/s", round(d["ms_per_step"],3), "ms; e2e", round(d["e2e"]["value"]/1e6,1), "M/s", round(d["e2e"]["ms_per_step"],2), d.get("parity"))
for k,v in d.get("configs",{}).items():
    if "error" in v: print(k, v); continue
    if k=="pieces":
        print(k, round(v["value"]/1e6,1), "M/s", round(v["ms_per_step"],3), "ms", (v.get("parity") or {}).get("sentences_mismatching")); continue
    if k=="estep":
        for n in ("final","seed"): print(k,n, round(v[n]["value"]/1e6,1),"Mwords/s", round(v[n]["ms_per_step"],3), "kernel", round(v[n]["kernel_ms"],3), "allreduce", round(v[n]["allreduce_ms"],3), (v[n].get("parity") or {}).get("ok"))
    else: print(k, round(v["value"]/1e6,1), "M/s", round(v["ms_per_step"],3), "ms; e2e", round(v["e2e"]["value"]/1e6,1), round(v["e2e"]["ms_per_step"],2), (v.get("parity") or {}).get("sentences_mismatching"))
PY
