# Multi-GPU pass without bench.py (the driver runs that itself): config 5 streamed and the E-step with its all-reduce.
V=${1:-vX}; N=${2:-8}; O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv,noheader | wc -l; nproc
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tools/config5.py --gpus $N --blocks 100 ) > $O/r2_${V}_config5_${N}gpu.json 2> $O/r2_${V}_config5_${N}gpu.err; echo "config5 rc=$?"; cut -c1-1700 $O/r2_${V}_config5_${N}gpu.json; grep -v "LOG(INFO)" $O/r2_${V}_config5_${N}gpu.err | grep -v "^$" | tail -4
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 tools/estep_bench.py --gpus $N ) > $O/r2_${V}_estep_${N}gpu.json 2> $O/r2_${V}_estep_${N}gpu.err; echo "estep rc=$?"; grep -v "LOG(INFO)" $O/r2_${V}_estep_${N}gpu.err | grep -v "^$" | tail -3
python - $O/r2_${V}_estep_${N}gpu.json <<'PY'
import json,sys
v=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
for n in ("final","seed"): print(n, v["n_gpus"], "GPUs", round(v[n]["value"]/1e6,1),"Mwords/s", round(v[n]["ms_per_step"],3), "kernel", round(v[n]["kernel_ms"],3), "allreduce", round(v[n]["allreduce_ms"],3), "share", round(v[n]["allreduce_share_of_step"],3), (v[n].get("parity") or {}).get("ok"))
PY
