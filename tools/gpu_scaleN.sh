#!/bin/bash
# One multi-GPU gpurun call (gpurun --gpus N): D2H ceiling of the box, then bench.py at N ranks.  usage: tools/gpu_scaleN.sh N TAG
N=${1:-2}; TAG=${2:-scale$N}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=index,name --format=csv > $OUT/smi.txt 2>&1; nproc > $OUT/nproc.txt
for n in ${D2H_NS-1 $N}; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 tools/d2h_ceiling.py 2>$OUT/d2h_$n.err | tee -a $OUT/d2h_ceiling.jsonl
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 3 > $OUT/bench_n$N.json 2> $OUT/bench_n$N.err; echo "bench N=$N rc=$?" | tee -a $OUT/rc.txt
python - $OUT/bench_n$N.json <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print({k: d[k] for k in ("n_gpus", "value", "ms_per_step")}, "e2e", round(d["e2e"]["value"], 1), "e2e_png", round(d["e2e_png"]["value"], 1), d["e2e_png"]["d2h_bytes_per_step"])
PY
tail -3 $OUT/bench_n$N.err
