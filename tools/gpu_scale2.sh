#!/bin/bash
# scaling check on one 2-GPU box: each GPU alone, then both (one rank per GPU)
OUT=gpurun_out/scale2; mkdir -p $OUT
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.mem,power.limit,temperature.gpu --format=csv > $OUT/smi.txt
CUDA_VISIBLE_DEVICES=0 python bench.py --steps 20 --warmup 3 --cpu-seconds 1 > $OUT/g0.json 2>/dev/null
CUDA_VISIBLE_DEVICES=1 python bench.py --steps 20 --warmup 3 --cpu-seconds 1 > $OUT/g1.json 2>/dev/null
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 > $OUT/n2.json 2>/dev/null
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 3 > $OUT/n2b.json 2>/dev/null
cat $OUT/smi.txt
for f in g0 g1 n2 n2b; do python - $OUT/$f.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms_step", round(d["ms_per_step"],3), d["roofline"]["kernels_ms"], d["clocks"])
PY
done
