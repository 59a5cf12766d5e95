#!/bin/bash
# scaling check on one multi-GPU box: topology, then N ranks with and without host binding
N=${1:-2}
OUT=gpurun_out/scale$N; mkdir -p $OUT
(nvidia-smi topo -m; lscpu | grep -i "numa\|socket\|^CPU(s)"; nproc; cat /sys/fs/cgroup/cpuset.cpus.effective 2>/dev/null) > $OUT/topo.txt 2>&1
cat $OUT/topo.txt | head -40
for bind in 0 1; do
  if [ $bind = 0 ]; then export VF_NO_BIND=1; else unset VF_NO_BIND; fi
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$bind bench.py --gpus $N --steps 20 --warmup 3 > $OUT/n$N.bind$bind.json 2>/dev/null
  python - $OUT/n$N.bind$bind.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms_step", round(d["ms_per_step"],3), d["config"].get("host_binding"))
PY
done
