#!/bin/bash
# workload E (or $WL) at $N GPUs with the records pulled in place vs exchanged (SHK_SHARD_PULL=0), same box
N=${N:-2}; W=${WL:-E}
A="bench.py --workload $W --gpus $N --steps 6 --warmup 3 --no-cpu-baseline --no-extras --no-e2e --no-cold"
for pull in 1 0 1 0; do
  SHK_SHARD_PULL=$pull python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 $A > gpurun_out/eab_${W}_n${N}_pull$pull.json 2> gpurun_out/eab.err
  python - gpurun_out/eab_${W}_n${N}_pull$pull.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d["ms_per_step"],3), d["config"].get("records"), {k:round(v,2) for k,v in d["roofline"]["stage_ms"].items()})
except Exception as e: print(sys.argv[1], "failed", e)
PY
done
