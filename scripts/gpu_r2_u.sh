#!/bin/bash
# Round 2, visit U (N GPUs): the bench exactly as the driver launches it, final code (clock sampler at 200 ms).
set -u
N=${1:-8}
OUT=gpurun_out/r2_u_n$N
mkdir -p $OUT
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29631 bench.py --gpus $N --steps 20 --warmup 5 > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"
python - $OUT <<'PY'
import json, sys
d = json.load(open(sys.argv[1] + "/bench.json"))
print("value %.4g e2e %.4g (%.3f ms e2e, %.3f device) parity %s clocks %s" % (d["value"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["ms_per_step"], d["sharded_parity"], d.get("clocks")))
for r in d["per_rank_ms"]["rows"]:
    print("   ", r)
PY
