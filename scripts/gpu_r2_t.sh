#!/bin/bash
# Round 2, visit T (N GPUs): strong-scaling bench with per-rank timings; clock sampler started before the warm-up.
set -u
N=${1:-8}
OUT=gpurun_out/r2_t_n$N
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for tag in a b; do
	timeout 600 $TR --master-port 2961$N bench.py --gpus $N --steps 20 --warmup 5 --no-secondary > $OUT/bench_$tag.json 2> $OUT/bench_$tag.err; echo "bench $tag rc=$?"
done
timeout 600 $TR --master-port 2962$N bench.py --gpus $N --steps 20 --warmup 5 --no-secondary --no-clocks > $OUT/bench_noclocks.json 2> $OUT/bench_noclocks.err; echo "bench noclocks rc=$?"
python - $OUT <<'PY'
import json, sys, glob
for f in sorted(glob.glob(sys.argv[1] + "/bench_*.json")):
    d = json.load(open(f))
    print(f.split("/")[-1], "value %.4g e2e %.4g (%.3f ms e2e, %.3f device) parity %s clocks %s" % (d["value"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["ms_per_step"], d["sharded_parity"], d.get("clocks")))
    for r in d["per_rank_ms"]["rows"]:
        print("   ", r)
PY
