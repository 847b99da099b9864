#!/bin/bash
# Round 2, visit Y (8 GPUs): ring all-gather vs grouped send / recv, one bench run each (no clock sampler).
set -u
N=${1:-8}
OUT=gpurun_out/r2_y_n$N
mkdir -p $OUT
TR="timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29651 bench.py --gpus $N --steps 60 --warmup 5 --no-secondary --no-clocks > $OUT/bench_ring.json 2> $OUT/bench_ring.err
VX_GATHER_P2P=1 $TR --master-port 29661 bench.py --gpus $N --steps 60 --warmup 5 --no-secondary --no-clocks > $OUT/bench_p2p.json 2> $OUT/bench_p2p.err
python - $OUT <<'PY'
import json, sys, glob
for f in sorted(glob.glob(sys.argv[1] + "/bench_*.json")):
    try:
        d = json.load(open(f))
        g = [r[1] for r in d["per_rank_ms"]["rows"]]
        print(f.split("/")[-1], "e2e %.4g (%.3f ms) device %.3f ms parity %s gather+merge per rank %.3f..%.3f" % (d["e2e"]["value"], d["e2e"]["ms_per_step"], d["ms_per_step"], d["sharded_parity"], min(g), max(g)))
    except Exception as e:
        print(f, "failed", e)
PY
