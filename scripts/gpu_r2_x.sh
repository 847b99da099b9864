#!/bin/bash
# Round 2, visit X (N GPUs): gather of the result blocks as grouped ncclSend / ncclRecv (VX_GATHER_P2P=1) vs ncclAllGather:
# parity on the hardware, then the bench without the clock sampler (its polling adds noise of the size of the difference).
set -u
N=${1:-2}
OUT=gpurun_out/r2_x_n$N
mkdir -p $OUT
TR="timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
VX_GATHER_P2P=1 $TR --master-port 29641 scripts/check_sharded.py > $OUT/check_p2p.log 2>&1; echo "check p2p rc=$?"; grep -E "sharded ==|MISMATCH|Error" $OUT/check_p2p.log | head -3
for rep in 1 2; do
	$TR --master-port 2965$rep bench.py --gpus $N --steps 40 --warmup 5 --no-secondary --no-clocks > $OUT/bench_ring_$rep.json 2> $OUT/bench_ring_$rep.err
	VX_GATHER_P2P=1 $TR --master-port 2966$rep bench.py --gpus $N --steps 40 --warmup 5 --no-secondary --no-clocks > $OUT/bench_p2p_$rep.json 2> $OUT/bench_p2p_$rep.err
done
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
