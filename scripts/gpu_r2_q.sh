#!/bin/bash
# Round 2, visit Q (N GPUs, final code): sharded == unsharded on the hardware, strong-scaling bench, C5 + C4 (full config).
set -u
N=${1:-8}
OUT=gpurun_out/r2_q_n$N
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29611 scripts/check_sharded.py > $OUT/check_sharded.log 2>&1; echo "check_sharded rc=$?"; grep -E "sharded ==|MISMATCH|Error|error" $OUT/check_sharded.log | head -5
timeout 900 $TR --master-port 29614 bench.py --gpus $N --steps 20 --warmup 5 > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"; tail -3 $OUT/bench.err
python - $OUT <<'PY'
import json, sys
d = json.load(open(sys.argv[1] + "/bench.json"))
print("N=%d value %.4g e2e %.4g (%.3f ms/step e2e, %.3f device) frac %.4f parity %s scaling %s\nphases %s" % (d["n_gpus"], d["value"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["ms_per_step"], d["roofline"]["frac"], d["sharded_parity"], d["scaling"], d["phases_ms"]))
PY
timeout 900 $TR --master-port 29615 scripts/bench_configs.py c5 c4full > $OUT/configs.jsonl 2> $OUT/configs.err; echo "configs rc=$?"; tail -2 $OUT/configs.err
python - $OUT <<'PY'
import json, sys
for l in open(sys.argv[1] + "/configs.jsonl"):
    d = json.loads(l)
    keys = [k for k in ("batch_wall_ms", "batch_device_ms", "sentence_alignments_per_s_e2e", "sentences_per_s_device", "sentences_per_s_e2e", "ms", "host_ms_rank0", "gemm_dp_hbm_frac_lower_bound") if k in d]
    print(d["config"][:60], {k: d[k] for k in keys})
PY
