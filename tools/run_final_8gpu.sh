#!/bin/bash
# Short 8-GPU confirmation of the current build: ResNet-50 and VGG-16 with the pull (p2p) and
# push (p2p_fused) exchange, the NCCL stand-in path, the all-reduce micro-benchmark and the
# P2P protocol check. Appends JSON lines to gpurun_out/final8.jsonl.
mkdir -p gpurun_out
OUT=gpurun_out/final8.jsonl
: > $OUT
: > gpurun_out/final8_arb.log
PORT=29700
NG=${1:-8}
run() {
  PORT=$((PORT+1))
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $NG --steps 20 --warmup 5 "$@" 2>gpurun_out/final8_err.log | tail -1 >> $OUT
  echo "done $* rc=$?"
}
run --no_e2e
run --no_e2e --comm_impl p2p_fused
run
run --no_e2e --model vgg16 --batch_size 64
run --no_e2e --model vgg16 --batch_size 64 --comm_impl p2p_fused
run --no_e2e --comm_impl nccl
for m in resnet50 vgg16; do for impl in p2p nccl; do
  echo "### all_reduce_benchmark $m $impl" >> gpurun_out/final8_arb.log
  timeout 200 python -m kf_benchmarks_b200.parallel.launcher -np $NG -- -m kf_benchmarks_b200.all_reduce_benchmark --model=$m --variable_update=replicated --comm_impl=$impl --use_fp16 --optimizer=momentum --num_batches=10 --num_warmup_batches=3 --iters_per_step=1 2>&1 | grep -E "Average|per all-reduce|Model:" >> gpurun_out/final8_arb.log
done; done
timeout 200 python -m kf_benchmarks_b200.parallel.launcher -np $NG -- tools/p2p_check.py 2>&1 | grep -E "P2P_CHECK|FAIL" | tail -3
python - <<'PY'
import json
for l in open("gpurun_out/final8.jsonl"):
    try:
        d = json.loads(l)
        print(d["config"].get("model"), d["config"].get("comm_impl"), d["n_gpus"], round(d["value"], 1),
              round(d["ms_per_step"], 2), "e2e", (d.get("e2e") or {}).get("value"))
    except Exception as e:
        print("bad line", l[:200])
PY
cat gpurun_out/final8_arb.log
