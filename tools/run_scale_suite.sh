#!/bin/bash
# BASELINE.json configs on one 8-GPU box: ours vs the PyTorch/NCCL stand-in. Appends JSON lines
# to gpurun_out/suite.jsonl. Usage: bash tools/run_scale_suite.sh [quick]
mkdir -p gpurun_out
OUT=gpurun_out/suite.jsonl
: > $OUT
PORT=29600
run() {  # n_gpus, extra args...
  local n=$1; shift
  PORT=$((PORT+1))
  if [ "$n" = "1" ]; then
    timeout 400 python bench.py --gpus 1 --steps 20 --warmup 5 "$@" 2>gpurun_out/suite_err.log | tail -1 >> $OUT
  else
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $n --steps 20 --warmup 5 "$@" 2>gpurun_out/suite_err.log | tail -1 >> $OUT
  fi
  echo "done n=$n $* rc=$?"
}
NG=$(nvidia-smi -L | wc -l)
# config 2: ResNet-50 replicated
for n in 1 2 4 8; do [ $n -le $NG ] && run $n --no_e2e; done
[ 8 -le $NG ] && run 8
[ 8 -le $NG ] && run 8 --impl standin
run 1 --impl standin
# config 3: KungFu S-SGD vs SMA
for n in 2 8; do [ $n -le $NG ] && run $n --no_e2e --variable_update kungfu --kungfu_option sync_sgd; done
for n in 2 8; do [ $n -le $NG ] && run $n --no_e2e --variable_update kungfu --kungfu_option sma; done
[ 8 -le $NG ] && run 8 --impl standin --variable_update kungfu --kungfu_option sma
# config 4: VGG-16 (138 M parameters)
run 1 --no_e2e --model vgg16 --batch_size 64
[ 8 -le $NG ] && run 8 --no_e2e --model vgg16 --batch_size 64
[ 8 -le $NG ] && run 8 --impl standin --model vgg16 --batch_size 64
# config 5: Inception-v3 with PairAveraging
run 1 --no_e2e --model inception3 --batch_size 64
[ 8 -le $NG ] && run 8 --no_e2e --model inception3 --batch_size 64 --variable_update kungfu --kungfu_option async_sgd
[ 8 -le $NG ] && run 8 --impl standin --model inception3 --batch_size 64 --variable_update kungfu --kungfu_option async_sgd
# all-reduce micro-benchmark, model-shaped
if [ 8 -le $NG ]; then
  for m in resnet50 vgg16; do for impl in p2p nccl; do
    echo "### all_reduce_benchmark $m $impl" >> gpurun_out/suite_arb.log
    timeout 300 python -m kf_benchmarks_b200.parallel.launcher -np 8 -- -m kf_benchmarks_b200.all_reduce_benchmark --model=$m --variable_update=replicated --comm_impl=$impl --use_fp16 --optimizer=momentum --num_batches=10 --num_warmup_batches=3 --iters_per_step=1 2>&1 | grep -E "Average|per all-reduce|Model:" >> gpurun_out/suite_arb.log
  done; done
fi
python - <<'PY'
import json
for l in open("gpurun_out/suite.jsonl"):
    try: d=json.loads(l)
    except Exception: print("BAD", l[:200]); continue
    c=d.get("config",{})
    print("%-8s %-10s n=%d vu=%-10s %-9s %10.1f img/s %7.2f ms e2e=%s" % (d.get("impl","ours"), c.get("model"), d.get("n_gpus",0), c.get("variable_update"), c.get("kungfu_option", ""), d.get("value",0), d.get("ms_per_step",0), (d.get("e2e") or {}).get("value")))
PY
cat gpurun_out/suite_arb.log 2>/dev/null
