#!/bin/bash
# One multi-GPU lease, everything that needs N ranks (default 8):
#   all_reduce_benchmark (ResNet-50 / VGG-16 variables; pull, NVLS, NCCL), the headline bench with
#   the stand-in, the NVLS exchange in training, VGG-16, KungFu S-SGD vs SMA.
# Results: gpurun_out/r2_<N>gpu_*.json(l) and gpurun_out/all_reduce_benchmark_<N>gpu_r2.txt
set -u
cd "$(dirname "$0")/.."
N=${1:-8}
mkdir -p gpurun_out
OUT=gpurun_out/all_reduce_benchmark_${N}gpu_r2.txt
: > $OUT
arb() {  # model impl extra...
  local model=$1 impl=$2; shift 2
  echo "### all_reduce_benchmark $model $impl $*" >> $OUT
  timeout 300 python -m kf_benchmarks_b200.parallel.launcher -np $N -- -m kf_benchmarks_b200.all_reduce_benchmark \
    --model=$model --variable_update=replicated --num_gpus=$N --use_fp16 --optimizer=momentum \
    --comm_impl=$impl --num_batches=20 --num_warmup_batches=5 "$@" 2>&1 \
    | grep -E "Model:|Average|per all-reduce|Error|error" >> $OUT
}
for model in resnet50 vgg16; do
  arb $model p2p
  arb $model p2p --compact_gradient_transfer=false
  arb $model nvls
  arb $model nccl
done
cat $OUT
run() {  # name, bench args...
  local name=$1; shift
  timeout 600 python bench.py --gpus $N --steps 30 --warmup 5 "$@" 2> gpurun_out/r2_${N}gpu_$name.err | tail -1 \
    > gpurun_out/r2_${N}gpu_$name.json
  python - "$name" "$N" <<'PY'
import json, sys
name, n = sys.argv[1], sys.argv[2]
try:
    d = json.loads(open("gpurun_out/r2_%sgpu_%s.json" % (n, name)).read())
    print("%-22s %9.0f img/s %7.2f ms/step e2e %s standin %s checksum_equal %s loss %.3f clocks %s" % (
        name, d["value"], d["ms_per_step"], (d.get("e2e") or {}).get("value"),
        (d.get("standin") or {}).get("value"), d.get("replica_checksum_equal"), d["final_loss"],
        (d.get("clocks") or {}).get("sm_mhz")))
except Exception as e:
    print(name, "no result", repr(e))
PY
}
run headline
run headline_nvls --comm_impl nvls --no_standin --no_e2e
run headline_nccl --comm_impl nccl --no_standin --no_e2e
run vgg16 --model vgg16 --batch_size 64 --no_standin --no_e2e
run vgg16_nvls --model vgg16 --batch_size 64 --comm_impl nvls --no_standin --no_e2e
run kungfu_ssgd --variable_update kungfu --kungfu_option sync_sgd --no_standin --no_e2e
run kungfu_sma --variable_update kungfu --kungfu_option sma --no_standin --no_e2e
