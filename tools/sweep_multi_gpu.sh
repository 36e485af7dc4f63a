#!/bin/bash
# Multi-GPU sweep of the gradient-exchange knobs on the headline step in one call:
#   bash tools/sweep_multi_gpu.sh 8 [full]
# KFB_AR_BLOCKS (CTAs of the fused all-reduce + update kernel), --bucket_bytes, --comm_impl.
# A multi-GPU call is charged N x its box time: the default list has 6 points (~1 min each),
# "full" 13.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-8}
MODE=${2:-short}
PORT=29800
run() {  # name, "ENV=.. ENV=..", bench args...
  local name=$1 envs=$2; shift 2
  PORT=$((PORT + 1))
  env $envs timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" \
    --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus "$N" --steps 15 --warmup 4 \
    --no_e2e "$@" 2> gpurun_out/sweepN_$name.err | tail -1 > gpurun_out/sweepN_$name.json
  python - "$name" <<'PY'
import json, sys
n = sys.argv[1]
try:
    d = json.loads(open("gpurun_out/sweepN_%s.json" % n).read())
    print("%-26s %9.0f img/s  %6.2f ms/step  loss %.3f" % (n, d["value"], d["ms_per_step"], d["final_loss"]))
except Exception as e:
    print("%-26s no result (%r)" % (n, e))
PY
}
run default "KFB_DUMMY=0"
if [ "$MODE" = full ]; then
  for b in 32 64 128; do run ar_blocks_$b "KFB_AR_BLOCKS=$b"; done
  for mb in 2 4 16 32; do run bucket_${mb}MB "KFB_DUMMY=0" --bucket_bytes $((mb << 20)); done
  run nccl "KFB_DUMMY=0" --comm_impl nccl
  run p2p_fused "KFB_DUMMY=0" --comm_impl p2p_fused
  run nvls "KFB_DUMMY=0" --comm_impl nvls
  run default_again "KFB_DUMMY=0"
else
  run ar_blocks_64 "KFB_AR_BLOCKS=64"
  run ar_blocks_128 "KFB_AR_BLOCKS=128"
  run bucket_4MB "KFB_DUMMY=0" --bucket_bytes $((4 << 20))
  run bucket_16MB "KFB_DUMMY=0" --bucket_bytes $((16 << 20))
  run nvls "KFB_DUMMY=0" --comm_impl nvls
fi
