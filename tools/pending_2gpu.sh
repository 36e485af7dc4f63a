# 2-GPU checks of the exchange kernels: host-simulation check (pull and NVLS forms), the
# multi-GPU tests and the isolated all-reduce benchmark for every implementation.
set -u
N=${1:-2}
MODEL=${2:-resnet50}
mkdir -p gpurun_out
echo "== p2p_check (pull) =="
timeout 600 python -m kf_benchmarks_b200.parallel.launcher -np $N -- tools/p2p_check.py > gpurun_out/p2p_check_pull.log 2>&1
grep -E "FAIL|P2P_CHECK|compact" gpurun_out/p2p_check_pull.log | head -20
echo "== p2p_check (nvls) =="
KFB_P2P_NVLS=1 timeout 600 python -m kf_benchmarks_b200.parallel.launcher -np $N -- tools/p2p_check.py > gpurun_out/p2p_check_nvls.log 2>&1
grep -E "FAIL|P2P_CHECK|compact" gpurun_out/p2p_check_nvls.log | head -20
echo "== all_reduce_benchmark =="
for impl in p2p nvls nccl; do
  for extra in "" "--compact_gradient_transfer=false"; do
    if [ "$impl" = nccl ] && [ -n "$extra" ]; then continue; fi
    timeout 600 python -m kf_benchmarks_b200.parallel.launcher -np $N -- -m kf_benchmarks_b200.all_reduce_benchmark \
      --model=$MODEL --variable_update=replicated --num_gpus=$N --use_fp16 --optimizer=momentum \
      --comm_impl=$impl --num_batches=20 --num_warmup_batches=5 $extra 2>&1 | grep -E "per all-reduce|Error|error" | sed "s/^/$impl $extra: /"
  done
done
KFB_AR_V2=0 timeout 600 python -m kf_benchmarks_b200.parallel.launcher -np $N -- -m kf_benchmarks_b200.all_reduce_benchmark \
      --model=$MODEL --variable_update=replicated --num_gpus=$N --use_fp16 --optimizer=momentum \
      --comm_impl=p2p --num_batches=20 --num_warmup_batches=5 2>&1 | grep -E "per all-reduce" | sed "s/^/p2p first generation: /"
