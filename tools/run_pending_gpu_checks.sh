#!/bin/bash
# Code paths that were written when no GPU time was left (NEXT.md) - run these first:
#   1 GPU : bash tools/run_pending_gpu_checks.sh
#   2 GPUs: bash tools/run_pending_gpu_checks.sh 2        (adds the NVLS all-reduce checks)
# Everything lands in gpurun_out/pending_*.log; nothing here is a bench value.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-1}
echo "== TMA-store epilogue: numerics of every GEMM / conv case =="
KFB_TMA_STORE=1 timeout 1500 python tools/kernel_check.py > gpurun_out/pending_tma_store_check.log 2>&1
tail -1 gpurun_out/pending_tma_store_check.log
echo "== TMA-store epilogue: A/B on the headline step (1 GPU) =="
# v=2: TMA stores + BN statistics fused into the conv epilogue (all layers)
for v in 0 1 2; do
  if [ "$v" = 2 ]; then export KFB_FUSE_STATS_MIN_K=0; fi
  KFB_TMA_STORE=$((v > 0 ? 1 : 0)) timeout 900 python bench.py --steps 20 --warmup 5 --no_e2e \
    > gpurun_out/pending_bench_tma_store_$v.json 2> gpurun_out/pending_bench_tma_store_$v.err
  python - "$v" <<'PY'
import json, sys
v = sys.argv[1]
try:
    d = json.loads(open("gpurun_out/pending_bench_tma_store_%s.json" % v).read().strip().splitlines()[-1])
    print("config %s (0 = default, 1 = TMA store, 2 = TMA store + fused BN stats): %.0f img/s, %.2f ms/step, loss %.4f" %
          (v, d["value"], d["ms_per_step"], d["final_loss"]))
except Exception as e:
    print("KFB_TMA_STORE=%s: no result (%r)" % (v, e))
PY
done
unset KFB_FUSE_STATS_MIN_K
echo "== fused BN statistics through the TMA-store layout: numerics =="
KFB_TMA_STORE=1 timeout 600 python -m pytest tests/test_gpu_gemm_conv.py -q -m gpu \
  -k fused_bn_statistics > gpurun_out/pending_fused_stats_tma.log 2>&1
tail -1 gpurun_out/pending_fused_stats_tma.log
if [ "$N" -ge 2 ]; then
  echo "== NVLS all-reduce (--comm_impl=nvls) on $N GPUs =="
  timeout 1800 python tools/nvls_check.py "$N" > gpurun_out/pending_nvls_check.log 2>&1
  grep -E "NVLS_CHECK|replicas|p2p_check|Average|per all-reduce|FAILED|==" gpurun_out/pending_nvls_check.log
  echo "== device-resident sequence offset: p2p_check section 8 =="
  KFB_P2P_DEVSEQ=1 timeout 900 python -m kf_benchmarks_b200.parallel.launcher -np "$N" -- \
    tools/p2p_check.py > gpurun_out/pending_p2p_devseq.log 2>&1
  grep -E "device sequence|device counter|P2P_CHECK" gpurun_out/pending_p2p_devseq.log
  echo "== multi-rank CUDA graph (device-resident sequence offset) on 2 GPUs =="
  KFB_MULTI_RANK_GRAPH=1 timeout 900 python -m pytest tests/test_gpu_multi.py -q -m gpu \
    -k multi_rank_cuda_graph > gpurun_out/pending_multi_rank_graph.log 2>&1
  tail -3 gpurun_out/pending_multi_rank_graph.log
  for g in "" "--cuda_graph"; do
    KFB_MULTI_RANK_GRAPH=1 timeout 900 python -m kf_benchmarks_b200.parallel.launcher -np "$N" -- \
      -m kf_benchmarks_b200.cli --model=mobilenet --batch_size=64 --use_fp16 --optimizer=momentum \
      --variable_update=replicated --num_gpus="$N" --num_batches=40 --num_warmup_batches=10 $g \
      2>&1 | grep -E "total images/sec|Captured" | sed "s/^/mobilenet $g: /"
  done
fi
