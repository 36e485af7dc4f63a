#!/bin/bash
# Sweep the environment launch knobs on one GPU in a single call (no rebuild between points):
#   KFB_APPLY_CTAS_PER_SM   grid cap of the BN apply kernels      (default 16)
#   KFB_REDUCE_BLOCKS_PER_SM grid cap of the channel reductions    (default 4)
#   KFB_GEMM_MAX_BN          widest GEMM tile (64 | 128 | 256)     (default 256)
#   KFB_SPLIT_ROUNDS         persistent-grid rounds split-K aims at (default 2)
#   KFB_TMA_STORE            TMA-store GEMM epilogue               (default 1)
#   KFB_FUSE_STATS_MIN_K     smallest reduction whose conv epilogue produces the BN statistics (default 0)
# Prints images/sec of the headline step per point; everything lands in gpurun_out/sweep_*.json.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() {  # name, env assignments...
  local name=$1; shift
  env "$@" timeout 600 python bench.py --steps 15 --warmup 4 --no_e2e \
    > gpurun_out/sweep_$name.json 2> gpurun_out/sweep_$name.err
  python - "$name" <<'PY'
import json, sys
n = sys.argv[1]
try:
    d = json.loads(open("gpurun_out/sweep_%s.json" % n).read().strip().splitlines()[-1])
    print("%-28s %8.0f img/s  %6.2f ms/step  loss %.3f  sm %s MHz %s" %
          (n, d["value"], d["ms_per_step"], d["final_loss"], (d.get("clocks") or {}).get("sm_mhz"),
           (d.get("clocks") or {}).get("reasons")))
except Exception as e:
    print("%-28s no result (%r)" % (n, e))
PY
}
run default KFB_DUMMY=0
for v in 4 8 32; do run apply_ctas_$v KFB_APPLY_CTAS_PER_SM=$v; done
for v in 2 6 8; do run reduce_blocks_$v KFB_REDUCE_BLOCKS_PER_SM=$v; done
run gemm_max_bn_128 KFB_GEMM_MAX_BN=128
for v in 1 3 4; do run split_rounds_$v KFB_SPLIT_ROUNDS=$v; done
run per_lane_store KFB_TMA_STORE=0
run no_fused_stats KFB_FUSE_STATS_MIN_K=1000000000
run fused_stats_k512 KFB_FUSE_STATS_MIN_K=512
run default_again KFB_DUMMY=0
