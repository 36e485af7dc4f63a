#!/bin/bash
# One-call refresh of the measured evidence under profiles/ for the current build (1 GPU):
#   1. plain bench run (must exit 0 before anything is profiled)       -> profiles/bench_1gpu_<tag>.json
#   2. launch list of one training step with device time + DRAM bytes  -> profiles/resnet50_bs256_launches_<tag>.txt
#   3. ncu --set full on the heaviest kernels (3 launches each)         -> gpurun_out/prof_<name>.ncu-rep,
#      selected raw metrics                                             -> profiles/ncu_<tag>_<name>.csv
#   4. per-class roofline against the new launch list                   -> profiles/roofline_resnet50_bs256_<tag>.md
# Usage (on the GPU box, e.g. through gpurun):  bash tools/capture_profiles.sh r2a
# Numbers printed by runs under ncu are never bench values.
set -u
cd "$(dirname "$0")/.."
TAG=${1:-new}
mkdir -p gpurun_out profiles
CMD="python tools/profile_step.py --model resnet50 --batch_size 256 --steps 3"

echo "== 1. bench (plain) =="
timeout 900 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err || {
  echo "bench failed"; tail -5 gpurun_out/bench_$TAG.err; exit 1; }
tail -1 gpurun_out/bench_$TAG.json > profiles/bench_1gpu_$TAG.json
python - "$TAG" <<'PY'
import json, sys
d = json.loads(open("profiles/bench_1gpu_%s.json" % sys.argv[1]).read())
print("%.0f img/s, %.2f ms/step, e2e %s, clocks %s" %
      (d["value"], d["ms_per_step"], (d.get("e2e") or {}).get("value"), d.get("clocks")))
PY

echo "== 2. launch list of one step =="
export KFB_CUDA_GRAPH=0   # profiled runs launch kernel by kernel (a graph replay hides them)
$CMD > gpurun_out/plain.log 2>&1 || { echo "profile_step failed"; tail -5 gpurun_out/plain.log; exit 1; }
PER=$(grep -o "launches per step (approx): [0-9]*" gpurun_out/plain.log | grep -o "[0-9]*$")
PER=${PER:-543}
# skip build-time launches + the first two steps; capture exactly one step
TOTAL=$(ncu --metrics gpu__time_duration.sum --clock-control none --csv \
          --log-file gpurun_out/launches_all_$TAG.csv $CMD > gpurun_out/ncu_count.log 2>&1;
        grep -c '"gpu__time_duration.sum"' gpurun_out/launches_all_$TAG.csv)
SKIP=$((TOTAL - PER))
echo "launches in the run: $TOTAL, per step: $PER, skipping $SKIP"
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum \
    --clock-control none -s "$SKIP" -c "$PER" --csv --log-file gpurun_out/launches_$TAG.csv \
    $CMD > gpurun_out/ncu_launches.log 2>&1
python tools/summarize_launches.py gpurun_out/launches_$TAG.csv \
    > profiles/resnet50_bs256_launches_$TAG.txt && head -12 profiles/resnet50_bs256_launches_$TAG.txt

echo "== 2b. the same launch list with kernels renamed after the enclosing NVTX range (layer) =="
KFB_NVTX=1 ncu --nvtx --print-nvtx-rename kernel \
    --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum \
    --clock-control none -s "$SKIP" -c "$PER" --csv --log-file gpurun_out/launches_nvtx_$TAG.csv \
    $CMD > gpurun_out/ncu_launches_nvtx.log 2>&1 \
  && python tools/summarize_launches.py gpurun_out/launches_nvtx_$TAG.csv 400 \
       > profiles/resnet50_bs256_launches_by_layer_$TAG.txt \
  && head -8 profiles/resnet50_bs256_launches_by_layer_$TAG.txt \
  || echo "(per-layer list not produced; see gpurun_out/ncu_launches_nvtx.log)"

echo "== 3. ncu --set full on the heaviest kernels =="
for spec in "gemm:gemm_tcgen05_kernel" "bnbwd:bn_bwd_fused_kernel" "bnapply:bn_apply_kernel" \
            "poolbwd:pool_bwd_max3s2_block_kernel"; do
  name=${spec%%:*}; pat=${spec##*:}
  skip=60; [ "$name" = poolbwd ] && skip=2
  ncu --set full --clock-control none --import-source on -k regex:$pat -s $skip -c 3 -f \
      -o gpurun_out/prof_${TAG}_$name $CMD > gpurun_out/ncu_$name.log 2>&1
  if [ -f gpurun_out/prof_${TAG}_$name.ncu-rep ]; then
    ncu -i gpurun_out/prof_${TAG}_$name.ncu-rep --page raw --csv > gpurun_out/raw_${TAG}_$name.csv 2>/dev/null
    python tools/ncu_select.py gpurun_out/raw_${TAG}_$name.csv gpurun_out/ncu_${TAG}_$name.csv
    cp gpurun_out/ncu_${TAG}_$name.csv profiles/ 2>/dev/null
  else
    echo "no report for $name"; tail -3 gpurun_out/ncu_$name.log
  fi
done

echo "== 4. roofline =="
# (only gpurun_out/ travels back from a gpurun box: everything under profiles/ is mirrored there)
MS=$(python -c "import json;print(json.loads(open('profiles/bench_1gpu_$TAG.json').read())['ms_per_step'])")
python tools/roofline_model.py --launches profiles/resnet50_bs256_launches_$TAG.txt --step_ms "$MS" \
    --by_layer profiles/resnet50_bs256_launches_by_layer_$TAG.txt \
    > profiles/roofline_resnet50_bs256_$TAG.md && tail -14 profiles/roofline_resnet50_bs256_$TAG.md
mkdir -p gpurun_out/profiles_$TAG && cp profiles/*_$TAG* gpurun_out/profiles_$TAG/ 2>/dev/null
