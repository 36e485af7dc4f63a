set -u
mkdir -p gpurun_out
timeout 600 python bench.py --gpus 8 --steps 30 --warmup 5 2> gpurun_out/r2_8gpu_headline.err | grep '^{"metric"' | tail -1 > gpurun_out/r2_8gpu_headline.json
timeout 300 python bench.py --gpus 8 --steps 30 --warmup 5 --no_standin --no_e2e --variable_update kungfu --kungfu_option sma 2> gpurun_out/r2_8gpu_kungfu_sma.err | grep '^{"metric"' | tail -1 > gpurun_out/r2_8gpu_kungfu_sma_graph.json
python - <<'PY'
import json
for n in ("headline", "kungfu_sma_graph"):
    try:
        d = json.loads(open("gpurun_out/r2_8gpu_%s.json" % n).read())
        print(n, d["value"], d["ms_per_step"], d.get("e2e"), d.get("standin"), d.get("replica_checksum_equal"), d["config"].get("comm_impl"), d["clocks"])
    except Exception as e:
        print(n, "no result", repr(e))
PY
