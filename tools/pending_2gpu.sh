set -u
mkdir -p gpurun_out
echo "== multi tests (all, incl. gated) =="
KFB_TEST_NVLS=1 KFB_MULTI_RANK_GRAPH=1 timeout 1200 python -m pytest tests/test_gpu_multi.py -q -m gpu -x 2>&1 | tail -15
echo "== nvls_check 2 =="
timeout 900 python tools/nvls_check.py 2 > gpurun_out/pending_nvls_check.log 2>&1
grep -E "NVLS_CHECK|replicas|p2p_check|Average|per all-reduce|FAILED|==|Error|error" gpurun_out/pending_nvls_check.log | head -40
echo "== devseq =="
KFB_P2P_DEVSEQ=1 timeout 600 python -m kf_benchmarks_b200.parallel.launcher -np 2 -- tools/p2p_check.py > gpurun_out/pending_p2p_devseq.log 2>&1
grep -E "device sequence|device counter|P2P_CHECK|FAIL" gpurun_out/pending_p2p_devseq.log | head
