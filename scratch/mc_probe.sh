set -x
timeout 90 python -m pytest tests/test_gpu_tc.py -x -q -m gpu -k multicast 2>&1 | tail -5
echo "--- mc on"; timeout 100 python profiles/gemm_probe.py 2>&1 | tail -7
echo "--- mc off"; NR_TC_MC=0 timeout 100 python profiles/gemm_probe.py 2>&1 | tail -7
