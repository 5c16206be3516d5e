set -x
cd /root/repo
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_aggregator.py -x -q 2>&1 | tail -3
timeout 600 python tools/run_configs.py --agg-once > $O/r02_agg3.log 2>&1
grep "^{" $O/r02_agg3.log | cut -c1-400
