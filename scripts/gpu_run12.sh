set -x
cd /root/repo
timeout 600 python -m pytest tests/test_gpu_aggregator.py -x -q 2>&1 | tail -3
timeout 600 python tools/run_configs.py --agg-once > gpurun_out/r02_agg2.log 2>&1
grep "^{" gpurun_out/r02_agg2.log | cut -c1-420
