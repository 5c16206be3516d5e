set -x
cd /root/repo
timeout 900 python -m pytest tests/test_gpu_aggregator.py -x -q 2>&1 | tail -25
timeout 600 python tools/run_configs.py --only-meter > gpurun_out/r2_meter3.log 2>&1
tail -4 gpurun_out/r2_meter3.log
