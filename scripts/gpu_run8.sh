set -x
cd /root/repo
timeout 300 python -m pytest tests/test_gpu_meter.py -x -q 2>&1 | tail -3
timeout 600 python tools/run_configs.py --sweep-meter > gpurun_out/r2_meter_sweep13.jsonl 2>&1
tail -2 gpurun_out/r2_meter_sweep13.jsonl
timeout 300 python -m pytest tests/test_gpu_aggregator.py -x -q 2>&1 | tail -3
timeout 300 python tools/run_configs.py --only-meter 2>&1 | tail -2
