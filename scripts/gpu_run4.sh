set -x
cd /root/repo
timeout 600 python -m pytest tests/test_gpu_meter.py tests/test_gpu_aggregator.py -x -q 2>&1 | tail -5
timeout 900 python tools/run_configs.py --sweep-meter > gpurun_out/r2_meter_sweep3.jsonl 2>&1
tail -3 gpurun_out/r2_meter_sweep3.jsonl
