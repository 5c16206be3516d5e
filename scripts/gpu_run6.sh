set -x
cd /root/repo
timeout 600 python -m pytest tests/test_gpu_meter.py -x -q 2>&1 | tail -3
timeout 900 python tools/run_configs.py --sweep-meter > gpurun_out/r2_meter_sweep5.jsonl 2>&1
tail -2 gpurun_out/r2_meter_sweep5.jsonl
