set -x
cd /root/repo
timeout 300 python -m pytest tests/test_gpu_meter.py tests/test_gpu_ingest.py tests/test_gpu_reserve.py -x -q 2>&1 | tail -5
timeout 120 python tools/run_configs.py --meter-once 960 --tune ws_consumers=4 2>&1 | tail -2
timeout 400 python tools/run_configs.py --sweep-meter > gpurun_out/r2_meter_sweep6.jsonl 2>&1
tail -2 gpurun_out/r2_meter_sweep6.jsonl
