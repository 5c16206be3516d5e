set -x
cd /root/repo
python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest1.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest1.log
tail -5 gpurun_out/r2_pytest1.log
python bench.py > gpurun_out/r2_bench1.json 2> gpurun_out/r2_bench1.err; echo "bench rc=$?"
python bench.py --sessions 131072 --no-e2e --no-cpu --sustained-seconds 0 --steps 200 > gpurun_out/r2_bench_s17.json 2>&1
python tools/run_configs.py --sweep-meter > gpurun_out/r2_meter_sweep1.jsonl 2>&1
python tools/run_configs.py --only-meter > gpurun_out/r2_meter1.log 2>&1
tail -3 gpurun_out/r2_meter1.log
