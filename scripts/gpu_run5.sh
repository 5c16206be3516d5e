set -x
cd /root/repo
timeout 600 python -m pytest tests/test_gpu_meter.py -x -q 2>&1 | tail -3
timeout 900 python tools/run_configs.py --sweep-meter > gpurun_out/r2_meter_sweep4.jsonl 2>&1
timeout 300 python tools/run_configs.py --meter-once 2944 > gpurun_out/r2_meter_once3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_frames_meter_ws -s 4 -c 1 -o gpurun_out/r2_prof_ws2944 python tools/run_configs.py --meter-once 2944 > gpurun_out/r2_ncu_ws2944.log 2>&1
timeout 300 python bench.py --sessions 131072 --no-e2e --no-cpu --sustained-seconds 0 --steps 200 --tune overlap_ticks=1 > gpurun_out/r2_bench_s17_pdl.json 2>&1
timeout 300 python bench.py --sessions 131072 --no-e2e --no-cpu --sustained-seconds 0 --steps 200 > gpurun_out/r2_bench_s17_b.json 2>&1
timeout 300 python bench.py --no-e2e --no-cpu --sustained-seconds 0 --steps 50 --tune overlap_ticks=1 > gpurun_out/r2_bench_s20_pdl.json 2>&1
tail -c 600 gpurun_out/r2_bench_s17_pdl.json
