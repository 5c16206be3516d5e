set -x
cd /root/repo
python -m pytest tests/test_gpu_meter.py -x -q 2>&1 | tail -3
python tools/run_configs.py --sweep-meter > gpurun_out/r2_meter_sweep2.jsonl 2>&1
python tools/run_configs.py --meter-once 960 > gpurun_out/r2_meter_once.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_frames_meter -s 4 -c 1 -o gpurun_out/r2_prof_meter960 python tools/run_configs.py --meter-once 960 > gpurun_out/r2_ncu_meter960.log 2>&1
python tools/run_configs.py --meter-once 2944 > gpurun_out/r2_meter_once2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_frames_meter -s 4 -c 1 -o gpurun_out/r2_prof_meter2944 python tools/run_configs.py --meter-once 2944 > gpurun_out/r2_ncu_meter2944.log 2>&1
tail -3 gpurun_out/r2_ncu_meter2944.log
