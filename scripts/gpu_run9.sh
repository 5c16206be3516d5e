set -x
cd /root/repo
timeout 300 python tools/run_configs.py --meter-once 2944 > gpurun_out/r2_meter_once5.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_frames_meter_ws -s 4 -c 1 -o gpurun_out/r2_prof_ws5_2944 python tools/run_configs.py --meter-once 2944 > gpurun_out/r2_ncu_ws5_2944.log 2>&1
tail -1 gpurun_out/r2_meter_once5.log
