set -x
cd /root/repo
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02_pytest_gpu.log; tail -4 $O/r02_pytest_gpu.log
timeout 600 python bench.py > $O/r02_bench_n1.json 2> $O/r02_bench_n1.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference > $O/r02_bench_reference_arm.json 2> $O/r02_bench_reference_arm.err; echo "ref rc=$?"
timeout 900 python tools/run_configs.py --out $O/r02_configs.json > $O/r02_configs.log 2>&1; echo "configs rc=$?"
# ncu: launch list of the bench command, then full captures (each after its plain run exited 0)
BENCH="python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu --sustained-seconds 0"
timeout 300 $BENCH > $O/r02_plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02_launches.csv $BENCH > $O/r02_ncu_launches.log 2>&1
timeout 300 $BENCH > $O/r02_plain_bench2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_decim_coal -s 3 -c 2 -o $O/r02_prof_coal3 $BENCH > $O/r02_ncu_coal3.log 2>&1
timeout 300 python tools/run_configs.py --meter-once 960 > $O/r02_plain_m960.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_frames_meter_ws -s 4 -c 1 -o $O/r02_prof_meter960 python tools/run_configs.py --meter-once 960 > $O/r02_ncu_m960.log 2>&1
timeout 300 python tools/run_configs.py --meter-once 2944 > $O/r02_plain_m2944.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_frames_meter_ws -s 4 -c 1 -o $O/r02_prof_meter2944 python tools/run_configs.py --meter-once 2944 > $O/r02_ncu_m2944.log 2>&1
timeout 300 python tools/run_configs.py --agg-once > $O/r02_plain_agg.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file $O/r02_launches_agg.csv python tools/run_configs.py --agg-once > $O/r02_ncu_agg_launches.log 2>&1
timeout 300 python tools/run_configs.py --agg-once > $O/r02_plain_agg2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_agg_convert_one -s 2 -c 1 -o $O/r02_prof_agg_convert python tools/run_configs.py --agg-order quantum > $O/r02_ncu_agg_convert.log 2>&1
timeout 300 python tools/run_configs.py --agg-order quantum > $O/r02_plain_agg3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_agg_walk -s 2 -c 1 -o $O/r02_prof_agg_walk python tools/run_configs.py --agg-order quantum > $O/r02_ncu_agg_walk.log 2>&1
timeout 300 python tools/run_configs.py --agg-stream 1048576 > $O/r02_plain_stream.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/r02_launches_agg_stream.csv python tools/run_configs.py --agg-stream 1048576 > $O/r02_ncu_agg_stream.log 2>&1
timeout 300 python tools/run_configs.py --agg-stream 1048576 > $O/r02_plain_stream2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_agg_append -s 6 -c 1 -o $O/r02_prof_agg_append python tools/run_configs.py --agg-stream 1048576 > $O/r02_ncu_agg_append.log 2>&1
timeout 300 python tools/run_configs.py --agg-stream 1048576 > $O/r02_plain_stream3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_agg_convert_warp -s 21 -c 1 -o $O/r02_prof_agg_convert_flush python tools/run_configs.py --agg-stream 1048576 > $O/r02_ncu_agg_convert_flush.log 2>&1
[ -x tools/scatter_probe ] && timeout 120 ./tools/scatter_probe > $O/r02_scatter_probe.jsonl 2>&1
timeout 300 python tools/run_configs.py --sweep-agg > $O/r02_sweep_aggregator.jsonl 2> $O/r02_sweep_aggregator.err
ls -la $O | tail -30
