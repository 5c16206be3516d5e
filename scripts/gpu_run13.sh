set -x
cd /root/repo
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02_pytest_gpu.log; tail -3 $O/r02_pytest_gpu.log
timeout 900 python tools/run_configs.py --out $O/r02_configs.json > $O/r02_configs.log 2>&1; echo "configs rc=$?"
timeout 600 python bench.py > $O/r02_bench_n1.json 2> $O/r02_bench_n1.err; echo "bench rc=$?"
python -c "import __graft_entry__ as g; g.smoke()"
timeout 300 python tools/run_configs.py --agg-once > $O/r02_plain_agg.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file $O/r02_launches_agg.csv python tools/run_configs.py --agg-once > $O/r02_ncu_agg_launches.log 2>&1
