# usage: bash scripts/collect_evidence.sh      (after `gpurun -- 'bash scripts/gpu_evidence.sh'` has merged gpurun_out/)
# Copies / summarises what gpu_evidence.sh brought back into the tracked files under profiles/.
set -e
cd "$(dirname "$0")/.."
O=gpurun_out; P=profiles
cp $O/r02_bench_n1.json $O/r02_bench_reference_arm.json $O/r02_configs.json $O/r02_sweep_aggregator.jsonl $P/
python $P/summarize.py launches $O/r02_launches.csv > $P/r02_launches.txt; gzip -nc $O/r02_launches.csv > $P/r02_launches.csv.gz
python $P/summarize.py launches $O/r02_launches_agg.csv > $P/r02_launches_aggregator.txt; gzip -nc $O/r02_launches_agg.csv > $P/r02_launches_aggregator.csv.gz
python $P/summarize.py launches $O/r02_launches_agg_stream.csv > $P/r02_launches_aggregator_stream.txt
for k in coal3:coal3 meter960:meter960 meter2944:meter2944 agg_convert:agg_convert agg_walk:agg_walk agg_append:agg_append_stream agg_convert_flush:agg_convert_stream_flush; do
  python $P/summarize.py raw $O/r02_prof_${k%%:*}.ncu-rep > $P/r02_kernel_${k##*:}_raw.txt
done
[ -f $O/r02_scatter_probe.jsonl ] && cp $O/r02_scatter_probe.jsonl $P/
tail -2 $O/r02_pytest_gpu.log
