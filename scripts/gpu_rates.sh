set -x
cd /root/repo
O=gpurun_out
for r in 44100 16000 32000 80000 96000 128000 192000; do
  timeout 300 python bench.py --rate $r --no-e2e --no-cpu --sustained-seconds 0 --steps 30 > $O/r02_rate_$r.json 2> $O/r02_rate_$r.err
done
timeout 300 python bench.py --frame-ms 60 --sessions 349525 --no-e2e --no-cpu --sustained-seconds 0 --steps 30 > $O/r02_rate_48000_60ms.json 2>&1
grep -h -o '"achieved": [0-9.]*, "peak": [0-9.]*, "unit": "GB/s", "frac": [0-9.]*' $O/r02_rate_*.json
