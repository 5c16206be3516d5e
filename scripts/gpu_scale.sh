# usage: bash scripts/gpu_scale.sh N   (run under gpurun --gpus N)
set -x
cd /root/repo
N=$1
O=gpurun_out
if [ "$N" = "1" ]; then
  timeout 600 python bench.py --gpus 1 > $O/r02_scale_n1.json 2> $O/r02_scale_n1.err
  timeout 300 python tools/pcie_probe.py > $O/r02_pcie_n1.json 2> $O/r02_pcie_n1.err
else
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N > $O/r02_scale_n$N.json 2> $O/r02_scale_n$N.err
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 tools/pcie_probe.py > $O/r02_pcie_n$N.json 2> $O/r02_pcie_n$N.err
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29535 tools/pcie_probe.py --scaling weak --reps 4 > $O/r02_pcie_weak_n$N.json 2> $O/r02_pcie_weak_n$N.err
  nvidia-smi topo -m > $O/r02_topo_n$N.txt 2>&1
fi
tail -c 400 $O/r02_scale_n$N.json; tail -c 300 $O/r02_pcie_n$N.json
