#!/bin/bash
# What bounds the end-to-end path when several GPUs of one box work at once: the host's copy bandwidth (tools/hostbw),
# the PCIe links driven together (tools/pcie_bw.py on k GPUs at a time), and bench.py's headline under torchrun.
#   gpurun --gpus 8 -- 'bash tools/scale_probe.sh "4 8" > gpurun_out/scale_probe.log 2>&1'
NS=${1:-"2 4 8"}
G=$(nvidia-smi -L | wc -l)
echo "# box: $(nproc) cpus, $G GPUs, $(free -g | awk '/Mem:/{print $2}') GiB RAM"
lscpu | grep -E "Model name|Socket|NUMA node\(s\)|Thread" | sed 's/^/# /'
nvidia-smi topo -m 2>/dev/null | head -12 | sed 's/^/# /'
g++ -O2 -pthread tools/hostbw.cpp -o /tmp/hostbw && /tmp/hostbw 2>&1 | sed 's/^/hostbw: /'
for k in 1 2 4 8; do
  [ $k -le $G ] || continue
  echo "## pcie_bw on $k GPU(s) at once"
  for ((i = 0; i < k; i++)); do CUDA_VISIBLE_DEVICES=$i python tools/pcie_bw.py 2>&1 | sed "s/^/gpu$i: /" & done
  wait
done
for n in $NS; do
  [ $n -le $G ] || continue
  echo "## bench.py --only-headline, $n GPU(s)"
  if [ $n -eq 1 ]; then python bench.py --only-headline --steps 3 --warmup 3 2>&1 | tail -1
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --only-headline --steps 3 --warmup 3 2>&1 | tail -1; fi
done
