#!/bin/bash
# round-2 GPU call 1: topology, integer-pipe microbenchmarks, parity suite, the default bench line
mkdir -p gpurun_out
{
  nvidia-smi topo -m; echo; lscpu | head -40; echo; nproc; cat /sys/devices/system/node/online 2>/dev/null
  for d in /sys/bus/pci/devices/*; do if [ -f $d/vendor ] && grep -q 0x10de $d/vendor 2>/dev/null; then echo "$d numa=$(cat $d/numa_node) cpus=$(cat $d/local_cpulist)"; fi; done
  free -g
} > gpurun_out/topo.txt 2>&1
./tools/bin/ubench_int > gpurun_out/ubench_int.jsonl 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python tools/h2d_probe.py --gb 4 > gpurun_out/h2d_probe_1gpu.jsonl 2>&1
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref rc=$?"
head -c 1500 gpurun_out/bench_default.json
