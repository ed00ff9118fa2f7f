#!/bin/bash
# round-2 multi-GPU call (run with gpurun --gpus 8): topology, the H2D fabric ceiling at 1/2/4/8 concurrent ranks with and
# without NUMA binding, and the headline bench at 8 (and 4) ranks with and without binding.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
{
  nvidia-smi topo -m; echo; lscpu | grep -i -E "model name|socket|numa|^cpu\(s\)|thread"; echo; nproc; cat /sys/devices/system/node/online
  for d in /sys/bus/pci/devices/*; do if [ -f $d/vendor ] && grep -q 0x10de $d/vendor 2>/dev/null && [ "$(cat $d/class)" = "0x030200" ]; then echo "$d numa=$(cat $d/numa_node) cpus=$(cat $d/local_cpulist)"; fi; done
  free -g | head -2
} > gpurun_out/topo_8gpu.txt 2>&1
: > gpurun_out/h2d_probe.jsonl
python tools/h2d_probe.py --gb 4 --seconds 2 >> gpurun_out/h2d_probe.jsonl 2>> gpurun_out/h2d_probe.err
for n in 2 4 8; do
  $TR --nproc-per-node $n --master-port $((29600+n)) tools/h2d_probe.py --gb 4 --seconds 2 >> gpurun_out/h2d_probe.jsonl 2>> gpurun_out/h2d_probe.err
done
cat gpurun_out/h2d_probe.jsonl
for n in 8 4; do
  $TR --nproc-per-node $n --master-port $((29700+n)) bench.py --gpus $n --steps 3 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_${n}gpu_bound.json 2> gpurun_out/bench_${n}gpu_bound.err; echo "bench $n bound rc=$?"
done
$TR --nproc-per-node 8 --master-port 29790 bench.py --gpus 8 --steps 3 --warmup 3 --no-secondary --no-cpu-baseline --no-numa > gpurun_out/bench_8gpu_unbound.json 2> gpurun_out/bench_8gpu_unbound.err; echo "bench 8 unbound rc=$?"
python - <<'PY'
import json
for f in ("bench_8gpu_bound","bench_4gpu_bound","bench_8gpu_unbound"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["value"], d["e2e"])
    except Exception as e: print(f, "ERR", e)
PY
