#!/bin/bash
# Multi-GPU checks on N GPUs of one box: the collective tier of the test suite, the headline bench at N,
# the all-reduce(+SGD) sweep with NVLink counters, and the topology probe.   bash tools/mgpu_round.sh N
N=${1:-2}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus_n$N.log 2>&1
nvidia-smi topo -m > gpurun_out/topo_n$N.log 2>&1
python tools/topo_probe.py > gpurun_out/topo_probe_n$N.log 2>&1
timeout 1500 python -m pytest tests/test_multigpu.py -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/test_multigpu_n$N.log 2>&1
echo "test_multigpu rc=$? $(tail -1 gpurun_out/test_multigpu_n$N.log)"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29611 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.log 2> gpurun_out/bench_n$N.err
echo "bench N=$N rc=$? $(tail -c 300 gpurun_out/bench_n$N.log)"
timeout 600 $TR --master-port 29612 bench/allreduce_sweep.py --max-mb ${SWEEP_MB:-256} --out gpurun_out/sweep_n$N.jsonl > gpurun_out/sweep_n$N.log 2>&1
echo "sweep N=$N rc=$? $(tail -1 gpurun_out/sweep_n$N.log | cut -c1-300)"
