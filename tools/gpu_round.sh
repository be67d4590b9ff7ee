#!/bin/bash
# One gpurun call's worth of checks on a single B200: every GPU test file in its own process (a trap in
# one kernel must not poison the others), smoke(), the topology probe and a short bench.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
nvidia-smi -L > gpurun_out/gpus.log 2>&1
python tools/topo_probe.py > gpurun_out/topo_probe.log 2>&1
rc_all=0
for f in tests/test_gpu_gemm.py tests/test_gpu_kernels.py tests/test_gpu_fused_sync.py tests/test_gpu_emulated_world.py tests/test_gpu_trainer.py tests/test_zz_no_alloc.py; do
  name=$(basename $f .py)
  timeout 900 python -m pytest $f -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/$name.log 2>&1
  rc=$?
  echo "$name rc=$rc $(tail -1 gpurun_out/$name.log)"
  [ $rc -ne 0 ] && rc_all=1
done
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$? $(tail -1 gpurun_out/smoke.log | cut -c1-200)"
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_n1.log 2> gpurun_out/bench_n1.err; echo "bench rc=$? $(tail -c 600 gpurun_out/bench_n1.log)"
if [ "$1" = "profile" ]; then
  timeout 300 python tools/step_timeline.py > gpurun_out/timeline_wide.md 2> gpurun_out/timeline.err; echo "timeline rc=$?"
  timeout 300 python tools/ncu_workload.py > gpurun_out/ncu_plain.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -f \
      -o gpurun_out/prof_step python tools/ncu_workload.py > gpurun_out/ncu_run.log 2>&1
  echo "ncu rc=$? $(tail -2 gpurun_out/ncu_run.log | cut -c1-200)"
fi
exit $rc_all
