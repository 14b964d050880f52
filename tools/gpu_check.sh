#!/bin/bash
# One gpurun call: parity tests, smoke, bench, then (only after the plain runs exited 0) the ncu
# launch list of the same bench command and one --set full capture of the dominant kernels.
# Outputs land in gpurun_out/.   usage: tools/gpu_check.sh [notests] [noncu] [micro]
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > gpurun_out/smi.csv 2>&1
nproc > gpurun_out/nproc.txt
if [[ " $* " == *" micro "* ]]; then
  for b in pipebench bflybench; do [ -x tools/$b ] && timeout 120 tools/$b > gpurun_out/$b.txt 2>&1; done
  cat gpurun_out/bflybench.txt
fi
if [[ " $* " != *" notests "* ]]; then
  timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
  tail -15 gpurun_out/pytest_gpu.log
  timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
  tail -3 gpurun_out/smoke.log
fi
timeout 600 python bench.py --steps 100 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
if [[ " $* " != *" noncu "* ]]; then
  PROF="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
  timeout 300 $PROF > gpurun_out/plain.log 2>&1 &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
      --log-file gpurun_out/launches.csv $PROF > gpurun_out/ncu_launches.log 2>&1
  echo "ncu launches exit $?"
  timeout 300 $PROF > gpurun_out/plain2.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_(row|col)_' -s 12 -c 3 \
      -o gpurun_out/prof -f $PROF > gpurun_out/ncu_full.log 2>&1
  echo "ncu full exit $?"
  tail -3 gpurun_out/ncu_full.log
fi
