#!/bin/bash
# One gpurun call: the bench's launch list, a full capture of the three config-4 kernels, and DRAM-traffic captures of
# the pipelined single-launch kernels against the per-pass kernels (30-bit limbs, config-4 shape).  Each ncu command runs
# only after the same command line exited 0 without ncu.
mkdir -p gpurun_out
PROF="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-sweep"
timeout 300 $PROF > gpurun_out/r02_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv $PROF > gpurun_out/r02_ncu_launches.log 2>&1
echo "launch list exit $?"
timeout 300 $PROF > gpurun_out/r02_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_(row_polymul|col_fwd|col_inv)' -s 12 -c 3 -o gpurun_out/r02_prof -f $PROF > gpurun_out/r02_ncu_full.log 2>&1
echo "full capture exit $?"
KT="python tools/kernel_times.py"
export BITS=30
GFHE_PIPE=1 timeout 300 $KT > gpurun_out/r02_kt_pipe30.log 2>&1 &&
GFHE_PIPE=1 timeout 600 ncu --set full --clock-control none -k regex:k_pipe_polymul -s 5 -c 1 -o gpurun_out/r02_pipe30_polymul -f $KT > gpurun_out/r02_ncu_p1.log 2>&1
echo "pipe polymul exit $?"
GFHE_PIPE=1 timeout 600 ncu --set full --clock-control none -k regex:k_pipe_ntt_fwd -s 5 -c 1 -o gpurun_out/r02_pipe30_ntt -f $KT > gpurun_out/r02_ncu_p2.log 2>&1
echo "pipe ntt exit $?"
GFHE_PIPE=0 timeout 300 $KT > gpurun_out/r02_kt_nopipe30.log 2>&1 &&
GFHE_PIPE=0 timeout 600 ncu --set full --clock-control none -k regex:'k_row_fwd|k_col_fwd' -s 160 -c 2 -o gpurun_out/r02_nopipe30_ntt -f $KT > gpurun_out/r02_ncu_p3.log 2>&1
echo "per-pass ntt exit $?"
export BITS=60
GFHE_PIPE=1 timeout 300 $KT > gpurun_out/r02_kt_pipe60.log 2>&1 &&
GFHE_PIPE=1 timeout 600 ncu --set full --clock-control none -k regex:k_pipe_polymul -s 5 -c 1 -o gpurun_out/r02_pipe60_polymul -f $KT > gpurun_out/r02_ncu_p4.log 2>&1
echo "pipe polymul 60 exit $?"
# the reports embed the library's 60 MB cubin: export what is read later (raw metrics per kernel, source-level counters of
# the main capture) and drop the .ncu-rep files, so that gpurun_out/ stays small enough to travel back
for r in gpurun_out/r02_*.ncu-rep; do
  b=${r%.ncu-rep}
  ncu -i $r --page raw --csv > ${b}_raw.csv 2>/dev/null
done
ncu -i gpurun_out/r02_prof.ncu-rep --page source --csv 2>/dev/null | gzip -9 > gpurun_out/r02_prof_source.csv.gz
ncu -i gpurun_out/r02_prof.ncu-rep --page details --csv > gpurun_out/r02_prof_details.csv 2>/dev/null
rm -f gpurun_out/r02_*.ncu-rep
ls -la gpurun_out/ | head -40
