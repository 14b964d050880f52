#!/bin/bash
# Build experiment variants of the library next to the real one (HERE, before gpurun) and time them on the GPU box:
#   tools/exp_variants.sh build "NOSYNC NOGLD NOTW"   then   gpurun -- 'bash tools/exp_variants.sh run "NOSYNC NOGLD NOTW"'
set -e
cd "$(dirname "$0")/../gpu-fhe_b200/csrc"
if [ "$1" = build ]; then
  for v in $2; do
    flags=""; for f in ${v//+/ }; do flags="$flags -DGFHE_EXP_$f"; done
    rm -rf /tmp/exp_$v && mkdir -p /tmp/exp_$v/csrc /tmp/exp_$v/include && cp *.cu *.cuh *.inl *.h Makefile /tmp/exp_$v/csrc/ && mkdir -p /tmp/exp_$v/x/include && cp ../../include/*.h /tmp/exp_$v/x/include/
    # same relative layout as the repo: csrc/../../include
    mkdir -p /tmp/exp_$v/r/g/csrc /tmp/exp_$v/r/include && cp *.cu *.cuh *.inl *.h Makefile /tmp/exp_$v/r/g/csrc/ && cp ../../include/*.h /tmp/exp_$v/r/include/
    make -C /tmp/exp_$v/r/g/csrc -j8 -s EXTRA="$flags" && cp /tmp/exp_$v/r/g/libgpufhe_b200.so ../exp_$v.so
  done
else
  cd ../..
  python tools/kernel_times.py
  for v in $2; do TAG=$v GFHE_B200_LIB=$PWD/gpu-fhe_b200/exp_$v.so python tools/kernel_times.py; done
fi
