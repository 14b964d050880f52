#!/bin/bash
# A/B of the pipelined single-launch kernels against one launch per pass, config-4 shape, every policy (run on the GPU box)
for bits in 60 50 30; do
  for pipe in 0 1; do BITS=$bits GFHE_PIPE=$pipe timeout 300 python tools/kernel_times.py; done
done
for pipe in 0 1; do BITS=30 PACKED=1 GFHE_PIPE=$pipe timeout 300 python tools/kernel_times.py; done
