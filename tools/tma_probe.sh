#!/bin/bash
# build first: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tma_probe tools/tma_probe.cu -lcuda
P=tools/tma_probe
# bk rows panels S hot ctas ld iters W gdesc
for hot in 1 0; do
 $P 32 128 1 3 $hot 148 4096 2000 1 0
 $P 32 128 1 3 $hot 148 4096 2000 1 1
 $P 32 128 1 3 $hot 148 4096 2000 2 0
 $P 32 128 1 3 $hot 148 4096 2000 4 0
 $P 32 128 1 2 $hot 148 4096 2000 6 0
 $P 32 64 1 3 $hot 148 4096 2000 4 0
 $P 32 64 1 3 $hot 148 4096 2000 8 0
 $P 16 128 1 3 $hot 148 4096 2000 4 0
 $P 16 128 1 3 $hot 148 4096 2000 8 0
 $P 32 128 1 3 $hot 1 4096 2000 4 0
done
