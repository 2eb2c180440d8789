#!/bin/bash
# Developer aid: disassemble the fp64 / one-quantile streaming kernel with line info into gpurun_out/cub/d1.sass
set -e
cd /root/repo/gpurun_out && mkdir -p cub && cd cub && rm -f *.cubin
cuobjdump -xelf all /root/repo/research*/build/cg_launch_warp_d1.o > /dev/null
nvdisasm -g -c cg_warp_inst.sm_100a.cubin > d1.sass 2>/dev/null
echo "static instructions: $(grep -c '^\s*/\*[0-9a-f]*\*/' d1.sass)"
