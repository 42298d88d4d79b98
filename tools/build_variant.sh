#!/bin/bash
# usage: tools/build_variant.sh <name> [-DEVX_WPG=.. -DEVX_CW=.. -DEVX_CC=..]  -> tools/variants/libevoxmo_<name>.so
# Builds nds.cu with other constants into a separate library so that several variants can be timed on ONE GPU box in one
# call (copy a variant over evox_b200/libevoxmo.so on the box, run tools/peel_times.py): profiles/r1_variants_constants.log.
set -e
name=$1; shift
cd "$(dirname "$0")/.."
mkdir -p tools/variants /tmp/evx_variants
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -diag-suppress 177 "$@" -c evox_b200/csrc/nds.cu -o /tmp/evx_variants/nds_$name.o
objs=$(ls evox_b200/build/*.o | grep -v '/nds\.o$')
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o tools/variants/libevoxmo_$name.so /tmp/evx_variants/nds_$name.o $objs -lcudart
echo built tools/variants/libevoxmo_$name.so
