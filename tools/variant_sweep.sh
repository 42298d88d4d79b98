#!/bin/bash
# usage (on a GPU box): tools/variant_sweep.sh <out> <sizes...>: tools/peel_times.py for the built library and for every tools/variants/*.so
out=$1; shift
cp evox_b200/libevoxmo.so /tmp/libevoxmo_default.so
echo "== default" >> $out; python tools/peel_times.py "$@" 2>&1 | grep -v Warning >> $out
for v in tools/variants/libevoxmo_*.so; do
  cp $v evox_b200/libevoxmo.so
  echo "== $v" >> $out; python tools/peel_times.py "$@" 2>&1 | grep -v Warning >> $out
done
cp /tmp/libevoxmo_default.so evox_b200/libevoxmo.so
