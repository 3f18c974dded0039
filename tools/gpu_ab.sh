#!/bin/bash
# A/B of two libraries on one box: $1 = variant library; recurrent pass times for every preset, twice
for rep in 1 2; do
for lib in muzero-cpp_b200/libmzb200.so $1; do
  for spec in "c4d 592" "c4d 1024" "c4r 1024" "ale 1024"; do set -- $spec
    MZB200_LIB=$PWD/$lib timeout 300 python tools/profile_recurrent.py $1 $2 200 2>&1 | grep recurrent | sed "s|^|$(basename $lib) $1 T=$2 |"
  done
done
done
