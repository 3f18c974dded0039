#!/bin/bash
# what a tower tile costs by games per tile: whole-launch times at cohorts that fill 148 SMs with 4-, 3-, 2-, 1-game tiles, and the in-kernel phase clocks
for T in 592 444 296 148 512 1024; do timeout 300 python tools/profile_recurrent.py c4d $T 50 2>&1 | grep recurrent | sed "s/^/T=$T /"; done
MZB_FUSED_TIMING=1 timeout 300 python tools/profile_recurrent.py c4d 1024 2 2>&1 | grep "fused" | tail -36
echo "--- weight stream out of the picture (dbg flag 2)"
MZB_FUSED_DBGFLAGS=2 MZB_FUSED_TIMING=1 timeout 300 python tools/profile_recurrent.py c4d 1024 2 2>&1 | grep "fused" | tail -3
for T in 592 444; do MZB_FUSED_DBGFLAGS=2 timeout 300 python tools/profile_recurrent.py c4d $T 50 2>&1 | grep recurrent | sed "s/^/noweights T=$T /"; done
