#!/bin/bash
# staggered half tiles: what the two sides contend for — phase clocks with the epilogue's stores into X switched off (dbg flag 8)
for fl in 4 12; do
echo "--- MZB_FUSED_DBGFLAGS=$fl split 2,2"
MZB_FUSED_DBGFLAGS=$fl MZB_TOWER_SPLIT=2,2 MZB_FUSED_TIMING=1 timeout 300 python tools/profile_recurrent.py c4d 592 2 2>&1 | grep "fused" | tail -24 | head -10
done
for fl in 0 8; do
echo "--- MZB_FUSED_DBGFLAGS=$fl split 0,0"
MZB_FUSED_DBGFLAGS=$fl MZB_TOWER_SPLIT=0,0 MZB_FUSED_TIMING=1 timeout 300 python tools/profile_recurrent.py c4d 592 2 2>&1 | grep "fused" | tail -24 | head -10
done
for sp in 0,0 1,1 2,2; do MZB_FUSED_DBGFLAGS=8 MZB_TOWER_SPLIT=$sp timeout 300 python tools/profile_recurrent.py c4d 592 200 2>&1 | grep recurrent | sed "s/^/nostores split=$sp /"; done
