#!/bin/bash
MZB_FUSED_TIMING=1 timeout 300 python tools/profile_recurrent.py c4r 1024 2 2>&1 | grep "fused" | tail -26
