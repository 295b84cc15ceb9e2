#!/bin/bash
# Config 3 (Go1 mixed gaits, H=20) and the reference preset (H=30) throughput on one GPU.
export LMPC_DEBUG_STATS=1
python tools_profile.py 16384 northstar_go1_h20 2 2>&1 | tail -2
python tools_profile.py 8192 reference_a1 2 2>&1 | tail -2
