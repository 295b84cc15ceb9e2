for z in 1 0.0001; do for g in trot tws crawl stand; do LMPC_Z0_SCALE=$z timeout 200 python tools_profile.py 4096 reference_a1 3 $g 2>&1 | tail -1 | cut -c1-330; done; done
