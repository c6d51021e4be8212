#!/bin/bash
# call J (1 GPU): ray-box switch on the other scene shapes (S4b random nodes, S5 lattice 100 rays/pair)
python tools/s4b_probe.py 2>&1 | tail -3
python tools/s5_probe.py 2>&1 | grep -E "connect|phases" | tail -2
SSPP_NVCC_EXTRA="-DSSPP_RAY_BOX=0" python -c "from sspp_b200 import build; build.build_gpu(force=True)" && echo "variant RAY_BOX=0"
python tools/s4b_probe.py 2>&1 | tail -3
python tools/s5_probe.py 2>&1 | grep -E "connect|phases" | tail -2
python tools/config2_full.py 2>&1 | cut -c1-900
