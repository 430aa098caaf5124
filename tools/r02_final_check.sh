#!/bin/bash
# Last GPU check of the round: smoke and the GPU suite on the final libraries (host layer changed since r02_s: glm half rounding,
# T*R*S zero signs, tinyobjloader triangulation, threaded builders).
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/r02_z_gpu_suite.txt
