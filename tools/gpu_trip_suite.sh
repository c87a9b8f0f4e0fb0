#!/bin/bash
# the whole GPU suite as the driver runs it at round end, then the smoke entry point
mkdir -p gpurun_out
exec > gpurun_out/suite.log 2>&1
date +%s
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -6
date +%s
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
