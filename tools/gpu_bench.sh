#!/bin/bash
# full bench (both arms) + the one failing test.  usage: tools/gpu_bench.sh <tag>
T=${1:-rb}; O=gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "corpus or container" > $O/${T}_pytest.txt 2>&1; tail -3 $O/${T}_pytest.txt
/usr/bin/env time -v true 2>/dev/null
SECONDS=0
timeout 1500 python bench.py --steps 5 > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench took $SECONDS s" | tee -a $O/${T}_bench.err; tail -3 $O/${T}_bench.err; head -c 400 $O/${T}_bench.json
