#!/bin/bash
# r2h: GPU tests + smoke + full bench (both arms)
T=${1:-r2h}; O=gpurun_out
set -x
timeout 900 python -m pytest tests -m gpu -x -q > $O/${T}_pytest.txt 2>&1; tail -5 $O/${T}_pytest.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.txt 2>&1; tail -3 $O/${T}_smoke.txt
timeout 1200 python bench.py --steps 5 > $O/${T}_bench.json 2> $O/${T}_bench.err; tail -3 $O/${T}_bench.err; head -c 600 $O/${T}_bench.json
AVR_BENCH_CORPUS=0 timeout 600 python bench.py --impl reference --steps 2 --warmup 0 > $O/${T}_bench_reference.json 2> $O/${T}_bench_reference.err; head -c 400 $O/${T}_bench_reference.json
