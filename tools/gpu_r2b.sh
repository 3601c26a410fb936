#!/bin/bash
# r2b: GPU tests + floors + a short bench (both arms) of the block-entry ring protocol
T=${1:-r2b}; O=gpurun_out
set -x
timeout 900 python -m pytest tests -m gpu -x -q > $O/${T}_pytest.txt 2>&1; tail -5 $O/${T}_pytest.txt
timeout 120 python tools/probe_case.py 20000 > $O/${T}_probes.txt 2>&1; cat $O/${T}_probes.txt
AVR_BENCH_CORPUS_CLIPS=32 timeout 900 python bench.py --steps 3 > $O/${T}_bench.json 2> $O/${T}_bench.err; tail -3 $O/${T}_bench.err; head -c 1500 $O/${T}_bench.json
AVR_BENCH_CORPUS=0 timeout 600 python bench.py --impl reference --steps 2 --warmup 0 > $O/${T}_bench_reference.json 2> $O/${T}_bench_reference.err; cat $O/${T}_bench_reference.json
