"""Prints the single-lane floors of the per-bin primitives (avr_probe) measured on cuda:0; run under gpurun."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import avrecode_b200 as A

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
for name, cpb, bins in A.probes(0, iters):
    print("%8.1f cycles/bin  (%9d bins)  %s" % (cpb, bins, name), flush=True)
