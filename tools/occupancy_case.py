"""Does a block run slower when every SM carries two?  N equal-sized intra slices (1080p, 8 slices per frame) in one
launch, N from a few (one block per SM at most, most SMs idle) to several per block slot; prints time and the cycles
one block spends per bin."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import avrecode_b200 as A
import helpers as H

data = H.gen_stream(width=1920, height=1088, frames=74, intra_period=1, slices=8, kbps=60000, fps=30, seed=3)
all_descs = H.batch_descs(A.index_h264(data), False)
CLK = 1.965e9
with A.Context() as ctx:
    for n in (16, 64, 148, 296, 592):
        descs = all_descs[:n]
        ctx.stage(descs, data)
        for _ in range(2):
            ctx.launch(0); ctx.sync()
        ms_c = ctx.last_kernel_ms()
        res, blob = ctx.collect()
        dd = H.coded_descs(descs, res)
        ctx.stage(dd, blob)
        for _ in range(2):
            ctx.launch(1); ctx.sync()
        ms_d = ctx.last_kernel_ms()
        bins = sum(r.n_bins for r in res); longest = max(r.n_bins for r in res)
        waves = max(1.0, n / 296.0)
        print("slices %4d bins %10d longest %7d  compress %7.2f ms (%.0f cycles/bin of the longest / wave)  decompress %7.2f ms (%.0f)" %
              (n, bins, longest, ms_c, ms_c * 1e-3 * CLK / longest / waves, ms_d, ms_d * 1e-3 * CLK / longest / waves), flush=True)
