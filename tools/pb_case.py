"""Cycles per bin of P and B slices (macroblock-level syntax is a quarter of their bins): up to 148 inter slices of
corpus-shaped clips in one launch, one block per SM at most; time = the longest slice."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import avrecode_b200 as A
import helpers as H
import bench as B

CLK = 1.965e9
with A.Context() as ctx:
    for sp in B.corpus_spec(8)[:3]:
        kw = dict(sp); kw["frames"] = min(kw["frames"], 60)
        data = H.gen_stream(**kw)
        descs = [d for d in H.batch_descs(A.index_h264(data), False) if d.slice_type != 2][:148]
        ctx.stage(descs, data)
        for _ in range(2):
            ctx.launch(0); ctx.sync()
        ms_c = ctx.last_kernel_ms()
        res, blob = ctx.collect()
        assert all(r.status == 0 for r in res)
        dd = H.coded_descs(descs, res)
        ctx.stage(dd, blob)
        for _ in range(2):
            ctx.launch(1); ctx.sync()
        ms_d = ctx.last_kernel_ms()
        r_d, b_d = ctx.collect()
        assert H.roundtrip_ok(descs, r_d, b_d, data)
        longest = max(r.n_bins for r in res)
        print("%dx%d %d kb/s: %d inter slices, longest %d bins: compress %.2f ms = %.0f cycles/bin, decompress %.2f ms = %.0f cycles/bin" %
              (kw["width"], kw["height"], kw["kbps"], len(descs), longest, ms_c, ms_c * 1e-3 * CLK / longest, ms_d, ms_d * 1e-3 * CLK / longest), flush=True)
