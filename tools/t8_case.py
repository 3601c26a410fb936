"""8x8-transform-heavy slices (cfg4-like: High profile, 85 % of the macroblocks with transform_size_8x8_flag): cycles per bin of the
longest slice, ordinary and wide (K = 2)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import avrecode_b200 as A
import helpers as H

CLK = 1.965e9
data = H.gen_stream(width=1920, height=1088, frames=12, high=True, p8x8=0.85, kbps=60000, fps=30, qp=20, seed=4)
with A.Context() as ctx:
    for k in (0, 2):
        descs = H.batch_descs(A.index_h264(data), False)
        for d in descs:
            d.n_streams = k
        ctx.stage(descs, data)
        for _ in range(2):
            ctx.launch(0); ctx.sync()
        ms_c = ctx.last_kernel_ms()
        res, blob = ctx.collect()
        assert all(r.status == 0 for r in res)
        ctx.stage(H.coded_descs(descs, res), blob)
        for _ in range(2):
            ctx.launch(1); ctx.sync()
        ms_d = ctx.last_kernel_ms()
        r_d, b_d = ctx.collect()
        assert H.roundtrip_ok(descs, r_d, b_d, data)
        longest = max(r.n_bins for r in res)
        print("K=%d: %d slices, longest %d bins: compress %.1f ms = %.0f cycles/bin, decompress %.1f ms = %.0f cycles/bin" %
              (k, len(descs), longest, ms_c, ms_c * 1e-3 * CLK / longest, ms_d, ms_d * 1e-3 * CLK / longest), flush=True)
