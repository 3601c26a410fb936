"""Diagnostic: wide slices of a seeded stream, per stream count, on the device (statuses and first difference)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import avrecode_b200 as A
import helpers as H

kw = dict(width=640, height=368, frames=5, high=True, seed=3, ragged=True, slices=3)
data = H.gen_stream(**kw)
with A.Context() as ctx:
    for K in (2, 3, 4, 5, 8, -1):
        for rep in range(3):
            descs = H.batch_descs(A.index_h264(data), False)
            for i, d in enumerate(descs):
                d.n_streams = H.WIDE_PATTERN[i % len(H.WIDE_PATTERN)] if K < 0 else K
            res, blob = ctx.compress(descs, data)
            dd = H.coded_descs(descs, res)
            r_d, b_d = ctx.decompress(dd, blob)
            got = H.slices_of(r_d, b_d)
            bad = []
            for i, d in enumerate(descs):
                want = data[d.in_off:d.in_off + d.payload_size]
                if r_d[i].status != 0 or H.fix_tail(got[i], want) != want:
                    n = next((k for k in range(min(len(got[i]), len(want))) if got[i][k] != want[k]), -1)
                    bad.append((i, d.n_streams, d.first_mb, r_d[i].status, r_d[i].n_mbs, res[i].n_mbs, len(got[i]), len(want), n))
            print("K", K, "rep", rep, "bad", bad, flush=True)
