"""Throughput batch: every slice of N corpus-shaped clips in one launch, estimators resident in shared memory (2 blocks
per SM) against the split placement (4 blocks per SM, 8x8 significance slab in global memory)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import avrecode_b200 as A
import helpers as H
import bench as B

n_clips = int(sys.argv[1]) if len(sys.argv) > 1 else 16
pieces, descs, offs = [], [], 0
for sp in B.corpus_spec(n_clips):
    kw = dict(sp); kw["frames"] = min(kw["frames"], 60)
    c = H.gen_stream(**kw)
    for d in H.batch_descs(A.index_h264(c), False):
        d.in_off += offs
        descs.append(d)
    pad = (-len(c)) % 16
    pieces += [c, bytes(pad)]
    offs += len(c) + pad
data = b"".join(pieces)
pay = sum(d.payload_size for d in descs)
for resident in [int(x) for x in os.environ.get("THROUGHPUT_RESIDENT", "1,0,1,0").split(",")]:
    with A.Context(resident=resident) as ctx:
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
        bins = sum(r.n_bins for r in res)
        print("resident=%d: %d slices %.1f MB payload %d bins: compress %.1f ms (%.0f MB/s) decompress %.1f ms (%.0f MB/s) round trip %.1f MB/s" %
              (resident, len(descs), pay / 1e6, bins, ms_c, pay / ms_c / 1e3, ms_d, pay / ms_d / 1e3, pay / (ms_c + ms_d) / 1e3), flush=True)
