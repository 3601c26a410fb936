"""Short profiling case (run under gpurun, plain first, then under ncu): one compress launch and one
decompress launch of a cfg2-shaped stream (1080p High, I+P, 1 slice/frame), parity-checked by round trip."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import avrecode_b200 as A
import helpers as H

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 30
REPS = int(os.environ.get("PROFILE_REPS", "2"))
FLUSH = int(os.environ.get("PROFILE_FLUSH", "0"))   # 1: L2 flush kernel queued right before each launch; 2: and waited for
per_stream = int(sys.argv[2]) if len(sys.argv) > 2 else 0   # > 0: slices of >= 2 * per_stream bytes become wide slices
data = H.gen_stream(width=1920, height=1088, frames=frames, high=True, kbps=30000, fps=60, intra_period=30, seed=2)
descs = H.batch_descs(A.index_h264(data), False)
if per_stream:
    for d in descs:
        k = min(int(os.environ.get("PROFILE_MAX_STREAMS", "2")), d.payload_size // per_stream)
        d.n_streams = k if k >= 2 else 0
with A.Context() as ctx:
    ctx.stage(descs, data)
    for _ in range(REPS):   # the first launch of a kernel also loads it (lazily), which can serialise two kernels
        if FLUSH: ctx.flush_l2()
        if FLUSH > 1: ctx.sync()
        ctx.launch(0); ctx.sync()
    ms_c = ctx.last_kernel_ms()
    res, blob = ctx.collect()
    dd = H.coded_descs(descs, res)
    ctx.stage(dd, blob)
    for _ in range(REPS):
        if FLUSH: ctx.flush_l2()
        if FLUSH > 1: ctx.sync()
        ctx.launch(1); ctx.sync()
    ms_d = ctx.last_kernel_ms()
    r_d, b_d = ctx.collect()
    pay = sum(d.payload_size for d in descs); coded = sum(r.out_size for r in res); bins = sum(r.n_bins for r in res)
    print("slices %d payload %d coded %d bins %d max_slice_bins %d compress %.2f ms decompress %.2f ms" %
          (len(descs), pay, coded, bins, max(r.n_bins for r in res), ms_c, ms_d), flush=True)
    big = max(range(len(res)), key=lambda i: res[i].n_bins)
    for name, rr, ms in (("compress", res, ms_c), ("decompress", r_d, ms_d)):
        w = [((rr[big].reserved >> (8 * i)) & 0xFF) * (1 << 22) / 1.965e6 for i in range(4)]   # units of 2^22 cycles
        print("  longest slice, %s (~%.0f ms): waits in ms -- parser for ring room %.0f, warp behind it for entries %.0f, "
              "estimator warp for record room %.0f, range-coder warp for records %.0f" % (name, ms, w[0], w[1], w[2], w[3]), flush=True)
    assert H.roundtrip_ok(descs, r_d, b_d, data)
