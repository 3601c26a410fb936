"""BASELINE config 5 shape (scaled): a corpus of independent 720p/1080p clips through the batch file API
(avr_files_compress / avr_files_decompress, host buffers).  Prints throughput; run under gpurun."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import avrecode_b200 as A
import helpers as H

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
t0 = time.time()
clips = []
for i in range(n):
    hd = i & 1
    clips.append(H.gen_stream(width=1920 if hd else 1280, height=1088 if hd else 720, frames=8 + (i * 7) % 17, seed=1000 + i,
                              high=bool(i % 3), kbps=5000 + (i * 3571) % 25000, fps=30, intra_period=12))
total = sum(len(c) for c in clips)
print("generated %d clips, %.1f MB in %.1f s" % (n, total / 1e6, time.time() - t0), flush=True)
with A.Context() as ctx:
    descs = []
    for c in clips:
        descs += H.batch_descs(A.index_h264(c), False)
    nsl = len(descs)
    for it in range(3):
        t0 = time.perf_counter(); recs = ctx.files_compress(clips, A.MODE_SLICE); tc = time.perf_counter() - t0
        t0 = time.perf_counter(); back = ctx.files_decompress(recs); td = time.perf_counter() - t0
    assert back == clips
    coded = sum(len(r) for r in recs)
    print("corpus %d clips %d slices %.1f MB: e2e compress %.1f ms (%.1f MB/s) decompress %.1f ms (%.1f MB/s) roundtrip %.1f MB/s saved %.3f%%"
          % (n, nsl, total / 1e6, tc * 1e3, total / tc / 1e6, td * 1e3, total / td / 1e6, total / (tc + td) / 1e6, 100 * (1 - coded / total)), flush=True)
    # kernel-only on the same batch
    big = b"".join(clips)
    offs, d2 = 0, []
    for c in clips:
        for d in H.batch_descs(A.index_h264(c), False):
            d.in_off += offs
            d2.append(d)
        offs += len(c)
    ctx.stage(d2, big)
    for it in range(3):
        ctx.launch(0); ctx.sync()
    ms_c = ctx.last_kernel_ms()
    res, blob = ctx.collect()
    ctx.stage(H.coded_descs(d2, res), blob)
    for it in range(3):
        ctx.launch(1); ctx.sync()
    ms_d = ctx.last_kernel_ms()
    bins = sum(r.n_bins for r in res)
    print("kernel only: compress %.2f ms (%.1f MB/s, %.0f Mbins/s) decompress %.2f ms (%.1f MB/s)" %
          (ms_c, total / ms_c / 1e3, bins / ms_c / 1e3, ms_d, total / ms_d / 1e3), flush=True)
