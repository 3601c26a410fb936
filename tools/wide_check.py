"""GPU check of wide slices (run under gpurun): device bytes == host emulation of the same code, exact round
trip, container round trip, and the kernel times of the 1080p profile case with and without wide slices."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import avrecode_b200 as A
import helpers as H

ctx = A.Context()
CASES = [
    dict(width=480, height=272, frames=6, slices=4, seed=2),
    dict(width=640, height=368, frames=5, high=True, seed=3, ragged=True, slices=3),
    dict(mbw=255, mbh=2, frames=2, seed=15, slices=2),
    dict(width=480, height=272, frames=12, bpics=0.8, seed=17, refs=4, high=True, p8x8=0.8, pskip=0.5),
    dict(width=1920, height=1088, frames=4, high=True, seed=31, kbps=30000, fps=60),
]
bad = 0
for kw in CASES:
    data = H.gen_stream(**kw)
    for k in (2,):
        descs = H.batch_descs(A.index_h264(data), False)
        for i, d in enumerate(descs):
            d.n_streams = k if i % 3 != 2 else 0    # wide and ordinary slices in one batch
        r_e, b_e = H.emu_compress(descs, data)
        r_g, b_g = ctx.compress(descs, data)
        same = H.slices_of(r_e, b_e) == H.slices_of(r_g, b_g) and all(r.status == 0 for r in r_g)
        r_d, b_d = ctx.decompress(H.coded_descs(descs, r_g), b_g)
        rt = H.roundtrip_ok(descs, r_d, b_d, data)
        print(kw, "wide", k, "device==emulation", same, "roundtrip", rt, flush=True)
        bad += (not same) + (not rt)
    os.environ["AVR_WIDE_STREAM_BYTES"] = "2048"
    rec = ctx.file_compress(data, A.MODE_SLICE)
    back = ctx.file_decompress(rec)
    os.environ.pop("AVR_WIDE_STREAM_BYTES")
    print("  container round trip with wide slices", back == data, len(data), len(rec), flush=True)
    bad += back != data
print("ALL OK" if not bad else "FAILURES %d" % bad)
