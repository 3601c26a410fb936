"""Ad-hoc GPU bring-up check (run under gpurun): parity vs oracle, roundtrip, first timings."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import avrecode_b200 as A
import helpers as H

ctx = A.Context()
cases = [
    dict(width=320, height=192, frames=3, intra_period=1, seed=1),
    dict(width=480, height=272, frames=6, slices=4, seed=2),
    dict(width=640, height=368, frames=5, high=True, seed=3, ragged=True, slices=3),
]
ok_all = True
for kw in cases:
    data = H.gen_stream(**kw)
    ents = A.index_h264(data)
    for mode_file in (True, False):
        descs = H.batch_descs(ents, mode_file)
        r_o, b_o = H.oracle_compress(descs, data)
        r_g, b_g = ctx.compress(descs, data)
        same = H.slices_of(r_o, b_o) == H.slices_of(r_g, b_g) and all(r.status == 0 for r in r_g)
        dd = H.coded_descs(descs, r_g)
        r_d, b_d = ctx.decompress(dd, b_g)
        rt = all(r.status == 0 for r in r_d) and all(
            H.fix_tail(g, data[d.in_off:d.in_off + d.payload_size]) == data[d.in_off:d.in_off + d.payload_size]
            for g, d in zip(H.slices_of(r_d, b_d), descs))
        print(kw, "file" if mode_file else "slice", "compress==oracle", same, "roundtrip", rt, [r.status for r in r_g][:4], flush=True)
        ok_all &= same and rt
    rec = ctx.file_compress(data, A.MODE_SLICE)
    back = ctx.file_decompress(rec)
    print("  file api slice mode roundtrip", back == data, len(data), len(rec))
    rec = ctx.file_compress(data, A.MODE_FILE)
    back = ctx.file_decompress(rec)
    print("  file api file mode roundtrip", back == data, len(data), len(rec))
    ok_all &= back == data
print("ALL OK" if ok_all else "FAILURES")

# timing: 1080p, 1 slice/frame
for kw in [dict(width=1920, height=1088, frames=60, high=True, seed=7, kbps=30000, fps=60),
           dict(width=1920, height=1088, frames=24, intra_period=1, slices=8, seed=8)]:
    t0 = time.time(); data = H.gen_stream(**kw); t1 = time.time()
    ents = A.index_h264(data)
    descs = H.batch_descs(ents, False)
    ctx.stage(descs, data)
    for it in range(3):
        ctx.launch(0); ctx.sync()
        ms = ctx.last_kernel_ms()
    res, blob = ctx.collect()
    bins = sum(r.n_bins for r in res); pay = sum(d.payload_size for d in descs)
    print(kw, "gen %.1fs" % (t1 - t0), "slices", len(descs), "payload", pay, "coded", sum(r.out_size for r in res),
          "bins", bins, "compress kernel ms %.2f" % ms, "MB/s %.1f" % (pay / ms / 1e3), "Mbins/s %.1f" % (bins / ms / 1e3),
          "bad", sum(1 for r in res if r.status))
    dd = H.coded_descs(descs, res)
    ctx.stage(dd, blob)
    for it in range(3):
        ctx.launch(1); ctx.sync()
        ms = ctx.last_kernel_ms()
    res2, blob2 = ctx.collect()
    print("   decompress kernel ms %.2f" % ms, "MB/s %.1f" % (pay / ms / 1e3), "bad", sum(1 for r in res2 if r.status))
