"""CPU check of the device logic (host emulation build) against the oracle on seeded streams wider than the golden set."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import avrecode_b200 as A
import helpers as H
CASES = [
    dict(width=480, height=272, frames=6, slices=4, seed=2),
    dict(width=640, height=368, frames=5, high=True, seed=3, ragged=True, slices=3),
    dict(width=720, height=480, frames=4, seed=12, high=True, p8x8=0.9, density=2.5, qp=18),
    dict(width=400, height=240, frames=12, seed=13, refs=4, pskip=0.7, pintra=0.2),
    dict(mbw=1, mbh=1, frames=9, seed=14),
    dict(mbw=255, mbh=2, frames=2, seed=15, slices=2),
    dict(width=1920, height=1088, frames=4, high=True, seed=31, kbps=30000, fps=60),
    dict(width=480, height=272, frames=12, bpics=0.8, seed=17, refs=4, high=True, p8x8=0.8, pskip=0.5),
    dict(width=640, height=368, frames=10, bpics=0.5, seed=18, refs=3, slices=3, ragged=True),
]
bad = 0
for kw in CASES:
    data = H.gen_stream(**kw)
    for mode_file in (True, False):
        descs = H.batch_descs(A.index_h264(data), mode_file)
        r_o, b_o = H.oracle_compress(descs, data)
        r_e, b_e = H.emu_compress(descs, data)
        same = H.slices_of(r_e, b_e) == H.slices_of(r_o, b_o) and [r.n_bins for r in r_e] == [r.n_bins for r in r_o]
        r_d, b_d = H.emu_decompress(H.coded_descs(descs, r_e), b_e)
        rt = H.roundtrip_ok(descs, r_d, b_d, data)
        print(kw, "file" if mode_file else "slice", "compress==oracle", same, "roundtrip", rt, flush=True)
        bad += (not same) + (not rt)
# wide slices (our own chunked format: K streams per slice): no oracle counterpart, exact round trip required,
# and the bins walked must be those of the ordinary path
for kw in CASES:
    data = H.gen_stream(**kw)
    base = H.batch_descs(A.index_h264(data), False)
    r_1, _ = H.emu_compress(base, data)
    for k in (2,):
        descs = H.batch_descs(A.index_h264(data), False)
        for d in descs:
            d.n_streams = k
        r_e, b_e = H.emu_compress(descs, data)
        same = [(r.status, r.n_bins, r.n_mbs) for r in r_e] == [(r.status, r.n_bins, r.n_mbs) for r in r_1]
        r_d, b_d = H.emu_decompress(H.coded_descs(descs, r_e), b_e)
        rt = H.roundtrip_ok(descs, r_d, b_d, data)
        grow = sum(r.out_size for r in r_e) / max(1, sum(r.out_size for r in r_1))
        print(kw, "wide", k, "bins==plain", same, "roundtrip", rt, "size x%.4f" % grow, flush=True)
        bad += (not same) + (not rt)
print("ALL OK" if not bad else "FAILURES %d" % bad)
sys.exit(1 if bad else 0)
