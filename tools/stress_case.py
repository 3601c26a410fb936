"""Randomised parity run on the device (run under gpurun): seeded streams of random shape, random stream counts per
slice (0, 2..8), with and without a random prior table, both estimator placements, repeated launches of the same batch
(the wide roles synchronise through global memory: a race would show as a difference between repetitions or against the
oracle).  usage: python tools/stress_case.py [seed] [streams]"""
import os, random, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import avrecode_b200 as A
import helpers as H

seed = int(sys.argv[1]) if len(sys.argv) > 1 else 1
n_streams = int(sys.argv[2]) if len(sys.argv) > 2 else 10
rng = random.Random(seed)
ctxs = {r: A.Context(resident=r) for r in (1, 0)}
bins = 0
for it in range(n_streams):
    w, h = rng.choice([(320, 192), (480, 272), (640, 368), (1280, 720), (1920, 1088)])
    kw = dict(width=w, height=h, frames=rng.randint(2, 5 if w < 1000 else 3), seed=rng.randint(1, 10 ** 6), high=rng.random() < 0.6,
              slices=rng.choice([1, 1, 2, 4]), bpics=rng.choice([0, 0, 0.5]), refs=rng.randint(1, 3), kbps=rng.choice([2000, 8000, 30000]),
              fps=30, intra_period=rng.choice([1, 3, 30]), ragged=rng.random() < 0.3)
    if kw["high"]:
        kw["p8x8"] = rng.choice([0.0, 0.5, 0.9])
    data = H.gen_stream(**kw)
    use_priors = rng.random() < 0.5
    table = H.random_priors(rng.randint(1, 10 ** 6)) if use_priors else None
    H.set_test_priors(table)
    descs = H.batch_descs(A.index_h264(data), False)
    for d in descs:
        d.n_streams = rng.choice([0, 0, 2, 2, 3, 4, 5, 8])
    if use_priors:
        H.with_priors(descs)
    r_o, b_o = H.oracle_compress(descs, data)
    want = H.slices_of(r_o, b_o)
    for resident, ctx in ctxs.items():
        ctx.set_priors(bytes(table) if use_priors else None)
        for rep in range(3):
            res, blob = ctx.compress(descs, data)
            assert [r.status for r in res] == [r.status for r in r_o], (it, resident, rep, "status")
            assert H.slices_of(res, blob) == want, (it, resident, rep, "compress != oracle")
            r_d, b_d = ctx.decompress(H.coded_descs(descs, res), blob)
            assert H.roundtrip_ok(descs, r_d, b_d, data), (it, resident, rep, "round trip")
    bins += sum(r.n_bins for r in r_o)
    print("stream %d ok: %dx%d %d slices priors=%s" % (it, w, h, len(descs), use_priors), flush=True)
H.set_test_priors(None)
print("STRESS OK: %d streams, %d bins, 6 device round trips each" % (n_streams, bins))
