"""Where the corpus call spends its time end to end (AVR_FILE_TIMING=1 prints the file layer's own laps)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import avrecode_b200 as A
import bench as B

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
paths = B.gen_corpus(B.corpus_spec(n))
clips = [open(p, "rb").read() for p in paths]
total = sum(map(len, clips))
with A.Context() as ctx:
    ctx.set_priors("default")
    recs = ctx.files_compress(clips, A.MODE_SLICE)
    assert ctx.files_decompress(recs) == clips
    for _ in range(2):
        t0 = time.perf_counter(); recs = ctx.files_compress(clips, A.MODE_SLICE); t1 = time.perf_counter()
        back = ctx.files_decompress(recs); t2 = time.perf_counter()
        print("clips %d bytes %.1f MB: compress %.2f s decompress %.2f s -> %.1f MB/s end to end" % (n, total / 1e6, t1 - t0, t2 - t1, total / (t2 - t0) / 1e6), flush=True)
