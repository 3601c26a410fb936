"""Writes tests/golden/bench_gate.json and tests/golden/corpus_gate.json (offline, CPU, ~2 minutes on 8 cores):

  bench_gate.json   for every single-file bench workload (cfg1..cfg4): sha256 of the stream, sha256 over the ORACLE's coded
                    bytes of ALL its ordinary slices (slice mode), every slice's bin count, and -- where oracle/_ref exists --
                    the size of the container the UNMODIFIED reference compressor writes for the whole file (its bytes saved
                    on the same input).  bench.py's pre-timing parity gate compares the CUDA path's bytes with these digests,
                    so the timed arm never executes anything under oracle/.
  corpus_gate.json  the same per clip for the 64-clip corpus of tests/test_gpu.py::test_corpus_every_slice_pinned.

Run here: python tests/make_bench_gate.py"""
import hashlib, json, os, subprocess, sys, tempfile
from concurrent.futures import ProcessPoolExecutor
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import avrecode_b200 as A
import bench as B
import helpers as H


def _oracle_part(args):
    path, lo, hi = args
    with open(path, "rb") as f:
        data = f.read()
    plain = H.batch_descs(A.index_h264(data), mode_file=False)[lo:hi]
    r_o, b_o = H.oracle_compress(plain, data)
    assert all(r.status == 0 for r in r_o), (path, lo)
    return [(int(r.n_bins), bytes(s)) for r, s in zip(r_o, H.slices_of(r_o, b_o))]


def digest_of(slices):
    h = hashlib.sha256()
    for s in slices:
        h.update(len(s).to_bytes(8, "little")); h.update(s)
    return h.hexdigest()


def oracle_all_slices(path, pool, parts=16):
    with open(path, "rb") as f:
        data = f.read()
    n = len(H.batch_descs(A.index_h264(data), mode_file=False))
    step = max(1, (n + parts - 1) // parts)
    res = []
    for part in pool.map(_oracle_part, [(path, lo, min(n, lo + step)) for lo in range(0, n, step)]):
        res += part
    return data, [r[0] for r in res], [r[1] for r in res]


def ref_container_bytes(path):
    ref = H.ref_binary("recode_ref")
    if not ref:
        return None
    with tempfile.TemporaryDirectory() as td:
        out = os.path.join(td, "x.rec")
        subprocess.run([ref, "compress", path, out], check=True, capture_output=True)
        return os.path.getsize(out)


def main():
    workers = os.cpu_count() or 1
    out = {"how": "tests/make_bench_gate.py: oracle (oracle/_build/libavr_oracle.so) slice-mode compress of every ordinary slice; "
                  "ref_rec_bytes = size of `recode_ref compress` of the whole file (unmodified reference, oracle/_ref)", "workloads": {}}
    with ProcessPoolExecutor(workers) as pool:
        refs = {}
        for name in ([] if "--corpus-only" in sys.argv else sorted(B.WORKLOADS)):
            _, path = B.gen_workload(name, 0)
            refs[name] = pool.submit(ref_container_bytes, path)
        for name in sorted(refs):
            _, path = B.gen_workload(name, 0)
            data, n_bins, slices = oracle_all_slices(path, pool)
            out["workloads"][name] = {"stream_sha256": hashlib.sha256(data).hexdigest(), "stream_bytes": len(data), "n_slices": len(slices),
                                      "coded_sha256": digest_of(slices), "coded_bytes": sum(len(s) for s in slices), "n_bins": n_bins}
            print(name, {k: v for k, v in out["workloads"][name].items() if k != "n_bins"}, flush=True)
        for name in sorted(refs):
            out["workloads"][name]["ref_rec_bytes"] = refs[name].result()
            print(name, "reference container:", out["workloads"][name]["ref_rec_bytes"], flush=True)
        if refs:
            with open(os.path.join(ROOT, "tests", "golden", "bench_gate.json"), "w") as f:
                json.dump(out, f, indent=1)
        # the GPU test's corpus: the first 64 clips of the cfg5 shape, a sixth of their length
        clips = B.corpus_spec(64, frame_div=6)
        paths = B.gen_corpus(clips, workers)
        gate = {"how": "first 64 clips of bench.corpus_spec(frame_div=6); per clip sha256 of the stream and of the oracle's coded slices",
                "clips": []}
        for spec, path in zip(clips, paths):
            data, n_bins, slices = oracle_all_slices(path, pool, parts=2)
            gate["clips"].append({"stream_sha256": hashlib.sha256(data).hexdigest(), "coded_sha256": digest_of(slices),
                                  "n_slices": len(slices), "n_bins": sum(n_bins)})
        with open(os.path.join(ROOT, "tests", "golden", "corpus_gate.json"), "w") as f:
            json.dump(gate, f, indent=1)
        print("corpus gate:", len(gate["clips"]), "clips,", sum(c["n_slices"] for c in gate["clips"]), "slices")


if __name__ == "__main__":
    main()
