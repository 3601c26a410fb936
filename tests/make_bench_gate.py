"""Writes tests/golden/bench_gate.json: sha256 of every bench workload's stream and of the ORACLE's coded bytes for its
first two slices (slice mode).  bench.py's pre-timing parity gate compares the CUDA path's bytes with these digests, so
the timed arm never executes anything under oracle/.  Run here (CPU): python tests/make_bench_gate.py"""
import hashlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import avrecode_b200 as A
import bench as B
import helpers as H

N_CHK = 2
out = {"n_slices": N_CHK, "how": "tests/make_bench_gate.py: oracle (oracle/_build/libavr_oracle.so) slice-mode compress of the first slices", "workloads": {}}
for name in sorted(B.WORKLOADS):
    data, _ = B.gen_workload(name, 0)
    plain = H.batch_descs(A.index_h264(data), mode_file=False)[:N_CHK]
    r_o, b_o = H.oracle_compress(plain, data)
    assert all(r.status == 0 for r in r_o), name
    h = hashlib.sha256()
    for s in H.slices_of(r_o, b_o):
        h.update(len(s).to_bytes(8, "little")); h.update(s)
    out["workloads"][name] = {"stream_sha256": hashlib.sha256(data).hexdigest(), "stream_bytes": len(data),
                              "coded_sha256": h.hexdigest(), "n_bins": [int(r.n_bins) for r in r_o]}
    print(name, out["workloads"][name], flush=True)
with open(os.path.join(ROOT, "tests", "golden", "bench_gate.json"), "w") as f:
    json.dump(out, f, indent=1)
