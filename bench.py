#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric: compress & decompress input MB/s, bit-exact, on synthetic CABAC streams.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg2]

One "step" = one pass of the hot path over one batch: every CABAC slice of the workload stream is
compressed (CABAC decode -> slice_data walk -> avrecode model -> range coder) and then decompressed
(the exact inverse) on the GPU.  `value` = stream bytes / (compress + decompress device time), inputs
resident in HBM, timed with CUDA events on the library's own stream; `e2e` = the same through the
C-ABI file calls (avr_file_compress / avr_file_decompress) with HOST buffers: slice index, H2D, kernel,
D2H and container assembly all inside the timed region.

What `value` is: the single cfg2 file (BASELINE.json configs[1]) on ONE GPU.  N > 1 (torchrun): every rank owns
one GPU and its own stream of the same shape (no collective in the data path; NCCL only for the barrier and the
max-over-ranks of the time) -> "scaling": "weak".  Beside it the same JSON line carries
  per_config   cfg1..cfg5 at N = 1: value, e2e, bit_exact, bytes saved, cycles/bin on the longest slice, issue fraction
  batched      B copies of the cfg2 file in one device batch (what the CPU arm does with one copy per core)
  corpus       BASELINE.json configs[4]: 256 independent 720p/1080p clips sharded BY FILE over the N ranks
               (avrecode_b200.shard.assign, longest-first) -> strong scaling; time = max over ranks, bytes = sum.

`--impl reference` times the UNMODIFIED reference CPU binary (oracle/_ref, built from /root/reference in the
development container), one process per host core.  Each step is a bounded sample of the SAME cfg2 file: the file
is cut at its intra pictures into closed groups of 30 pictures (byte ranges of the file, SPS/PPS in front), every
core recodes one group per step and the groups rotate, so that the steps cover the whole file evenly.  The same
measurement feeds `cpu_baseline` of the default arm (rank 0, N = 1 only)."""
import argparse, hashlib
import json
import os
import subprocess
import sys
import tempfile
import threading
import time
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "roundtrip_input_MBps"   # input bytes / (compress time + decompress time)
UNIT = "MB/s"

# BASELINE.json configs (synthetic stand-ins); cfg2 is the one the metric is quoted on.
WORKLOADS = {
    "cfg1": dict(width=1280, height=720, frames=30, kbps=4000, fps=30, seed=1),
    "cfg2": dict(width=1920, height=1088, frames=300, high=True, kbps=30000, fps=60, intra_period=30, seed=2),
    "cfg3": dict(width=1920, height=1088, frames=60, intra_period=1, slices=8, kbps=60000, fps=30, seed=3),
    "cfg4": dict(width=3840, height=2160, frames=60, high=True, p8x8=0.85, kbps=100000, fps=30, qp=20, seed=4),
}
DESCR = {
    "cfg1": "1280x720 30f Main 1 slice/frame ~4 Mb/s",
    "cfg2": "1080p60 High 30 Mb/s 300 frames 1 slice/frame (BASELINE.json configs[1])",
    "cfg3": "1080p intra 8 slices/frame 60 frames",
    "cfg4": "2160p High 8x8 100 Mb/s 60 frames",
    "cfg5": "corpus of 256 independent 720p/1080p clips, 30-120 frames, 5-30 Mb/s (BASELINE.json configs[4])",
}
CACHE = os.environ.get("AVR_BENCH_CACHE", "/tmp/avr_bench_cache")
N_CORPUS = int(os.environ.get("AVR_BENCH_CORPUS_CLIPS", "256"))


def _gen(kw, path):
    gen = os.path.join(ROOT, "build", "h264gen")
    if not os.path.exists(gen):
        from avrecode_b200 import build as b
        b.build(with_ref=False)
    tmp = path + ".tmp%d_%d" % (os.getpid(), threading.get_ident())
    cmd = [gen, "-o", tmp]
    for k, v in kw.items():
        opt = "--" + k.replace("_", "-")
        if v is True:
            cmd.append(opt)
        elif v is not False and v is not None:
            cmd += [opt, str(v)]
    subprocess.run(cmd, check=True, capture_output=True)
    os.replace(tmp, path)


def _cache_path(name, kw):
    os.makedirs(CACHE, exist_ok=True)
    tag = "_".join("%s%s" % (k, v) for k, v in sorted(kw.items()))
    return os.path.join(CACHE, "%s_%s.h264" % (name, tag))


def gen_workload(name, seed_offset=0, frames=None):
    """Synthesises the stream with build/h264gen (test infrastructure) and caches it on local disk."""
    kw = dict(WORKLOADS[name])
    kw["seed"] = kw["seed"] + 1000 * seed_offset
    if frames is not None:
        kw["frames"] = frames
    path = _cache_path(name, kw)
    if not os.path.exists(path):
        _gen(kw, path)
    with open(path, "rb") as f:
        return f.read(), path


def corpus_spec(n=256, frame_div=1):
    """BASELINE.json configs[4]: clips with seeds 1000.., alternating 720p / 1080p, 30-120 frames,
    5-30 Mb/s, Main and High mixed, I+P.  frame_div > 1 shortens every clip (tests)."""
    out = []
    for i in range(n):
        hd = i & 1
        out.append(dict(width=1920 if hd else 1280, height=1088 if hd else 720, frames=max(2, (30 + (i * 37) % 91) // frame_div),
                        seed=1000 + i, high=bool(i % 3), kbps=5000 + (i * 3571) % 25001, fps=30, intra_period=30))
    return out


def gen_corpus(specs, workers=None):
    """Generates (or finds in the cache) one stream per spec, `workers` generator processes at a time; -> paths."""
    paths = [_cache_path("cfg5", kw) for kw in specs]
    todo = [(kw, p) for kw, p in zip(specs, paths) if not os.path.exists(p)]
    if todo:
        with ThreadPoolExecutor(max(1, workers or os.cpu_count() or 1)) as ex:
            list(ex.map(lambda a: _gen(*a), todo))
    return paths


def gop_segments(data, entries):
    """Cuts a stream at its intra pictures: [bytes]; every piece but the first gets the stream's SPS/PPS in front.
    The pieces are byte ranges of the file and tile it."""
    def nal_start(off):
        k = data.rfind(b"\x00\x00\x01", 0, off)
        return k - 1 if k > 0 and data[k - 1] == 0 else k
    head = data[:nal_start(entries[0].desc.in_off)]
    cuts = [nal_start(e.desc.in_off) for e in entries if e.desc.slice_type == 2 and e.desc.first_mb == 0]
    if not cuts or cuts[0] != len(head):
        cuts = [len(head)] + cuts
    cuts.append(len(data))
    return [(head if i else b"") + data[(0 if i == 0 else cuts[i]):cuts[i + 1]] for i in range(len(cuts) - 1)]


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.rows = []
        self.stop_flag = False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                for line in out.strip().splitlines():
                    self.rows.append([c.strip() for c in line.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        self.stop_flag = True
        self.join(timeout=6)
        sm = sorted(int(float(r[1])) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit())
        mx = [int(float(r[2])) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) < 8:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def load_gate():
    try:
        with open(os.path.join(ROOT, "tests", "golden", "bench_gate.json")) as f:
            return json.load(f)["workloads"]
    except Exception:
        return {}


def config_of(workload, data_len, n_slices):
    """The SAME dict in both arms (the driver compares them)."""
    return {"workload": "%s: %s" % (workload, DESCR[workload]), "stream_bytes": data_len, "cabac_slices": n_slices}


# ---- the reference's own CPU implementation, all host cores -------------------------------------------
def _ref_binaries(has_8x8):
    """`recode compress` is always the stock binary.  The stock DEcompressor cannot decode 8x8-transform macroblocks
    (SURVEY hazard H8), so for High-profile input the decompress leg uses the copy with the one-line fix (its compressor
    output is byte-identical to the stock binary's)."""
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    comp = os.path.join(ref_dir, "recode_ref")
    decomp = os.path.join(ref_dir, "recode_ref_h8" if has_8x8 else "recode_ref")
    return (comp, decomp) if os.path.exists(comp) and os.path.exists(decomp) else None


def reference_pass(paths, has_8x8):
    """One process per entry of `paths`, all at once: compress, then decompress, then compare one output.
    -> (wall seconds compress, wall seconds decompress, ok, [container bytes])."""
    bins = _ref_binaries(has_8x8)
    if bins is None:
        return None
    comp, decomp = bins
    with tempfile.TemporaryDirectory() as td:
        def run_all(argv_of):
            t0 = time.perf_counter()
            procs = [subprocess.Popen(argv_of(i), stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL) for i in range(len(paths))]
            rcs = [p.wait() for p in procs]
            return time.perf_counter() - t0, rcs
        tc, rc1 = run_all(lambda i: [comp, "compress", paths[i], os.path.join(td, "%d.rec" % i)])
        td_, rc2 = run_all(lambda i: [decomp, "decompress", os.path.join(td, "%d.rec" % i), os.path.join(td, "%d.out" % i)])
        ok = all(r == 0 for r in rc1 + rc2)
        if ok:
            with open(os.path.join(td, "0.out"), "rb") as f, open(paths[0], "rb") as g:
                ok = f.read() == g.read()
        recs = [os.path.getsize(os.path.join(td, "%d.rec" % i)) if os.path.exists(os.path.join(td, "%d.rec" % i)) else 0
                for i in range(len(paths))]
    return tc, td_, ok, recs


def reference_segments(workload):
    """The workload file cut into its closed groups of pictures, written to the cache; -> (paths, sizes, n_slices, bytes)."""
    import avrecode_b200 as A
    data, path = gen_workload(workload, 0)
    entries = A.index_h264(data)
    segs = gop_segments(data, entries)
    paths = []
    for i, s in enumerate(segs):
        p = path + ".seg%d" % i
        if not os.path.exists(p) or os.path.getsize(p) != len(s):
            with open(p + ".tmp%d" % os.getpid(), "wb") as f:
                f.write(s)
            os.replace(p + ".tmp%d" % os.getpid(), p)
        paths.append(p)
    return paths, [len(s) for s in segs], len([e for e in entries if e.recodable]), len(data)


def cpu_reference_measure(workload, steps=1, warmup=0):
    cores = os.cpu_count() or 1
    if _ref_binaries(True) is None:
        return None
    seg_paths, seg_sizes, n_slices, data_len = reference_segments(workload)
    has_8x8 = bool(WORKLOADS[workload].get("high"))
    tc = td = 0.0
    ok = True
    total = rec_total = 0
    for step in range(-warmup, steps):
        pick = [((step + warmup) * cores + c) % len(seg_paths) for c in range(cores)]
        r = reference_pass([seg_paths[i] for i in pick], has_8x8)
        if step < 0:
            continue
        tc += r[0]; td += r[1]; ok = ok and r[2]
        total += sum(seg_sizes[i] for i in pick); rec_total += sum(r[3])
    pin = load_gate().get(workload) or {}
    full = (1.0 - pin["ref_rec_bytes"] / pin["stream_bytes"]) if pin.get("ref_rec_bytes") else None
    return {
        "value": total / (tc + td) / 1e6, "unit": UNIT, "cores": cores, "kind": "reference",
        "sample": "per step: %d processes (one per host core), each recodes one closed group of pictures of the %s file "
                  "(%d groups of ~%.2f MB, byte ranges of the file with SPS/PPS in front, rotating so that the steps cover the "
                  "file evenly); recode_ref -O2 -DNDEBUG compress + %s decompress"
                  % (cores, workload, len(seg_paths), sum(seg_sizes) / len(seg_sizes) / 1e6, "H8-patched" if has_8x8 else "stock"),
        "compress_MBps": total / tc / 1e6, "decompress_MBps": total / td / 1e6, "bit_exact": ok,
        "bytes_saved_frac_sample": 1.0 - rec_total / total if rec_total else None,
        "bytes_saved_frac": full, "bytes_saved_note": "whole file through the unmodified reference compressor (pinned offline in "
                                                     "tests/golden/bench_gate.json: ref_rec_bytes)",
        "ms_per_step": (tc + td) * 1e3 / max(1, steps), "config": config_of(workload, data_len, n_slices),
    }


def cpu_reference_corpus(cores):
    """The reference on a bounded sample of the corpus: its first `cores` clips, one per core, once."""
    specs = corpus_spec(N_CORPUS)[:cores]
    paths = gen_corpus(specs, cores)
    r = reference_pass(paths, True)
    if r is None:
        return None
    total = sum(os.path.getsize(p) for p in paths)
    return {"value": total / (r[0] + r[1]) / 1e6, "unit": UNIT, "cores": cores, "bit_exact": r[2], "clips": len(paths), "bytes": total,
            "bytes_saved_frac": 1.0 - sum(r[3]) / total, "seconds": r[0] + r[1],
            "sample": "first %d clips of the corpus, one process per host core, once (compress + decompress)" % len(paths)}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    m = cpu_reference_measure(args.workload, steps=max(1, args.steps), warmup=max(0, args.warmup))
    if m is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref (reference binaries) not present on this machine"}))
        return 0
    line = {
        "impl": "reference", "metric": METRIC, "value": m["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": m["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8/u64 integer", "data": "synthetic", "config": m["config"],
        "cpu_baseline": {k: m[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": m["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "compress_MBps": m["compress_MBps"], "decompress_MBps": m["decompress_MBps"], "bit_exact": m["bit_exact"],
        "bytes_saved_frac": m["bytes_saved_frac"], "bytes_saved_frac_sample": m["bytes_saved_frac_sample"],
        "bytes_saved_note": m["bytes_saved_note"], "gpu_launches": 0,
    }
    if os.environ.get("AVR_BENCH_CORPUS", "1") != "0":
        try:
            line["corpus"] = cpu_reference_corpus(m["cores"])
        except Exception as e:
            line["corpus"] = {"error": str(e)}
    print(json.dumps(line))
    return 0


# ---- our arm ---------------------------------------------------------------------------------------
def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


_OUT_FD = None


def load_traffic():
    """dram bytes / warp instructions per launch of the dominant kernel from the committed ncu --set full capture, if any."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            return json.load(f)
    except Exception:
        return {}


class Gpu:
    """One rank's device side: two contexts (one per direction) and the single-file measurement."""

    def __init__(self, local):
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import avrecode_b200 as A
        import helpers as H   # desc helpers only (the oracle binding in there is never called from this arm)
        self.A, self.H = A, H
        self.cc = A.Context(device=local)
        self.cd = A.Context(device=local)
        # chunked containers of the file calls start fresh models from the shipped prior table (what the CLI does by
        # default); the slice-level batches that are timed and pinned against the oracle do not carry AVR_SLICE_PRIORS
        self.priors = os.path.exists(A.PRIORS_PATH) and os.environ.get("AVR_BENCH_PRIORS", "1") != "0"
        if self.priors:
            self.cc.set_priors("default"); self.cd.set_priors("default")
        self.gate = load_gate()

    def plain_container_bytes(self, data):
        """Size of the chunked container WITHOUT the prior table (the ratio the format has on its own)."""
        if not self.priors:
            return None
        self.cc.set_priors(None)
        try:
            return len(self.cc.file_compress(data, self.A.MODE_SLICE))
        finally:
            self.cc.set_priors("default")

    def descs_of(self, data, wide=True, n_batch=0):
        """n_batch: CABAC slices of the whole batch these descriptors go into (0: just this file's)."""
        A, H = self.A, self.H
        descs = H.batch_descs(A.index_h264(data), mode_file=False)
        if wide:
            # what the file API does in chunked mode: long slices become wide slices -- while the batch is small enough to
            # be a latency job (avr_config.wide_batch_limit); a throughput batch gets none
            for d in descs:
                d.n_streams = self.cc.slice_streams(d, n_batch or len(descs))
        return descs

    def prepare(self, name, data, copies=1):
        """Stages both directions (inputs resident in HBM), checks the round trip and -- for a pinned stream -- that ALL
        ordinary slices' coded bytes equal the oracle's (digest in tests/golden/bench_gate.json).  Nothing under oracle/
        runs here."""
        H, cc, cd = self.H, self.cc, self.cd
        descs = self.descs_of(data, n_batch=copies * len(self.A.index_h264(data)) if copies > 1 else 0)
        arena = data
        if copies > 1:   # the same file `copies` times in one batch
            pad = (-len(data)) % 16
            arena = (data + bytes(pad)) * copies
            one = descs
            descs = []
            for c in range(copies):
                for d in one:
                    d2 = self.A._abi.copy_desc(d)
                    d2.in_off += c * (len(data) + pad)
                    descs.append(d2)
        cc.stage(descs, arena)
        cc.launch(0); cc.sync()
        res, blob = cc.collect()
        if not all(r.status == 0 for r in res):
            raise SystemExit("bench.py: compress failed on %d slices of %s" % (sum(1 for r in res if r.status), name))
        try:   # per-class bin counts of the launch (the reference's per-CodingType bills, recode.cpp:635-655, as counts)
            hist = [sum(h[k] for h in cc.collect_hist()) for k in range(8)]
        except Exception:
            hist = None
        dd = H.coded_descs(descs, res)
        cd.stage(dd, blob)
        cd.launch(1); cd.sync()
        res_d, blob_d = cd.collect()
        ok = H.roundtrip_ok(descs, res_d, blob_d, arena) and [r.n_bins for r in res_d] == [r.n_bins for r in res]
        gate = "round trip only"
        pin = self.gate.get(name)
        if copies == 1 and pin and hashlib.sha256(data).hexdigest() == pin["stream_sha256"]:
            plain = self.descs_of(data, wide=False)   # ordinary-slice bytes are the ones the oracle restates
            r_g, b_g = cc.compress(plain, data)
            h = hashlib.sha256()
            for sl in H.slices_of(r_g, b_g):
                h.update(len(sl).to_bytes(8, "little")); h.update(sl)
            ok = ok and len(plain) == pin["n_slices"] and h.hexdigest() == pin["coded_sha256"] \
                and [int(r.n_bins) for r in r_g] == pin["n_bins"] and [r.n_bins for r in r_g] == [r.n_bins for r in res]
            gate = "round trip + all %d slices equal the oracle's pinned bytes" % len(plain)
            cc.stage(descs, arena)
        if not ok:
            raise SystemExit("bench.py: parity check failed on %s -- refusing to time a wrong kernel" % name)
        return dict(descs=descs, res=res, gate=gate, hist=hist, payload=sum(d.payload_size for d in descs), coded=sum(r.out_size for r in res),
                    bins=sum(r.n_bins for r in res), longest=max(r.n_bins for r in res), n_wide=sum(1 for d in descs if (d.n_streams & 15) >= 2),
                    k_wide=max([d.n_streams & 15 for d in descs] + [0]), nbytes=len(arena))

    def step(self):
        cc, cd = self.cc, self.cd
        cc.flush_l2(); cc.launch(0); cc.sync()
        mc = cc.last_kernel_ms()
        cd.flush_l2(); cd.launch(1); cd.sync()
        return mc, cd.last_kernel_ms()

    def timed(self, steps, warmup, barrier=None):
        for _ in range(warmup):
            self.step()
        if barrier:
            barrier()
        n0 = self.cc.launch_count() + self.cd.launch_count()
        t0 = time.perf_counter()
        ms_c = ms_d = 0.0
        for _ in range(steps):
            a, b = self.step()
            ms_c += a; ms_d += b
        if barrier:
            barrier()
        wall_ms = (time.perf_counter() - t0) * 1e3 / steps
        launches = (self.cc.launch_count() + self.cd.launch_count() - n0) + 2 * steps   # job kernels + L2 flush kernels
        return ms_c / steps, ms_d / steps, wall_ms, launches

    def e2e_file(self, data, steps, barrier=None):
        A, cc = self.A, self.cc
        rec = cc.file_compress(data, A.MODE_SLICE)
        if cc.file_decompress(rec) != data:
            raise SystemExit("bench.py: file round trip is not the identity")
        if barrier:
            barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            rec = cc.file_compress(data, A.MODE_SLICE)
            back = cc.file_decompress(rec)
        if barrier:
            barrier()
        s = (time.perf_counter() - t0) / steps
        assert back == data
        return s, len(rec)

    def file_mode(self, data):
        """Throughput of the reference's own container format, end to end through the file calls with host buffers: a
        file is ONE job (its slices share one model), so one file runs on one thread block and a set of files on as
        many.  (a) one file: cfg1; (b) the headline file as its closed-GOP segments -- the very files the reference arm
        times on the host cores -- in one avr_files_compress(AVR_MODE_FILE) call."""
        A, cc = self.A, self.cc
        out = {"unit": UNIT, "api": "avr_file[s]_compress(AVR_MODE_FILE) + avr_file[s]_decompress, host buffers, wall clock",
               "note": "byte-identical to the reference compressor's containers (tests/golden, bench_gate ref_rec_bytes); "
                       "parallelism = number of files"}
        one, _ = gen_workload("cfg1", 0)
        rec = cc.file_compress(one, A.MODE_FILE)                      # warm-up + check
        if cc.file_decompress(rec) != one:
            raise SystemExit("bench.py: file-mode round trip is not the identity")
        pin = self.gate.get("cfg1") or {}
        t0 = time.perf_counter()
        rec = cc.file_compress(one, A.MODE_FILE)
        back = cc.file_decompress(rec)
        s1 = time.perf_counter() - t0
        assert back == one
        out["one_file"] = {"workload": DESCR["cfg1"], "value": len(one) / s1 / 1e6, "bytes_saved_frac": 1.0 - len(rec) / len(one),
                           "container_bytes_equal_reference": (len(rec) == pin["ref_rec_bytes"]) if pin.get("ref_rec_bytes") else None}
        segs = gop_segments(data, A.index_h264(data))
        recs = cc.files_compress(segs, A.MODE_FILE)
        if cc.files_decompress(recs) != segs:
            raise SystemExit("bench.py: file-mode segment round trip is not the identity")
        t0 = time.perf_counter()
        recs = cc.files_compress(segs, A.MODE_FILE)
        back = cc.files_decompress(recs)
        s2 = time.perf_counter() - t0
        assert back == segs
        total = sum(len(x) for x in segs)
        out["segments"] = {"workload": "the headline file as %d closed-GOP segment files (what the reference arm times)" % len(segs),
                           "files": len(segs), "value": total / s2 / 1e6, "bytes_saved_frac": 1.0 - sum(len(r) for r in recs) / total}
        return out

    def per_config(self, name, steps=2, warmup=1, f_hz=0.0, inst_per_bin=None):
        data, _ = gen_workload(name, 0)
        p = self.prepare(name, data)
        ms_c, ms_d, _, _ = self.timed(steps, warmup)
        e2e_s, rec_len = self.e2e_file(data, 1)
        rec_plain = self.plain_container_bytes(data)
        dom_ms = max(ms_c, ms_d)
        pin = self.gate.get(name) or {}
        ref_saved = (1.0 - pin["ref_rec_bytes"] / pin["stream_bytes"]) if pin.get("ref_rec_bytes") else None
        saved = 1.0 - rec_len / len(data)
        bins_s = p["bins"] / dom_ms * 1e3
        issue_peak = (148 * 4 * f_hz / inst_per_bin) if (inst_per_bin and f_hz) else None
        return {"workload": DESCR[name], "stream_bytes": len(data), "value": len(data) / (ms_c + ms_d) / 1e3, "e2e": len(data) / e2e_s / 1e6, "unit": UNIT,
                "compress_MBps": len(data) / ms_c / 1e3, "decompress_MBps": len(data) / ms_d / 1e3,
                "bit_exact": True, "parity_gate": p["gate"], "bytes_saved_frac": saved, "bytes_saved_frac_reference": ref_saved,
                "bytes_saved_frac_no_priors": (1.0 - rec_plain / len(data)) if rec_plain else None,
                "delta_saved_vs_ref": (saved - ref_saved) if ref_saved is not None else None,
                "saved_frac_of_reference": (saved / ref_saved) if ref_saved else None,
                "slices": len(p["descs"]), "wide_slices": p["n_wide"], "longest_slice_bins": int(p["longest"]),
                "cycles_per_bin_longest": (dom_ms * 1e-3 * f_hz / p["longest"]) if f_hz else None,
                "issue_frac": (bins_s / issue_peak) if issue_peak else None, "steps": steps}

    def corpus(self, clips, steps, barrier):
        """My shard of the corpus: every slice of every clip in one device batch.  -> (kernel ms compress, decompress, e2e
        seconds compress + decompress, container bytes, slices, bins)."""
        A, H, cc, cd = self.A, self.H, self.cc, self.cd
        recs = cc.files_compress(clips, A.MODE_SLICE)
        if cc.files_decompress(recs) != clips:
            raise SystemExit("bench.py: corpus round trip is not the identity")
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            recs = cc.files_compress(clips, A.MODE_SLICE)
            back = cc.files_decompress(recs)
        barrier()
        e2e_s = (time.perf_counter() - t0) / steps
        assert back == clips
        del back
        # kernel only, inputs resident in HBM
        pieces, descs, offs = [], [], 0
        n_total = sum(len(self.A.index_h264(c)) for c in clips)
        for c in clips:
            for d in self.descs_of(c, n_batch=n_total):
                d.in_off += offs
                descs.append(d)
            pad = (-len(c)) % 16
            pieces.append(c); pieces.append(bytes(pad))
            offs += len(c) + pad
        big = b"".join(pieces)
        cc.stage(descs, big)
        cc.launch(0); cc.sync()
        res, blob = cc.collect()
        assert all(r.status == 0 for r in res)
        cd.stage(H.coded_descs(descs, res), blob)
        cd.launch(1); cd.sync()
        ms_c, ms_d, _, launches = self.timed(steps, 1, barrier)
        return ms_c, ms_d, e2e_s, sum(len(r) for r in recs), len(descs), sum(r.n_bins for r in res), launches


def run_ours(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    from avrecode_b200 import shard

    def max_over_ranks(x):
        return shard.max_over_ranks(x, device="cuda")

    def sum_over_ranks(x):
        return shard.sum_over_ranks(x, device="cuda")

    g = Gpu(local)
    A = g.A
    steps, warmup = max(1, args.steps), max(3, args.warmup)
    # ---- headline: one stream of the workload shape per GPU (rank r: seed offset r) ------------------------------
    stream_ids = shard.my_items([1] * world, rank, world)
    assert len(stream_ids) == 1
    data, _ = gen_workload(args.workload, stream_ids[0])
    p = g.prepare(args.workload, data)
    sampler = ClockSampler(local)
    sampler.start()
    ms_c, ms_d, wall_ms, launches = g.timed(steps, warmup, barrier)
    dev_ms = max_over_ranks(ms_c + ms_d)
    dev_ms_c, dev_ms_d = max_over_ranks(ms_c), max_over_ranks(ms_d)
    wall_ms = max_over_ranks(wall_ms)
    total_bytes = sum_over_ranks(float(len(data)))
    total_bins = sum_over_ranks(float(p["bins"]))
    e2e_steps = max(1, min(steps, 3))
    e2e_s, rec_len = g.e2e_file(data, e2e_steps, barrier)
    e2e_s = max_over_ranks(e2e_s)
    clocks = sampler.summary()
    f_hz = float((clocks or {}).get("sm_mhz") or 0) * 1e6
    tr = load_traffic()
    extra_launches = 0

    # ---- the same file B times in one batch (the CPU arm runs one copy per core) -- N = 1 only -------------------
    batched = None
    if world == 1 and os.environ.get("AVR_BENCH_BATCHED", "1") != "0":
        try:
            B = int(os.environ.get("AVR_BENCH_BATCH", "8"))
            pb = g.prepare(args.workload, data, copies=B)
            b_c, b_d, _, bl = g.timed(2, 1)
            extra_launches += bl
            batched = {"copies": B, "value": B * len(data) / (b_c + b_d) / 1e3, "unit": UNIT, "compress_MBps": B * len(data) / b_c / 1e3,
                       "decompress_MBps": B * len(data) / b_d / 1e3, "slices": len(pb["descs"]), "steps": 2,
                       "note": "throughput figure: %d copies of the file in one device batch, device-timed like `value`; "
                               "`value` itself stays the single file" % B}
        except Exception as e:
            batched = {"error": str(e)}

    # ---- per-primitive floors measured on this device (avr_probe) ------------------------------------------------
    floors = None
    if rank == 0:
        try:
            floors = {name: round(cpb, 1) for name, cpb, _ in A.probes(local, 20000)}
        except Exception as e:
            floors = {"error": str(e)}

    # ---- corpus (BASELINE.json configs[4]), sharded by file over the ranks: strong scaling -----------------------
    def corpus_section():
        specs = corpus_spec(N_CORPUS)
        weights = [s["frames"] * s["kbps"] for s in specs]   # bytes ~ frames x bitrate: known without generating anything
        mine = shard.my_items(weights, rank, world)
        t0 = time.perf_counter()
        paths = gen_corpus([specs[i] for i in mine], max(1, (os.cpu_count() or 1) // world))
        clips = []
        for pth in paths:
            with open(pth, "rb") as f:
                clips.append(f.read())
        gen_s = time.perf_counter() - t0
        c_steps = max(1, min(steps, int(os.environ.get("AVR_BENCH_CORPUS_STEPS", "1"))))
        k_c, k_d, ce2e_s, rec_bytes, n_sl, c_bins, cl = g.corpus(clips, c_steps, barrier)
        my_bytes = float(sum(len(c) for c in clips))
        c_total = sum_over_ranks(my_bytes)
        c_kernel_ms = max_over_ranks(k_c + k_d)
        c_e2e_s = max_over_ranks(ce2e_s)
        return cl, {"workload": DESCR["cfg5"], "clips": len(specs), "bytes": c_total, "slices": sum_over_ranks(float(n_sl)),
                    "scaling": "strong", "sharding": "by file, longest first (avrecode_b200.shard.assign), no collective",
                    "value": c_total / c_kernel_ms / 1e3, "e2e": c_total / c_e2e_s / 1e6, "unit": UNIT,
                    "compress_MBps": c_total / max_over_ranks(k_c) / 1e3, "decompress_MBps": c_total / max_over_ranks(k_d) / 1e3,
                    "bins_per_s": sum_over_ranks(float(c_bins)) * 2 / c_kernel_ms * 1e3,
                    "bytes_saved_frac": 1.0 - sum_over_ranks(float(rec_bytes)) / c_total, "bit_exact": True, "steps": c_steps,
                    "ms_per_step": c_kernel_ms, "e2e_api": "avr_files_compress + avr_files_decompress (host buffers)",
                    "max_rank_share": max_over_ranks(my_bytes) / c_total, "generate_s": max_over_ranks(gen_s)}

    corpus = None
    if os.environ.get("AVR_BENCH_CORPUS", "1") != "0":
        if world > 1:
            cl, corpus = corpus_section()   # every rank must reach the collectives: no catching here
            extra_launches += cl
        else:
            try:
                cl, corpus = corpus_section()
                extra_launches += cl
            except BaseException as e:   # a secondary figure must never sink the headline
                corpus = {"error": "%s: %s" % (type(e).__name__, e)}

    # ---- the other BASELINE shapes, N = 1 only ------------------------------------------------------------------
    # At N > 1 every rank runs the same file on its own GPU (weak scaling, like the headline): rates = N x bytes / the
    # slowest rank's time.  Every rank takes part in every collective, whatever went wrong on it.
    per_config = None
    if os.environ.get("AVR_BENCH_PER_CONFIG", "1") != "0":
        per_config = {}
        for name in sorted(WORKLOADS):
            if name == args.workload:
                continue
            try:
                pc = g.per_config(name, steps=2 if world == 1 else 1, f_hz=f_hz)
                extra_launches += 2 * 3 * 2
            except BaseException as e:   # a secondary figure must never sink the headline
                pc = {"error": str(e)}
            if world > 1:
                failed = max_over_ranks(1.0 if "error" in pc else 0.0)
                nbytes = float(pc.get("stream_bytes", 0))
                t_c = max_over_ranks(nbytes / pc["compress_MBps"] / 1e3 if not failed else 0.0)      # ms
                t_d = max_over_ranks(nbytes / pc["decompress_MBps"] / 1e3 if not failed else 0.0)
                t_e = max_over_ranks(nbytes / pc["e2e"] / 1e6 if not failed else 0.0)                 # s
                if failed:
                    pc = {"error": pc.get("error", "failed on another rank")}
                else:
                    pc.update({"value": world * nbytes / (t_c + t_d) / 1e3, "compress_MBps": world * nbytes / t_c / 1e3,
                               "decompress_MBps": world * nbytes / t_d / 1e3, "e2e": world * nbytes / t_e / 1e6,
                               "n_gpus": world, "scaling": "weak (the same file on every GPU)"})
            per_config[name] = pc

    # ---- the reference's own format (AVR_MODE_FILE: one model per file = one serial job per file), N = 1 only ----
    file_mode = None
    if world == 1 and os.environ.get("AVR_BENCH_FILE_MODE", "1") != "0":
        try:
            file_mode = g.file_mode(data)
            extra_launches += 8
        except BaseException as e:
            file_mode = {"error": str(e)}

    if rank != 0:
        return 0
    peak, peak_src = load_peaks()
    dom = "decompress" if ms_d >= ms_c else "compress"
    dom_ms = max(ms_c, ms_d)
    algo_bytes = p["payload"] + p["coded"]     # N + C per direction
    achieved = algo_bytes / dom_ms / 1e6
    trw = tr if args.workload == "cfg2" else {}   # the committed capture is of the cfg2 stream
    traffic = trw.get(dom)
    pin = g.gate.get(args.workload) or {}
    ref_saved = (1.0 - pin["ref_rec_bytes"] / pin["stream_bytes"]) if pin.get("ref_rec_bytes") else None
    saved = 1.0 - rec_len / len(data)
    rec_plain = g.plain_container_bytes(data) if rank == 0 else None
    saved_plain = (1.0 - rec_plain / len(data)) if rec_plain else None
    # the bounds that mean something for a serial integer recurrence, from live numbers
    try:
        inst = (trw.get("warp_inst_per_launch") or {}).get(dom)   # smsp__inst_executed.sum of the committed --set full capture
        bins = p["bins"]
        bins_s_dom = bins / dom_ms * 1e3
        issue_peak = (148 * 4 * f_hz / (inst / bins)) if (inst and f_hz) else None
        cpb = (dom_ms * 1e-3 * f_hz / p["longest"]) if f_hz else None
        # the serial stage of the dominant direction: the CABAC decoder + walk (compress) or the CABAC re-encoder (decompress)
        floor = None
        if isinstance(floors, dict) and "error" not in floors:
            # the serial stage that bounds the longest slice: the parser's inner loop; for a wide slice being decompressed
            # (K parsers share the slice) the one warp that re-encodes the CABAC stream
            want = "compress significance loop" if dom == "compress" else ("decompress coder warp" if p["n_wide"] else "decompress significance loop")
            key = [k for k in floors if want in k]
            floor = floors[key[0]] if key else None
        bounds = {
            "latency": {"longest_slice_bins": int(p["longest"]), "mean_slice_bins": bins / max(1, len(p["descs"])),
                        "cycles_per_bin_on_longest_slice": cpb, "floor_cycles_per_bin": floor,
                        "frac_of_floor": (floor / cpb) if (floor and cpb) else None,
                        "floor_note": "avr_probe on this device: one lane looping the serial primitive of the dominant direction "
                                      "(compress: the parser's CABAC decode of a significance map; decompress: the parser's range decode of one, or -- when the "
                                      "longest slices are wide, K parsers each -- the CABAC re-encoder warp over a ring of entries)",
                        "note": "a launch lasts as long as its longest slice; cycles/bin = kernel time x SM clock / its bins"},
            "issue": {"achieved_bins_per_s": bins_s_dom, "peak_bins_per_s": issue_peak,
                      "frac": (bins_s_dom / issue_peak) if issue_peak else None,
                      "warp_inst_per_bin": (inst / bins) if inst else None,
                      "note": "148 SM x 4 schedulers x f_SM / warp instructions per bin (ncu, spin-wait included)"},
            "probes_cycles_per_bin": floors,
        }
    except Exception as e:   # explanatory only: must never sink the measured line
        bounds = {"error": str(e)}
    if per_config is not None:
        per_config[args.workload] = {
            "workload": DESCR[args.workload], "value": total_bytes / dev_ms / 1e3, "e2e": total_bytes / e2e_s / 1e6, "unit": UNIT,
            "compress_MBps": total_bytes / dev_ms_c / 1e3, "decompress_MBps": total_bytes / dev_ms_d / 1e3, "bit_exact": True, "parity_gate": p["gate"],
            "bytes_saved_frac": saved, "bytes_saved_frac_reference": ref_saved, "bytes_saved_frac_no_priors": saved_plain,
            "delta_saved_vs_ref": (saved - ref_saved) if ref_saved is not None else None,
            "saved_frac_of_reference": (saved / ref_saved) if ref_saved else None, "slices": len(p["descs"]),
            "wide_slices": p["n_wide"], "longest_slice_bins": int(p["longest"]),
            "cycles_per_bin_longest": bounds.get("latency", {}).get("cycles_per_bin_on_longest_slice"),
            "issue_frac": bounds.get("issue", {}).get("frac"), "steps": steps}
        if corpus is not None and "error" not in corpus:
            per_config["cfg5"] = {k: corpus[k] for k in ("workload", "value", "e2e", "unit", "compress_MBps", "decompress_MBps", "bit_exact",
                                                         "bytes_saved_frac", "slices", "steps")}
    line = {
        "metric": METRIC, "value": total_bytes / dev_ms / 1e3, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": warmup, "ms_per_step": dev_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8/u64 integer", "data": "synthetic",
        "config": config_of(args.workload, len(data), len(p["descs"])),
        "value_is": "one %s file per GPU, device-timed compress + decompress (N > 1: one such file per rank, weak scaling); "
                    "`corpus` is the strong-scaled cfg5 figure" % args.workload,
        "details": {"chunking": "slice (model reset per slice); %d slices of >= 64 KiB as wide slices (up to %d streams each)" % (p["n_wide"], p["k_wide"]),
                    "sharding": "one stream per GPU, no collective", "l2": "flushed (256 MB write) before every kernel",
                    "e2e_steps": e2e_steps},
        "compress_MBps": total_bytes / dev_ms_c / 1e3, "decompress_MBps": total_bytes / dev_ms_d / 1e3,
        "bins_per_s": total_bins * 2 / dev_ms * 1e3, "bins_per_byte": p["bins"] / p["payload"],
        "bytes_saved_frac": saved, "bytes_saved_frac_reference": ref_saved, "bytes_saved_frac_no_priors": saved_plain,
        "bytes_saved_note": "chunked container through the file API, fresh models started from avrecode_b200/priors_v1.bin (trained by "
                            "tools/train_priors.py on h264gen streams of OTHER seeds: synthetic like this workload, so on real video expect "
                            "less); _no_priors: the same without the table; _reference: one model per file, the reference's format" if g.priors else None,
        "delta_saved_vs_ref": (saved - ref_saved) if ref_saved is not None else None,
        "saved_frac_of_reference": (saved / ref_saved) if ref_saved else None,
        "saved_frac_of_reference_no_priors": (saved_plain / ref_saved) if (ref_saved and saved_plain is not None) else None,
        "bin_classes": (dict(zip(("mb", "cbf", "sig", "level", "bypass", "nnz_side_channel", "pred_mode", "terminate"),
                                 [h / max(1, p["bins"]) for h in p["hist"]])) if p.get("hist") else None),
        "bit_exact": True, "parity_gate": p["gate"], "wall_ms_per_step": wall_ms,
        "gpu_launches": int(launches), "gpu_launches_other_sections": int(extra_launches),
        "e2e": {"value": total_bytes / e2e_s / 1e6, "unit": UNIT, "h2d_bytes_per_step": len(data) + p["coded"] + 80 * len(p["descs"]),
                "d2h_bytes_per_step": p["coded"] + p["payload"] + 64 * len(p["descs"]), "steps": e2e_steps,
                "api": "avr_file_compress + avr_file_decompress (host buffers)"},
        "roofline": {"bound": "hbm", "kernel": "avr_job_kernel<%s, true> (one persistent launch per direction; wide slices are roles of the same kernel)" % dom.capitalize(),
                     "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src, "bounds": bounds,
                     "note": "serial integer recurrence per slice: latency-bound, not HBM-bound; see DESIGN.md for the per-bin latency bound"},
        "clocks": clocks,
    }
    if batched is not None:
        line["batched"] = batched
    if corpus is not None:
        line["corpus"] = corpus
    if per_config is not None:
        line["per_config"] = per_config
    if file_mode is not None:
        line["file_mode"] = file_mode
    if world == 1:
        try:
            m = cpu_reference_measure(args.workload, steps=2)
            if m is not None:
                line["cpu_baseline"] = {k: m[k] for k in ("value", "unit", "cores", "kind", "sample")}
                line["cpu_baseline"]["bytes_saved_frac"] = m["bytes_saved_frac"]
        except Exception as e:   # the baseline must never sink the GPU number
            line["cpu_baseline"] = {"error": str(e)}
    sys.stdout.flush()
    os.write(_OUT_FD if _OUT_FD is not None else 1, (json.dumps(line) + "\n").encode())
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    # stdout carries the ONE JSON line and nothing else: libraries that write to fd 1 (NCCL prints its version there)
    # are sent to stderr for the length of the run
    global _OUT_FD
    sys.stdout.flush()
    _OUT_FD = os.dup(1)
    os.dup2(2, 1)
    rc = run_ours(args)
    try:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()
    except Exception:
        pass
    return rc


if __name__ == "__main__":
    sys.exit(main())
