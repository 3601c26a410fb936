#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric: compress & decompress input MB/s, bit-exact, on synthetic CABAC streams.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg2]

One "step" = one pass of the hot path over one batch: every CABAC slice of the workload stream is
compressed (CABAC decode -> slice_data walk -> avrecode model -> range coder) and then decompressed
(the exact inverse) on the GPU.  `value` = stream bytes / (compress + decompress device time), inputs
resident in HBM, timed with CUDA events on the library's own stream; `e2e` = the same through the
C-ABI file calls (avr_file_compress / avr_file_decompress) with HOST buffers: slice index, H2D, kernel,
D2H and container assembly all inside the timed region.  N > 1 (torchrun): every rank owns one GPU
and its own stream of the same shape (file-level sharding, no collective in the data path; NCCL only
for the barrier and the max-over-ranks of the time) -> "scaling": "weak".

`--impl reference` times the UNMODIFIED reference CPU binary (oracle/_ref, built from
/root/reference in the development container) on all host cores on a bounded sample of the same
workload.  The same measurement feeds `cpu_baseline` of the default arm (rank 0, N = 1 only)."""
import argparse, hashlib
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

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
}
CACHE = os.environ.get("AVR_BENCH_CACHE", "/tmp/avr_bench_cache")


def gen_workload(name, seed_offset=0, frames=None):
    """Synthesises the stream with build/h264gen (test infrastructure) and caches it on local disk."""
    kw = dict(WORKLOADS[name])
    kw["seed"] = kw["seed"] + 1000 * seed_offset
    if frames is not None:
        kw["frames"] = frames
    os.makedirs(CACHE, exist_ok=True)
    tag = "_".join("%s%s" % (k, v) for k, v in sorted(kw.items()))
    path = os.path.join(CACHE, "%s_%s.h264" % (name, tag))
    if not os.path.exists(path):
        gen = os.path.join(ROOT, "build", "h264gen")
        if not os.path.exists(gen):
            from avrecode_b200 import build as b
            b.build(with_ref=False)
        cmd = [gen, "-o", path + ".tmp%d" % os.getpid()]
        for k, v in kw.items():
            opt = "--" + k.replace("_", "-")
            cmd += [opt] if v is True else [opt, str(v)]
        subprocess.run(cmd, check=True, capture_output=True)
        os.replace(path + ".tmp%d" % os.getpid(), path)
    with open(path, "rb") as f:
        return f.read(), path


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


# ---- the reference's own CPU implementation, all host cores -------------------------------------------
def reference_pass(sample_path, sample_bytes, cores, has_8x8):
    """One process per core: `recode compress` then `recode decompress` of the sample (reference binary,
    -O2 -DNDEBUG).  The stock decompressor cannot decode 8x8-transform macroblocks (SURVEY hazard H8), so
    for High-profile samples the decompress leg uses the copy with the one-line fix (its compressor output
    is byte-identical to the stock binary's).  Returns (wall seconds compress, wall seconds decompress)."""
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    comp = os.path.join(ref_dir, "recode_ref")
    decomp = os.path.join(ref_dir, "recode_ref_h8" if has_8x8 else "recode_ref")
    if not (os.path.exists(comp) and os.path.exists(decomp)):
        return None
    with tempfile.TemporaryDirectory() as td:
        def run_all(argv_of):
            t0 = time.perf_counter()
            procs = [subprocess.Popen(argv_of(i), stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL) for i in range(cores)]
            rcs = [p.wait() for p in procs]
            return time.perf_counter() - t0, rcs
        tc, rc1 = run_all(lambda i: [comp, "compress", sample_path, os.path.join(td, "%d.rec" % i)])
        td_, rc2 = run_all(lambda i: [decomp, "decompress", os.path.join(td, "%d.rec" % i), os.path.join(td, "%d.out" % i)])
        ok = all(r == 0 for r in rc1 + rc2)
        if ok:
            with open(os.path.join(td, "0.out"), "rb") as f, open(sample_path, "rb") as g:
                ok = f.read() == g.read()
        rec_bytes = os.path.getsize(os.path.join(td, "0.rec")) if os.path.exists(os.path.join(td, "0.rec")) else 0
    return tc, td_, ok, rec_bytes


def cpu_reference_measure(workload, steps=1, warmup=0):
    cores = os.cpu_count() or 1
    sample_frames = {"cfg1": 6, "cfg2": 12, "cfg3": 3, "cfg4": 2}[workload]   # ~0.5-1 MB -> ~2-3 s per leg per core
    _, path = gen_workload(workload, 0, frames=sample_frames)
    nbytes = os.path.getsize(path)
    has_8x8 = bool(WORKLOADS[workload].get("high"))
    for _ in range(warmup):
        reference_pass(path, nbytes, cores, has_8x8)
    tc = td = 0.0
    ok = True
    rec_bytes = 0
    for _ in range(steps):
        r = reference_pass(path, nbytes, cores, has_8x8)
        if r is None:
            return None
        tc += r[0]; td += r[1]; ok = ok and r[2]; rec_bytes = r[3]
    total = nbytes * cores * steps
    return {
        "value": total / (tc + td) / 1e6, "unit": UNIT, "cores": cores, "kind": "reference",
        "sample": "%s first %d frames (%d bytes) x %d processes (one per host core), recode_ref -O2 -DNDEBUG compress + %s decompress"
                  % (workload, sample_frames, nbytes, cores, "H8-patched" if has_8x8 else "stock"),
        "compress_MBps": total / tc / 1e6, "decompress_MBps": total / td / 1e6, "bit_exact": ok,
        "bytes_saved_frac": 1.0 - rec_bytes / nbytes if rec_bytes else None,
        "ms_per_step": (tc + td) * 1e3 / steps,
    }


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    m = cpu_reference_measure(args.workload, steps=max(1, args.steps), warmup=min(args.warmup, 1))
    if m is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref (reference binaries) not present on this machine"}))
        return 0
    line = {
        "impl": "reference", "metric": METRIC, "value": m["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": m["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8/u64 integer", "data": "synthetic",
        "config": {"workload": "%s: %s" % (args.workload, DESCR[args.workload])},
        "cpu_baseline": {k: m[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": m["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "compress_MBps": m["compress_MBps"], "decompress_MBps": m["decompress_MBps"], "bit_exact": m["bit_exact"],
        "bytes_saved_frac": m["bytes_saved_frac"], "gpu_launches": 0,
    }
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
    """dram bytes per launch of the dominant kernel from the committed ncu --set full capture, if any."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            return json.load(f)
    except Exception:
        return {}


def run_ours(args):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import avrecode_b200 as A
    import helpers as H   # desc helpers only (the oracle binding in there is never called from this arm)

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

    # file-level sharding: rank r owns stream r (same shape, different seed) -- per-GPU work is fixed
    #   (avrecode_b200.shard.assign: the same LPT table on every rank, nothing exchanged)
    stream_ids = shard.my_items([1] * world, rank, world)
    assert len(stream_ids) == 1
    data, _ = gen_workload(args.workload, stream_ids[0])
    entries = A.index_h264(data)
    descs = H.batch_descs(entries, mode_file=False)
    for d in descs:   # what the file API does in chunked mode: long slices become wide slices (two streams)
        d.n_streams = A.slice_streams(d)
    n_wide = sum(1 for d in descs if d.n_streams >= 2)
    payload = sum(d.payload_size for d in descs)
    cc = A.Context(device=local)
    cd = A.Context(device=local)

    # stage both directions once: inputs are resident in HBM when the timed region starts
    cc.stage(descs, data)
    cc.launch(0); cc.sync()
    res, blob = cc.collect()
    assert all(r.status == 0 for r in res), "compress failed on %d slices" % sum(1 for r in res if r.status)
    coded = sum(r.out_size for r in res)
    bins = sum(r.n_bins for r in res)
    dd = H.coded_descs(descs, res)
    cd.stage(dd, blob)
    cd.launch(1); cd.sync()
    res_d, blob_d = cd.collect()
    # parity gate before any timing: round trip is the identity, and the first slices' coded bytes equal the ORACLE's,
    #   pinned as digests in tests/golden/bench_gate.json (tests/make_bench_gate.py) -- nothing under oracle/ runs here
    ok = H.roundtrip_ok(descs, res_d, blob_d, data)
    gate = "round trip only"
    try:
        with open(os.path.join(ROOT, "tests", "golden", "bench_gate.json")) as f:
            pin = json.load(f)["workloads"].get(args.workload)
    except Exception:
        pin = None
    if pin and hashlib.sha256(data).hexdigest() == pin["stream_sha256"]:   # rank 0's stream (other ranks: other seeds)
        n_chk = min(len(pin["n_bins"]), len(descs))
        plain = H.batch_descs(entries, mode_file=False)[:n_chk]   # ordinary-slice bytes are the ones the oracle restates
        r_g, b_g = cc.compress(plain, data)
        h = hashlib.sha256()
        for sl in H.slices_of(r_g, b_g):
            h.update(len(sl).to_bytes(8, "little")); h.update(sl)
        ok = ok and h.hexdigest() == pin["coded_sha256"] and [int(r.n_bins) for r in r_g] == pin["n_bins"][:n_chk] \
            and [r.n_bins for r in r_g] == [r.n_bins for r in res[:n_chk]]
        gate = "round trip + first %d slices equal the oracle's pinned bytes" % n_chk
    cc.stage(descs, data)
    if not ok:
        raise SystemExit("bench.py: parity check failed -- refusing to time a wrong kernel")

    def step():
        cc.flush_l2(); cc.launch(0); cc.sync()
        mc = cc.last_kernel_ms()
        cd.flush_l2(); cd.launch(1); cd.sync()
        return mc, cd.last_kernel_ms()

    for _ in range(max(3, args.warmup) if args.warmup >= 0 else 3):
        step()
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    n0 = cc.launch_count() + cd.launch_count()
    t0 = time.perf_counter()
    ms_c = ms_d = 0.0
    for _ in range(args.steps):
        a, b = step()
        ms_c += a; ms_d += b
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    launches = (cc.launch_count() + cd.launch_count() - n0) + 2 * args.steps   # job kernels + L2 flush kernels
    ms_c /= args.steps; ms_d /= args.steps
    dev_ms = max_over_ranks(ms_c + ms_d)
    wall_ms = max_over_ranks(wall_ms)
    total_bytes = sum_over_ranks(float(len(data)))
    total_bins = sum_over_ranks(float(bins))

    # e2e: the C-ABI file calls with host buffers (index + H2D + kernels + D2H + container inside the clock)
    e2e_steps = max(1, min(args.steps, 3))
    rec = cc.file_compress(data, A.MODE_SLICE)
    assert cc.file_decompress(rec) == data
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        rec = cc.file_compress(data, A.MODE_SLICE)
        back = cc.file_decompress(rec)
    barrier()
    e2e_s = max_over_ranks((time.perf_counter() - t0) / e2e_steps)
    clocks = sampler.summary()
    assert back == data

    if rank != 0:
        return 0
    peak, peak_src = load_peaks()
    dom = "decompress" if ms_d >= ms_c else "compress"
    dom_ms = max(ms_c, ms_d)
    algo_bytes = payload + coded     # N + C per direction
    achieved = algo_bytes / dom_ms / 1e6
    tr = load_traffic() if args.workload == "cfg2" else {}   # the committed capture is of the cfg2 stream
    traffic = tr.get(dom)
    # the bounds that mean something for a serial integer recurrence, from live numbers
    try:
        f_hz = float((clocks or {}).get("sm_mhz") or 0) * 1e6
        longest = max(r.n_bins for r in res)
        inst = (tr.get("warp_inst_per_launch") or {}).get(dom)   # smsp__inst_executed.sum of the committed --set full capture
        bins_s_dom = bins / dom_ms * 1e3
        issue_peak = (148 * 4 * f_hz / (inst / bins)) if (inst and f_hz) else None
        bounds = {
            "latency": {"longest_slice_bins": int(longest), "mean_slice_bins": bins / max(1, len(descs)),
                        "cycles_per_bin_on_longest_slice": (dom_ms * 1e-3 * f_hz / longest) if f_hz else None,
                        "note": "a launch lasts as long as its longest slice; cycles/bin = kernel time x SM clock / its bins"},
            "issue": {"achieved_bins_per_s": bins_s_dom, "peak_bins_per_s": issue_peak,
                      "frac": (bins_s_dom / issue_peak) if issue_peak else None,
                      "warp_inst_per_bin": (inst / bins) if inst else None,
                      "note": "148 SM x 4 schedulers x f_SM / warp instructions per bin (ncu, spin-wait included)"},
        }
    except Exception as e:   # explanatory only: must never sink the measured line
        bounds = {"error": str(e)}
    line = {
        "metric": METRIC, "value": total_bytes / dev_ms / 1e3, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": dev_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8/u64 integer", "data": "synthetic",
        "config": {"workload": "%s: %s" % (args.workload, DESCR[args.workload]), "stream_bytes_per_gpu": len(data),
                   "slices_per_gpu": len(descs), "chunking": "slice (model reset per slice); %d slices of >= 64 KiB as wide slices (2 streams each)" % n_wide, "sharding": "one stream per GPU, no collective",
                   "l2": "flushed (256 MB write) before every kernel"},
        "compress_MBps": total_bytes / max_over_ranks(ms_c) / 1e3 if world == 1 else None,
        "decompress_MBps": total_bytes / max_over_ranks(ms_d) / 1e3 if world == 1 else None,
        "bins_per_s": total_bins * 2 / dev_ms * 1e3, "bins_per_byte": bins / payload,
        "bytes_saved_frac": 1.0 - len(rec) / len(data), "bit_exact": True, "parity_gate": gate, "wall_ms_per_step": wall_ms,
        "gpu_launches": int(launches),
        "e2e": {"value": total_bytes / e2e_s / 1e6, "unit": UNIT, "h2d_bytes_per_step": len(data) + coded + 80 * len(descs),
                "d2h_bytes_per_step": coded + payload + 64 * len(descs), "api": "avr_file_compress + avr_file_decompress (host buffers)"},
        "roofline": {"bound": "hbm", "kernel": ("avr_wide_kernel<%s> + avr_job_kernel<%s> (one launch pair, concurrent)" if n_wide else "avr_job_kernel<%s>%s") % (dom.capitalize(), "" if not n_wide else dom.capitalize()),
                     "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src, "bounds": bounds,
                     "note": "serial integer recurrence per slice: latency-bound, not HBM-bound; see DESIGN.md for the per-bin latency bound"},
        "clocks": clocks,
    }
    if world == 1:
        try:
            m = cpu_reference_measure(args.workload)
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
