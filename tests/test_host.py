"""CPU-side checks: the C-ABI library loads and exports what include/avrecode_b200.h declares, fails
loudly without a device, the host slice indexer agrees with the reference's segmentation, the device
code compiled as plain C++ (tests/emu, debugging aid) matches the oracle, and the product never
touches oracle/.  No GPU needed."""
import ctypes as C
import os
import re
import subprocess

import pytest

import helpers as H
import recproto
import avrecode_b200 as A
from avrecode_b200 import _abi
from test_oracle_pin import CASES, load_case, MANIFEST

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_library_exports_every_declared_symbol(built):
    hdr = open(os.path.join(ROOT, "include", "avrecode_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = set(re.findall(r"\b(avr_[a-z0-9_]+)\s*\(", hdr))
    assert {"avr_create", "avr_compress", "avr_decompress", "avr_stage", "avr_launch", "avr_collect",
            "avr_index_h264", "avr_file_compress", "avr_file_decompress"} <= names
    lib = A.load_library()
    for n in sorted(names):
        assert hasattr(lib, n), "missing export " + n
    assert lib.avr_abi_version() == 2 and C.sizeof(_abi.Config) == 32
    assert b"no CPU fallback" in lib.avr_strerror(_abi.AVR_ERR_NO_DEVICE)


def test_struct_layouts_match_header(built):
    assert C.sizeof(_abi.SliceDesc) == 40 and C.sizeof(_abi.SliceResult) == 32 and C.sizeof(_abi.IndexEntry) == 48


@pytest.mark.skipif(_has_gpu(), reason="only meaningful on a machine without a CUDA device")
def test_no_device_fails_loudly(built):
    with pytest.raises(A.NoDeviceError):
        A.Context()
    with pytest.raises(A.NoDeviceError):
        A.MultiContext([0, 1])
    cli = os.path.join(ROOT, "avrecode_b200", "recode")
    src = os.path.join(ROOT, "tests", "golden", CASES[0] + ".h264")
    p = subprocess.run([cli, "roundtrip", src], capture_output=True)
    assert p.returncode == 1 and b"Exception" in p.stderr and b"no CPU fallback" in p.stderr


def test_cli_usage_matches_reference(built):
    cli = os.path.join(ROOT, "avrecode_b200", "recode")
    p = subprocess.run([cli], capture_output=True)
    assert p.returncode == 1
    assert b"[compress|decompress|roundtrip] <input> [output]" in p.stderr   # recode.cpp:1631-1635


def test_product_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "avrecode_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if not fn.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                continue
            text = open(os.path.join(dirpath, fn), errors="replace").read()
            for line in text.splitlines():
                code = line.split("//")[0].split("#")[0] if not line.lstrip().startswith(("#include", "#if", "#define", "#else", "#endif", "#pragma")) else line
                assert "avr_oracle" not in code and "oracle/" not in code or fn == "build.py", (fn, line)
    out = subprocess.run(["nm", "-D", "--defined-only", A.lib_path()], capture_output=True, text=True).stdout
    assert "oracle" not in out and "emu" not in out


@pytest.mark.parametrize("name", CASES)
def test_index_agrees_with_reference_segmentation(built, name):
    data, ref_rec = load_case(name)
    entries = A.index_h264(data)
    _, blocks = recproto.parse(ref_rec)
    coded = [b for b in blocks if "cabac" in b or "skip_coded" in b]
    assert len(coded) == len(entries)
    pos = 0
    k = 0
    for b in blocks:
        if "literal" in b:
            pos += len(b["literal"])
        elif "cabac" in b:
            e = entries[k]
            k += 1
            assert e.recodable and e.desc.in_off == pos and e.desc.payload_size == b["size"]
            pos += b["size"]
        else:
            assert not entries[k].recodable and entries[k].desc.payload_size == b["size"]
            k += 1
    assert pos == len(data)


def test_index_edge_cases(built):
    assert A.index_h264(b"") == []
    assert A.index_h264(b"\x00" * 64) == []
    assert A.index_h264(os.urandom(4096)) is not None        # garbage must not crash
    data, _ = load_case(CASES[0])
    for cut in (5, 40, len(data) // 2, len(data) - 1):       # truncated streams must not crash
        A.index_h264(data[:cut])
    ents = A.index_h264(data)
    assert ents[0].desc.flags & _abi.AVR_SLICE_NEW_PICTURE
    assert all(e.desc.chroma_format_idc == 1 for e in ents)


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("mode_file", [True, False])
def test_device_logic_on_host_matches_oracle(built, name, mode_file):
    """avr_core.cuh / avr_job.cuh compiled as plain C++ (one emulated lane): same bytes as the oracle."""
    data, _ = load_case(name)
    descs = H.batch_descs(A.index_h264(data), mode_file)
    r_o, b_o = H.oracle_compress(descs, data)
    r_e, b_e = H.emu_compress(descs, data)
    assert [r.status for r in r_e] == [0] * len(descs)
    assert H.slices_of(r_e, b_e) == H.slices_of(r_o, b_o)
    assert [r.n_bins for r in r_e] == [r.n_bins for r in r_o]
    r_d, b_d = H.emu_decompress(H.coded_descs(descs, r_e), b_e)
    assert H.roundtrip_ok(descs, r_d, b_d, data)


@pytest.mark.parametrize("kw", [dict(width=480, height=272, frames=4, slices=2, seed=2),
                                dict(width=352, height=288, frames=6, bpics=0.6, seed=19, refs=3, high=True),
                                dict(mbw=3, mbh=1, frames=2, seed=20)], ids=["ip", "b", "one-row"])
def test_wide_slices_on_host_round_trip(built, kw):
    """Wide slices (two range-coder streams per slice, rows interleaved) in the host emulation of the device
    code: exact round trip, the bins of the ordinary path, and a size within a few percent of it."""
    data = H.gen_stream(**kw)
    plain = H.batch_descs(A.index_h264(data), False)
    r_1, _ = H.emu_compress(plain, data)
    descs = H.batch_descs(A.index_h264(data), False)
    for i, d in enumerate(descs):
        d.n_streams = (2, 3, 8)[i % 3]
    r_e, b_e = H.emu_compress(descs, data)
    assert [(r.status, r.n_bins, r.n_mbs) for r in r_e] == [(r.status, r.n_bins, r.n_mbs) for r in r_1]
    r_d, b_d = H.emu_decompress(H.coded_descs(descs, r_e), b_e)
    assert H.roundtrip_ok(descs, r_d, b_d, data)
    assert sum(r.out_size for r in r_e) < 1.5 * sum(r.out_size for r in r_1) + 128 * len(descs)


ESCAPED_STREAM = dict(width=640, height=368, frames=6, high=True, seed=108, kbps=9000)   # slice 4 has a 00 00 03 in its payload


def test_escaped_slice_on_host_round_trips(built):
    """Chunked mode recodes slices with emulation-prevention bytes (the reference leaves them literal): the
    coder sees the unescaped slice_data -- same bytes as the oracle gives for it -- and decompression puts the
    0x03 bytes back."""
    data = H.gen_stream(**ESCAPED_STREAM)
    ents = A.index_h264(data)
    case = H.escaped_slice_case(data, ents)
    assert case is not None and sum(1 for e in ents if not e.recodable) >= 1
    e, d, raw, lay = case
    assert not e.recodable                                    # file mode keeps the reference's behaviour
    assert len(raw) < e.desc.payload_size
    r_e, b_e = H.emu_compress([d], raw)
    r_o, b_o = H.oracle_compress([d], raw)
    assert r_e[0].status == 0 and H.slices_of(r_e, b_e) == H.slices_of(r_o, b_o)
    dd = H.coded_descs([d], r_e)
    dd[0].n_streams = lay                                     # header zeros + AVR_SLICE_ESCAPED
    r_d, b_d = H.emu_decompress(dd, b_e)
    want = data[e.desc.in_off:e.desc.in_off + e.desc.payload_size]
    assert r_d[0].status == 0
    assert H.fix_tail(H.slices_of(r_d, b_d)[0], want) == want
    # and as a wide slice
    d.n_streams = 2
    r_e, b_e = H.emu_compress([d], raw)
    dd = H.coded_descs([d], r_e)
    dd[0].n_streams = lay | 2
    r_d, b_d = H.emu_decompress(dd, b_e)
    assert H.fix_tail(H.slices_of(r_d, b_d)[0], want) == want


def test_manifest_matches_files():
    import hashlib
    for name, m in MANIFEST.items():
        data, rec = load_case(name)
        assert hashlib.sha256(data).hexdigest() == m["sha256"] and hashlib.sha256(rec).hexdigest() == m["rec_sha256"]


def test_bench_gate_digests_are_the_oracles(built):
    """bench.py's pre-timing parity gate compares the CUDA bytes of ALL ordinary slices with tests/golden/bench_gate.json
    instead of running the oracle; here (CPU) the pinned digest of the small workload is re-derived from the oracle."""
    import hashlib, json, sys
    sys.path.insert(0, H.ROOT)
    import bench as B
    with open(os.path.join(H.ROOT, "tests", "golden", "bench_gate.json")) as f:
        pin = json.load(f)
    assert set(pin["workloads"]) == set(B.WORKLOADS)
    w = pin["workloads"]["cfg1"]
    data, _ = B.gen_workload("cfg1", 0)
    assert hashlib.sha256(data).hexdigest() == w["stream_sha256"]
    plain = H.batch_descs(A.index_h264(data), mode_file=False)
    assert len(plain) == w["n_slices"] == len(w["n_bins"])
    r_o, b_o = H.oracle_compress(plain, data)
    h = hashlib.sha256()
    for s in H.slices_of(r_o, b_o):
        h.update(len(s).to_bytes(8, "little")); h.update(s)
    assert h.hexdigest() == w["coded_sha256"] and [int(r.n_bins) for r in r_o] == w["n_bins"]
    for name, ww in pin["workloads"].items():   # every workload is pinned on all of its slices, and against the reference's size
        assert ww["n_slices"] == len(ww["n_bins"]) > 0 and ww["ref_rec_bytes"] and ww["ref_rec_bytes"] < ww["stream_bytes"], name


def test_reference_arm_segments_tile_the_file(built):
    """bench.py --impl reference times closed groups of pictures of the workload file: they must be byte ranges that tile it
    (same-config with the GPU arm), each a stream the indexer reads on its own."""
    import sys
    sys.path.insert(0, H.ROOT)
    import bench as B
    data = H.gen_stream(width=320, height=192, frames=10, intra_period=4, seed=9)
    ents = A.index_h264(data)
    segs = B.gop_segments(data, ents)
    assert len(segs) == 3
    # pieces after the first start with the SPS/PPS of the stream; without that prefix they tile the file
    first_slice = data.index(b"\x00\x00\x01", data.index(b"\x00\x00\x01", 4) + 3)   # third start code = first slice NAL (4-byte codes)
    prefix = data[:first_slice - 1] if data[first_slice - 1] == 0 else data[:first_slice]
    assert all(s.startswith(prefix) for s in segs)
    assert segs[0] + b"".join(s[len(prefix):] for s in segs[1:]) == data
    n = [len(A.index_h264(s)) for s in segs]
    assert n == [4, 4, 2] and sum(n) == len(ents)
    specs = B.corpus_spec(8)
    assert [s["seed"] for s in specs] == list(range(1000, 1008)) and {s["width"] for s in specs} == {1280, 1920}
    assert all(30 <= s["frames"] <= 120 and 5000 <= s["kbps"] <= 30000 for s in B.corpus_spec(256))


@pytest.mark.parametrize("kw", [
    dict(width=480, height=272, frames=6, slices=4, seed=2),
    dict(width=640, height=368, frames=5, high=True, seed=3, ragged=True, slices=3),     # slices that start mid-row
    dict(mbw=255, mbh=2, frames=2, seed=15, slices=2),                                  # fewer rows than streams
    dict(width=480, height=272, frames=8, bpics=0.8, seed=17, refs=4, high=True, p8x8=0.8, pskip=0.5),
], ids=lambda k: "seed%d" % k["seed"])
def test_wide_format_restatement_equals_device_logic(built, kw):
    """Wide slices have no reference counterpart: the oracle restates the format's DEFINITION serially with its own
    per-bin primitives (WideCompressIO / WideDecompressIO), the device code (here: its host emulation) implements it
    as a wavefront over rings.  Same bytes, and each decodes the other's output."""
    data = H.gen_stream(**kw)
    descs = H.batch_descs(A.index_h264(data), False)
    for i, d in enumerate(descs):
        d.n_streams = H.WIDE_PATTERN[i % len(H.WIDE_PATTERN)]
    r_e, b_e = H.emu_compress(descs, data)
    r_o, b_o = H.oracle_compress(descs, data)
    assert [r.status for r in r_o] == [0] * len(descs)
    assert H.slices_of(r_e, b_e) == H.slices_of(r_o, b_o)
    assert [(r.n_bins, r.n_mbs) for r in r_e] == [(r.n_bins, r.n_mbs) for r in r_o]
    dd = H.coded_descs(descs, r_o)
    for dec in (H.oracle_decompress, H.emu_decompress):
        r_d, b_d = dec(dd, b_o)
        assert H.roundtrip_ok(descs, r_d, b_d, data)


# ---- the product's file layer (avr_file.cpp + h264_index.cpp) on the host emulation: container codec, multi-device ----
FILE_LAYER_CASES = ["main_ip_4slices", "high_b_8x8_slices", "main_ip_mp4", "one_mb_stop_byte", "high_8x8_ragged"]


@pytest.fixture(scope="module")
def emu_ctx(built):
    lib = H.emu_file_lib()
    h = C.c_void_p()
    assert lib.avr_create(None, C.byref(h)) == 0
    yield h
    lib.avr_destroy(h)


@pytest.fixture(scope="module")
def emu_multi(built):
    lib = H.emu_file_lib()
    devs = (C.c_int * 3)(0, 1, 2)
    m = C.c_void_p()
    assert lib.avr_multi_create(None, devs, 3, C.byref(m)) == 0
    yield m
    lib.avr_multi_destroy(m)


def test_file_layer_writes_the_references_containers(emu_ctx):
    """AVR_MODE_FILE through the product's own host code (index -> batch -> Recoded container): byte-identical to the
    .rec files the reference binary wrote for the golden cases (tests/golden), raw streams and mp4."""
    lib = H.emu_file_lib()
    files, want = zip(*[load_case(n) for n in FILE_LAYER_CASES])
    rc, recs = H.emu_files_call(lib.avr_files_compress, emu_ctx, list(files), A.MODE_FILE)
    assert rc == 0
    assert [hashlib_sha(r) for r in recs] == [hashlib_sha(w) for w in want]
    rc, back = H.emu_files_call(lib.avr_files_decompress, emu_ctx, list(want))
    assert rc == 0 and back == list(files)


def hashlib_sha(b):
    import hashlib
    return hashlib.sha256(b).hexdigest()


@pytest.mark.parametrize("mode", [A.MODE_FILE, A.MODE_SLICE], ids=["file", "slice"])
def test_multi_device_sharding_is_invisible_in_the_output(emu_ctx, emu_multi, mode):
    """avr_multi_*: the units of a call (files in the reference's mode, slices in chunked mode) are dealt out over three
    "devices" (emulation contexts), each with a compact arena of its own and a host thread of its own.  Same containers
    as one device, byte for byte; every device took part; either side decompresses the other's output."""
    lib = H.emu_file_lib()
    files = [load_case(n)[0] for n in FILE_LAYER_CASES] + [H.gen_stream(**ESCAPED_STREAM), b"", b"\x00\x00\x01\x09\x10"]
    rc, one = H.emu_files_call(lib.avr_files_compress, emu_ctx, files, mode)
    assert rc == 0
    before = [lib.avr_emu_ctx_calls(lib.avr_multi_ctx(emu_multi, i)) for i in range(3)]
    rc, many = H.emu_files_call(lib.avr_multi_files_compress, emu_multi, files, mode)
    assert rc == 0 and many == one
    after = [lib.avr_emu_ctx_calls(lib.avr_multi_ctx(emu_multi, i)) for i in range(3)]
    assert all(a > b for a, b in zip(after, before))
    rc, back = H.emu_files_call(lib.avr_multi_files_decompress, emu_multi, one)
    assert rc == 0 and back == files
    rc, back = H.emu_files_call(lib.avr_files_decompress, emu_ctx, many)
    assert rc == 0 and back == files


def test_multi_device_argument_errors(built):
    lib = H.emu_file_lib()
    m = C.c_void_p()
    assert lib.avr_multi_create(None, (C.c_int * 2)(0, 0), 2, C.byref(m)) != 0      # the same device twice
    assert lib.avr_multi_create(None, (C.c_int * 1)(0), 0, C.byref(m)) != 0
    assert lib.avr_multi_create(None, None, 2, C.byref(m)) != 0
    assert lib.avr_multi_ctx(None, 0) is None


# ---- static priors (chunked mode): fresh models start from a table instead of {1, 1} everywhere ----
def test_shipped_prior_table_is_valid(built):
    t = A.load_priors("default")
    assert len(t) == _abi.AVR_PRIOR_ENTRIES == 47892
    assert all(1 <= (v & 0xFF) < (v >> 8) <= 0x50 for v in t)
    assert sum(1 for v in t if v != 0x0201) > 4000


@pytest.mark.parametrize("kw", [
    dict(width=480, height=272, frames=6, slices=4, seed=2),
    dict(width=640, height=368, frames=5, high=True, seed=3, ragged=True, slices=3),
    dict(width=480, height=272, frames=8, bpics=0.8, seed=17, refs=4, high=True, p8x8=0.8, pskip=0.5),
], ids=lambda k: "seed%d" % k["seed"])
def test_prior_table_layout_restatement_equals_device_logic(built, kw):
    """The oracle maps a model KEY (recode.cpp:676-815, :884) to its table index from the documented layout
    (oracle/avr_oracle.hpp prior_index), the device code indexes the table with its own arithmetic: with a random table
    -- no entry {1, 1} -- every estimator either side would look up elsewhere changes the bytes.  Ordinary and wide
    slices; each decodes the other's output."""
    data = H.gen_stream(**kw)
    H.set_test_priors(H.random_priors(kw["seed"]))
    try:
        descs = H.with_priors(H.batch_descs(A.index_h264(data), False))
        for i, d in enumerate(descs):
            d.n_streams |= H.WIDE_PATTERN[i % len(H.WIDE_PATTERN)]
        r_e, b_e = H.emu_compress(descs, data)
        r_o, b_o = H.oracle_compress(descs, data)
        assert [r.status for r in r_o] == [0] * len(descs)
        assert H.slices_of(r_e, b_e) == H.slices_of(r_o, b_o)
        dd = H.coded_descs(descs, r_o)
        for dec in (H.oracle_decompress, H.emu_decompress):
            r_d, b_d = dec(dd, b_o)
            assert H.roundtrip_ok(descs, r_d, b_d, data)
        plain = H.batch_descs(A.index_h264(data), False)
        r_p, b_p = H.emu_compress(plain, data)                       # without the flag the table is not used
        H.set_test_priors(None)
        r_q, b_q = H.emu_compress(plain, data)
        assert H.slices_of(r_p, b_p) == H.slices_of(r_q, b_q) != H.slices_of(r_e, b_e)
        r_x, _ = H.oracle_compress(descs, data)                      # the flag without a table: refused
        assert all(r.status != 0 for r in r_x)
    finally:
        H.set_test_priors(None)


def test_priors_in_the_container(built, emu_multi):
    """File layer: chunked containers record the digest of the table they were written from (Metadata field 2) and are
    refused by a context that holds another (or none); the reference's mode never uses the table; the shipped table
    makes the synthetic test streams smaller."""
    lib = H.emu_file_lib()
    lib.avr_set_priors.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
    lib.avr_multi_set_priors.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
    lib.avr_priors_digest_of.restype = C.c_uint32
    lib.avr_priors_digest_of.argtypes = [C.c_void_p, C.c_uint32]
    table = A.load_priors("default")
    files = [load_case(n)[0] for n in FILE_LAYER_CASES]
    want_file = [load_case(n)[1] for n in FILE_LAYER_CASES]
    plain, primed = C.c_void_p(), C.c_void_p()
    assert lib.avr_create(None, C.byref(plain)) == 0 and lib.avr_create(None, C.byref(primed)) == 0
    try:
        assert lib.avr_set_priors(primed, table, len(table)) == 0
        assert lib.avr_set_priors(primed, table, len(table) - 1) == _abi.AVR_ERR_ARG
        bad = (C.c_uint16 * len(table))(*table)
        bad[5] = 0x0200                                               # pos = 0
        assert lib.avr_set_priors(plain, bad, len(bad)) == _abi.AVR_ERR_ARG
        rc, rec_0 = H.emu_files_call(lib.avr_files_compress, plain, files, A.MODE_SLICE)
        rc1, rec_p = H.emu_files_call(lib.avr_files_compress, primed, files, A.MODE_SLICE)
        assert rc == 0 and rc1 == 0
        digest = lib.avr_priors_digest_of(table, len(table))
        assert digest != 0 and [recproto.priors_digest(r) for r in rec_p] == [digest] * len(files)
        assert [recproto.priors_digest(r) for r in rec_0] == [0] * len(files)
        assert sum(map(len, rec_p)) < sum(map(len, rec_0))
        rc, back = H.emu_files_call(lib.avr_files_decompress, primed, rec_p)
        assert rc == 0 and back == files
        rc, back = H.emu_files_call(lib.avr_files_decompress, primed, rec_0)      # a table does no harm to plain containers
        assert rc == 0 and back == files
        rc, _ = H.emu_files_call(lib.avr_files_decompress, plain, rec_p)
        assert rc == _abi.AVR_ERR_PRIORS
        assert lib.avr_set_priors(plain, H.random_priors(1), len(table)) == 0     # another table: refused as well
        rc, _ = H.emu_files_call(lib.avr_files_decompress, plain, rec_p)
        assert rc == _abi.AVR_ERR_PRIORS
        rc, rec_f = H.emu_files_call(lib.avr_files_compress, primed, files, A.MODE_FILE)
        assert rc == 0 and rec_f == list(want_file)                   # the reference's format, byte for byte
        # several devices: every one holds the table, same containers
        assert lib.avr_multi_set_priors(emu_multi, table, len(table)) == 0
        rc, many = H.emu_files_call(lib.avr_multi_files_compress, emu_multi, files, A.MODE_SLICE)
        assert rc == 0 and many == rec_p
        rc, back = H.emu_files_call(lib.avr_multi_files_decompress, emu_multi, rec_p)
        assert rc == 0 and back == files
    finally:
        lib.avr_multi_set_priors(emu_multi, None, 0)
        lib.avr_destroy(plain); lib.avr_destroy(primed)


# ---- damaged inputs: refused or lossless, never a crash (file layer on the host emulation) ----
def _mutate(rng, b):
    d = bytearray(b)
    for _ in range(rng.randint(1, 4)):
        if len(d) < 2:
            break
        op = rng.randint(0, 3)
        if op == 0:
            d[rng.randrange(len(d))] ^= 1 << rng.randrange(8)
        elif op == 1 and len(d) > 8:
            i = rng.randrange(len(d))
            del d[i:i + rng.randint(1, 8)]
        elif op == 2:
            i = rng.randrange(len(d) + 1)
            d[i:i] = bytes(rng.randrange(256) for _ in range(rng.randint(1, 4)))
        else:
            d = d[:rng.randrange(len(d) + 1)]
    return bytes(d)


def test_damaged_streams_are_refused_or_recoded_losslessly(emu_ctx):
    """Any byte string is a legal input of `recode compress` (what cannot be parsed stays literal): a mutated stream
    either is refused with an error code or round-trips to exactly the mutated bytes, in both container modes."""
    import random
    lib = H.emu_file_lib()
    rng = random.Random(11)
    srcs = [load_case(n)[0] for n in ("main_ip_4slices", "main_ip_mp4", "rich_headers_b", "tiny_1mb_rows")]
    for it in range(24):
        d = _mutate(rng, rng.choice(srcs))
        mode = (A.MODE_FILE, A.MODE_SLICE)[it & 1]
        rc, rec = H.emu_files_call(lib.avr_files_compress, emu_ctx, [d], mode)
        if rc != 0:
            continue
        rc, back = H.emu_files_call(lib.avr_files_decompress, emu_ctx, rec)
        assert rc == 0 and back == [d], "mutation %d is not lossless" % it


def test_damaged_containers_are_refused_not_crashed_on(emu_ctx):
    """A damaged container decodes to an error or to some byte string -- the coded payloads carry their own check (the
    rebuilt slice must have the recorded size) -- but never takes the process down or hangs."""
    import random
    lib = H.emu_file_lib()
    rng = random.Random(5)
    data = load_case("main_ip_4slices")[0]
    recs = []
    for mode in (A.MODE_FILE, A.MODE_SLICE):
        rc, r = H.emu_files_call(lib.avr_files_compress, emu_ctx, [data], mode)
        assert rc == 0
        recs.append(r[0])
    errors = 0
    for it in range(120):
        rc, out = H.emu_files_call(lib.avr_files_decompress, emu_ctx, [_mutate(rng, recs[it & 1])])
        errors += rc != 0
        assert rc != 0 or len(out) == 1
    assert errors > 60
