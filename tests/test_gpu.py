"""Parity tests proper: the CUDA path, called through the C ABI, against the reference compressor's
golden containers, against the CPU oracle on seeded streams, and -- at BASELINE sizes -- through the
round-trip property.  Bit-exact everywhere (this is integer/byte work)."""
import os
import subprocess
import tempfile

import pytest

import helpers as H
import recproto
import avrecode_b200 as A
from avrecode_b200 import _abi
from test_oracle_pin import CASES, load_case, container_from

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def ctx(built):
    c = A.Context()
    yield c
    c.close()


def check_roundtrip(ctx, descs, data, res, blob):
    r_d, b_d = ctx.decompress(H.coded_descs(descs, res), blob)
    assert H.roundtrip_ok(descs, r_d, b_d, data)
    assert [r.n_bins for r in r_d] == [r.n_bins for r in res]


@pytest.mark.parametrize("name", CASES)
def test_file_mode_equals_reference_compressor(ctx, name):
    """Block.cabac of every slice and the whole container == bytes written by the unmodified reference."""
    data, ref_rec = load_case(name)
    entries = A.index_h264(data)
    descs = H.batch_descs(entries, True)
    res, blob = ctx.compress(descs, data)
    assert [r.status for r in res] == [0] * len(descs)
    _, ref_blocks = recproto.parse(ref_rec)
    assert [s for s, d in zip(H.slices_of(res, blob), descs) if not d.flags & _abi.AVR_SLICE_SKIP] == \
        [b["cabac"] for b in ref_blocks if "cabac" in b]
    assert container_from(data, entries, res, blob) == ref_rec
    assert ctx.file_compress(data, A.MODE_FILE) == ref_rec
    check_roundtrip(ctx, descs, data, res, blob)
    # the reference's own container decodes on the device
    assert ctx.file_decompress(ref_rec) == data


@pytest.mark.parametrize("name", CASES)
def test_slice_mode_equals_oracle(ctx, name):
    data, _ = load_case(name)
    descs = H.batch_descs(A.index_h264(data), False)
    res, blob = ctx.compress(descs, data)
    r_o, b_o = H.oracle_compress(descs, data)
    assert [r.status for r in res] == [0] * len(descs)
    assert H.slices_of(res, blob) == H.slices_of(r_o, b_o)
    assert [(r.n_bins, r.n_mbs) for r in res] == [(r.n_bins, r.n_mbs) for r in r_o]
    check_roundtrip(ctx, descs, data, res, blob)
    rec = ctx.file_compress(data, A.MODE_SLICE)
    assert recproto.parse(rec)[0] == b"avr-b200/slice"
    assert ctx.file_decompress(rec) == data


SEEDED = [
    dict(width=480, height=272, frames=6, slices=4, seed=2),
    dict(width=640, height=368, frames=5, high=True, seed=3, ragged=True, slices=3),
    dict(width=1280, height=720, frames=3, seed=11, kbps=4000),
    dict(width=720, height=480, frames=4, seed=12, high=True, p8x8=0.9, density=2.5, qp=18),
    dict(width=400, height=240, frames=12, seed=13, refs=4, pskip=0.7, pintra=0.2),
    dict(mbw=1, mbh=1, frames=9, seed=14),
    dict(mbw=255, mbh=2, frames=2, seed=15, slices=2),
    dict(width=480, height=272, frames=12, bpics=0.8, seed=17, refs=4, high=True, p8x8=0.8, pskip=0.5),   # B slices
    dict(width=640, height=368, frames=10, bpics=0.5, seed=18, refs=3, slices=3, ragged=True),
]


@pytest.mark.parametrize("kw", SEEDED, ids=lambda k: "seed%d" % k["seed"])
@pytest.mark.parametrize("mode_file", [True, False], ids=["file", "slice"])
def test_seeded_streams_equal_oracle(ctx, kw, mode_file):
    data = H.gen_stream(**kw)
    descs = H.batch_descs(A.index_h264(data), mode_file)
    assert descs
    res, blob = ctx.compress(descs, data)
    r_o, b_o = H.oracle_compress(descs, data)
    assert [r.status for r in res] == [r.status for r in r_o] == [0] * len(descs)
    assert H.slices_of(res, blob) == H.slices_of(r_o, b_o)
    check_roundtrip(ctx, descs, data, res, blob)


def test_decompress_equals_oracle_decompress(ctx):
    data, _ = load_case("high_8x8_ragged")
    descs = H.batch_descs(A.index_h264(data), True)
    res, blob = ctx.compress(descs, data)
    dd = H.coded_descs(descs, res)
    r_g, b_g = ctx.decompress(dd, blob)
    r_o, b_o = H.oracle_decompress(dd, blob)
    assert H.slices_of(r_g, b_g) == H.slices_of(r_o, b_o)


def test_empty_and_degenerate_batches(ctx):
    res, blob = ctx.compress([], b"\0")
    assert res == [] and blob == b""
    data, _ = load_case(CASES[0])
    ents = A.index_h264(data)
    # unsupported slice kinds are reported per slice, never crash, and cost nothing else
    d = _abi.copy_desc(ents[0].desc)
    d.flags |= _abi.AVR_SLICE_NEW_MODEL | _abi.AVR_SLICE_MBAFF
    res, _ = ctx.compress([d], data)
    assert res[0].status == _abi.AVR_SLICE_UNSUPPORTED and res[0].out_size == 0
    d = _abi.copy_desc(ents[0].desc)
    d.flags |= _abi.AVR_SLICE_NEW_MODEL
    d.slice_type = 3   # SP / SI slices are not walked
    res, _ = ctx.compress([d], data)
    assert res[0].status == _abi.AVR_SLICE_UNSUPPORTED
    # a payload of garbage must end as CORRUPT/OK-with-different-size, not hang or fault
    junk = os.urandom(4096)
    d = _abi.copy_desc(ents[0].desc)
    d.flags |= _abi.AVR_SLICE_NEW_MODEL
    d.in_off, d.in_size, d.payload_size = 0, len(junk), len(junk)
    res, _ = ctx.compress([d], junk)
    assert res[0].status in (_abi.AVR_SLICE_OK, _abi.AVR_SLICE_CORRUPT, _abi.AVR_SLICE_UNSUPPORTED)
    # truncated coded input on decompress: status, not a fault
    descs = H.batch_descs(ents, False)[:1]
    res, blob = ctx.compress(descs, data)
    dd = H.coded_descs(descs, res)
    dd[0].in_size = max(1, dd[0].in_size // 3)
    r_d, _ = ctx.decompress(dd, blob)
    # (garbage may also decode to an I_PCM macroblock, which the device reports as UNSUPPORTED)
    assert r_d[0].status in (_abi.AVR_SLICE_OK, _abi.AVR_SLICE_CORRUPT, _abi.AVR_SLICE_OVERFLOW, _abi.AVR_SLICE_UNSUPPORTED)
    # bad arguments
    bad = _abi.copy_desc(ents[0].desc)
    bad.in_off = len(data)
    with pytest.raises(A.AvrError):
        ctx.compress([bad], data)


def test_staged_pipeline_and_counters(ctx):
    data = H.gen_stream(width=640, height=368, frames=8, seed=21, slices=2)
    descs = H.batch_descs(A.index_h264(data), False)
    n0 = ctx.launch_count()
    ctx.stage(descs, data)
    for _ in range(2):          # re-launchable on the same staged input
        ctx.flush_l2()
        ctx.launch(0)
        ctx.sync()
    assert ctx.last_kernel_ms() > 0
    assert ctx.launch_count() == n0 + 2
    res, blob = ctx.collect()
    r_o, b_o = H.oracle_compress(descs, data)
    assert H.slices_of(res, blob) == H.slices_of(r_o, b_o)


def test_cli_roundtrip_verbs(ctx):
    cli = os.path.join(ROOT, "avrecode_b200", "recode")
    src = os.path.join(ROOT, "tests", "golden", "main_ip_4slices.h264")
    data, ref_rec = load_case("main_ip_4slices")
    with tempfile.TemporaryDirectory() as td:
        p = subprocess.run([cli, "roundtrip", src, os.path.join(td, "rt.rec")], capture_output=True, text=True)
        assert p.returncode == 0, p.stderr
        assert "Compress-decompress roundtrip succeeded:" in p.stdout and " compression ratio: " in p.stdout \
            and " protobuf overhead: " in p.stdout           # recode.cpp:1616-1618
        env = dict(os.environ, AVR_MODE="file")
        rec = os.path.join(td, "f.rec")
        subprocess.run([cli, "compress", src, rec], check=True, env=env)
        assert open(rec, "rb").read() == ref_rec
        out = os.path.join(td, "f.h264")
        subprocess.run([cli, "decompress", rec, out], check=True)
        assert open(out, "rb").read() == data
        p = subprocess.run([cli, "decompress", src], capture_output=True)   # not a container
        assert p.returncode == 1 and b"Exception" in p.stderr


@pytest.mark.parametrize("kw", [
    dict(width=1920, height=1088, frames=12, high=True, seed=31, kbps=30000, fps=60),          # cfg2 shape
    dict(width=1920, height=1088, frames=4, intra_period=1, slices=8, seed=32, kbps=60000),   # cfg3 shape
    dict(width=3840, height=2160, frames=2, high=True, p8x8=0.85, seed=33, density=2.0, qp=20),  # cfg4 shape
], ids=["cfg2", "cfg3", "cfg4"])
def test_full_size_roundtrip_property(ctx, kw):
    """BASELINE-sized pictures: size-independent properties (file round trip is the identity, bins agree
    in both directions, first slice equals the oracle's bytes)."""
    data = H.gen_stream(**kw)
    rec = ctx.file_compress(data, A.MODE_SLICE)
    assert ctx.file_decompress(rec) == data
    assert len(rec) < len(data)
    descs = H.batch_descs(A.index_h264(data), False)
    res, blob = ctx.compress(descs, data)
    r_o, b_o = H.oracle_compress(descs[:1], data)
    assert H.slices_of(res, blob)[0] == H.slices_of(r_o, b_o)[0]


def test_corpus_batch_equals_per_file(ctx):
    """BASELINE config 5 shape: many independent clips in one device batch == each clip alone, both modes."""
    clips = [H.gen_stream(width=320 + 64 * (i % 3), height=192 + 48 * (i % 2), frames=3 + i % 4, seed=400 + i,
                          high=bool(i & 1), slices=1 + i % 3) for i in range(9)]
    clips.append(b"")                                    # an empty file is a legal corpus member
    clips.append(b"\x00\x00\x01\x09\x10not a video")     # no CABAC slice at all: one literal block
    for mode in (A.MODE_SLICE, A.MODE_FILE):
        recs = ctx.files_compress(clips, mode)
        assert recs == [ctx.file_compress(c, mode) for c in clips]
        assert ctx.files_decompress(recs) == clips
    assert ctx.files_compress([], A.MODE_SLICE) == []
    # containers written by the reference decode in a batch too
    names = CASES[:3]
    pairs = [load_case(n) for n in names]
    assert ctx.files_decompress([r for _, r in pairs]) == [d for d, _ in pairs]


def test_corpus_every_slice_pinned(built):
    """BASELINE config 5 shape at a size the box generates in seconds: the first 64 clips of the bench corpus, a sixth of
    their length (~760 slices), through the kernel variant a large corpus selects -- estimators in GLOBAL memory, forced
    here by avr_config.resident = 0 -- and every slice's coded bytes against the ORACLE's, pinned per clip in
    tests/golden/corpus_gate.json (tests/make_bench_gate.py).  Then the batch file calls on the same corpus."""
    import hashlib, json, sys
    sys.path.insert(0, ROOT)
    import bench as B
    with open(os.path.join(ROOT, "tests", "golden", "corpus_gate.json")) as f:
        gate = json.load(f)["clips"]
    specs = B.corpus_spec(64, frame_div=6)
    clips = []
    for path in B.gen_corpus(specs):
        with open(path, "rb") as f:
            clips.append(f.read())
    assert [hashlib.sha256(c).hexdigest() for c in clips] == [g["stream_sha256"] for g in gate]
    with A.Context(resident=0) as c:
        descs, spans, offs, pieces = [], [], 0, []
        for clip in clips:
            ds = H.batch_descs(A.index_h264(clip), False)
            for d in ds:
                d.in_off += offs
            spans.append((len(descs), len(descs) + len(ds)))
            descs += ds
            pad = (-len(clip)) % 16
            pieces += [clip, bytes(pad)]
            offs += len(clip) + pad
        arena = b"".join(pieces)
        assert len(descs) == sum(g["n_slices"] for g in gate) > 700
        res, blob = c.compress(descs, arena)
        assert all(r.status == 0 for r in res)
        coded = H.slices_of(res, blob)
        for (lo, hi), g in zip(spans, gate):
            h = hashlib.sha256()
            for sl in coded[lo:hi]:
                h.update(len(sl).to_bytes(8, "little")); h.update(sl)
            assert h.hexdigest() == g["coded_sha256"] and sum(r.n_bins for r in res[lo:hi]) == g["n_bins"]
        r_d, b_d = c.decompress(H.coded_descs(descs, res), blob)
        assert H.roundtrip_ok(descs, r_d, b_d, arena)
        recs = c.files_compress(clips, A.MODE_SLICE)
        assert c.files_decompress(recs) == clips
    # a batch this big is a throughput job: no wide slices in its containers (avr_config.wide_batch_limit), while one
    # clip on its own gets them; both decode to the same bytes
    big = max(range(len(clips)), key=lambda i: len(clips[i]))
    with A.Context() as c:
        alone = c.file_compress(clips[big], A.MODE_SLICE)
        many = c.files_compress(clips, A.MODE_SLICE)
        assert not any("streams" in b for r in many for b in recproto.parse(r)[1])
        assert c.file_decompress(alone) == clips[big] == c.file_decompress(many[big])


@pytest.mark.parametrize("resident", [1, 0], ids=["shared-estimators", "global-estimators"])
def test_both_estimator_placements_equal_oracle(built, resident):
    """The kernel has two variants (estimators resident in shared memory / in the global workspace); the library
    picks per batch by estimated makespan, avr_config.resident forces one.  Both must give the oracle's bytes."""
    with A.Context(resident=resident) as c:
        for kw in (dict(width=320, height=192, frames=3, seed=1, slices=2),
                   dict(width=352, height=288, frames=4, high=True, seed=3, p8x8=0.7, refs=2)):
            data = H.gen_stream(**kw)
            for mode_file in (True, False):
                descs = H.batch_descs(A.index_h264(data), mode_file)
                res, blob = c.compress(descs, data)
                r_o, b_o = H.oracle_compress(descs, data)
                assert H.slices_of(res, blob) == H.slices_of(r_o, b_o)
                check_roundtrip(c, descs, data, res, blob)


# ---- wide slices (chunked mode only: a slice over K range-coder streams, decompressed as a wavefront) ----
WIDE = [
    dict(width=480, height=272, frames=6, slices=4, seed=2),
    dict(width=640, height=368, frames=5, high=True, seed=3, ragged=True, slices=3),     # slices that start mid-row
    dict(mbw=255, mbh=2, frames=2, seed=15, slices=2),                                  # fewer rows than streams
    dict(mbw=1, mbh=1, frames=3, seed=14),
    dict(width=480, height=272, frames=12, bpics=0.8, seed=17, refs=4, high=True, p8x8=0.8, pskip=0.5),
    dict(width=1920, height=1088, frames=3, high=True, seed=31, kbps=30000, fps=60),
]


@pytest.mark.parametrize("kw", WIDE, ids=lambda k: "seed%d" % k["seed"])
def test_wide_slices_equal_oracle_and_round_trip(ctx, kw):
    """The format is this library's own (no reference counterpart).  Its definition is restated independently in the
    oracle (oracle/avr_oracle.hpp: WideCompressIO / WideDecompressIO, serial, built from the oracle's own per-bin
    primitives): the device must write exactly those bytes, walk the bins of the ordinary path, round-trip, and the
    oracle must decode what the device wrote."""
    data = H.gen_stream(**kw)
    plain = H.batch_descs(A.index_h264(data), False)
    r_p, _ = ctx.compress(plain, data)
    descs = H.batch_descs(A.index_h264(data), False)
    for i, d in enumerate(descs):
        d.n_streams = H.WIDE_PATTERN[i % len(H.WIDE_PATTERN)]   # wide (2..8 streams) and ordinary slices in one batch
    res, blob = ctx.compress(descs, data)
    r_o, b_o = H.oracle_compress(descs, data)
    assert [r.status for r in res] == [r.status for r in r_o] == [0] * len(descs)
    assert H.slices_of(res, blob) == H.slices_of(r_o, b_o)
    assert [(r.n_bins, r.n_mbs) for r in res] == [(r.n_bins, r.n_mbs) for r in r_p] == [(r.n_bins, r.n_mbs) for r in r_o]
    check_roundtrip(ctx, descs, data, res, blob)
    r_d, b_d = H.oracle_decompress(H.coded_descs(descs, res), blob)
    assert H.roundtrip_ok(descs, r_d, b_d, data)


def test_wide_slices_in_the_container(built):
    data = H.gen_stream(width=640, height=368, frames=6, high=True, seed=41, kbps=6000)
    with A.Context(wide_stream_bytes=2048) as cw, A.Context(wide_stream_bytes=0) as c0, \
            A.Context(wide_stream_bytes=2048, wide_max_streams=A.MAX_STREAMS) as c8:
        rec_w = cw.file_compress(data, A.MODE_SLICE)
        rec_0 = c0.file_compress(data, A.MODE_SLICE)
        rec_8 = c8.file_compress(data, A.MODE_SLICE)
        _, blocks_w = recproto.parse(rec_w)
        _, blocks_0 = recproto.parse(rec_0)
        _, blocks_8 = recproto.parse(rec_8)
        assert {b.get("streams", 0) for b in blocks_w} == {0, 2} and not any("streams" in b for b in blocks_0)
        assert max(b.get("streams", 0) for b in blocks_8) > 2
        assert c0.file_decompress(rec_w) == data and cw.file_decompress(rec_0) == data
        assert c0.file_decompress(rec_8) == data and c8.file_decompress(rec_w) == data
        assert len(rec_w) < len(rec_0) * 1.02        # two models instead of one cost little
        assert len(rec_8) < len(rec_0) * 1.10        # a model per 2 KB of payload costs more: the default stays at two
        # file mode (the reference's format) never uses them
        assert not any("streams" in b for b in recproto.parse(cw.file_compress(data, A.MODE_FILE))[1])
    # a batch beyond avr_config.wide_batch_limit is a throughput job: no wide slices
    with A.Context(wide_stream_bytes=2048, wide_batch_limit=3) as c:
        assert not any("streams" in b for b in recproto.parse(c.file_compress(data, A.MODE_SLICE))[1])


def test_wide_slice_errors(ctx):
    data = H.gen_stream(width=480, height=272, frames=2, seed=2)
    descs = H.batch_descs(A.index_h264(data), False)
    descs[0].n_streams = A.MAX_STREAMS + 1                   # more than this build runs
    with pytest.raises(A.AvrError):
        ctx.compress(descs, data)
    descs = H.batch_descs(A.index_h264(data), True)          # file mode: the slices share one model
    descs[0].n_streams = 2
    with pytest.raises(A.AvrError):
        ctx.compress(descs, data)
    # damaged blob headers are reported per slice, not crashed on
    descs = H.batch_descs(A.index_h264(data), False)
    for d in descs:
        d.n_streams = 2
    res, blob = ctx.compress(descs, data)
    dd = H.coded_descs(descs, res)
    for damage in (lambda b: b[:4] + bytes([b[4] ^ 1]) + b[5:],       # stream size
                   lambda b: bytes(4) + b[4:],                        # zero macroblocks
                   lambda b: b[:8]):                                  # header cut short
        bad = damage(blob[res[0].out_off:res[0].out_off + res[0].out_size])
        d0 = _abi.copy_desc(dd[0])
        d0.in_off, d0.in_size = 0, len(bad)
        r_d, _ = ctx.decompress([d0], bad)
        assert r_d[0].status != 0


def test_escaped_slices_are_recoded_in_chunked_mode(ctx):
    """slice_data with emulation-prevention bytes goes to the device unescaped and comes back
    escaped; the reference's own format (file mode) still leaves such slices as literals."""
    from test_host import ESCAPED_STREAM
    data = H.gen_stream(**ESCAPED_STREAM)
    ents = A.index_h264(data)
    e, d, raw, lay = H.escaped_slice_case(data, ents)
    res, blob = ctx.compress([d], raw)
    r_o, b_o = H.oracle_compress([d], raw)
    assert res[0].status == 0 and H.slices_of(res, blob) == H.slices_of(r_o, b_o)
    want = data[e.desc.in_off:e.desc.in_off + e.desc.payload_size]
    for k in (0, 2):                                           # ordinary and wide
        d.n_streams = k
        res, blob = ctx.compress([d], raw)
        dd = H.coded_descs([d], res)
        dd[0].n_streams = lay | k
        r_d, b_d = ctx.decompress(dd, blob)
        assert r_d[0].status == 0 and H.fix_tail(H.slices_of(r_d, b_d)[0], want) == want
    rec_s = ctx.file_compress(data, A.MODE_SLICE)
    rec_f = ctx.file_compress(data, A.MODE_FILE)
    blocks_s = recproto.parse(rec_s)[1]
    assert sum(1 for b in blocks_s if b.get("escaped")) == 1 and not any(b.get("skip_coded") for b in blocks_s)
    assert sum(1 for b in recproto.parse(rec_f)[1] if b.get("skip_coded")) == 1
    assert ctx.file_decompress(rec_s) == data and ctx.file_decompress(rec_f) == data
    assert ctx.files_decompress(ctx.files_compress([data, data[:1000], data], A.MODE_SLICE)) == [data, data[:1000], data]


# ---- several devices behind one handle (avr_multi_*): every device this box has; one device is the degenerate case ----
def test_multi_context_equals_one_device(ctx):
    import torch
    from test_oracle_pin import load_case
    n_dev = torch.cuda.device_count()
    files = [load_case(n)[0] for n in ("main_ip_4slices", "high_b_8x8_slices", "main_ip_mp4", "rich_headers_b")]
    files += [H.gen_stream(width=1280, height=720, frames=8, seed=70 + i, kbps=6000) for i in range(4)]
    with A.MultiContext(list(range(n_dev))) as m:
        assert m.n_devices == n_dev
        for mode in (A.MODE_FILE, A.MODE_SLICE):
            one = ctx.files_compress(files, mode)
            many = m.files_compress(files, mode)
            assert many == one
            assert m.files_decompress(one) == files and ctx.files_decompress(many) == files
    with pytest.raises(A.AvrError):
        A.MultiContext([0, 0])
    with pytest.raises(A.AvrError):
        A.MultiContext([n_dev])


# ---- static priors (chunked mode): fresh models start from a table instead of {1, 1} everywhere ----
@pytest.mark.parametrize("table", ["random", "shipped"])
def test_priors_equal_oracle_and_round_trip(built, table):
    """Slices that carry AVR_SLICE_PRIORS, ordinary and wide, both kernel variants: the device writes the oracle's bytes
    (the oracle restates the table layout from the model keys, oracle/avr_oracle.hpp prior_index) and round-trips."""
    kw = dict(width=640, height=368, frames=6, high=True, seed=3, ragged=True, slices=3, bpics=0.4, refs=2, p8x8=0.6)
    data = H.gen_stream(**kw)
    t = H.random_priors(7) if table == "random" else A.load_priors("default")
    H.set_test_priors(t)
    try:
        for resident in (1, 0):
            with A.Context(resident=resident) as c:
                descs = H.with_priors(H.batch_descs(A.index_h264(data), False))
                with pytest.raises(A.AvrError):                     # no table in this context yet
                    c.compress(descs, data)
                c.set_priors(bytes(t))
                for i, d in enumerate(descs):
                    d.n_streams |= H.WIDE_PATTERN[i % len(H.WIDE_PATTERN)]
                res, blob = c.compress(descs, data)
                r_o, b_o = H.oracle_compress(descs, data)
                assert [r.status for r in res] == [r.status for r in r_o] == [0] * len(descs)
                assert H.slices_of(res, blob) == H.slices_of(r_o, b_o)
                check_roundtrip(c, descs, data, res, blob)
                plain = H.batch_descs(A.index_h264(data), False)  # without the flag: the reference's {1, 1}
                r_p, b_p = c.compress(plain, data)
                r_q, b_q = H.oracle_compress(plain, data)
                assert H.slices_of(r_p, b_p) == H.slices_of(r_q, b_q)
    finally:
        H.set_test_priors(None)


def test_priors_in_the_container_on_the_device(ctx):
    data = H.gen_stream(width=640, height=368, frames=6, high=True, seed=41, kbps=6000)
    name = CASES[0]
    gold, gold_rec = load_case(name)
    with A.Context() as primed:
        primed.set_priors("default")
        assert primed.priors_digest == A.priors_digest_of("default") != 0 and ctx.priors_digest == 0
        rec_p = primed.file_compress(data, A.MODE_SLICE)
        rec_0 = ctx.file_compress(data, A.MODE_SLICE)
        assert recproto.priors_digest(rec_p) == primed.priors_digest and recproto.priors_digest(rec_0) == 0
        assert len(rec_p) < len(rec_0)
        assert primed.file_decompress(rec_p) == data and primed.file_decompress(rec_0) == data
        with pytest.raises(A.AvrError) as e:
            ctx.file_decompress(rec_p)
        assert e.value.code == _abi.AVR_ERR_PRIORS
        assert primed.file_compress(gold, A.MODE_FILE) == gold_rec     # the reference's format does not use the table
        primed.set_priors(None)
        assert primed.priors_digest == 0 and primed.file_compress(data, A.MODE_SLICE) == rec_0
