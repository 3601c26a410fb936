"""Pins the CPU oracle (oracle/avr_oracle.hpp) to the reference.

 * known answers of the reference's own unit test (test/arithmetic_code.cpp:48-112, :14-47);
 * tests/golden/*.ref.rec -- containers written by the UNMODIFIED reference compressor
   (tests/make_golden.py), compared block for block and byte for byte;
 * when oracle/_ref holds the reference binaries (development container), freshly generated
   streams against the live reference.
No GPU needed."""
import ctypes as C
import json
import os
import subprocess
import tempfile

import pytest

import helpers as H
import recproto
import avrecode_b200 as A

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
with open(os.path.join(GOLD, "manifest.json")) as _f:
    MANIFEST = json.load(_f)
CASES = sorted(MANIFEST)


def load_case(name):
    with open(os.path.join(GOLD, name + "." + MANIFEST[name].get("ext", "h264")), "rb") as f:
        data = f.read()
    with open(os.path.join(GOLD, name + ".ref.rec"), "rb") as f:
        rec = f.read()
    return data, rec


def container_from(data, entries, results, blob, version=None):
    """Re-creates compressor::run's block sequence (recode.cpp:1115-1125, 1275-1297) from slice results."""
    blocks = []
    prev = 0
    k = 0
    for e in entries:   # file mode: one result per entry (skip_coded ones ride along as AVR_SLICE_SKIP)
        d = e.desc
        r = results[k]
        k += 1
        if e.recodable and r.status == 0:
            blocks.append(recproto.block_literal(data[prev:d.in_off]))
            payload = data[d.in_off:d.in_off + d.payload_size]
            blocks.append(recproto.block_cabac(d.payload_size, blob[r.out_off:r.out_off + r.out_size], payload[-1]))
            prev = d.in_off + d.payload_size
        else:
            blocks.append(recproto.block_skip(d.payload_size))
    blocks.append(recproto.block_literal(data[prev:]))
    return recproto.serialize(blocks, version)


def test_kat_half_probability(built):
    # test/arithmetic_code.cpp:93-110 with p = 1/2: N bits -> N/8 + 1 bytes, decodes back
    import random
    rnd = random.Random(7)
    for n, expect in ((100000, 12501), (1000, 126), (8, 2)):
        bits = bytes(rnd.getrandbits(1) for _ in range(n))
        dec = C.create_string_buffer(n)
        fn = H.oracle_lib().avr_oracle_kat_half
        fn.restype = C.c_uint64
        size = fn(bits, C.c_uint64(n), dec)
        assert size == expect
        assert dec.raw == bits


def test_kat_cabac_regression_vector(built):
    # test/arithmetic_code.cpp:16-17 (disabled branch): cabac::encoder output must decode with the CABAC decoder
    states = bytes([15, 17, 106, 28, 16, 0, 10, 26, 33, 22, 35, 58, 44, 0, 0, 1, 3, 5])
    bits = bytes([1, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 0, 1, 1, 1, 1, 1, 1])
    out = C.create_string_buffer(256)
    n = C.c_int(256)
    rc = H.oracle_lib().avr_oracle_kat_cabac(states, bits, len(states), 16, out, C.byref(n))
    assert rc == 0
    assert 0 < n.value < 32


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_compressor_bytes(built, name):
    data, ref_rec = load_case(name)
    entries = A.index_h264(data)
    descs = H.batch_descs(entries, mode_file=True)
    res, blob = H.oracle_compress(descs, data)
    assert all(r.status == 0 for r in res)
    # block-level: every Block.cabac equals the reference's
    _, ref_blocks = recproto.parse(ref_rec)
    ref_cabac = [b["cabac"] for b in ref_blocks if "cabac" in b]
    assert [s for s, d in zip(H.slices_of(res, blob), descs) if not d.flags & 0x80] == ref_cabac
    # container-level: literal segmentation, sizes, parity/last_byte, field order
    assert container_from(data, entries, res, blob) == ref_rec


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("mode_file", [True, False])
def test_oracle_roundtrip(built, name, mode_file):
    data, _ = load_case(name)
    entries = A.index_h264(data)
    descs = H.batch_descs(entries, mode_file)
    res, blob = H.oracle_compress(descs, data)
    res2, blob2 = H.oracle_decompress(H.coded_descs(descs, res), blob)
    assert H.roundtrip_ok(descs, res2, blob2, data)
    assert sum(r.n_bins for r in res) == sum(r.n_bins for r in res2)


def test_first_slice_of_chunk_mode_equals_reference(built):
    # with a cold model the first coded slice is byte-identical in every chunking mode
    for name in CASES:
        data, ref_rec = load_case(name)
        descs = H.batch_descs(A.index_h264(data), mode_file=False)
        res, blob = H.oracle_compress(descs[:1], data)
        _, ref_blocks = recproto.parse(ref_rec)
        assert H.slices_of(res, blob)[0] == [b["cabac"] for b in ref_blocks if "cabac" in b][0]


LIVE = [
    dict(width=208, height=160, frames=4, seed=101, slices=2),
    dict(width=192, height=128, frames=5, seed=102, high=True, p8x8=0.7),
    dict(width=176, height=144, frames=6, seed=103, refs=2, pskip=0.6, ragged=True, slices=3),
    dict(width=160, height=96, frames=3, seed=104, intra_period=1, pdense=0.1, density=2.0),
]


@pytest.mark.parametrize("kw", LIVE, ids=lambda k: "seed%d" % k["seed"])
def test_oracle_matches_live_reference(built, kw):
    ref = H.ref_binary("recode_ref")
    if ref is None:
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")
    data = H.gen_stream(**kw)
    with tempfile.TemporaryDirectory() as td:
        src, rec = os.path.join(td, "s.h264"), os.path.join(td, "s.rec")
        with open(src, "wb") as f:
            f.write(data)
        subprocess.run([ref, "compress", src, rec], check=True, capture_output=True)
        with open(rec, "rb") as f:
            ref_rec = f.read()
    entries = A.index_h264(data)
    descs = H.batch_descs(entries, True)
    res, blob = H.oracle_compress(descs, data)
    assert container_from(data, entries, res, blob) == ref_rec
