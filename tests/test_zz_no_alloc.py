"""avr_recode_noalloc -- the entry point that allocates nothing (include/avrecode_b200.h) -- through the C ABI, against
the existing entry points (avr_compress / avr_decompress): same bytes, same results, on calls in a row that share the
static scratch array.  No counterpart in the reference.

Collected last on purpose.  Device buffers come from torch: plumbing, as everywhere."""
import ctypes as C

import pytest

import helpers as H
import avrecode_b200 as A
from avrecode_b200 import _abi

pytestmark = pytest.mark.gpu

STREAMS = [   # call 1: many slices; call 2: fewer, other pictures (stale tables / results / counter must not show)
    dict(width=640, height=368, frames=6, high=True, seed=3, ragged=True, slices=3),
    dict(width=480, height=272, frames=3, seed=21, bpics=0.5, refs=2),
]


def _lib():
    lib = A.load_library()
    lib.avr_recode_noalloc.restype = C.c_int
    lib.avr_recode_noalloc.argtypes = [C.c_int, C.c_int, C.POINTER(_abi.SliceDesc), C.c_int, C.c_void_p, C.c_uint64,
                                       C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(_abi.SliceResult), C.c_void_p]
    return lib


def _noalloc(lib, torch, direction, descs, data):
    """-> (results, [bytes per slice]) from one avr_recode_noalloc call with torch-owned device buffers."""
    n = len(descs)
    arr = (_abi.SliceDesc * n)(*descs)
    res = (_abi.SliceResult * n)()
    need = C.c_uint64()
    assert lib.avr_recode_noalloc(0, direction, arr, n, None, len(data), None, 0, C.byref(need), res, None) == 0   # query
    d_in = torch.frombuffer(bytearray(data), dtype=torch.uint8).cuda()
    d_out = torch.full((need.value + 64,), 0xA5, dtype=torch.uint8, device="cuda")   # dirty on purpose
    torch.cuda.synchronize()
    rc = lib.avr_recode_noalloc(0, direction, arr, n, d_in.data_ptr(), len(data), d_out.data_ptr(), need.value, None, res, None)
    assert rc == 0, lib.avr_strerror(rc)
    out = d_out.cpu().numpy().tobytes()
    assert out[need.value:] == b"\xA5" * 64                      # nothing written past what it asked for
    return list(res), [out[r.out_off:r.out_off + r.out_size] for r in res]


def test_noalloc_equals_the_allocating_entry_points_on_calls_in_a_row(built):
    import torch
    lib = _lib()
    with A.Context() as ctx:
        for rep in range(2):                                     # twice over both streams: four calls per direction
            for kw in STREAMS:
                data = H.gen_stream(**kw)
                descs = H.batch_descs(A.index_h264(data), False)
                r_ref, blob = ctx.compress(descs, data)
                want = H.slices_of(r_ref, blob)
                r_na, got = _noalloc(lib, torch, 0, descs, data)
                assert [(r.status, r.out_size, r.n_mbs, r.n_bins) for r in r_na] == [(r.status, r.out_size, r.n_mbs, r.n_bins) for r in r_ref]
                assert [r.status for r in r_na] == [0] * len(descs)
                assert got == want                               # bit-exact: integer / byte work
                # and back: the coded bytes through the no-allocation decompress == the allocating one == the file
                dd = H.coded_descs(descs, r_ref)
                r_dref, b_dref = ctx.decompress(dd, blob)
                r_dna, got_d = _noalloc(lib, torch, 1, dd, blob)
                assert [(r.status, r.out_size, r.n_bins) for r in r_dna] == [(r.status, r.out_size, r.n_bins) for r in r_dref]
                assert got_d == H.slices_of(r_dref, b_dref)
                assert H.roundtrip_ok(descs, r_dref, b_dref, data)


def test_noalloc_limits_and_argument_errors(built):
    import torch
    lib = _lib()
    data = H.gen_stream(**STREAMS[1])
    descs = H.batch_descs(A.index_h264(data), False)
    n = len(descs)
    arr = (_abi.SliceDesc * n)(*descs)
    res = (_abi.SliceResult * n)()
    need = C.c_uint64()
    d_in = torch.frombuffer(bytearray(data), dtype=torch.uint8).cuda()
    d_out = torch.zeros(1 << 20, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    call = lambda a, k, cap: lib.avr_recode_noalloc(0, 0, a, k, d_in.data_ptr(), len(data), d_out.data_ptr(), cap, C.byref(need), res, None)
    assert call(arr, n, 16) == _abi.AVR_ERR_LIMIT and need.value > 16          # the caller's buffer is too small
    assert b"avr_recode_noalloc" in lib.avr_strerror(_abi.AVR_ERR_LIMIT)
    many = (_abi.SliceDesc * (_abi.AVR_NOALLOC_MAX_SLICES + 1))()
    assert lib.avr_recode_noalloc(0, 0, many, len(many), d_in.data_ptr(), len(data), d_out.data_ptr(), 1 << 20, None, None, None) == _abi.AVR_ERR_LIMIT
    shared = H.batch_descs(A.index_h264(data), True)                           # the reference's mode: one model for the file
    if len(shared) > 1:
        assert call((_abi.SliceDesc * n)(*shared), n, 1 << 20) == _abi.AVR_ERR_ARG
    wide = [_abi.copy_desc(d) for d in descs]
    wide[0].n_streams = 2
    assert call((_abi.SliceDesc * n)(*wide), n, 1 << 20) == _abi.AVR_ERR_ARG
    assert lib.avr_recode_noalloc(0, 2, arr, n, d_in.data_ptr(), len(data), d_out.data_ptr(), 1 << 20, None, res, None) == _abi.AVR_ERR_ARG
    assert lib.avr_recode_noalloc(99, 0, arr, n, d_in.data_ptr(), len(data), d_out.data_ptr(), 1 << 20, None, res, None) == _abi.AVR_ERR_ARG
    assert call(arr, n, 1 << 20) == 0 and all(r.status == 0 for r in res)      # still fine after the refusals
