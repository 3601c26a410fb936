"""Small cases for compute-sanitizer --tool memcheck (run under gpurun): every kernel variant (compress / decompress,
resident / split estimators, file / slice mode, I / P / B / 8x8 / multi-slice), wide slices of 2..8 streams, prior
tables, the no-allocation entry point -- each checked against the oracle."""
import ctypes as C
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import avrecode_b200 as A
from avrecode_b200 import _abi
import helpers as H

cases = [dict(width=320, height=192, frames=3, seed=1, slices=2),
         dict(width=352, height=288, frames=4, high=True, seed=3, p8x8=0.7, refs=2, bpics=0.4)]
table = A.load_priors("default")
for resident in (1, 0):
    with A.Context(resident=resident, wide_stream_bytes=1024, wide_max_streams=8) as ctx:
        ctx.set_priors("default")
        for kw in cases:
            data = H.gen_stream(**kw)
            for mode_file in (True, False):
                descs = H.batch_descs(A.index_h264(data), mode_file)
                res, blob = ctx.compress(descs, data)
                r_o, b_o = H.oracle_compress(descs, data)
                assert H.slices_of(res, blob) == H.slices_of(r_o, b_o)
                r_d, b_d = ctx.decompress(H.coded_descs(descs, res), blob)
                assert H.roundtrip_ok(descs, r_d, b_d, data)
            # wide slices of 2..8 streams with priors
            H.set_test_priors(table)
            descs = H.with_priors(H.batch_descs(A.index_h264(data), False))
            for i, d in enumerate(descs):
                d.n_streams |= H.WIDE_PATTERN[i % len(H.WIDE_PATTERN)]
            res, blob = ctx.compress(descs, data)
            r_o, b_o = H.oracle_compress(descs, data)
            assert [r.status for r in res] == [0] * len(descs) and H.slices_of(res, blob) == H.slices_of(r_o, b_o)
            r_d, b_d = ctx.decompress(H.coded_descs(descs, res), blob)
            assert H.roundtrip_ok(descs, r_d, b_d, data)
            H.set_test_priors(None)
            recs = ctx.files_compress([data, data[: len(data) // 2]], A.MODE_SLICE)
            assert ctx.files_decompress(recs)[0] == data
        print("resident=%d ok" % resident, flush=True)

# the no-allocation entry point, device buffers from torch
import torch
lib = A.load_library()
lib.avr_recode_noalloc.argtypes = [C.c_int, C.c_int, C.POINTER(_abi.SliceDesc), C.c_int, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64,
                                   C.POINTER(C.c_uint64), C.POINTER(_abi.SliceResult), C.c_void_p]
data = H.gen_stream(**cases[1])
descs = H.batch_descs(A.index_h264(data), False)
n = len(descs)
arr = (_abi.SliceDesc * n)(*descs); res = (_abi.SliceResult * n)(); need = C.c_uint64()
assert lib.avr_recode_noalloc(0, 0, arr, n, None, len(data), None, 0, C.byref(need), res, None) == 0
d_in = torch.frombuffer(bytearray(data), dtype=torch.uint8).cuda()
d_out = torch.zeros(need.value, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()
assert lib.avr_recode_noalloc(0, 0, arr, n, d_in.data_ptr(), len(data), d_out.data_ptr(), need.value, None, res, None) == 0
out = d_out.cpu().numpy().tobytes()
r_o, b_o = H.oracle_compress(descs, data)
assert [out[r.out_off:r.out_off + r.out_size] for r in res] == H.slices_of(r_o, b_o)
print("SANITIZE CASE OK")
