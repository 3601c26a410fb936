"""Small cases for compute-sanitizer --tool memcheck (run under gpurun): every kernel variant (compress / decompress,
resident / global estimators, file / slice mode, I / P / 8x8 / multi-slice), checked against the oracle."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import avrecode_b200 as A
import helpers as H

cases = [dict(width=320, height=192, frames=3, seed=1, slices=2), dict(width=352, height=288, frames=4, high=True, seed=3, p8x8=0.7, refs=2)]
for resident in ("1", "0"):
    os.environ["AVR_RESIDENT"] = resident
    with A.Context() as ctx:
        for kw in cases:
            data = H.gen_stream(**kw)
            for mode_file in (True, False):
                descs = H.batch_descs(A.index_h264(data), mode_file)
                res, blob = ctx.compress(descs, data)
                r_o, b_o = H.oracle_compress(descs, data)
                assert H.slices_of(res, blob) == H.slices_of(r_o, b_o)
                r_d, b_d = ctx.decompress(H.coded_descs(descs, res), blob)
                assert H.roundtrip_ok(descs, r_d, b_d, data)
            recs = ctx.files_compress([data, data[: len(data) // 2]], A.MODE_SLICE)
            assert ctx.files_decompress(recs)[0] == data
        print("resident=%s ok" % resident, flush=True)
print("SANITIZE CASE OK")
