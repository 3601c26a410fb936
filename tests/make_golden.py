"""Regenerates tests/golden/: small synthetic CABAC streams + what the UNMODIFIED reference makes of them.

Run in the development container (needs /root/reference built into oracle/_ref by
oracle/build_ref.sh and build/h264gen):   python tests/make_golden.py

For every case it writes
  <name>.h264      the input stream (h264gen, fixed seed -- committed so the GPU box needs no generator parity)
  <name>.ref.rec   `recode_ref compress <name>.h264` -- the reference COMPRESSOR's container, byte for byte
and records in manifest.json whether the stock reference round-trips the stream (`roundtrip` verb),
whether the one-line H8-patched copy does (SURVEY.md hazard H8), sizes, and sha256 of both files.
The reference compressor's bytes are the normative format definition."""
import hashlib
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
import helpers as H  # noqa: E402

CASES = {
    # name: h264gen options
    "main_i_320x192": dict(width=320, height=192, frames=3, intra_period=1, seed=1),
    "main_ip_4slices": dict(width=320, height=192, frames=5, slices=4, seed=2),
    "high_8x8_ragged": dict(width=320, height=192, frames=4, high=True, ragged=True, slices=3, seed=3),
    "main_dense_h1": dict(width=320, height=192, frames=3, pdense=0.05, seed=4, intra_period=2),
    "main_p_3refs": dict(width=352, height=288, frames=8, refs=3, seed=5, density=0.5),
    "high_no8x8": dict(width=256, height=144, frames=4, high=True, p8x8=0.0, seed=6),
    "tiny_1mb_rows": dict(mbw=3, mbh=2, frames=6, slices=2, seed=7, density=1.5),
    # 1x1-macroblock pictures: slices whose rbsp_stop_one_bit sits alone in the next byte (AVR_SLICE_STOP_BYTE)
    "one_mb_stop_byte": dict(mbw=1, mbh=1, frames=9, seed=14),
    # B slices (kept as reference pictures with increasing frame_num, SURVEY hazard H2): B mb/sub-mb types,
    # direct, both reference lists, 8x8 transform under direct_8x8_inference
    "main_b_2refs": dict(width=320, height=192, frames=8, bpics=0.6, refs=2, seed=5),
    "high_b_8x8_slices": dict(width=352, height=288, frames=8, bpics=0.6, refs=3, high=True, slices=2, seed=16, pskip=0.3),
    # slice headers with ref_pic_list_modification, pred_weight_table and memory management control operations
    # 1, 2, 4, 6: the host indexer must land on the same payload byte and slice QP as libavcodec does
    "rich_headers_b": dict(width=320, height=192, frames=12, refs=3, bpics=0.4, rich_headers=True, slices=2, seed=77),
    # the same kind of stream inside an mp4 container (avcC parameter sets, 4-byte NAL lengths, 3 samples per chunk):
    # the reference demuxes it with libavformat (recode.cpp:92, 118), the indexer walks the boxes itself
    "main_ip_mp4": dict(width=320, height=192, frames=7, slices=2, seed=21, mp4=dict(nal_len=4, chunk=3)),
    "high_p_mp4_co64": dict(width=352, height=288, frames=5, high=True, refs=2, seed=22, mp4=dict(nal_len=2, chunk=2, co64=True)),
}


def sha(b):
    return hashlib.sha256(b).hexdigest()


def main():
    out_dir = os.path.join(HERE, "golden")
    os.makedirs(out_dir, exist_ok=True)
    ref = H.ref_binary("recode_ref")
    ref_dbg = H.ref_binary("recode_ref_dbg")
    ref_h8 = H.ref_binary("recode_ref_h8")
    assert ref and ref_dbg and ref_h8, "build oracle/_ref first (oracle/build_ref.sh)"
    manifest = {}
    for name, kw in CASES.items():
        gen_kw = {k: v for k, v in kw.items() if k != "mp4"}
        data = H.gen_stream(**gen_kw)
        ext = "h264"
        if "mp4" in kw:
            import mp4mux
            data = mp4mux.mux(data, kw["width"], kw["height"], **kw["mp4"])
            ext = "mp4"
        src = os.path.join(out_dir, name + "." + ext)
        rec = os.path.join(out_dir, name + ".ref.rec")
        with open(src, "wb") as f:
            f.write(data)
        subprocess.run([ref, "compress", src, rec], check=True, capture_output=True)
        stock = subprocess.run([ref_dbg, "roundtrip", src], capture_output=True).returncode == 0
        h8 = subprocess.run([ref_h8, "roundtrip", src], capture_output=True).returncode == 0
        with open(rec, "rb") as f:
            rec_bytes = f.read()
        manifest[name] = {
            "gen": kw, "ext": ext, "bytes": len(data), "rec_bytes": len(rec_bytes), "sha256": sha(data), "rec_sha256": sha(rec_bytes),
            "stock_reference_roundtrips": stock, "h8_patched_reference_roundtrips": h8,
        }
        print(name, manifest[name])
    with open(os.path.join(out_dir, "manifest.json"), "w") as f:
        json.dump(manifest, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
