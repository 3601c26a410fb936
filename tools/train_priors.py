"""Makes a prior table for the chunked container mode (include/avrecode_b200.h, "static priors").

    python tools/train_priors.py -o avrecode_b200/priors_v1.bin [--tmax 24] [file.h264 ...]

For every training stream the compressor runs once in the reference's mode (one model for the whole file) on the HOST
EMULATION of the device code (build/libavr_emu.so: test infrastructure, used here as an offline tool; the product never
loads it), and the estimators are read as that run left them.  An entry of the table is the pooled probability of the
files that touched the estimator, with a total of at most --tmax so that a fresh model still adapts quickly; an
estimator no training file touched stays {1, 1}.

Without file arguments the training set is synthetic (build/h264gen, seeds 5000..: none of the test / bench seeds): this
box has no real video.  A table trained on real streams is made by passing them on the command line."""
import argparse
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import avrecode_b200 as A
from avrecode_b200 import _abi
import helpers as H

N = _abi.AVR_PRIOR_ENTRIES


def training_specs():
    out = []
    for i in range(24):
        hd = i % 3
        out.append(dict(width=(640, 1280, 1920)[hd], height=(368, 720, 1088)[hd], frames=(10, 6, 4)[hd], seed=5000 + i,
                        high=bool(i % 2), kbps=(1500, 6000, 14000)[hd] * (1 + i % 4), fps=30, intra_period=(4, 6, 30)[i % 3],
                        bpics=(0, 0, 0.5)[(i // 3) % 3], refs=1 + i % 3, p8x8=(None, 0.5)[(i // 2) % 2], slices=1 + i % 2))
    return out


def final_counters(data):
    """Estimators after `data` went through the compressor with one model for the file -> uint16[N] (or None)."""
    descs = H.batch_descs(A.index_h264(data), True)
    if not descs:
        return None
    res, _ = H.emu_compress(descs, data)
    if any(r.status != 0 for r in res):
        return None
    lib = H.emu_lib()
    buf = (C.c_uint16 * N)()
    if lib.avr_emu_last_counters(buf, N) != 0:
        return None
    return np.frombuffer(buf, dtype=np.uint16).copy()


def make_table(counters, tmax):
    pos = np.zeros(N, dtype=np.float64)
    tot = np.zeros(N, dtype=np.float64)
    for c in counters:
        p = (c & 0xFF).astype(np.float64)
        t = (c >> 8).astype(np.float64)
        touched = t > 2                                      # {1, 1} = never coded in this file
        pos += np.where(touched, p - 1, 0)                   # observations only (the initial {1, 1} is not evidence)
        tot += np.where(touched, t - 2, 0)
    prob = np.where(tot > 0, (pos + 0.5) / (tot + 1.0), 0.5)
    total = np.clip(np.minimum(tot + 2, tmax), 2, 0x50).astype(np.int64)
    p_i = np.clip(np.rint(prob * total), 1, total - 1).astype(np.int64)
    return (p_i | (total << 8)).astype(np.uint16)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("-o", "--out", required=True)
    ap.add_argument("--tmax", type=int, default=24)
    ap.add_argument("files", nargs="*")
    a = ap.parse_args()
    streams = [open(f, "rb").read() for f in a.files] if a.files else [H.gen_stream(**{k: v for k, v in kw.items()}) for kw in training_specs()]
    counters = [c for c in (final_counters(d) for d in streams) if c is not None]
    if not counters:
        raise SystemExit("train_priors: no training stream could be coded")
    table = make_table(counters, a.tmax)
    with open(a.out, "wb") as f:
        f.write(table.tobytes())
    touched = int(np.count_nonzero(table != 0x0201))
    print("train_priors: %d streams, %d of %d estimators set, tmax %d -> %s" % (len(counters), touched, N, a.tmax, a.out))


if __name__ == "__main__":
    main()
