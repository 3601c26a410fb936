"""Test-side helpers: stream synthesis (build/h264gen), oracle + emulation bindings, slice batches.

Only tests/, bench.py's cpu_baseline leg and __graft_entry__.smoke() import this; the product
package never does."""
import ctypes as C
import os
import subprocess
import tempfile

from avrecode_b200 import _abi
from avrecode_b200._abi import SliceDesc, SliceResult, copy_desc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
H264GEN = os.path.join(ROOT, "build", "h264gen")
REF_DIR = os.path.join(ROOT, "oracle", "_ref")

_oracle = None
_emu = None


def oracle_lib():
    global _oracle
    if _oracle is None:
        _oracle = C.CDLL(os.path.join(ROOT, "oracle", "_build", "libavr_oracle.so"))
    return _oracle


def emu_lib():
    global _emu
    if _emu is None:
        _emu = C.CDLL(os.path.join(ROOT, "build", "libavr_emu.so"))
    return _emu


def gen_stream(**kw):
    """Runs build/h264gen; keyword names are its long options (width=..., high=True, ...)."""
    fd, path = tempfile.mkstemp(suffix=".h264")
    os.close(fd)
    cmd = [H264GEN, "-o", path]
    for k, v in kw.items():
        opt = "--" + k.replace("_", "-")
        if v is True:
            cmd.append(opt)
        elif v is not False and v is not None:
            cmd += [opt, str(v)]
    try:
        subprocess.run(cmd, check=True, capture_output=True)
        with open(path, "rb") as f:
            return f.read()
    finally:
        os.unlink(path)


def batch_descs(entries, mode_file):
    """Index entries -> slice descs for one file in file (shared model) or slice (chunk) mode."""
    descs = []
    for e in entries:
        if not e.recodable and not mode_file:
            continue
        d = copy_desc(e.desc)
        if not e.recodable:   # skip_coded: only its frame_spec reaches the shared model (recode.cpp:1139-1145)
            d.flags |= _abi.AVR_SLICE_SKIP
            d.in_off, d.in_size = 0, 0
        if not descs or not mode_file:
            d.flags |= _abi.AVR_SLICE_NEW_MODEL
        descs.append(d)
    return descs


def _call(fn, descs, data, *extra):
    n = len(descs)
    arr = (SliceDesc * n)(*descs)
    res = (SliceResult * n)()
    out = C.POINTER(C.c_uint8)()
    nb = C.c_uint64()
    buf = (C.c_uint8 * max(1, len(data))).from_buffer_copy(data if data else b"\0")
    rc = fn(arr, n, buf, C.c_uint64(len(data)), res, C.byref(out), C.byref(nb), *extra)
    assert rc == 0, rc
    blob = C.string_at(out, nb.value)
    return list(res), blob


def oracle_compress(descs, data):
    return _call(oracle_lib().avr_oracle_compress, descs, data)


def oracle_decompress(descs, data, mirror_compressor=1):
    return _call(oracle_lib().avr_oracle_decompress, descs, data, C.c_int(mirror_compressor))


WIDE_PATTERN = (2, 3, 0, 4, 8, 2, 5)   # avr_slice_desc.n_streams over the slices of a test batch (0 = ordinary)


def emu_compress(descs, data):
    return _call(emu_lib().avr_emu_compress, descs, data)


def emu_decompress(descs, data):
    return _call(emu_lib().avr_emu_decompress, descs, data)


def coded_descs(descs, results):
    """Descs for the decompress direction: input = the coded bytes described by `results`."""
    out = []
    for d, r in zip(descs, results):
        d2 = copy_desc(d)
        d2.in_off = r.out_off
        d2.in_size = r.out_size
        out.append(d2)
    return out


def slices_of(results, blob):
    return [blob[r.out_off:r.out_off + r.out_size] for r in results]


def roundtrip_ok(descs, results, blob, data):
    """True when every rebuilt payload (after the x264 fix-up) equals the original slice_data bytes."""
    for d, r in zip(descs, results):
        if d.flags & _abi.AVR_SLICE_SKIP:
            if r.status != 0 or r.out_size != 0:
                return False
            continue
        payload = data[d.in_off:d.in_off + d.payload_size]
        if r.status != 0 or fix_tail(blob[r.out_off:r.out_off + r.out_size], payload) != payload:
            return False
    return True


def fix_tail(got, payload):
    """x264 padding fix-up of recode.cpp:1347-1354 using the original payload's parity / last byte."""
    if (len(got) & 1) != (len(payload) & 1):
        return got + payload[-1:]
    return got[:-1] + payload[-1:]


def ref_binary(name="recode_ref"):
    p = os.path.join(REF_DIR, name)
    return p if os.path.exists(p) else None


def unescape(payload, zeros=0):
    """H.264 7.4.1: drop the 0x03 of every 00 00 03; zeros = zero bytes right before the payload."""
    out = bytearray()
    z = zeros
    for b in payload:
        if z >= 2 and b == 3:
            z = 0
            continue
        out.append(b)
        z = z + 1 if b == 0 else 0
    return bytes(out)


def escaped_slice_case(data, entries):
    """For the first slice whose payload contains emulation-prevention bytes: (entry, compress desc over a buffer
    that holds the unescaped payload, that buffer, decompress flags byte)."""
    for e in entries:
        if e.desc.n_streams & _abi.AVR_SLICE_ESCAPED:
            zeros = (e.desc.n_streams >> _abi.AVR_ESC_ZEROS_SHIFT) & 3
            raw = unescape(data[e.desc.in_off:e.desc.in_off + e.desc.payload_size], zeros)
            d = copy_desc(e.desc)
            d.in_off, d.in_size = 0, len(raw)
            d.n_streams = 0
            d.flags |= _abi.AVR_SLICE_NEW_MODEL
            return e, d, raw, e.desc.n_streams & 0xF0
    return None


# ---- the product's host file layer on top of the emulation (build/libavr_emu_file.so, tests/emu/avr_emu_file.cpp) ----
_emu_file = None


def emu_file_lib():
    global _emu_file
    if _emu_file is None:
        lib = C.CDLL(os.path.join(ROOT, "build", "libavr_emu_file.so"))
        u8p = C.POINTER(C.c_uint8)
        lib.avr_create.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
        lib.avr_destroy.argtypes = [C.c_void_p]
        lib.avr_destroy.restype = None
        lib.avr_multi_create.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_void_p)]
        lib.avr_multi_destroy.argtypes = [C.c_void_p]
        lib.avr_multi_destroy.restype = None
        lib.avr_multi_ctx.argtypes = [C.c_void_p, C.c_int]
        lib.avr_multi_ctx.restype = C.c_void_p
        lib.avr_emu_ctx_calls.argtypes = [C.c_void_p]
        many = [C.c_void_p, C.POINTER(u8p), C.POINTER(C.c_uint64), C.c_int]
        lib.avr_files_compress.argtypes = many + [C.c_int, C.POINTER(u8p), C.POINTER(C.c_uint64)]
        lib.avr_multi_files_compress.argtypes = many + [C.c_int, C.POINTER(u8p), C.POINTER(C.c_uint64)]
        lib.avr_files_decompress.argtypes = many + [C.POINTER(u8p), C.POINTER(C.c_uint64)]
        lib.avr_multi_files_decompress.argtypes = many + [C.POINTER(u8p), C.POINTER(C.c_uint64)]
        lib.avr_free.argtypes = [C.c_void_p]
        lib.avr_free.restype = None
        _emu_file = lib
    return _emu_file


def emu_files_call(fn, handle, blobs, mode=None):
    """avr_[multi_]files_compress / _decompress of the emulation-backed file layer: list of bytes -> (rc, list of bytes)."""
    lib = emu_file_lib()
    u8p = C.POINTER(C.c_uint8)
    n = len(blobs)
    keep = [(C.c_uint8 * max(1, len(b))).from_buffer_copy(b if b else b"\0") for b in blobs]
    ptrs = (u8p * max(1, n))(*[C.cast(k, u8p) for k in keep])
    sizes = (C.c_uint64 * max(1, n))(*[len(b) for b in blobs])
    outs = (u8p * max(1, n))()
    out_sizes = (C.c_uint64 * max(1, n))()
    args = [handle, ptrs, sizes, n] + ([mode] if mode is not None else []) + [outs, out_sizes]
    rc = fn(*args)
    got = []
    if rc == 0:
        for i in range(n):
            got.append(C.string_at(outs[i], out_sizes[i]))
            lib.avr_free(outs[i])
    return rc, got


# ---- prior tables (chunked mode; include/avrecode_b200.h "static priors") ----
_priors_keep = {}


def set_test_priors(table):
    """Gives the host emulation and the oracle the table that slices with AVR_SLICE_PRIORS start from (None: none).
    table: a ctypes uint16 array (A.load_priors) or None."""
    _priors_keep["t"] = table                       # both libraries only keep the pointer
    emu_lib().avr_emu_set_priors.argtypes = [C.c_void_p]
    oracle_lib().avr_oracle_set_priors.argtypes = [C.c_void_p]
    emu_lib().avr_emu_set_priors(table)
    oracle_lib().avr_oracle_set_priors(table)


def random_priors(seed):
    """A valid table with every entry different from {1, 1}: any index the device and the oracle disagree on shows."""
    import random
    rng = random.Random(seed)
    t = (C.c_uint16 * _abi.AVR_PRIOR_ENTRIES)()
    for i in range(_abi.AVR_PRIOR_ENTRIES):
        tot = rng.randint(3, 0x50)
        t[i] = rng.randint(1, tot - 1) | tot << 8
    return t


def with_priors(descs):
    for d in descs:
        d.n_streams |= _abi.AVR_SLICE_PRIORS
    return descs
