"""ctypes binding of include/avrecode_b200.h.

Mirrors the reference's driver surface for this path: Context.file_compress / file_decompress are
`recode compress` / `recode decompress` (recode.cpp:1102-1309, 1312-1591) on in-memory files;
Context.compress / decompress are the slice-batch calls that replace the per-slice
compressor::cabac_decoder / decompressor::cabac_decoder objects (recode.cpp:1134-1268, 1411-1520)."""
import ctypes as C
import os

from . import _abi
from ._abi import SliceDesc, SliceResult, Config, IndexEntry

MODE_FILE, MODE_SLICE = _abi.AVR_MODE_FILE, _abi.AVR_MODE_SLICE
MAX_STREAMS = _abi.AVR_MAX_STREAMS
_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None


class AvrError(RuntimeError):
    def __init__(self, code, what):
        self.code = code
        super().__init__("%s: %s" % (what, _strerror(code)))


class NoDeviceError(AvrError):
    pass


def lib_path():
    return os.path.join(_HERE, "libavrecode_b200.so")


def load_library():
    """Loads libavrecode_b200.so (built in-tree by avrecode_b200.build / __graft_entry__.build)."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise FileNotFoundError("%s is missing: run `python -m avrecode_b200.build` (there is no fallback path)" % path)
    lib = C.CDLL(path)
    u8p = C.POINTER(C.c_uint8)
    lib.avr_strerror.restype = C.c_char_p
    lib.avr_strerror.argtypes = [C.c_int]
    lib.avr_abi_version.restype = C.c_int
    lib.avr_create.argtypes = [C.POINTER(Config), C.POINTER(C.c_void_p)]
    lib.avr_destroy.argtypes = [C.c_void_p]
    lib.avr_destroy.restype = None
    batch = [C.c_void_p, C.POINTER(SliceDesc), C.c_int, C.c_void_p, C.c_uint64, C.POINTER(SliceResult),
             C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
    lib.avr_compress.argtypes = batch
    lib.avr_decompress.argtypes = batch
    lib.avr_stage.argtypes = [C.c_void_p, C.POINTER(SliceDesc), C.c_int, C.c_void_p, C.c_uint64]
    lib.avr_launch.argtypes = [C.c_void_p, C.c_int]
    lib.avr_sync.argtypes = [C.c_void_p]
    lib.avr_collect.argtypes = [C.c_void_p, C.POINTER(SliceResult), C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
    lib.avr_last_kernel_ms.argtypes = [C.c_void_p, C.POINTER(C.c_float)]
    lib.avr_launch_count.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
    lib.avr_flush_l2.argtypes = [C.c_void_p]
    lib.avr_index_h264.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(C.POINTER(IndexEntry)), C.POINTER(C.c_int)]
    many = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.c_int]
    lib.avr_files_compress.argtypes = many + [C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
    lib.avr_files_decompress.argtypes = many + [C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
    lib.avr_file_compress.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
    lib.avr_file_decompress.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
    lib.avr_free.argtypes = [C.c_void_p]
    lib.avr_free.restype = None
    lib.avr_probe_name.argtypes = [C.c_int]
    lib.avr_probe_name.restype = C.c_char_p
    lib.avr_probe.argtypes = [C.c_int, C.c_int, C.c_uint32, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    _lib = lib
    return lib


def _strerror(code):
    try:
        return load_library().avr_strerror(code).decode()
    except Exception:  # pragma: no cover
        return "error %d" % code


def _check(rc, what):
    if rc == _abi.AVR_ERR_NO_DEVICE:
        raise NoDeviceError(rc, what)
    if rc != 0:
        raise AvrError(rc, what)


def _string_at(ptr, n):
    """ctypes.string_at for any size (it takes a C int: outputs of a corpus batch pass 2 GiB)."""
    if not n:
        return b""
    addr = ptr.value if hasattr(ptr, "value") else ptr
    if n < (1 << 31):
        return C.string_at(addr, n)
    return bytes((C.c_char * n).from_address(addr))


def _buf(data):
    """-> (address, length, keepalive) for bytes / bytearray / numpy / torch (CPU) buffers."""
    if isinstance(data, bytes):      # the bytes object's own storage: no copy (the library only reads its inputs)
        return C.cast(C.c_char_p(data), C.c_void_p).value or 0, len(data), data
    if isinstance(data, bytearray):
        arr = (C.c_uint8 * len(data)).from_buffer(data)
        return C.addressof(arr), len(data), arr
    if hasattr(data, "data_ptr"):  # torch tensor (pinned host memory for the e2e path)
        return data.data_ptr(), data.numel() * data.element_size(), data
    if hasattr(data, "ctypes"):  # numpy
        return data.ctypes.data, data.nbytes, data
    raise TypeError("unsupported buffer type %r" % type(data))


def slice_streams(desc):
    """n_streams the file calls would give this slice in MODE_SLICE in a small batch (0 = ordinary slice, 2 = wide slice)."""
    lib = load_library()
    lib.avr_slice_streams.argtypes = [C.POINTER(SliceDesc)]
    lib.avr_slice_streams.restype = C.c_int
    return int(lib.avr_slice_streams(C.byref(desc)))


def probes(device=0, iters=20000):
    """Single-lane microbenchmarks of the per-bin primitives: [(name, cycles per bin, bins)] measured on `device`."""
    lib = load_library()
    out = []
    which = 0
    while True:
        name = lib.avr_probe_name(which)
        if not name:
            break
        cyc, bins, chk = C.c_uint64(), C.c_uint64(), C.c_uint64()
        _check(lib.avr_probe(device, which, iters, C.byref(cyc), C.byref(bins), C.byref(chk)), "avr_probe")
        out.append((name.decode(), cyc.value / max(1, bins.value), bins.value))
        which += 1
    return out


def index_h264(data):
    """Slice index of an Annex-B H.264 stream (host only, no GPU): list of IndexEntry."""
    lib = load_library()
    addr, n, keep = _buf(data)
    ents = C.POINTER(IndexEntry)()
    cnt = C.c_int()
    _check(lib.avr_index_h264(addr, n, C.byref(ents), C.byref(cnt)), "avr_index_h264")
    out = []
    for i in range(cnt.value):
        e = IndexEntry()
        C.memmove(C.byref(e), C.byref(ents[i]), C.sizeof(IndexEntry))
        out.append(e)
    if cnt.value:
        lib.avr_free(C.cast(ents, C.c_void_p))
    return out


PRIORS_PATH = os.path.join(_HERE, "priors_v1.bin")


def load_priors(source="default"):
    """A prior table as a ctypes uint16 array: "default" = the table that ships (priors_v1.bin), a path, or bytes."""
    if source == "default":
        source = PRIORS_PATH
    raw = open(source, "rb").read() if isinstance(source, str) else bytes(source)
    if len(raw) != 2 * _abi.AVR_PRIOR_ENTRIES:
        raise ValueError("a prior table is %d bytes" % (2 * _abi.AVR_PRIOR_ENTRIES))
    return (C.c_uint16 * _abi.AVR_PRIOR_ENTRIES).from_buffer_copy(raw)


def priors_digest_of(source="default"):
    lib = load_library()
    lib.avr_priors_digest_of.restype = C.c_uint32
    lib.avr_priors_digest_of.argtypes = [C.POINTER(C.c_uint16), C.c_uint32]
    return lib.avr_priors_digest_of(load_priors(source), _abi.AVR_PRIOR_ENTRIES)


class Context:
    """One CUDA device.  Raises NoDeviceError when there is none: no CPU fallback exists."""

    def __init__(self, device=0, blocks_per_sm=0, wide_stream_bytes=-1, resident=-1, wide_batch_limit=-1, wide_max_streams=0):
        self._lib = load_library()
        cfg = Config()
        cfg.abi_version = self._lib.avr_abi_version()
        cfg.device, cfg.blocks_per_sm, cfg.wide_stream_bytes = device, blocks_per_sm, wide_stream_bytes
        cfg.resident, cfg.wide_batch_limit = resident, wide_batch_limit
        cfg.wide_max_streams = wide_max_streams
        h = C.c_void_p()
        _check(self._lib.avr_create(C.byref(cfg), C.byref(h)), "avr_create")
        self._h = h
        self._keep = None

    def set_priors(self, source="default"):
        """Fresh models of chunked-mode slices start from this table (avr_set_priors); None clears."""
        self._lib.avr_set_priors.argtypes = [C.c_void_p, C.POINTER(C.c_uint16), C.c_uint32]
        if source is None:
            _check(self._lib.avr_set_priors(self._h, None, 0), "avr_set_priors")
        else:
            _check(self._lib.avr_set_priors(self._h, load_priors(source), _abi.AVR_PRIOR_ENTRIES), "avr_set_priors")
        return self

    @property
    def priors_digest(self):
        self._lib.avr_priors_digest.restype = C.c_uint32
        self._lib.avr_priors_digest.argtypes = [C.c_void_p]
        return self._lib.avr_priors_digest(self._h)

    def close(self):
        if getattr(self, "_h", None):
            self._lib.avr_destroy(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def slice_streams(self, desc, n_batch_slices):
        """n_streams the file calls of THIS context give the slice in a batch of n_batch_slices CABAC slices."""
        self._lib.avr_ctx_slice_streams.argtypes = [C.c_void_p, C.POINTER(SliceDesc), C.c_int]
        return int(self._lib.avr_ctx_slice_streams(self._h, C.byref(desc), n_batch_slices))

    # ---- slice batches -------------------------------------------------------------------------
    def _batch(self, fn, what, descs, data):
        n = len(descs)
        arr = (SliceDesc * n)(*descs)
        res = (SliceResult * n)()
        addr, nbytes, keep = _buf(data)
        out = C.c_void_p()
        nout = C.c_uint64()
        _check(fn(self._h, arr, n, addr, nbytes, res, C.byref(out), C.byref(nout)), what)
        blob = _string_at(out, nout.value)
        return list(res), blob

    def compress(self, descs, data):
        """-> (results, blob); Block.cabac of slice i = blob[r.out_off : r.out_off + r.out_size]."""
        return self._batch(self._lib.avr_compress, "avr_compress", descs, data)

    def decompress(self, descs, data):
        return self._batch(self._lib.avr_decompress, "avr_decompress", descs, data)

    # ---- staged pipeline (inputs resident in HBM) ----------------------------------------------
    def stage(self, descs, data):
        n = len(descs)
        arr = (SliceDesc * n)(*descs)
        addr, nbytes, keep = _buf(data)
        self._keep = (arr, keep)
        self._n = n
        _check(self._lib.avr_stage(self._h, arr, n, addr, nbytes), "avr_stage")

    def launch(self, direction):
        _check(self._lib.avr_launch(self._h, direction), "avr_launch")

    def sync(self):
        _check(self._lib.avr_sync(self._h), "avr_sync")

    def collect(self, copy=True):
        res = (SliceResult * self._n)()
        out = C.c_void_p()
        nout = C.c_uint64()
        _check(self._lib.avr_collect(self._h, res, C.byref(out), C.byref(nout)), "avr_collect")
        blob = _string_at(out, nout.value) if copy else b""
        return list(res), blob

    def collect_hist(self):
        """Per-class bin counts of the last launch: one list of 8 per slice (see avr_collect_hist)."""
        self._lib.avr_collect_hist.argtypes = [C.c_void_p, C.POINTER(C.c_uint32), C.c_int]
        buf = (C.c_uint32 * (8 * max(1, self._n)))()
        _check(self._lib.avr_collect_hist(self._h, buf, self._n), "avr_collect_hist")
        return [list(buf[8 * i:8 * i + 8]) for i in range(self._n)]

    def last_kernel_ms(self):
        ms = C.c_float()
        _check(self._lib.avr_last_kernel_ms(self._h, C.byref(ms)), "avr_last_kernel_ms")
        return ms.value

    def launch_count(self):
        n = C.c_uint64()
        _check(self._lib.avr_launch_count(self._h, C.byref(n)), "avr_launch_count")
        return n.value

    def flush_l2(self):
        _check(self._lib.avr_flush_l2(self._h), "avr_flush_l2")

    # ---- whole files (recode compress / decompress) ---------------------------------------------
    def file_compress(self, data, mode=MODE_SLICE):
        addr, n, keep = _buf(data)
        out = C.c_void_p()
        nout = C.c_uint64()
        _check(self._lib.avr_file_compress(self._h, addr, n, mode, C.byref(out), C.byref(nout)), "avr_file_compress")
        blob = _string_at(out, nout.value)
        self._lib.avr_free(out)
        return blob

    def file_decompress(self, rec):
        addr, n, keep = _buf(rec)
        out = C.c_void_p()
        nout = C.c_uint64()
        _check(self._lib.avr_file_decompress(self._h, addr, n, C.byref(out), C.byref(nout)), "avr_file_decompress")
        blob = _string_at(out, nout.value)
        self._lib.avr_free(out)
        return blob

    # ---- corpora: many independent files in one device batch (BASELINE config 5) ------------------
    def _many(self, fn, name, blobs, *extra):
        n = len(blobs)
        bufs = [_buf(b) for b in blobs]
        ptrs = (C.c_void_p * max(1, n))(*[b[0] for b in bufs])
        sizes = (C.c_uint64 * max(1, n))(*[b[1] for b in bufs])
        outs = (C.c_void_p * max(1, n))()
        nouts = (C.c_uint64 * max(1, n))()
        _check(fn(self._h, ptrs, sizes, n, *extra, outs, nouts), name)
        res = []
        for i in range(n):
            res.append(_string_at(outs[i], nouts[i]))
            self._lib.avr_free(outs[i])
        return res

    def files_compress(self, files, mode=MODE_SLICE):
        return self._many(self._lib.avr_files_compress, "avr_files_compress", files, C.c_int(mode))

    def files_decompress(self, recs):
        return self._many(self._lib.avr_files_decompress, "avr_files_decompress", recs)


class MultiContext:
    """Several CUDA devices of one node behind one handle (avr_multi_*): the slices (MODE_SLICE) or files (MODE_FILE) of
    a call are dealt out over the devices, longest first; the containers are those one device writes."""

    def __init__(self, devices, **config):
        self._lib = load_library()
        lib = self._lib
        lib.avr_multi_create.argtypes = [C.POINTER(Config), C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_void_p)]
        lib.avr_multi_destroy.argtypes = [C.c_void_p]
        lib.avr_multi_destroy.restype = None
        lib.avr_multi_devices.argtypes = [C.c_void_p]
        many = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.c_int]
        lib.avr_multi_files_compress.argtypes = many + [C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
        lib.avr_multi_files_decompress.argtypes = many + [C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
        cfg = Config()
        lib.avr_config_init(C.byref(cfg))
        for k, v in config.items():
            if not hasattr(cfg, k) or k in ("abi_version", "device", "reserved"):
                raise TypeError("MultiContext: unknown option %r" % k)
            setattr(cfg, k, v)
        devs = (C.c_int * max(1, len(devices)))(*devices)
        h = C.c_void_p()
        _check(lib.avr_multi_create(C.byref(cfg), devs, len(devices), C.byref(h)), "avr_multi_create")
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            self._lib.avr_multi_destroy(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def set_priors(self, source="default"):
        self._lib.avr_multi_set_priors.argtypes = [C.c_void_p, C.POINTER(C.c_uint16), C.c_uint32]
        if source is None:
            _check(self._lib.avr_multi_set_priors(self._h, None, 0), "avr_multi_set_priors")
        else:
            _check(self._lib.avr_multi_set_priors(self._h, load_priors(source), _abi.AVR_PRIOR_ENTRIES), "avr_multi_set_priors")
        return self

    @property
    def n_devices(self):
        return self._lib.avr_multi_devices(self._h)

    _many = Context._many

    def files_compress(self, files, mode=MODE_SLICE):
        return self._many(self._lib.avr_multi_files_compress, "avr_multi_files_compress", files, C.c_int(mode))

    def files_decompress(self, recs):
        return self._many(self._lib.avr_multi_files_decompress, "avr_multi_files_decompress", recs)
