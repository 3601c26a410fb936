"""ctypes mirror of include/avrecode_b200.h (struct layouts + constants).

Shared by the product binding (avrecode_b200/api.py) and by the test-side oracle binding
(tests/oracle_binding.py); contains no algorithm."""
import ctypes as C

AVR_SLICE_P, AVR_SLICE_B, AVR_SLICE_I = 0, 1, 2
AVR_SLICE_TRANSFORM_8X8 = 0x01
AVR_SLICE_DIRECT_8X8_INFERENCE = 0x02
AVR_SLICE_NEW_PICTURE = 0x04
AVR_SLICE_MBAFF = 0x08
AVR_SLICE_FIELD = 0x10
AVR_SLICE_NEW_MODEL = 0x20
AVR_STREAMS_MASK = 0x0F      # avr_slice_desc.n_streams: bits 0-3 streams, 4-5 header zeros, 7 escaped
AVR_ESC_ZEROS_SHIFT = 4
AVR_SLICE_ESCAPED = 0x80
AVR_SLICE_STOP_BYTE = 0x40
AVR_SLICE_SKIP = 0x80

AVR_SLICE_OK, AVR_SLICE_UNSUPPORTED, AVR_SLICE_CORRUPT, AVR_SLICE_OVERFLOW = 0, 1, 2, 3
AVR_MODE_FILE, AVR_MODE_SLICE = 0, 1
AVR_MAX_STREAMS = 8          # include/avrecode_b200.h

AVR_OK = 0
AVR_ERR_NO_DEVICE, AVR_ERR_CUDA, AVR_ERR_ARG, AVR_ERR_NOMEM, AVR_ERR_FORMAT, AVR_ERR_STATE = -1, -2, -3, -4, -5, -6
AVR_ERR_PRIORS = -7
AVR_SLICE_PRIORS = 0x40      # avr_slice_desc.n_streams bit 6
AVR_PRIOR_ENTRIES = 47892
AVR_ERR_LIMIT = -8           # avr_recode_noalloc only
AVR_NOALLOC_MAX_SLICES, AVR_NOALLOC_SCRATCH_BYTES = 4096, 512 * 1024


class SliceDesc(C.Structure):
    _fields_ = [
        ("in_off", C.c_uint64),
        ("in_size", C.c_uint32),
        ("payload_size", C.c_uint32),
        ("first_mb", C.c_uint32),
        ("mb_width", C.c_uint16),
        ("mb_height", C.c_uint16),
        ("frame_num", C.c_int32),
        ("slice_type", C.c_uint8),
        ("cabac_init_idc", C.c_uint8),
        ("slice_qp", C.c_int8),
        ("num_ref_idx_l0", C.c_uint8),
        ("num_ref_idx_l1", C.c_uint8),
        ("chroma_format_idc", C.c_uint8),
        ("flags", C.c_uint8),
        ("n_streams", C.c_uint8),
    ]


class SliceResult(C.Structure):
    _fields_ = [
        ("out_off", C.c_uint64),
        ("out_size", C.c_uint32),
        ("status", C.c_int32),
        ("n_mbs", C.c_uint32),
        ("reserved", C.c_uint32),
        ("n_bins", C.c_uint64),
    ]


class Config(C.Structure):
    _fields_ = [("abi_version", C.c_int32), ("device", C.c_int32), ("blocks_per_sm", C.c_int32), ("wide_stream_bytes", C.c_int32),
                ("resident", C.c_int32), ("wide_batch_limit", C.c_int32), ("wide_max_streams", C.c_int32), ("reserved", C.c_int32 * 1)]


class IndexEntry(C.Structure):
    _fields_ = [("desc", SliceDesc), ("recodable", C.c_int32), ("nal_type", C.c_int32)]


assert C.sizeof(SliceDesc) == 40, C.sizeof(SliceDesc)
assert C.sizeof(SliceResult) == 32
assert C.sizeof(IndexEntry) == 48


def copy_desc(d):
    out = SliceDesc()
    C.memmove(C.byref(out), C.byref(d), C.sizeof(SliceDesc))
    return out
