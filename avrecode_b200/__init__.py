"""avrecode_b200 -- B200-native implementation of avrecode's CABAC decode / model / re-encode path.

Python is only a thin ctypes binding over libavrecode_b200.so (include/avrecode_b200.h); all
per-bin work runs in sm_100a CUDA kernels.  There is no CPU fallback: creating a Context without a
CUDA device raises NoDeviceError."""
from .api import (Context, MultiContext, NoDeviceError, AvrError, index_h264, slice_streams, probes, lib_path, load_library,
                  MODE_FILE, MODE_SLICE, MAX_STREAMS, load_priors, priors_digest_of, PRIORS_PATH)
from . import _abi

__all__ = ["Context", "MultiContext", "NoDeviceError", "AvrError", "index_h264", "slice_streams", "probes", "lib_path", "load_library",
           "MODE_FILE", "MODE_SLICE", "MAX_STREAMS", "load_priors", "priors_digest_of", "PRIORS_PATH", "_abi"]
