"""ctypes loader for libvaeclust.so (the C-ABI declared in include/vaeclust.h).

The prototypes are parsed from the header itself so that the Python side can never drift from the C side, and so the
CPU test-suite can check that the library exports every declared symbol.  There is NO fallback: if the shared
library is missing, importing an op raises with the build command.
"""
from __future__ import annotations

import ctypes
import os
import re
from pathlib import Path
from typing import Dict, List, Tuple

_PKG = Path(__file__).resolve().parent
_ROOT = _PKG.parent.parent
HEADER = _ROOT / "include" / "vaeclust.h"
LIB_PATH = Path(os.environ.get("VAECLUST_LIB", _PKG / "libvaeclust.so"))

VC_OK = 0
VC_LOSS_MEAN_MEAN = 0
VC_LOSS_SUM_MEAN = 1

_SCALARS = {
    "int": ctypes.c_int,
    "int64_t": ctypes.c_int64,
    "size_t": ctypes.c_size_t,
    "float": ctypes.c_float,
    "double": ctypes.c_double,
}


def parse_header(path: Path = HEADER) -> Dict[str, Tuple[object, List[object]]]:
    """{name: (restype, [argtypes])} for every function prototype in the header."""
    text = path.read_text()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    text = re.sub(r"//[^\n]*", " ", text)
    text = re.sub(r"typedef\s+struct\s+\w+\s*\{.*?\}\s*\w+\s*;", " ", text, flags=re.S)
    protos: Dict[str, Tuple[object, List[object]]] = {}
    for m in re.finditer(r"(const\s+char\s*\*|int|size_t)\s+(vc_\w+)\s*\(([^)]*)\)\s*;", text):
        ret, name, args = m.group(1), m.group(2), m.group(3).strip()
        restype = ctypes.c_char_p if "char" in ret else _SCALARS[ret]
        argtypes: List[object] = []
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                if "*" in a:
                    argtypes.append(ctypes.c_void_p)
                else:
                    argtypes.append(_SCALARS[a.replace("const ", "").split()[0]])
        protos[name] = (restype, argtypes)
    return protos


class VaeclustError(RuntimeError):
    pass


# kernels launched per entry point (for bench.py's gpu_launches claim); anything absent launches one kernel
_LAUNCHES = {
    "vc_recon_loss_fwd_bwd": 2, "vc_conv2d_s2k3_wgrad": 2, "vc_convT2d_s2k4_wgrad": 2, "vc_convT2d_s2k4_fwd": 4,
    "vc_kmeans_assign_update": 2, "vc_cluster_sums": 2, "vc_kmeans_inertia": 2, "vc_dbi_intra": 2, "vc_col_stats": 2,
    "vc_silhouette_finalize": 2, "vc_bias_grad": 2, "vc_relu_bwd_bias": 2, "vc_dense_op": 2,
    "vc_relu_bwd_bias_nhwc": 2, "vc_conv_dgrad_nhwc": 2, "vc_mlpvae_step": 4, "vc_kmeans_lloyd_estep": 2,
    "vc_conv_nhwc": 2, "vc_conv_nhwc_wgrad": 2, "vc_nhwc_bias_grad": 2, "vc_nhwc_tap_wgrad": 2, "vc_conv1_wgrad_bias": 3, "vc_convT1_wgrad": 2, "vc_nhwc_conv1_wgrad_bias": 3, "vc_nhwc_convT1_dgrad_bias": 2,
}


class Profiler:
    """Brackets every C-ABI call with CUDA events on the launching stream (bench.py's per-kernel durations)."""

    def __init__(self):
        self.records = []  # (name, int-args signature, start event, end event)

    def run(self, name, args, fn):
        import torch

        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        rc = fn(*args)
        b.record()
        self.records.append((name, tuple(x for x in args[:-1] if isinstance(x, int) and not isinstance(x, bool)
                                         and abs(x) < (1 << 31)), a, b))
        return rc

    def summary(self):
        """{(name, signature): (count, total_ms)} -- call after torch.cuda.synchronize()."""
        out = {}
        for name, sig, a, b in self.records:
            c, t = out.get((name, sig), (0, 0.0))
            out[(name, sig)] = (c + 1, t + a.elapsed_time(b))
        return out


class _Lib:
    def __init__(self) -> None:
        if not LIB_PATH.is_file():
            raise VaeclustError(
                f"{LIB_PATH} not found: build it with `make -C {_PKG.parent}` (or python -c 'import "
                f"__graft_entry__ as g; g.build()'); vaeclust has no CPU or PyTorch fallback."
            )
        self.cdll = ctypes.CDLL(str(LIB_PATH))
        self.profiler = None      # set to a Profiler to time every entry point with CUDA events
        self.n_calls = 0          # C-ABI calls made
        self.n_launches = 0       # kernels those calls launched (table below)
        self.protos = parse_header()
        for name, (restype, argtypes) in self.protos.items():
            fn = getattr(self.cdll, name)
            fn.restype = restype
            fn.argtypes = argtypes

    def last_error(self) -> str:
        return (self.cdll.vc_last_error_string() or b"").decode()

    def call(self, name: str, *args):
        prof = self.profiler
        if prof is None:
            rc = getattr(self.cdll, name)(*args)
        else:
            rc = prof.run(name, args, getattr(self.cdll, name))
        self.n_calls += 1
        self.n_launches += _LAUNCHES.get(name, 1)
        if rc != VC_OK:
            raise VaeclustError(f"{name} failed ({rc}): {self.last_error()}")

    def size(self, name: str, *args) -> int:
        return int(getattr(self.cdll, name)(*args))


_lib = None


def lib() -> _Lib:
    global _lib
    if _lib is None:
        _lib = _Lib()
    return _lib


def ptr(t) -> int:
    """Device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def stream_ptr():
    import torch

    return torch.cuda.current_stream().cuda_stream
