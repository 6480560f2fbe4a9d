"""ctypes binding of libpenroz_b200.so.

Prototypes are taken from include/penroz_b200.h itself (one source of truth for the C ABI); every call checks the
return code and raises RuntimeError with penroz_last_error(). There is NO fallback: if the library is missing or the
device is not sm_100a, the product path fails loudly.
"""
import ctypes
import os
import re
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpenroz_b200.so")
HEADER_PATH = os.path.normpath(os.path.join(_HERE, "..", "..", "include", "penroz_b200.h"))

F32, BF16 = 0, 1
EPI_STORE, EPI_GELU, EPI_RESID_F32, EPI_RESID_BF16, EPI_ACCUM_F32, EPI_DGELU, EPI_F32, EPI_ROWDOT = range(8)
FLAG_TANH, FLAG_ACCUM, FLAG_BN128, FLAG_BN256 = 1, 2, 4, 8

_SCALARS = {
    "int": ctypes.c_int,
    "long long": ctypes.c_longlong,
    "unsigned long long": ctypes.c_ulonglong,
    "size_t": ctypes.c_size_t,
    "float": ctypes.c_float,
    "double": ctypes.c_double,
}


class PenrozError(RuntimeError):
    pass


def parse_header(path: str = HEADER_PATH) -> dict:
    """Return {function name: (restype, [argtypes])} for every prototype in the header."""
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    text = "\n".join(l for l in text.splitlines() if not l.strip().startswith("#"))
    text = text.replace('extern "C" {', " ")
    protos = {}
    for m in re.finditer(r"([A-Za-z_][\w\s\*]*?)\b(penroz_\w+)\s*\(([^)]*)\)\s*;", text):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        if "*" in ret:
            restype = ctypes.c_char_p if "char" in ret else ctypes.c_void_p
        else:
            restype = _SCALARS[ret.replace("const", "").strip()]
        argtypes = []
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                if "*" in a:
                    argtypes.append(ctypes.c_void_p)
                else:
                    ty = re.sub(r"\b\w+$", "", a.replace("const", "")).strip()
                    argtypes.append(_SCALARS[ty])
        protos[name] = (restype, argtypes)
    return protos


class _Lib:
    def __init__(self):
        if not os.path.exists(LIB_PATH):
            raise PenrozError(
                f"{LIB_PATH} is missing: build it with `python -m penroz.build` (or __graft_entry__.build()). "
                "penroz_b200 has no CPU or PyTorch fallback.")
        self._cdll = ctypes.CDLL(LIB_PATH)
        self.protos = parse_header()
        self._status_fns = set()
        for name, (restype, argtypes) in self.protos.items():
            fn = getattr(self._cdll, name)  # AttributeError if the .so does not export a declared symbol
            fn.restype = restype
            fn.argtypes = argtypes
            if restype is ctypes.c_int and name not in ("penroz_abi_version", "penroz_kv_page_size"):
                self._status_fns.add(name)

    def __getattr__(self, name):
        full = "penroz_" + name
        fn = getattr(self._cdll, full)
        if full in self._status_fns:
            def checked(*args, _fn=fn, _name=full):
                rc = _fn(*args)
                if rc != 0:
                    msg = self._cdll.penroz_last_error()
                    raise PenrozError(f"{_name} failed ({rc}): {msg.decode() if msg else ''}")
                return rc
            setattr(self, name, checked)
            return checked
        setattr(self, name, fn)
        return fn


_lib = None
_lock = threading.Lock()


def lib() -> _Lib:
    """The loaded library (loads on first use; raises PenrozError when it cannot)."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                _lib = _Lib()
    return _lib


def require_device():
    """Raise unless the current CUDA device can run the sm_100a kernels."""
    import torch
    if not torch.cuda.is_available():
        raise PenrozError("penroz_b200 needs a CUDA device (B200, sm_100a); none is visible and there is no CPU path")
    lib().query(None, None, None)


def dt(t) -> int:
    """dtype code of a torch tensor."""
    import torch
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.bfloat16:
        return BF16
    raise PenrozError(f"unsupported dtype {t.dtype}: penroz_b200 computes in float32 or bfloat16")


def ptr(t):
    return None if t is None else t.data_ptr()


_raw_stream = None


def stream():
    """cudaStream_t of torch's current stream on the current device (what every entry point launches on).
    torch.cuda.current_stream() builds a Stream object per call (~14 us, more than some of the kernels it is asked
    about); the raw getter behind it costs ~1 us."""
    global _raw_stream
    import torch
    if _raw_stream is None:
        _raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", False)
    if _raw_stream:
        return _raw_stream(torch.cuda.current_device())
    return torch.cuda.current_stream().cuda_stream
