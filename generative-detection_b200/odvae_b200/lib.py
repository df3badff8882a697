"""ctypes binding of libodvae_b200.so (the C ABI declared in include/odvae_b200.h).

The prototypes are parsed from the header itself, so the Python side cannot drift from the C ABI.
There is no fallback: if the shared library is missing, or the device is not sm_100, calls raise.
"""
import ctypes
import os
import re

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(os.path.dirname(_HERE))
HEADER = os.path.join(REPO_ROOT, "include", "odvae_b200.h")
# ODVAE_LIB_PATH: A/B runs of two builds of the same C ABI on one box (tools/gpu_ab_lib.sh)
LIB_PATH = os.environ.get("ODVAE_LIB_PATH") or os.path.join(_HERE, "libodvae_b200.so")

F32, BF16, F16 = 0, 1, 2
CONV_3x3_S1, CONV_1x1, CONV_3x3_S2 = 0, 1, 2
IMPL_SIMT, IMPL_TC = 0, 1

_CTYPE = {
    "int": ctypes.c_int, "float": ctypes.c_float, "int64_t": ctypes.c_int64, "size_t": ctypes.c_size_t,
}


def parse_header(path=HEADER):
    """-> {name: (restype, [(argtype, argname), ...])} for every prototype in the header."""
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = re.sub(r"//[^\n]*", "", text)
    protos = {}
    for m in re.finditer(r"(const char\*|int64_t|int|size_t)\s+(odvae_\w+)\s*\(([^)]*)\)\s*;", text):
        ret, name, args = m.group(1), m.group(2), m.group(3).strip()
        arglist = []
        if args and args != "void":
            for a in args.split(","):
                a = " ".join(a.split())
                t, n = a.rsplit(" ", 1)
                if n.startswith("*"):
                    t, n = t + "*", n[1:]
                arglist.append((t.replace(" *", "*"), n))
        protos[name] = (ret, arglist)
    return protos


def _ctype_of(t):
    if t.endswith("*"):
        return ctypes.c_void_p
    return _CTYPE[t.replace("const ", "")]


class _Lib:
    def __init__(self):
        self._dll = None
        self.protos = parse_header()

    def load(self):
        if self._dll is None:
            if not os.path.exists(LIB_PATH):
                raise RuntimeError(
                    "libodvae_b200.so is not built (%s). Run `python -c 'import __graft_entry__ as g; g.build()'` "
                    "or `make -C generative-detection_b200/csrc`. There is no CPU/PyTorch fallback." % LIB_PATH)
            dll = ctypes.CDLL(LIB_PATH)
            for name, (ret, args) in self.protos.items():
                fn = getattr(dll, name)  # raises AttributeError if the symbol is not exported
                fn.restype = {"int": ctypes.c_int, "int64_t": ctypes.c_int64, "size_t": ctypes.c_size_t, "const char*": ctypes.c_char_p}[ret]
                fn.argtypes = [_ctype_of(t) for t, _ in args]
            self._dll = dll
        return self._dll

    def status_string(self, st):
        dll = self.load()
        return "%s [%s]" % (dll.odvae_status_string(st).decode(), dll.odvae_last_cuda_error().decode())


LIB = _Lib()
_RESOLVED = {}  # name -> (ctypes function, #args, trailing stream?, status-checked?, full name)


def _ptr(x):
    if x is None:
        return None
    if isinstance(x, torch.Tensor):
        return x.data_ptr()
    return x


def stream_handle(device_index=None):
    """Raw cudaStream_t of torch's current stream. torch.cuda.current_stream() builds a Stream object through three
    layers of Python (13 us; 850 calls per train step = 11 ms of host time, tools/host_profile.py); the private C
    getters are what it ends up calling."""
    try:
        idx = torch._C._cuda_getDevice() if device_index is None else device_index
        return torch._C._cuda_getCurrentRawStream(idx)
    except AttributeError:  # other torch builds: the public (slow) path
        return torch.cuda.current_stream(device_index).cuda_stream


def call(name, *args):
    """Call `odvae_<name>`; the trailing `void* stream` argument is supplied automatically when the
    caller passes one argument fewer than the prototype. Non-zero status raises RuntimeError."""
    ent = _RESOLVED.get(name)
    if ent is None:
        dll = LIB.load()
        full = name if name.startswith("odvae_") else "odvae_" + name
        ret, proto = LIB.protos[full]
        ent = (getattr(dll, full), len(proto), bool(proto) and proto[-1][1] == "stream",
               ret == "int" and not full.endswith(("_supported", "_version")), full)
        _RESOLVED[name] = ent
    fn, nproto, has_stream, checked, full = ent
    conv = [a.data_ptr() if isinstance(a, torch.Tensor) else a for a in args]
    if len(conv) == nproto - 1 and has_stream:
        conv.append(stream_handle())
    if len(conv) != nproto:
        raise TypeError("%s expects %d arguments, got %d" % (full, nproto, len(conv)))
    out = fn(*conv)
    if checked:
        if out != 0:
            raise RuntimeError("%s failed: %s" % (full, LIB.status_string(out)))
        return None
    return out


def require_device():
    if not torch.cuda.is_available():
        raise RuntimeError("odvae_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    if not LIB.load().odvae_device_supported():
        raise RuntimeError("odvae_b200 kernels are built for sm_100a only; current device is unsupported")
