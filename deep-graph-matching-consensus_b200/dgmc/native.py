"""ctypes binding of libdgmc_b200.so — the C ABI declared in include/dgmc_b200.h.

The prototypes are read from the header itself, so the Python side cannot drift from the ABI.
There is NO fallback: if the shared library is missing the import fails, and every call checks
that its tensors live on a CUDA device.
"""
import ctypes
import os
import re

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
PKG_ROOT = os.path.dirname(_HERE)
REPO_ROOT = os.path.dirname(PKG_ROOT)
LIB_PATH = os.environ.get('DGMC_B200_LIB', os.path.join(PKG_ROOT, 'lib', 'libdgmc_b200.so'))
HEADER_PATH = os.path.join(REPO_ROOT, 'include', 'dgmc_b200.h')

_CTYPES = {
    'int': ctypes.c_int, 'int64_t': ctypes.c_int64, 'size_t': ctypes.c_size_t,
    'float': ctypes.c_float, 'dgmc_stream_t': ctypes.c_void_p, 'uint64_t': ctypes.c_uint64,
}


def _ctype(decl):
    decl = decl.replace('const', ' ').strip()
    if '*' in decl:
        return ctypes.c_void_p
    return _CTYPES[decl.split()[0]]


def parse_header(path=HEADER_PATH):
    """-> {name: (restype, [argtypes])} for every prototype in the header."""
    text = open(path).read()
    text = re.sub(r'/\*.*?\*/', ' ', text, flags=re.S)
    text = re.sub(r'^\s*#.*$', ' ', text, flags=re.M)
    protos = {}
    for m in re.finditer(r'([A-Za-z_][\w\s\*]*?)\b(dgmc_\w+)\s*\(([^)]*)\)\s*;', text):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        if ret.startswith('typedef'):
            continue
        argtypes = []
        if args and args != 'void':
            for a in args.split(','):
                a = a.strip()
                a = re.sub(r'\b\w+$', '', a).strip() if not a.endswith('*') else a
                argtypes.append(_ctype(a))
        restype = ctypes.c_char_p if 'char' in ret else _ctype(ret)
        protos[name] = (restype, argtypes)
    return protos


def _try_build():
    """The shared library is built in-tree by __graft_entry__.build(); if it is absent (fresh
    checkout) and nvcc is available, build it now.  This compiles the extension — it is not a
    fallback path: without the library the import fails."""
    import importlib.util
    spec = importlib.util.spec_from_file_location('dgmc_b200_build', os.path.join(PKG_ROOT, 'build.py'))
    mod = importlib.util.module_from_spec(spec)
    try:
        spec.loader.exec_module(mod)
        mod.build()
    except Exception as exc:  # pragma: no cover
        raise ImportError('dgmc (B200): could not build %s: %s' % (LIB_PATH, exc))


if not os.path.exists(LIB_PATH) and 'DGMC_B200_LIB' not in os.environ:
    _try_build()

if not os.path.exists(LIB_PATH):
    raise ImportError(
        'dgmc (B200): %s not found. Build it with `python -c "import __graft_entry__ as g; '
        'g.build()"` from the repository root. There is no CPU or PyTorch fallback.' % LIB_PATH)

lib = ctypes.CDLL(LIB_PATH)
PROTOTYPES = parse_header()
for _name, (_res, _args) in PROTOTYPES.items():
    _fn = getattr(lib, _name)  # AttributeError here == header/library mismatch: fail loudly
    _fn.restype, _fn.argtypes = _res, _args


class DgmcNativeError(RuntimeError):
    pass


def last_error():
    msg = lib.dgmc_last_error()
    return msg.decode() if msg else ''


class RowStrided(object):
    """Wraps a 2-D view whose rows are strided (e.g. a column slice of a wider buffer) for entry
    points that take an explicit leading dimension."""
    def __init__(self, t):
        assert t.dim() == 2 and t.stride(1) == 1
        self.t = t


def _ptr(t):
    if t is None:
        return None
    if t.__class__ is RowStrided:
        t = t.t
        if not t.is_cuda:
            raise DgmcNativeError('dgmc (B200): tensor on %s; there is no CPU fallback' % t.device)
        return t.data_ptr()
    if not t.is_cuda:
        raise DgmcNativeError('dgmc (B200): tensor on %s; the kernels need CUDA tensors and there '
                              'is no CPU fallback' % t.device)
    if not t.is_contiguous():
        raise DgmcNativeError('dgmc (B200): non-contiguous tensor passed to a kernel')
    return t.data_ptr()


_raw_stream = getattr(torch._C, '_cuda_getCurrentRawStream', None)
_raw_device = getattr(torch._C, '_cuda_getDevice', None)


def stream_ptr():
    """The current CUDA stream of the current device as a raw cudaStream_t (fast path through torch._C:
    this is called once per kernel launch)."""
    if _raw_stream is not None and _raw_device is not None and torch.cuda.is_initialized():
        return _raw_stream(_raw_device())
    if not torch.cuda.is_available():
        raise DgmcNativeError('dgmc (B200): no CUDA device; the kernels have no CPU fallback')
    return torch.cuda.current_stream().cuda_stream


LAUNCHES = [0]  # number of native entry points called (bench.py reports it as gpu_launches)
_Tensor = torch.Tensor


def call(name, *args):
    """Call an `int dgmc_*` entry point; tensors become device pointers; raises on failure.  The
    kernels run on the CURRENT device and stream: tensors that live on another GPU are an error (wrap
    the call in `torch.cuda.device(tensor.device)`), not a launch on the wrong device."""
    fn = getattr(lib, name)
    conv = []
    first = None
    for a in args:
        if a is None:
            conv.append(None)
        elif isinstance(a, _Tensor):
            if first is None:
                first = a
            conv.append(_ptr(a))
        elif a.__class__ is RowStrided:
            if first is None:
                first = a.t
            conv.append(_ptr(a))
        else:
            conv.append(a)
    if first is not None and first.is_cuda:
        # torch ops already require all tensors of a call on one device: check the first one
        cur = _raw_device() if _raw_device is not None else torch.cuda.current_device()
        if first.device.index != cur:
            raise DgmcNativeError('%s: tensor on %s but the current device is cuda:%d; use '
                                  'torch.cuda.set_device / torch.cuda.device(...)' % (name, first.device, cur))
    LAUNCHES[0] += 1
    rc = fn(*conv)
    if rc != 0:
        raise DgmcNativeError('%s failed (%d): %s' % (name, rc, last_error()))


def query(name, *args):
    """Call a size_t-returning helper (`*_workspace`)."""
    return getattr(lib, name)(*args)
