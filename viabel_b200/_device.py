"""Device plumbing shared by the Python surface: torch owns memory and streams, nothing else."""
import numpy as np
import torch

from . import _lib

_DT = {_lib.F32: torch.float32, _lib.F64: torch.float64}
_NP = {_lib.F32: np.float32, _lib.F64: np.float64}


def dtype_code(dtype):
    """Accepts 'f64'/'f32', numpy or torch dtypes; returns the C-ABI code."""
    if dtype in ('f64', 'float64', np.float64, torch.float64, _lib.F64, None):
        return _lib.F64
    if dtype in ('f32', 'float32', np.float32, torch.float32, _lib.F32):
        return _lib.F32
    raise ValueError('dtype must be float64 or float32')


def torch_dtype(code):
    return _DT[code]


def device():
    if not torch.cuda.is_available():
        raise RuntimeError('viabel_b200 needs a CUDA device (sm_100a); there is no CPU fallback')
    return torch.device('cuda', torch.cuda.current_device())


def stream_ptr():
    return torch.cuda.current_stream().cuda_stream


def is_host(a):
    return not (isinstance(a, torch.Tensor) and a.is_cuda)


def to_dev(a, code=_lib.F64):
    """numpy / torch (host or device) -> contiguous CUDA tensor of the requested dtype, no copy if possible."""
    dt = _DT[code]
    if isinstance(a, torch.Tensor):
        t = a
    else:
        t = torch.from_numpy(np.ascontiguousarray(np.asarray(a, dtype=_NP[code])))
    return t.to(device=device(), dtype=dt, non_blocking=True).contiguous()


def like_input(t, template):
    """Return `t` (a CUDA tensor) as numpy if the caller handed us host data, else as is."""
    if is_host(template):
        return t.cpu().numpy()
    return t


def ptr(t):
    return 0 if t is None else t.data_ptr()


class Workspace:
    """Grow-only scratch buffers keyed by name (zero-filled on allocation)."""

    def __init__(self):
        self._bufs = {}

    def get(self, key, nbytes):
        nbytes = max(int(nbytes), 16)
        buf = self._bufs.get(key)
        if buf is None or buf.numel() < nbytes or buf.device != device():
            buf = torch.zeros(nbytes, dtype=torch.uint8, device=device())
            self._bufs[key] = buf
        return buf


WS = Workspace()


def all_gather_rows(t, world, group=None):
    """One all_gather of a small vector: returns (world, numel) with row r = rank r's vector.
    Flat output buffer so that the same call works on NCCL and on gloo (CPU tests)."""
    import torch.distributed as dist
    flat = t.contiguous().reshape(-1)
    out = torch.empty(world * flat.numel(), dtype=flat.dtype, device=flat.device)
    dist.all_gather_into_tensor(out, flat, group=group)
    return out.view(world, flat.numel())
