"""Feature-side pieces of the path: column statistics (K3), the exact in-place standardisation
pass, padded feature storage, and the lazily standardised feature matrix.

Reference behaviour being replaced: /root/reference/utils.py:118-119
    features = torch.FloatTensor(np.array(features))
    features = (features - features.mean(dim=0)) / features.std(dim=0)
which materialises a standardised copy on the CPU.  Here `load_reddit_data` returns a
`StandardizedFeatures`: the raw fp32 matrix in HBM plus fp32 mean/std vectors.  It behaves like the
standardised tensor for every consumer (any torch op on it sees ``(raw-mean)/std``), row selection
(``features[idx_train]``, reddit.py:45) stays lazy, and ``sgc_precompute`` either standardises the
storage in place with one exact streaming pass before the first hop (default) or hands raw +
statistics to the first hop, which then standardises the gathered values in registers.
"""
import torch

from . import _ffi
from .graph import require_cuda, stream_ptr


def col_stats(x, want64=False):
    """Column mean and unbiased std of a CUDA fp32 [N, F] matrix, computed by sgc_col_stats with
    fp64 accumulation.  Returns (mean32, std32) or (mean32, std32, mean64, std64)."""
    dev = require_cuda(x.device)
    if x.dtype != torch.float32 or x.dim() != 2:
        raise TypeError("col_stats expects a 2-D float32 tensor")
    if x.stride(1) != 1:
        x = x.contiguous()
    n, f = x.shape
    lib = _ffi.lib()
    with torch.cuda.device(dev):
        mean32 = torch.empty(f, dtype=torch.float32, device=dev)
        std32 = torch.empty(f, dtype=torch.float32, device=dev)
        mean64 = torch.empty(f, dtype=torch.float64, device=dev) if want64 else None
        std64 = torch.empty(f, dtype=torch.float64, device=dev) if want64 else None
        ws = torch.empty(lib.sgc_col_stats_workspace_bytes(n, f), dtype=torch.uint8, device=dev)
        _ffi.call("sgc_col_stats", x.data_ptr(), n, f, x.stride(0), _ffi.ptr(mean32), _ffi.ptr(std32),
                  _ffi.ptr(mean64), _ffi.ptr(std64), _ffi.ptr(ws), ws.numel(), stream_ptr())
    return (mean32, std32, mean64, std64) if want64 else (mean32, std32)


def pad_ld(f):
    """Leading dimension (floats) of feature storage for f columns: the next multiple of 8 floats,
    so that every row starts on a 32-byte sector and is a run of whole 16-byte vectors (602 -> 608,
    100 -> 104).  Padding rows further, to 128-byte lines (80 -> 96, 100 -> 128), was measured on
    B200 and does not pay: an 80-column block of the Reddit-shaped X costs 0.846 ms per hop at a
    320-byte stride and 0.884 ms at 384 bytes (larger L2 footprint), products-shaped hops are
    unchanged (8.1 ms)."""
    return (int(f) + 7) // 8 * 8


def padded_rows(n, f, device):
    """fp32 [n, f] view of zero-padded row storage with leading dimension `pad_ld(f)`: every row
    is a run of whole, aligned 16-byte vectors, which the propagation kernel's fast path needs."""
    ld = pad_ld(f)
    base = torch.empty((n, ld), dtype=torch.float32, device=device)
    if ld != f:
        base[:, f:].zero_()
    return base[:, :f]


def to_padded(x, device=None):
    """Copy (host or device) features into padded device storage; no-op if already laid out so."""
    device = x.device if device is None else device
    n, f = x.shape
    if x.device == torch.device(device) and x.dtype == torch.float32 and x.stride(1) == 1 \
            and x.stride(0) == pad_ld(f) and x.data_ptr() % 16 == 0:
        return x
    out = padded_rows(n, f, device)
    out.copy_(x, non_blocking=True)
    return out


def standardize_(x, mean, std):
    """In place x <- (x - mean) / std (utils.py:119), exact fp32 expression, on the GPU."""
    dev = require_cuda(x.device)
    if x.dtype != torch.float32 or x.dim() != 2 or x.stride(1) != 1:
        raise TypeError("standardize_ expects a 2-D float32 tensor with unit column stride")
    with torch.cuda.device(dev):
        _ffi.call("sgc_standardize_inplace", x.data_ptr(), x.shape[0], x.shape[1], x.stride(0),
                  _ffi.ptr(mean), _ffi.ptr(std), stream_ptr())
    return x


def _getter(name):
    return getattr(torch.Tensor, name).__get__


_METADATA = {
    torch.Tensor.size, torch.Tensor.dim, torch.Tensor.stride, torch.Tensor.numel, torch.Tensor.__len__,
    torch.Tensor.is_contiguous, torch.Tensor.data_ptr, torch.Tensor.element_size, torch.Tensor.get_device,
    torch.Tensor.is_floating_point, torch.Tensor.type,
    _getter("shape"), _getter("device"), _getter("dtype"), _getter("is_cuda"), _getter("ndim"),
    _getter("layout"), _getter("requires_grad"), _getter("is_sparse"), _getter("grad"), _getter("grad_fn"),
    _getter("is_leaf"),
}
_IDENTITY = {torch.Tensor.cuda, torch.Tensor.float, torch.Tensor.contiguous, torch.Tensor.detach}


class StandardizedFeatures(torch.Tensor):
    """Raw features + column statistics, semantically equal to ``(raw - mean) / std``."""

    @staticmethod
    def __new__(cls, raw, mean, std):
        t = torch.Tensor._make_subclass(cls, raw.detach(), False)
        t._sgc_mean, t._sgc_std, t._sgc_applied = mean, std, False
        return t

    def unwrap(self):
        """(storage, mean, std); mean/std are None once the storage holds standardised values."""
        with torch._C.DisableTorchFunctionSubclass():
            raw = self.as_subclass(torch.Tensor)
        if self._sgc_applied:
            return raw, None, None
        return raw, self._sgc_mean, self._sgc_std

    def apply_(self):
        """Standardise the storage in place (one streaming pass, 8 B/element) and remember it."""
        if not self._sgc_applied:
            raw, mean, std = self.unwrap()
            standardize_(raw, mean, std)
            self._sgc_applied = True
        return self

    @staticmethod
    def materialize(raw, mean, std):
        """The reference's own expression (utils.py:119) on the GPU, in fp32."""
        return (raw - mean) / std

    def standardized(self):
        raw, mean, std = self.unwrap()
        return raw if mean is None else StandardizedFeatures.materialize(raw, mean, std)

    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        kwargs = kwargs or {}
        me = args[0] if args and isinstance(args[0], StandardizedFeatures) else None
        if me is not None and func in _METADATA:
            with torch._C.DisableTorchFunctionSubclass():
                return func(*args, **kwargs)
        if me is not None and func in _IDENTITY:
            return me
        if me is not None and func is torch.Tensor.__getitem__ and len(args) == 2 \
                and not isinstance(args[1], tuple):
            raw, mean, std = me.unwrap()             # row selection keeps the statistics
            idx = args[1]
            if not isinstance(idx, (torch.Tensor, slice, int)):
                idx = torch.as_tensor(idx, device=raw.device)
            rows = raw[idx]
            if mean is None:
                return rows
            if rows.untyped_storage().data_ptr() == raw.untyped_storage().data_ptr():
                # a slice is a view: parent and child each carry their own "applied" flag, so an in-place
                # apply_() on one would standardise the shared storage twice for the other -> own copy
                rows = rows.clone()
            if rows.dim() != 2:
                return (rows - mean) / std
            return StandardizedFeatures(to_padded(rows) if rows.is_cuda else rows, mean, std)
        # anything else sees the standardised values
        conv = lambda a: a.standardized() if isinstance(a, StandardizedFeatures) else a
        args = tuple(conv(a) for a in args)
        kwargs = {k: conv(v) for k, v in kwargs.items()}
        with torch._C.DisableTorchFunctionSubclass():
            return func(*args, **kwargs)
