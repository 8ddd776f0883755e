"""Row gather for the generator API (mappo_gather_rows)."""
import ctypes as C

import torch

from . import _lib
from ._lib import check, ptr, stream_ptr

_lib._EXTRA_SIGS["mappo_gather_rows"] = [C.c_void_p, C.c_longlong, C.c_int, C.c_int, C.c_void_p, C.c_longlong,
                                         C.c_void_p, C.c_void_p]


_lib._EXTRA_SIGS["mappo_onehot_to_ids"] = [C.c_void_p, C.c_int, C.c_longlong, C.c_int, C.POINTER(C.c_int), C.c_void_p,
                                           C.c_int, C.c_void_p]
_ONEHOT_DTYPES = {torch.float32: 0, torch.float64: 1, torch.uint8: 2}


def onehot_to_ids(onehot, head_dims, out):
    """onehot: CUDA [rows, sum(head_dims)] float32 / float64 / uint8 (contiguous); out: float32 [rows, >= heads]."""
    assert onehot.is_cuda and onehot.is_contiguous() and onehot.dtype in _ONEHOT_DTYPES, "one-hot must be f32/f64/u8 CUDA"
    rows = onehot.shape[0]
    hd = (C.c_int * len(head_dims))(*[int(d) for d in head_dims])
    check(_lib.lib().mappo_onehot_to_ids(ptr(onehot), _ONEHOT_DTYPES[onehot.dtype], rows, len(head_dims), hd, ptr(out),
                                         int(out.shape[-1]), stream_ptr()), "onehot_to_ids")
    return out


def gather_rows(src, lead_dims, index):
    """src: float32 CUDA tensor whose first `lead_dims` dims enumerate rows (row-major) and whose remaining dims
    are a contiguous row; may be a broadcast (stride-0) view over trailing lead dims.  -> [len(index), width]."""
    assert src.dtype == torch.float32 and src.is_cuda
    shape, stride = src.shape, src.stride()
    if all(st == 0 or sh == 1 for st, sh in zip(stride, shape)):          # a constant (fully broadcast) tensor
        w = 1
        for d in shape[lead_dims:]:
            w *= d
        return torch.empty(index.numel(), w, dtype=torch.float32, device=src.device).fill_(src[(0,) * src.dim()])
    width = 1
    for d in shape[lead_dims:]:
        width *= d
    tail_contig = True
    exp = 1
    for d in range(src.dim() - 1, lead_dims - 1, -1):
        if shape[d] != 1 and stride[d] != exp:
            tail_contig = False
        exp *= shape[d]
    row_div, row_stride, ok = 1, width, tail_contig
    if ok:
        # lead dims: allow "contiguous" or "trailing lead dims broadcast with stride 0"
        div, exp_stride = 1, None
        d = lead_dims - 1
        while d >= 0 and (stride[d] == 0 or shape[d] == 1):
            div *= shape[d]
            d -= 1
        if d < 0:
            row_div, row_stride = 1, 0            # every row is the same row
        else:
            rs = stride[d]
            e = rs
            for dd in range(d, -1, -1):
                if shape[dd] != 1 and stride[dd] != e:
                    ok = False
                e *= shape[dd]
            row_div, row_stride = div, rs
    if not ok:
        src = src.contiguous()
        row_div, row_stride = 1, width
    index = index.to(device=src.device, dtype=torch.int64).contiguous()
    out = torch.empty(index.numel(), width, dtype=torch.float32, device=src.device)
    check(_lib.lib().mappo_gather_rows(C.c_void_p(src.data_ptr()), int(row_stride), int(row_div), int(width),
                                       ptr(index), index.numel(), ptr(out), stream_ptr()), "gather_rows")
    return out


class _CopyList(C.Structure):
    _fields_ = [("n", C.c_int32), ("reserved", C.c_int32), ("src", C.c_void_p * 8), ("dst", C.c_void_p * 8),
                ("bytes", C.c_longlong * 8)]


_lib._EXTRA_SIGS["mappo_copy_segments"] = [C.POINTER(_CopyList), C.c_void_p]


def copy_segments(pairs):
    """pairs: [(dst, src)] of contiguous CUDA tensors of equal dtype and element count (4-byte elements) -> the copies in
    one launch per 8 pairs (mappo_copy_segments)."""
    fn = _lib.lib().mappo_copy_segments
    for i in range(0, len(pairs), 8):
        cl = _CopyList()
        chunk = pairs[i:i + 8]
        cl.n = len(chunk)
        for k, (dst, src) in enumerate(chunk):
            cl.src[k], cl.dst[k], cl.bytes[k] = src.data_ptr(), dst.data_ptr(), dst.numel() * dst.element_size()
        check(fn(C.byref(cl), stream_ptr()), "copy_segments")


_lib._EXTRA_SIGS["mappo_ids_to_onehot_u8"] = [C.c_void_p, C.c_longlong, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_void_p, C.c_void_p,
                                              C.c_void_p]


def ids_to_onehot(ids, head_dims, out, bad=None):
    """ids: CUDA float32 [rows, >= heads] (contiguous); out: CUDA uint8 [rows, sum(head_dims)]; bad: CUDA int32 [1]
    the kernel sets to 1 if an id lies outside its head (np.eye(n)[id] raises there), or None."""
    assert ids.is_cuda and ids.dtype == torch.float32 and ids.is_contiguous() and out.dtype == torch.uint8 and out.is_contiguous()
    hd = (C.c_int * len(head_dims))(*[int(d) for d in head_dims])
    check(_lib.lib().mappo_ids_to_onehot_u8(ptr(ids), ids.shape[0], int(ids.shape[-1]), len(head_dims), hd, ptr(out),
                                            ptr(bad), stream_ptr()), "ids_to_onehot_u8")
    return out
