"""The reference's operator interface for this path, on torch CUDA tensors (device memory only; no arithmetic here).

Names, argument meaning and error behaviour follow ggml:
  quantize_row_q8_1(x)            quantize_row_q8_1_ref          ggml/src/ggml-quants.c:260-293
  mul_mat(type, W, x)             ggml_mul_mat   (ggml.c:3240-3255) / ggml_compute_forward_mul_mat    (ggml-cpu.c:1241-1439)
  mul_mat_id(type, as, b, ids)    ggml_mul_mat_id(ggml.c:3291-3315) / ggml_compute_forward_mul_mat_id (ggml-cpu.c:1521-1694)
Shapes are written torch-style (outermost first), i.e. reversed ggml `ne`: a ggml tensor [k, n, ne2, ne3] is a torch tensor
of shape (ne3, ne2, n, k).  Weight tensors are uint8 with the last dim = ggml_row_size(type, k) bytes.
"""
from __future__ import annotations

import ctypes as C

from .binding import F32, I32, Context, Tensor, blck_size, check, lib, row_size, type_size


def _desc(t, ggml_type: int, k_elems: int | None = None) -> Tensor:
    """b200_tensor for a torch tensor of up to 4 dims.  For quantized weights pass k_elems (elements per row)."""
    import torch
    assert t.is_cuda, "device tensors only"
    d = Tensor()
    d.data = t.data_ptr()
    d.type = ggml_type
    shape = list(t.shape)
    strides = [s * t.element_size() for s in t.stride()]
    while len(shape) < 4:
        shape.insert(0, 1)
        strides.insert(0, strides[0] * shape[1] if strides else t.element_size())
    # torch (ne3, ne2, ne1, ne0) -> ggml ne[0..3]
    for i in range(4):
        d.ne[i] = shape[3 - i]
        d.nb[i] = strides[3 - i]
    if k_elems is not None:
        assert t.dtype == torch.uint8 and t.stride(-1) == 1
        assert shape[3] == row_size(ggml_type, k_elems), "last dim must be ggml_row_size(type, k)"
        d.ne[0] = k_elems
        d.nb[0] = type_size(ggml_type)
    return d


def quantize_row_q8_1(ctx: Context, x):
    """x: f32 (nrows, k) with contiguous rows -> uint8 (nrows, k/32*36) of block_q8_1"""
    import torch
    assert x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1
    nrows, k = x.shape
    if k % 32 != 0:
        raise ValueError("k must be a multiple of 32 (QK8_1)")
    y = torch.empty((nrows, k // 32 * 36), dtype=torch.uint8, device=x.device)
    check(lib().b200_quantize_row_q8_1(ctx.handle, x.data_ptr(), x.stride(0) * 4, y.data_ptr(), k, nrows), "b200_quantize_row_q8_1")
    return y


def mul_mat(ctx: Context, wtype: int, w, x, out=None):
    """w: uint8 ([ne03, ne02,] m, row_size); x: f32 ([ne13, ne12,] n, k) -> f32 ([ne13, ne12,] n, m)"""
    import torch
    k = x.shape[-1]
    if k % blck_size(wtype) != 0:
        raise ValueError("k is not a whole number of blocks")
    a = _desc(w, wtype, k_elems=k)
    b = _desc(x, F32)
    m = w.shape[-2]
    if out is None:
        out = torch.empty(tuple(x.shape[:-1]) + (m,), dtype=torch.float32, device=x.device)
    d = _desc(out, F32)
    check(lib().b200_mul_mat(ctx.handle, C.byref(a), C.byref(b), C.byref(d)), "b200_mul_mat")
    return out


def mul_mat_noalloc(ctx: Context, wtype: int, w, x, out):
    """b200_mul_mat_noalloc: same contract as mul_mat, the caller passes `out`; nothing is allocated inside the call (the
    activation planes live in the library's static device scratch, B200_NOALLOC_SCRATCH_BYTES)"""
    a = _desc(w, wtype, k_elems=x.shape[-1])
    b = _desc(x, F32)
    d = _desc(out, F32)
    check(lib().b200_mul_mat_noalloc(ctx.handle, C.byref(a), C.byref(b), C.byref(d)), "b200_mul_mat_noalloc")
    return out


def mul_mat_multi(ctx: Context, wtype: int, ws, x, outs=None):
    """outs[j] = ws[j] x x for several weight matrices sharing the activations (q/k/v, gate/up) — one launch when fusable"""
    import torch
    k = x.shape[-1]
    if outs is None:
        outs = [torch.empty(tuple(x.shape[:-1]) + (w.shape[-2],), dtype=torch.float32, device=x.device) for w in ws]
    a = [_desc(w, wtype, k_elems=k) for w in ws]
    d = [_desc(o, F32) for o in outs]
    b = _desc(x, F32)
    T = C.POINTER(Tensor)
    ap = (T * len(ws))(*[C.pointer(t) for t in a])
    dp = (T * len(ws))(*[C.pointer(t) for t in d])
    check(lib().b200_mul_mat_multi(ctx.handle, len(ws), ap, C.byref(b), dp), "b200_mul_mat_multi")
    return outs


def mul_mat_fused(ctx: Context, wtype: int, w_up, x, w_gate=None, bias_up=None, bias_gate=None, glu_op: int = 2, out=None):
    """N2: out = w_up x (+ bias_up), or glu(w_gate x (+ bias_gate), w_up x (+ bias_up)), one activation column per batch, ONE launch
    (b200_mul_mat_fused; glu_op: enum ggml_glu_op, 2 = SWIGLU)"""
    import torch
    k = x.shape[-1]
    if out is None:
        out = torch.empty(tuple(x.shape[:-1]) + (w_up.shape[-2],), dtype=torch.float32, device=x.device)
    a = _desc(w_up, wtype, k_elems=k)
    g = _desc(w_gate, wtype, k_elems=k) if w_gate is not None else None
    bu = _desc(bias_up, F32) if bias_up is not None else None
    bg = _desc(bias_gate, F32) if bias_gate is not None else None
    b, d = _desc(x, F32), _desc(out, F32)
    ref = lambda t: C.byref(t) if t is not None else None
    check(lib().b200_mul_mat_fused(ctx.handle, C.byref(a), ref(g), C.byref(b), ref(bu), ref(bg), glu_op, C.byref(d)), "b200_mul_mat_fused")
    return out


def mul_mat_add(ctx: Context, wtype: int, w, x, addend, out=None):
    """out = w x + addend through the fast GEMV kernels (b200_mul_mat_add; n <= 8; out may be the addend itself)"""
    import torch
    if out is None:
        out = torch.empty(tuple(x.shape[:-1]) + (w.shape[-2],), dtype=torch.float32, device=x.device)
    a, b, c, d = _desc(w, wtype, k_elems=x.shape[-1]), _desc(x, F32), _desc(addend, F32), _desc(out, F32)
    check(lib().b200_mul_mat_add(ctx.handle, C.byref(a), C.byref(b), C.byref(c), C.byref(d)), "b200_mul_mat_add")
    return out


_TORCH_TYPES = {"torch.float32": 0, "torch.float16": 1, "torch.bfloat16": 30, "torch.int32": 26, "torch.int64": 27}


def _fdesc(t):
    return _desc(t, _TORCH_TYPES[str(t.dtype)])


def graph_op(ctx: Context, name: str, *tensors, **params):
    """The other graph ops of the C ABI on torch tensors (ggml shapes: torch (ne3, ne2, ne1, ne0)).  name in {"add","sub","mul","div",
    "add_id","glu","unary","scale","rms_norm","rope","soft_max","cpy","get_rows","set_rows","mul_mat_f"}; the LAST tensor is the destination,
    None stands for an absent optional source."""
    L = lib()
    d = [C.byref(_fdesc(t)) if t is not None else None for t in tensors]
    h = ctx.handle
    if name in ("add", "sub", "mul", "div"):
        rc = L.b200_op_binary(h, ("add", "sub", "mul", "div").index(name), *d)
    elif name == "add_id":
        rc = L.b200_op_add_id(h, *d)
    elif name == "glu":
        rc = L.b200_op_glu(h, params["op"], d[0], d[1], d[2], params.get("swapped", 0), params.get("alpha", 0.0), params.get("limit", 0.0))
    elif name == "unary":
        rc = L.b200_op_unary(h, params["op"], *d)
    elif name == "scale":
        rc = L.b200_op_scale(h, d[0], d[1], params["scale"], params.get("bias", 0.0))
    elif name == "rms_norm":
        rc = L.b200_op_rms_norm(h, d[0], d[1], params["eps"])
    elif name == "rope":
        import struct
        words = [0, params["n_dims"], params.get("mode", 0), 0, params.get("n_ctx_orig", 0)]
        words += [struct.unpack("<i", struct.pack("<f", params.get(k_, dv)))[0] for k_, dv in (("freq_base", 10000.0), ("freq_scale", 1.0), ("ext_factor", 0.0),
                                                                                             ("attn_factor", 1.0), ("beta_fast", 32.0), ("beta_slow", 1.0))]
        words += [0, 0, 0, 0]
        rc = L.b200_op_rope(h, d[0], d[1], d[2], d[3], (C.c_int32 * 15)(*words))
    elif name == "soft_max":
        rc = L.b200_op_soft_max(h, d[0], d[1], d[2], d[3], params.get("scale", 1.0), params.get("max_bias", 0.0))
    elif name == "cpy":
        rc = L.b200_op_cpy(h, *d)
    elif name == "get_rows":
        rc = L.b200_op_get_rows(h, *d)
    elif name == "set_rows":
        rc = L.b200_op_set_rows(h, *d)
    elif name == "mul_mat_f":
        rc = L.b200_mul_mat_f(h, *d)
    else:
        raise ValueError(name)
    check(rc, "b200 graph op " + name)
    return tensors[-1]


class MatmulBatch:
    """A prepared list of matmuls for b200_mul_mat_batch: d[i] = ws[i] x xs[i], executed as if in list order.  The descriptor
    arrays are built once, run() is one C-ABI call (what the backend shell does for the MUL_MAT nodes of a cgraph)."""

    def __init__(self, wtype: int, ws, xs, outs):
        self.keep = (ws, xs, outs)
        n = len(ws)
        self.a = [_desc(w, wtype, k_elems=x.shape[-1]) for w, x in zip(ws, xs)]
        self.b = [_desc(x, F32) for x in xs]
        self.d = [_desc(o, F32) for o in outs]
        T = C.POINTER(Tensor)
        self.ap = (T * n)(*[C.pointer(t) for t in self.a])
        self.bp = (T * n)(*[C.pointer(t) for t in self.b])
        self.dp = (T * n)(*[C.pointer(t) for t in self.d])
        self.n = n

    def run(self, ctx: Context):
        check(lib().b200_mul_mat_batch(ctx.handle, self.n, self.ap, self.bp, self.dp), "b200_mul_mat_batch")
        return self.keep[2]


def stream_comm_init(ctx: Context, rank: int, world: int, buf_ptrs, slot_entries: int):
    """row split of b200_mul_mat_batch_gather: buf_ptrs[p] = rank p's zeroed, peer-mapped buffer of 2 * slot_entries * 8 bytes"""
    arr = (C.c_void_p * max(1, len(buf_ptrs)))(*buf_ptrs)
    check(lib().b200_stream_comm_init(ctx.handle, rank, world, arr, slot_entries), "b200_stream_comm_init")


class MatmulBatchGather(MatmulBatch):
    """b200_mul_mat_batch_gather: ws[i] = this rank's row shard (rows [row0s[i], ...)), outs[i] = the full-height result"""

    def __init__(self, wtype: int, ws, xs, outs, row0s):
        super().__init__(wtype, ws, xs, outs)
        self.r0 = (C.c_int64 * self.n)(*row0s)

    def run(self, ctx: Context):
        check(lib().b200_mul_mat_batch_gather(ctx.handle, self.n, self.ap, self.bp, self.dp, self.r0), "b200_mul_mat_batch_gather")
        return self.keep[2]


def mul_mat_batch(ctx: Context, wtype: int, ws, xs, outs=None):
    """outs[i] = ws[i] x xs[i] for a list of matmuls, in list order, through one C-ABI call"""
    import torch
    if outs is None:
        outs = [torch.empty(tuple(x.shape[:-1]) + (w.shape[-2],), dtype=torch.float32, device=x.device) for w, x in zip(ws, xs)]
    return MatmulBatch(wtype, ws, xs, outs).run(ctx)


def mul_mat_supported(wtype: int, w, x) -> bool:
    import torch
    a = _desc(w, wtype, k_elems=x.shape[-1])
    b = _desc(x, F32)
    out = torch.empty(tuple(x.shape[:-1]) + (w.shape[-2],), dtype=torch.float32, device="meta")
    d = Tensor()
    d.type = F32
    shape = [1] * (4 - out.dim()) + list(out.shape)
    nb = 4
    for i in range(4):
        d.ne[i] = shape[3 - i]
        d.nb[i] = nb
        nb *= shape[3 - i]
    return bool(lib().b200_mul_mat_supported(C.byref(a), C.byref(b), C.byref(d)))


def mul_mat_id(ctx: Context, wtype: int, as_, b, ids, out=None):
    """as_: uint8 (n_expert, m, row_size); b: f32 (n_tokens, b_ne1, k); ids: int32 (n_tokens, n_used), rows may be strided
    -> f32 (n_tokens, n_used, m)"""
    import torch
    assert as_.dim() == 3 and b.dim() == 3 and ids.dim() == 2 and ids.dtype == torch.int32
    k = b.shape[-1]
    a = _desc(as_, wtype, k_elems=k)
    bd = _desc(b, F32)
    idd = _desc(ids, I32)
    n_tokens, n_used = ids.shape
    m = as_.shape[-2]
    if out is None:
        out = torch.empty((n_tokens, n_used, m), dtype=torch.float32, device=b.device)
    cd = _desc(out, F32)
    check(lib().b200_mul_mat_id(ctx.handle, C.byref(a), C.byref(bd), C.byref(idd), C.byref(cd)), "b200_mul_mat_id")
    return out


def mm_ids(ctx: Context, ids, n_expert: int):
    """expert gather: ids int32 (n_tokens, n_used) -> (ids_src1 int32 [n_pairs], expert_bounds int32 [n_expert+1])"""
    import torch
    n_tokens, n_used = ids.shape
    src1 = torch.full((n_tokens * n_used,), -1, dtype=torch.int32, device=ids.device)
    bounds = torch.zeros((n_expert + 1,), dtype=torch.int32, device=ids.device)
    check(lib().b200_mm_ids(ctx.handle, ids.data_ptr(), ids.stride(0) * 4, n_expert, n_used, n_tokens, src1.data_ptr(), bounds.data_ptr()),
          "b200_mm_ids")
    return src1, bounds


def mul_mat_gather(ctx: Context, wtype: int, w_shard, x, out_shard, dst_ptrs, dst_m: int, dst_row0: int):
    """row split, one process per GPU: out_shard = w_shard x x, and every result is also stored into the full-height
    outputs `dst_ptrs` (peer-mapped device pointers, e.g. torch symmetric memory) at rows [dst_row0, dst_row0 + rows)"""
    a = _desc(w_shard, wtype, k_elems=x.shape[-1])
    b = _desc(x, F32)
    d = _desc(out_shard, F32)
    arr = (C.c_void_p * len(dst_ptrs))(*dst_ptrs)
    check(lib().b200_mul_mat_gather(ctx.handle, C.byref(a), C.byref(b), C.byref(d), len(dst_ptrs), arr, dst_m, dst_row0), "b200_mul_mat_gather")
    return out_shard


def gather_ll_init(ctx: Context, rank: int, world: int, buf_ptrs, slot_entries: int):
    """LL exchange: buf_ptrs[p] = rank p's zeroed, peer-mapped buffer of 2 * slot_entries * 8 bytes"""
    arr = (C.c_void_p * max(1, len(buf_ptrs)))(*buf_ptrs)
    check(lib().b200_gather_ll_init(ctx.handle, rank, world, arr, slot_entries), "b200_gather_ll_init")


def mul_mat_gather_ll(ctx: Context, wtype: int, w_shards, x, outs_full, row0s):
    """row split, decode: outs_full[j] (full height, local) = all ranks' rows of matrix j x x; w_shards[j] = this rank's rows
    [row0s[j], row0s[j] + rows) — one GEMV launch + one receiver launch, no barrier"""
    k = x.shape[-1]
    a = [_desc(w, wtype, k_elems=k) for w in w_shards]
    d = [_desc(o, F32) for o in outs_full]
    b = _desc(x, F32)
    T = C.POINTER(Tensor)
    n = len(w_shards)
    ap = (T * n)(*[C.pointer(t) for t in a])
    dp = (T * n)(*[C.pointer(t) for t in d])
    r0 = (C.c_int64 * n)(*row0s)
    check(lib().b200_mul_mat_gather_ll(ctx.handle, n, ap, C.byref(b), dp, r0), "b200_mul_mat_gather_ll")
    return outs_full


def row_shard(m: int, rank: int, world: int, rounding: int = 128):
    """(row0, rows) of rank's shard of an m-row weight matrix: even cut points rounded DOWN to `rounding` rows, the last
    rank takes the remainder — the rule of get_row_split (ggml/src/ggml-cuda/ggml-cuda.cu:855-881) and b200_rowsplit_rows"""
    def cut(i):
        if i >= world:
            return m
        c = m * i // world
        return c - c % rounding
    lo, hi = cut(rank), cut(rank + 1)
    return lo, hi - lo
