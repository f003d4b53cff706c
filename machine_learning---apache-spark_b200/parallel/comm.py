"""Symmetric (peer-mapped) buffers and the fused all-reduce + optimizer launcher (csrc/comm/allreduce.cu).

Set-up (once): every rank allocates its flat gradient / weight / working-copy buffers and a signal pad with
``b200ml_comm_alloc`` (plain cudaMalloc, so the CUDA-IPC handle maps the whole allocation at offset 0), exports the IPC
handles, exchanges them through the process group (an object all-gather — control plane only) and imports every
peer's handles.  From then on the data plane is our own kernels dereferencing peer pointers over NVLink; NCCL is not
involved.  This is the B200 counterpart of gloo's TCP pairs + c10d's Reducer buckets in the reference's DDP setup
(distributed_cnn.py:152-156).
"""
from __future__ import annotations

import ctypes as C
import os

import torch
import torch.distributed as dist

from .. import _native, _native_ext

MAX_RANKS = 8
SIGNAL_WORDS = 2 * 320 * 8


class CommStruct(C.Structure):
    _fields_ = [("world", C.c_int), ("rank", C.c_int), ("grads", C.c_void_p * 8), ("params", C.c_void_p * 8),
                ("w16", C.c_void_p * 8), ("signals", C.c_void_p * 8), ("counter", C.c_void_p),
                ("mc_grads", C.c_void_p), ("mc_params", C.c_void_p), ("mc_w16", C.c_void_p)]


vp, i, ll, f = C.c_void_p, C.c_int, C.c_longlong, C.c_float
_native_ext.register("b200ml_comm_alloc", [C.c_size_t, C.POINTER(C.c_void_p)])
_native_ext.register("b200ml_comm_free", [vp])
_native_ext.register("b200ml_ipc_export", [vp, C.c_char_p])
_native_ext.register("b200ml_ipc_import", [C.c_char_p, C.POINTER(C.c_void_p)])
_native_ext.register("b200ml_ipc_close", [vp])
_native_ext.register("b200ml_allreduce_update", [C.POINTER(CommStruct), ll, ll, i, i, f, f, f, f, f, vp, vp, vp, vp, i,
                                                 i, vp])


class _RawCuda:
    """Minimal __cuda_array_interface__ holder so torch can view memory we allocated ourselves."""

    def __init__(self, ptr: int, numel: int, typestr: str):
        self.__cuda_array_interface__ = {"shape": (numel,), "typestr": typestr, "data": (ptr, False), "version": 3}


class SymmetricBuffer:
    """Device buffer allocated by libb200ml and exportable over CUDA IPC; ``.tensor`` is a torch view of it."""

    _DT = {torch.float32: ("<f4", 4), torch.uint8: ("|u1", 1), torch.int32: ("<i4", 4), torch.float64: ("<f8", 8)}

    def __init__(self, numel: int, dtype, device):
        lib = _native.load()
        self.device = torch.device(device)
        esize = 2 if dtype == torch.bfloat16 else self._DT[dtype][1]
        self.nbytes = max(numel * esize, 16)
        p = C.c_void_p()
        with torch.cuda.device(self.device):
            rc = lib.b200ml_comm_alloc(self.nbytes, C.byref(p))
        if rc != 0:
            raise _native.NativeError("comm_alloc: " + lib.b200ml_last_error().decode())
        self.ptr = p.value
        if dtype == torch.bfloat16:
            raw = torch.as_tensor(_RawCuda(self.ptr, numel, "<i2"), device=self.device)
            self.tensor = raw.view(torch.bfloat16)
        else:
            self.tensor = torch.as_tensor(_RawCuda(self.ptr, numel, self._DT[dtype][0]), device=self.device)
        self._keep = self  # tensors created from the interface do not own the memory

    def export(self) -> bytes:
        buf = C.create_string_buffer(64)
        lib = _native.load()
        if lib.b200ml_ipc_export(self.ptr, buf) != 0:
            raise _native.NativeError("ipc_export: " + lib.b200ml_last_error().decode())
        return buf.raw

    def free(self) -> None:
        if self.ptr:
            _native.load().b200ml_comm_free(self.ptr)
            self.ptr = None


def _import(handle: bytes) -> int:
    lib = _native.load()
    p = C.c_void_p()
    if lib.b200ml_ipc_import(handle, C.byref(p)) != 0:
        raise _native.NativeError("ipc_import: " + lib.b200ml_last_error().decode())
    return p.value


class SymmetricAllocator:
    """Hands FlatParams its three flat buffers as symmetric allocations."""

    def __init__(self, device):
        self.device = device
        self.buffers: dict[str, SymmetricBuffer] = {}

    def __call__(self, name: str, numel: int, dtype) -> torch.Tensor:
        b = SymmetricBuffer(numel, dtype, self.device)
        self.buffers[name] = b
        return b.tensor


class SymmMemAllocator:
    """Flat buffers from torch.distributed._symmetric_memory (CUDA VMM allocations that can additionally be bound to an
    NVSwitch multicast object).  Used when NVLS is wanted; peer and multicast addresses come from the rendezvous
    handles instead of CUDA IPC."""

    def __init__(self, device):
        import torch.distributed._symmetric_memory as symm_mem

        self._sm = symm_mem
        self.device = torch.device(device)
        self.tensors: dict[str, torch.Tensor] = {}
        self.handles: dict = {}

    def __call__(self, name: str, numel: int, dtype) -> torch.Tensor:
        t = self._sm.empty(numel, dtype=dtype, device=self.device)
        t.zero_()
        self.tensors[name] = t
        return t

    def rendezvous(self, group) -> None:
        for name, t in self.tensors.items():
            self.handles[name] = self._sm.rendezvous(t, group if group is not None else dist.group.WORLD)


def make_allocator(device, group=None):
    """NVLS-capable symmetric memory when B200ML_NVLS != 0 and torch can provide it, else cudaMalloc + CUDA IPC."""
    import os

    if os.environ.get("B200ML_NVLS", "0") == "1":  # opt-in (see profiles/allreduce_microbench.md for the comparison)
        try:
            alloc = SymmMemAllocator(device)
            probe = alloc("_probe", 1024, torch.float32)
            h = alloc._sm.rendezvous(probe, group if group is not None else dist.group.WORLD)
            ok = bool(getattr(h, "multicast_ptr", 0))
            del alloc.tensors["_probe"]
            if ok:
                return alloc
        except Exception as e:  # noqa: BLE001
            if os.environ.get("B200ML_DEBUG"):
                print(f"[b200ml] symmetric-memory/NVLS unavailable: {e}")
    return SymmetricAllocator(device)


class SparseRowsStruct(C.Structure):
    """b200ml_sparse_rows (include/b200ml.h)."""
    _fields_ = [("start", C.c_longlong), ("rows", C.c_int), ("row_elems", C.c_int), ("touched", C.c_void_p * 8)]


_native_ext.register("b200ml_allreduce_update_sparse", [C.POINTER(CommStruct), ll, ll, i, i, f, f, f, f, f, vp, vp, vp, vp,
                                                        i, i, C.POINTER(SparseRowsStruct), vp])
_native_ext.register("b200ml_allreduce_update_tail", [C.POINTER(CommStruct), ll, ll, ll, ll, i, i, f, f, f, f, f, vp, vp, vp,
                                                      vp, i, i, C.POINTER(SparseRowsStruct), vp])
_native_ext.register("b200ml_allreduce_sum_f64", [C.POINTER(CommStruct), C.POINTER(C.c_void_p * 8), ll, vp, vp])


class F64AllReduce:
    """Sum of one float64 vector over the ranks of a node with the one-shot peer-memory kernel
    (b200ml_allreduce_sum_f64): every rank owns a symmetric buffer ``self.buf`` that producers write straight into,
    ``reduce()`` returns the rank-order sum (identical bits on every rank) in ``self.out``.  Used by the MLlib-style
    L-BFGS trainer (spark/ml.py) in place of Spark's treeAggregate (mllib_multilayer_perceptron_classifier.py:35-39)."""

    def __init__(self, n: int, device, group=None):
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        assert self.world <= MAX_RANKS
        self.group = group
        self.n = (n + 1) // 2 * 2
        self._buf = SymmetricBuffer(self.n, torch.float64, device)
        self.buf = self._buf.tensor
        self.buf.zero_()
        self.out = torch.zeros(self.n, dtype=torch.float64, device=device)
        self._signals = SymmetricBuffer(SIGNAL_WORDS, torch.int32, device)
        self._counter = torch.zeros(2, dtype=torch.int32, device=device)
        mine = {"buf": self._buf.export(), "signals": self._signals.export()}
        gathered = [None] * self.world
        dist.all_gather_object(gathered, mine, group=group)
        self._imported = []
        self.cs = CommStruct()
        self.cs.world, self.cs.rank = self.world, self.rank
        self._bufs = (C.c_void_p * 8)()
        for r in range(self.world):
            if r == self.rank:
                self._bufs[r], self.cs.signals[r] = self._buf.ptr, self._signals.ptr
            else:
                self._bufs[r] = _import(gathered[r]["buf"])
                self.cs.signals[r] = _import(gathered[r]["signals"])
                self._imported += [self._bufs[r], self.cs.signals[r]]
        self.cs.counter = self._counter.data_ptr()
        torch.cuda.synchronize(device)
        dist.barrier(group=group)

    def reduce(self) -> torch.Tensor:
        _native.check(_native.load().b200ml_allreduce_sum_f64(C.byref(self.cs), C.byref(self._bufs), self.n,
                                                              self.out.data_ptr(), _native.stream_ptr()),
                      "allreduce_sum_f64")
        return self.out

    def close(self) -> None:
        try:
            torch.cuda.synchronize(self.buf.device)
            if dist.is_initialized():
                dist.barrier(group=self.group)
        except Exception:  # noqa: BLE001
            pass
        lib = _native.load()
        for p in self._imported:
            lib.b200ml_ipc_close(p)
        self._imported = []


class FusedAllReduce:
    """Peer-pointer tables + launcher for b200ml_allreduce_update."""

    transport = "p2p-ipc"

    def __init__(self, flat, m, v, group=None, allocator: SymmetricAllocator | None = None):
        self.flat, self.m, self.v, self.group = flat, m, v, group
        if isinstance(allocator, SymmMemAllocator):
            self._init_symm_mem(allocator)
            return
        assert allocator is not None and {"flat_p", "flat_g", "flat_w16"} <= set(allocator.buffers), \
            "fused all-reduce needs FlatParams built on a SymmetricAllocator"
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        assert self.world <= MAX_RANKS, f"fused all-reduce supports up to {MAX_RANKS} ranks per node"
        dev = flat.device
        self.alloc = allocator
        self.signals = SymmetricBuffer(SIGNAL_WORDS, torch.int32, dev)
        self.counter = torch.zeros(2, dtype=torch.int32, device=dev)
        mine = {k: b.export() for k, b in allocator.buffers.items()}
        mine["signals"] = self.signals.export()
        gathered = [None] * self.world
        dist.all_gather_object(gathered, mine, group=group)
        self._imported = []
        self.cs = CommStruct()
        self.cs.world, self.cs.rank = self.world, self.rank
        table = {"flat_g": self.cs.grads, "flat_p": self.cs.params, "flat_w16": self.cs.w16, "signals": self.cs.signals}
        local = {"flat_g": allocator.buffers["flat_g"].ptr, "flat_p": allocator.buffers["flat_p"].ptr,
                 "flat_w16": allocator.buffers["flat_w16"].ptr, "signals": self.signals.ptr}
        for r in range(self.world):
            for key, arr in table.items():
                if r == self.rank:
                    arr[r] = local[key]
                else:
                    p = _import(gathered[r][key])
                    self._imported.append(p)
                    arr[r] = p
        self.cs.counter = self.counter.data_ptr()
        # torch views of the peers' fp32 master buffers (for the on-demand pull of slices this rank does not own)
        self.peer_p = [flat.flat_p if r == self.rank else
                       torch.as_tensor(_RawCuda(self.cs.params[r], flat.numel, "<f4"), device=dev)
                       for r in range(self.world)]
        self.push_master = 1
        self.max_blocks = int(os.environ.get("B200ML_AR_BLOCKS", "0"))  # 0 = library default (depends on world / algo)
        self.default_algo = int(os.environ.get("B200ML_AR_ALGO", "0"))  # 0 auto; 2 / 4 / 5 force a two-shot transport
        torch.cuda.synchronize(dev)
        dist.barrier(group=group)  # nobody launches before every mapping exists

    def _init_symm_mem(self, allocator: "SymmMemAllocator") -> None:
        """Peer + multicast addresses from torch symmetric-memory rendezvous handles (NVLS path)."""
        flat, group = self.flat, self.group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        assert self.world <= MAX_RANKS
        dev = flat.device
        self.alloc = allocator
        self.signals_t = allocator("signals", SIGNAL_WORDS, torch.int32)
        self.counter = torch.zeros(2, dtype=torch.int32, device=dev)
        allocator.rendezvous(group)
        h = allocator.handles
        self._imported = []
        self.cs = CommStruct()
        self.cs.world, self.cs.rank = self.world, self.rank
        table = {"flat_g": self.cs.grads, "flat_p": self.cs.params, "flat_w16": self.cs.w16, "signals": self.cs.signals}
        for key, arr in table.items():
            ptrs = list(h[key].buffer_ptrs)
            off = int(getattr(h[key], "offset", 0) or 0)
            for r in range(self.world):
                arr[r] = ptrs[r] + off
            assert arr[self.rank] == allocator.tensors[key].data_ptr(), "symmetric-memory offset mismatch"
        self.cs.counter = self.counter.data_ptr()

        def mc(key):
            p = int(getattr(h[key], "multicast_ptr", 0) or 0)
            return (p + int(getattr(h[key], "offset", 0) or 0)) if p else None

        self.cs.mc_grads, self.cs.mc_params, self.cs.mc_w16 = mc("flat_g"), mc("flat_p"), mc("flat_w16")
        self.transport = "symm-mem+nvls" if self.cs.mc_grads and self.cs.mc_w16 else "symm-mem"
        self.peer_p = [flat.flat_p if r == self.rank else
                       torch.as_tensor(_RawCuda(self.cs.params[r], flat.numel, "<f4"), device=dev)
                       for r in range(self.world)]
        self.push_master = 1
        self.max_blocks = int(os.environ.get("B200ML_AR_BLOCKS", "0"))
        self.default_algo = int(os.environ.get("B200ML_AR_ALGO", "0"))
        torch.cuda.synchronize(dev)
        dist.barrier(group=group)

    def rider_capable(self) -> bool:
        """Can a second range ride in the launch of a two-shot bucket (bulk-async transport, no NVLS)?"""
        return self.default_algo in (0, 4, 5) and not (self.cs.mc_grads and self.cs.mc_w16)

    def allreduce_update(self, start, n, kind, lr, betas, eps, weight_decay, step_tensor, out=None, algo: int = 0,
                         sparse: "SparseRowsStruct | None" = None, tail: "tuple[int, int] | None" = None):
        """``tail`` = (start2, n2): a second range that rides in the same launch (sliced and updated like the bucket,
        fp32 masters always pushed) - only on the bulk-async two-shot path."""
        lib = _native.load()
        n4 = (n + 3) // 4 * 4  # slots are 64-element aligned, so rounding up stays inside the bucket
        if algo == 0 and kind != 0 and n4 * 4 > (512 << 10):
            algo = self.default_algo
        start2, n2 = (0, 0) if tail is None else (tail[0], (tail[1] + 7) // 8 * 8)
        _native.check(lib.b200ml_allreduce_update_tail(
            C.byref(self.cs), start, n4, start2, n2, algo, kind, lr, betas[0], betas[1], eps, weight_decay,
            step_tensor.data_ptr(),
            None if self.m is None else self.m.data_ptr(), None if self.v is None else self.v.data_ptr(),
            None if out is None else out.data_ptr(), self.push_master, self.max_blocks,
            None if sparse is None else C.byref(sparse), _native.stream_ptr()), "allreduce_update")

    def make_sparse_rows(self, start: int, rows: int, row_elems: int):
        """Peer-mapped touched-row flags for a row-sparse table at flat offset ``start``: returns (descriptor, local
        uint8 tensor of ``rows`` flags, initialised to 1 = "every row may be non-zero").  Collective."""
        dev = self.flat.device
        buf = SymmetricBuffer(rows, torch.uint8, dev)
        buf.tensor.fill_(1)
        torch.cuda.synchronize(dev)
        gathered = [None] * self.world
        dist.all_gather_object(gathered, buf.export(), group=self.group)
        d = SparseRowsStruct()
        d.start, d.rows, d.row_elems = start, rows, row_elems
        for r in range(self.world):
            if r == self.rank:
                d.touched[r] = buf.ptr
            else:
                p = _import(gathered[r])
                self._imported.append(p)
                d.touched[r] = p
        self._sparse_bufs = getattr(self, "_sparse_bufs", []) + [buf]
        dist.barrier(group=self.group)
        return d, buf.tensor

    def owned_slice(self, start: int, n: int, sliced: bool | None = None, rank: int | None = None):
        """(lo, hi) relative to ``start``: the part of bucket [start, start+n) whose optimizer state (fp32 master under
        bf16 push, Adam moments always) rank ``rank`` keeps, or None when the bucket takes the one-shot path and every
        rank updates all of it.  Mirrors the slice arithmetic of csrc/comm/allreduce.cu."""
        n4 = (n + 3) // 4 * 4
        if sliced is None:
            sliced = n4 * 4 > (512 << 10)
        if not sliced:
            return None
        gran = 8 if (start % 8 == 0 and n4 % 8 == 0) else 4   # as owned_range() in csrc/comm/allreduce.cu
        per = ((n4 // gran + self.world - 1) // self.world) * gran
        r = self.rank if rank is None else rank
        lo = min(per * r, n4)
        return lo, min(lo + per, n4)

    def pull_master(self, start: int, n: int, sliced: bool | None = None) -> None:
        """Refresh the fp32 master values of [start, start+n) that other ranks own (two-shot buckets updated with
        push_master = 0).  One-sided peer reads; the owners' values are final once our own last all-reduce kernel
        has returned, because that kernel's closing barrier waits for every rank's update."""
        for r in range(self.world):
            if r == self.rank:
                continue
            own = self.owned_slice(start, n, sliced, rank=r)
            if own is None:
                return  # such a bucket takes the one-shot path: every rank already updated all of it locally
            lo, hi = own
            if hi > lo:
                self.flat.flat_p[start + lo:start + hi].copy_(self.peer_p[r][start + lo:start + hi])

    def close(self) -> None:
        lib = _native.load()
        try:
            torch.cuda.synchronize(self.flat.device)
            if dist.is_initialized():
                dist.barrier(group=self.group)
        except Exception:  # noqa: BLE001
            pass
        for p in self._imported:
            lib.b200ml_ipc_close(p)
        self._imported = []
