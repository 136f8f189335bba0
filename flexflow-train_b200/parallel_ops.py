"""Multi-GPU forms of the four parallel operators: one process per GPU,
`torch.distributed` (NCCL over NVLink 5 / NVSwitch) moves the bytes, the local halves (slice,
place, accumulate) run on the sm_100a kernels through the C ABI.

    Replicate   fwd: broadcast from the owning rank (bit-exact)       bwd: all-reduce(sum) of the replica gradients
    Reduction   fwd: all-reduce(sum) of the partial tensors            bwd: copy (bit-exact)
    Repartition fwd: every rank keeps its slab along `dim` (bit-exact) bwd: all-gather the slab gradients, accumulate
    Combine     fwd: all-gather the slabs along `dim` (bit-exact)      bwd: accumulate this rank's slab of the gradient

The single-device forms (`kernels.Replicate` ...) are what the reference's kernels compute
(`lib/kernels/src/cuda/ops/{replicate,reduction,partition,combine}_kernels.cu`); the movement
between devices was Legion's job there and is NCCL's here. The reference sums gradients
without dividing by the replica count (optimizer_kernel.cu:83-89), and so does this.

`local` supplies copy / accumulate / strided copy. The default is the GPU kernels; there is
no CPU implementation in the product — the CPU tests inject one built on the test oracle to
exercise the rank arithmetic over gloo.
"""
from __future__ import annotations

import os
from typing import List, Optional, Sequence

import torch
import torch.distributed as dist


class GpuLocalOps:
    """copy / accumulate / strided block copy on the device through libffk.so."""

    def __init__(self):
        from . import kernels as K  # raises if libffk.so is missing: no fallback

        self.K = K

    def copy(self, dst, src):
        self.K.copy(None, dst, src)

    def accumulate(self, dst, src):
        self.K.accumulate(None, dst, src)

    def copy_blocks(self, out, out_off, inp, in_off, num_blocks, out_blk, in_blk, add=False):
        name = "ffk_add_with_stride" if add else "ffk_copy_with_stride"
        es = out.element_size()
        self.K._call(name, self.K._stream(None), out.data_ptr() + out_off * es,
                     inp.data_ptr() + in_off * es, num_blocks, out_blk, in_blk, self.K.dtype_code(out))


def _blocks(shape: Sequence[int], dim: int):
    outer = 1
    for s in shape[:dim]:
        outer *= s
    inner = 1
    for s in shape[dim:]:
        inner *= s
    return outer, inner


class ParallelOps:
    def __init__(self, group: Optional[dist.ProcessGroup] = None, local=None):
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.local = local if local is not None else GpuLocalOps()

    # ------------------------------------------------------------------ Replicate
    def replicate_forward(self, tensor: torch.Tensor, src: int = 0) -> torch.Tensor:
        dist.broadcast(tensor, src, group=self.group)
        return tensor

    def replicate_backward(self, input_grad: torch.Tensor, output_grad: torch.Tensor) -> None:
        """input_grad += sum over ranks of output_grad. `output_grad` is used as the
        communication buffer and holds the sum afterwards."""
        dist.all_reduce(output_grad, op=dist.ReduceOp.SUM, group=self.group)
        self.local.accumulate(input_grad, output_grad)

    # ------------------------------------------------------------------ Reduction
    def reduction_forward(self, partial: torch.Tensor, output: torch.Tensor) -> None:
        self.local.copy(output, partial)
        dist.all_reduce(output, op=dist.ReduceOp.SUM, group=self.group)

    def reduction_backward(self, input_grad: torch.Tensor, output_grad: torch.Tensor) -> None:
        self.local.copy(input_grad, output_grad)

    # ------------------------------------------------------------------ Repartition
    def slab_shape(self, full_shape: Sequence[int], dim: int) -> List[int]:
        assert full_shape[dim] % self.world == 0, "dimension must divide evenly over the ranks"
        s = list(full_shape)
        s[dim] //= self.world
        return s

    def repartition_forward(self, input: torch.Tensor, output: torch.Tensor, dim: int) -> None:
        """output = this rank's slab of the (replicated) input along `dim`."""
        assert list(output.shape) == self.slab_shape(input.shape, dim)
        outer, full_blk = _blocks(input.shape, dim)
        _, slab_blk = _blocks(output.shape, dim)
        self.local.copy_blocks(output, 0, input, self.rank * slab_blk, outer, slab_blk, full_blk)

    def repartition_backward(self, input_grad: torch.Tensor, output_grad: torch.Tensor, dim: int) -> None:
        """input_grad (full) += concatenation over ranks of the slab gradients along `dim`."""
        gathered = torch.empty((self.world,) + tuple(output_grad.shape), dtype=output_grad.dtype,
                               device=output_grad.device)
        dist.all_gather_into_tensor(gathered.view(-1), output_grad.contiguous().view(-1), group=self.group)
        outer, full_blk = _blocks(input_grad.shape, dim)
        _, slab_blk = _blocks(output_grad.shape, dim)
        n = output_grad.numel()
        for r in range(self.world):
            self.local.copy_blocks(input_grad, r * slab_blk, gathered, r * n, outer, full_blk, slab_blk, add=True)

    # ------------------------------------------------------------------ Combine
    def combine_forward(self, input: torch.Tensor, output: torch.Tensor, dim: int) -> None:
        """output (full) = concatenation over ranks of the slabs `input` along `dim`."""
        assert list(input.shape) == self.slab_shape(output.shape, dim)
        outer, full_blk = _blocks(output.shape, dim)
        _, slab_blk = _blocks(input.shape, dim)
        if dim == 0 or outer == 1:
            dist.all_gather_into_tensor(output.view(-1), input.contiguous().view(-1), group=self.group)
            return
        gathered = torch.empty((self.world,) + tuple(input.shape), dtype=input.dtype, device=input.device)
        dist.all_gather_into_tensor(gathered.view(-1), input.contiguous().view(-1), group=self.group)
        n = input.numel()
        for r in range(self.world):
            self.local.copy_blocks(output, r * slab_blk, gathered, r * n, outer, full_blk, slab_blk)

    def combine_backward(self, output_grad: torch.Tensor, input_grad: torch.Tensor, dim: int) -> None:
        """input_grad (slab) += this rank's slab of output_grad (full)."""
        outer, full_blk = _blocks(output_grad.shape, dim)
        _, slab_blk = _blocks(input_grad.shape, dim)
        self.local.copy_blocks(input_grad, 0, output_grad, self.rank * slab_blk, outer, slab_blk, full_blk, add=True)

    # ------------------------------------------------------------------ optimizer sync
    def allreduce_gradients(self, flat_buffers: Sequence[torch.Tensor]) -> None:
        """The `*_nccl_update_task_gpu` gradient sync (optimizer_kernel.cu:83-89,191-197):
        SUM over ranks, no division; one collective per flat bucket."""
        for b in flat_buffers:
            dist.all_reduce(b, op=dist.ReduceOp.SUM, group=self.group)


class _DeviceBytes:
    """Raw device memory presented through __cuda_array_interface__ (so torch can view it)."""

    def __init__(self, ptr: int, nbytes: int, owner):
        self.owner = owner
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 3}


class PeerMemoryGroup:
    """Gradient sums over NVLink peer memory (ffk_p2p_* in include/ffk_parallel.h): every rank of
    one node maps the other ranks' regions through CUDA IPC and one kernel per collective reads
    them directly. Tensors that take part are views into the region (`tensor`). The handle
    exchange rides on torch.distributed (any backend); the data never touches NCCL."""

    def __init__(self, nbytes: int, group: Optional[dist.ProcessGroup] = None):
        import ctypes

        from . import kernels as K

        self.K, self.ctypes = K, ctypes
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        nbytes = (int(nbytes) + 255) // 256 * 256
        self.handle = ctypes.c_void_p()
        mine = (ctypes.c_ubyte * 64)()
        # Every rank runs every exchange below even when its own step failed, so that a failure
        # anywhere becomes the same exception on all ranks instead of a rank waiting forever.
        err = None
        try:
            K._call("ffk_p2p_create", ctypes.byref(self.handle), self.rank, self.world, nbytes, mine)
        except Exception as e:  # noqa: BLE001 - reported to every rank below
            err, self.handle = f"rank {self.rank}: {e}", None
        if self.world > 1:
            gathered: List[Optional[tuple]] = [None] * self.world
            dist.all_gather_object(gathered, (err, bytes(mine)), group=group)
            errors = [g[0] for g in gathered if g[0]]
            if not errors:
                try:
                    blob = (ctypes.c_ubyte * (64 * self.world)).from_buffer_copy(b"".join(g[1] for g in gathered))
                    K._call("ffk_p2p_connect", self.handle, blob)
                except Exception as e:  # noqa: BLE001
                    err = f"rank {self.rank}: {e}"
            second: List[Optional[str]] = [None] * self.world
            dist.all_gather_object(second, err, group=group)
            errors += [e for e in second if e and e not in errors]
            if errors:
                self.close()
                raise RuntimeError("peer-memory group could not be set up: " + "; ".join(errors))
        elif err:
            raise RuntimeError("peer-memory group could not be set up: " + err)
        base, size = ctypes.c_void_p(), ctypes.c_size_t()
        K._call("ffk_p2p_local_buffer", self.handle, ctypes.byref(base), ctypes.byref(size))
        self.nbytes = size.value
        self.bytes = torch.as_tensor(_DeviceBytes(base.value, self.nbytes, self), device="cuda")
        if self.world > 1:
            dist.barrier(group=group)  # nobody starts a collective before every rank has mapped its peers

    def tensor(self, offset_bytes: int, shape, dtype) -> torch.Tensor:
        n = 1
        for s in shape:
            n *= s
        nb = n * torch.empty(0, dtype=dtype).element_size()
        assert offset_bytes % 16 == 0 and offset_bytes + nb <= self.nbytes
        return self.bytes[offset_bytes:offset_bytes + nb].view(dtype).view(tuple(shape))

    def all_reduce_sum(self, in_offset: int, out_offset: int, count: int, dtype, stream=None,
                       background_blocks: int = 0) -> None:
        """background_blocks > 0 selects the slim kernel that shares SMs with a running GEMM
        (ffk_p2p_all_reduce_sum_background); every rank must pass the same value."""
        K = self.K
        if background_blocks > 0:
            K._call("ffk_p2p_all_reduce_sum_background", K._stream(stream), self.handle, in_offset, out_offset,
                    count, K._DT[dtype], int(background_blocks))
        else:
            K._call("ffk_p2p_all_reduce_sum", K._stream(stream), self.handle, in_offset, out_offset, count,
                    K._DT[dtype])

    def check(self) -> None:
        self.K._call("ffk_p2p_status", self.handle)

    def close(self) -> None:
        if self.handle:
            self.K._call("ffk_p2p_destroy", self.handle)
            self.handle = None


class PeerGradSync:
    """Data-parallel gradient sums (the Replicate-backward / `*_nccl_update` exchange: SUM over
    ranks, no division, optimizer_kernel.cu:83-89) over NVLink peer memory, issued slice by slice
    WHILE the backward pass is still running.

    The model's flat gradient buckets are re-homed into a peer-mapped region ([grads | sums] per
    bucket). Backward visits the layers in reverse, so the tail of each bucket is final first:
    whenever `min_bytes` more of it are complete, one background all-reduce of that slice goes out
    on a high-priority side stream behind an event; `finish()` sends the remainder and makes the
    main stream wait. The optimizer then reads `summed(g)`."""

    def __init__(self, model, group: Optional[dist.ProcessGroup] = None, min_bytes: int = 48 << 20,
                 background_blocks: int = 32):
        self.model, self.min_bytes, self.blocks = model, int(min_bytes), int(background_blocks)
        buckets = model.grad_buckets
        sizes = [(b.numel() * b.element_size() + 255) // 256 * 256 for b in buckets]
        self.peer = PeerMemoryGroup(2 * sum(sizes), group)
        self.world = self.peer.world
        self.slots = []   # per bucket: (grad offset, sum offset, dtype, element size, numel)
        off = 0
        flats, self.sums = [], []
        for b, sz in zip(buckets, sizes):
            g = self.peer.tensor(off, (b.numel(),), b.dtype).zero_()
            sm = self.peer.tensor(off + sz, (b.numel(),), b.dtype).zero_()
            self.slots.append((off, off + sz, b.dtype, b.element_size(), b.numel()))
            flats.append(g)
            self.sums.append(sm)
            off += 2 * sz
        model.rehome_gradients(flats)
        # where each layer's gradients start inside its bucket (bytes), for the progress tracking
        self._layer_lo = {}
        for l in model.layers:
            for g in l.grads:
                for bi, flat in enumerate(model.grad_buckets):
                    if g.untyped_storage().data_ptr() == flat.untyped_storage().data_ptr() and g.dtype == flat.dtype:
                        lo = (g.data_ptr() - flat.data_ptr()) // 16 * 16
                        key = (id(l), bi)
                        self._layer_lo[key] = min(self._layer_lo.get(key, lo), lo)
        self.side = torch.cuda.Stream(priority=-1)
        self.ready = torch.cuda.Event()
        self.done = torch.cuda.Event()
        self.launches = 0
        self.summed_grads = [self.summed(g) for g in model.grads]
        # The K-cut GEMM schedule waits on CTAs of its own grid, and so does the two-shot collective that
        # now runs beside the backward GEMMs. The side stream's priority orders the two (DESIGN 3.1,
        # "Residency"), but that pairing has not been measured on a multi-GPU box yet: whole tiles while
        # this object lives. FFK_PEER_SYNC_KEEP_KCUT=1 keeps the cut schedule on.
        self._kcut_off = os.environ.get("FFK_PEER_SYNC_KEEP_KCUT") is None
        if self._kcut_off:
            self.peer.K.set_gemm_kcut(False)
        self.begin()

    def begin(self):
        self._sent_lo = [n * es for (_, _, _, es, n) in self.slots]  # bytes [sent_lo, end) are on their way
        self._done_lo = list(self._sent_lo)

    def summed(self, grad: torch.Tensor) -> torch.Tensor:
        """The view of the summed bucket that corresponds to the gradient view `grad`."""
        for flat, sm in zip(self.model.grad_buckets, self.sums):
            if grad.dtype == flat.dtype and flat.data_ptr() <= grad.data_ptr() < flat.data_ptr() + flat.numel() * flat.element_size():
                o = (grad.data_ptr() - flat.data_ptr()) // flat.element_size()
                return sm[o:o + grad.numel()].view(grad.shape)
        raise ValueError("gradient does not live in a synchronised bucket")

    def _send(self, bi: int, lo: int, hi: int):
        if hi <= lo:
            return
        goff, soff, dt, es, _ = self.slots[bi]
        self.ready.record()
        self.side.wait_event(self.ready)
        self.peer.all_reduce_sum(goff + lo, soff + lo, (hi - lo) // es, dt, stream=self.side,
                                 background_blocks=self.blocks)
        self.launches += 1

    def after_layer(self, layer):
        """Called by Model.backward after `layer`'s backward has been queued."""
        for bi in range(len(self.slots)):
            lo = self._layer_lo.get((id(layer), bi))
            if lo is None:
                continue
            self._done_lo[bi] = min(self._done_lo[bi], lo)
            if self._sent_lo[bi] - self._done_lo[bi] >= self.min_bytes:
                self._send(bi, self._done_lo[bi], self._sent_lo[bi])
                self._sent_lo[bi] = self._done_lo[bi]

    def finish(self):
        for bi in range(len(self.slots)):
            self._send(bi, 0, self._sent_lo[bi])
            self._sent_lo[bi] = 0
        self.done.record(self.side)
        torch.cuda.current_stream().wait_event(self.done)
        self.begin()

    def check(self):
        self.peer.check()

    def close(self):
        if self._kcut_off:
            self.peer.K.set_gemm_kcut(True)
            self._kcut_off = False
        self.peer.close()
