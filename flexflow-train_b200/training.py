"""Step driver over the kernel API — the glue the reference's local-execution backend lacks
(SURVEY §8f-1): `LocalTrainingBacking::execute_update` is NOT_IMPLEMENTED
(lib/local-execution/src/local_training_backing.cc:107-109) and no loss task is wired, so
nothing in the reference can run a full step. This driver walks a layer list exactly the
way `execute_forward` / `execute_backward` do (topological order forward, reversed order
backward: local_training_backing.cc:69-105), then calls the loss and optimizer kernels.

Every arithmetic step is a C-ABI call into libffk.so (through `kernels.py`); torch owns the
device buffers. Gradients follow the reference's convention: backward kernels accumulate, so
the driver zeroes all gradient buffers at the start of a step (`ff.zero_gradients()` in the
legacy loop, examples/cpp/Transformer/transformer.cc:192-197).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional

import torch
import torch.distributed as dist

from . import kernels as K


class Tensor:
    """An activation with its gradient buffer (allocated on demand)."""

    def __init__(self, shape, dtype, device, requires_grad=True):
        self.data = torch.zeros(shape, dtype=dtype, device=device)
        self.grad = torch.zeros(shape, dtype=dtype, device=device) if requires_grad else None

    @property
    def shape(self):
        return self.data.shape


def _fill0(t: torch.Tensor):
    K._call("ffk_fill", K._stream(None), t.data_ptr(), t.numel(), 0.0, K.dtype_code(t))


class Layer:
    params: List[torch.Tensor] = []
    grads: List[torch.Tensor] = []
    # fused-operator execution (Model(fuse=True)): a layer that is the FIRST writer of an input's
    # gradient in backward order may write it instead of adding to it, which removes the zero fill
    # of that gradient and the read of its old value
    supports_overwrite = False
    # backward finishes reading y.grad before it writes x.grad, so the two may share one buffer
    safe_inplace = False
    _first: Dict[int, bool] = {}

    def input_tensors(self) -> List["Tensor"]:
        """Every activation this layer reads (and sends a gradient to)."""
        out = [getattr(self, n, None) for n in ("x", "a", "b", "residual")]
        out += list(getattr(self, "inputs", []) or [])
        seen, uniq = set(), []
        for t in out:
            if isinstance(t, Tensor) and id(t) not in seen:
                seen.add(id(t))
                uniq.append(t)
        return uniq

    def output_tensors(self) -> List["Tensor"]:
        y = getattr(self, "y", None)
        return [y] if isinstance(y, Tensor) else []

    def first(self, t: "Tensor") -> bool:
        """This layer is the first writer of t.grad in backward order (set by Model when fusing)."""
        return self._first.get(id(t), False)

    def forward(self):
        raise NotImplementedError

    def backward(self):
        raise NotImplementedError


_scratch_pool: Dict = {}


def _scratch_like(t: torch.Tensor, tag: str) -> torch.Tensor:
    """Backward scratch that is dead between layers: one buffer per (device, shape, dtype, tag)."""
    key = (str(t.device), tuple(t.shape), t.dtype, tag)
    buf = _scratch_pool.get(key)
    if buf is None:
        buf = _scratch_pool[key] = torch.empty_like(t)
    return buf


class Linear(Layer):
    supports_overwrite = True

    def __init__(self, x: Tensor, out_dim: int, activation=K.ACT_NONE, bias=True, seed=0,
                 shard_out=None, shard_in=None, dropout=None, residual: Optional[Tensor] = None):
        """shard_out = (index, count): this rank holds rows [index*out_dim, ...) of a weight with
        count*out_dim outputs (column-parallel layer); shard_in = (index, count): it holds the
        matching slice of the input features (row-parallel layer, partial sums out). The full
        weight is drawn from the seed and sliced, so that a sharded model equals the unsharded one.

        Fused-operator form: dropout = (rate, seed) and / or residual make this layer
        y = dropout(act(x W^T + b)) + residual, ONE GEMM whose epilogue does the rest
        (ffk_linear_forward_fused) instead of Linear, Dropout and ElementBinary layers."""
        self.x = x
        dt, dev = x.data.dtype, x.data.device
        self.in_dim = x.shape[-1]
        self.batch = x.data.numel() // self.in_dim
        self.out_dim = out_dim
        g = torch.Generator(device=dev).manual_seed(4321 + seed)
        full_out = out_dim * (shard_out[1] if shard_out else 1)
        full_in = self.in_dim * (shard_in[1] if shard_in else 1)
        bound = (6.0 / (full_in + full_out)) ** 0.5  # glorot uniform, the reference's default initializer
        full = (torch.rand(full_out, full_in, device=dev, generator=g) * 2 - 1) * bound
        if shard_out:
            full = full[shard_out[0] * out_dim:(shard_out[0] + 1) * out_dim]
        if shard_in:
            full = full[:, shard_in[0] * self.in_dim:(shard_in[0] + 1) * self.in_dim]
        self.w = full.to(dt).contiguous()
        self.b = torch.zeros(out_dim, dtype=dt, device=dev) if bias else None
        self.dw = torch.zeros_like(self.w)
        self.db = torch.zeros_like(self.b) if bias else None
        self.y = Tensor(tuple(x.shape[:-1]) + (out_dim,), dt, dev)
        self.state = K.Linear.init_kernel(activation, 0.0, bias, K.GEMM_AUTO, out_dim=out_dim, device=dev)
        self.preact = torch.zeros_like(self.y.data) if activation == K.ACT_GELU else None
        self.plain = K.Linear.init_kernel(K.ACT_NONE, 0.0, bias, K.GEMM_AUTO)
        self.params = [self.w] + ([self.b] if bias else [])
        self.grads = [self.dw] + ([self.db] if bias else [])
        self.drop = K.Dropout.init_kernel(dropout[0], dropout[1], self.y.data.numel(), device=dev) if dropout else None
        self.residual = residual
        self.residual_mode = "acc"       # "alias" | "copy" | "acc": how y.grad reaches residual.grad (Model)
        self.gate_src: Optional["Linear"] = None  # producer of x whose ReLU(+dropout) gradient this dgrad applies
        self.grad_is_preact = False      # y.grad arrives as the pre-activation gradient (a consumer gated it)
        self.fused_preact = False        # the pre-activation is stored by the same GEMM (Model(fuse=True))

    def forward(self):
        if self.drop is not None or self.residual is not None or (self.preact is not None and self.fused_preact):
            K.FusedLinear.forward_kernel(None, self.state, self.x.data, self.y.data, self.w, self.b, self.in_dim,
                                         self.out_dim, self.batch, preact_out=self.preact,
                                         residual=self.residual.data if self.residual is not None else None,
                                         dropout=self.drop)
            return
        if self.preact is not None:
            # GELU's gradient needs the pre-activation: one extra plain projection kept beside
            # the activated output (the reference's Linear cannot differentiate GELU at all)
            K.Linear.forward_kernel(None, self.plain, self.x.data, self.preact, self.w, self.b,
                                    self.in_dim, self.out_dim, self.batch)
        K.Linear.forward_kernel(None, self.state, self.x.data, self.y.data, self.w, self.b, self.in_dim,
                                self.out_dim, self.batch)

    def backward(self):
        G = self.y.grad
        if self.residual is not None and self.residual.grad is not None:
            if self.residual_mode == "copy":
                K.copy(None, self.residual.grad, G)
            elif self.residual_mode == "acc":
                K.accumulate(None, self.residual.grad, G)
            # "alias": residual.grad IS y.grad
        dy, fused_drop = G, None
        act = self.state.activation
        shared = self.residual is not None          # G also serves as the residual's gradient: keep it intact
        if self.drop is not None and not self.grad_is_preact:
            if shared or act == K.ACT_NONE:
                dy = _scratch_like(G, "dz")
                K.Dropout.backward_kernel(None, self.drop, G, dy)
            else:
                fused_drop = self.drop               # dropout + activation gradient, one pass in place on G
        elif shared and act != K.ACT_NONE and not self.grad_is_preact:
            dy = _scratch_like(G, "dz")
            K.copy(None, dy, G)
        gate = self.gate_src
        gate_scale = 1.0
        if gate is not None and gate.drop is not None:
            gate_scale = 1.0 / (1.0 - gate.drop.rate) if gate.drop.rate < 1.0 else 0.0
        K.FusedLinear.backward_kernel(None, self.state, self.x.data, self.x.grad, self.y.data, dy, self.w, self.dw,
                                      self.db, self.in_dim, self.out_dim, self.batch, self.preact,
                                      accumulate_input_grad=not self.first(self.x),
                                      output_grad_is_preactivation=self.grad_is_preact, dropout=fused_drop,
                                      input_gate=gate.y.data if gate is not None else None,
                                      input_gate_scale=gate_scale)


class LayerNorm(Layer):
    supports_overwrite = True

    def __init__(self, x: Tensor, eps=1e-5):
        self.x = x
        dt, dev = x.data.dtype, x.data.device
        N = x.shape[-1]
        M = x.data.numel() // N
        self.gamma = torch.ones(N, dtype=dt, device=dev)
        self.beta = torch.zeros(N, dtype=dt, device=dev)
        self.dgamma = torch.zeros_like(self.gamma)
        self.dbeta = torch.zeros_like(self.beta)
        self.y = Tensor(x.shape, dt, dev)
        self.state = K.LayerNorm.init_kernel(True, M, N, eps, device=dev)
        self.params = [self.gamma, self.beta]
        self.grads = [self.dgamma, self.dbeta]

    def forward(self):
        K.LayerNorm.forward_kernel(None, self.state, self.x.data, self.y.data, self.gamma, self.beta)

    def backward(self):
        K.LayerNorm.backward_kernel(None, self.state, self.y.grad, self.x.data, self.x.grad, self.gamma,
                                    self.dgamma, self.dbeta, accumulate_param_grads=True,
                                    accumulate_input_grad=not self.first(self.x))


class Add(Layer):
    supports_overwrite = True

    def __init__(self, a: Tensor, b: Tensor):
        self.a, self.b = a, b
        self.y = Tensor(a.shape, a.data.dtype, a.data.device)
        # set by Model(fuse=True) per operand: "alias" (operand.grad IS y.grad), "copy", or "acc"
        self.mode = {"a": "acc", "b": "acc"}

    def forward(self):
        K.ElementBinary.forward_kernel(None, K.OperatorType.EW_ADD, self.a.data, self.b.data, self.y.data)

    def backward(self):
        if self.mode["a"] == "acc" and self.mode["b"] == "acc":
            K.ElementBinary.backward_kernel(None, K.OperatorType.EW_ADD, self.y.grad, self.a.data, self.b.data,
                                            self.a.grad, self.b.grad)
            return
        for name, t in (("a", self.a), ("b", self.b)):
            if t.grad is None or self.mode[name] == "alias":
                continue
            (K.copy if self.mode[name] == "copy" else K.accumulate)(None, t.grad, self.y.grad)


class Dropout(Layer):
    supports_overwrite = True

    def __init__(self, x: Tensor, rate: float, seed: int):
        self.x = x
        self.y = Tensor(x.shape, x.data.dtype, x.data.device)
        self.state = K.Dropout.init_kernel(rate, seed, x.data.numel(), device=x.data.device)
        self.scratch = None
        self.sole_reader = False  # set by Model once the whole graph is known

    def forward(self):
        K.Dropout.forward_kernel(None, self.state, self.x.data, self.y.data)

    def backward(self):
        # the kernel overwrites (reference semantics). When this layer is the only reader of x the
        # result IS x's gradient; otherwise the driver accumulates it into x.grad.
        if self.sole_reader or self.first(self.x):
            K.Dropout.backward_kernel(None, self.state, self.y.grad, self.x.grad)
            return
        if self.scratch is None:
            self.scratch = torch.zeros_like(self.x.data)
        K.Dropout.backward_kernel(None, self.state, self.y.grad, self.scratch)
        K.accumulate(None, self.x.grad, self.scratch)


class SelfAttention(Layer):
    """MultiHeadAttention with q = k = v = x (what lib/models' encoder layer builds:
    transformer.cc:32-45). The reference's wrapper passes NULL for dK/dV in this case and
    thereby drops two thirds of the input gradient; the driver asks for all three and sums them.

    residual=True is the fused-operator form y = MHA(x, x, x) + x (the residual add that follows
    the attention block in lib/models runs in the output projection's epilogue)."""

    supports_overwrite = True
    safe_inplace = True   # y.grad is read by the first GEMMs of the backward pass, x.grad written by the last
    _shared_ws: Dict = {}  # backward scratch is dead between layers: one buffer per (device, size)

    def __init__(self, x: Tensor, heads: int, kdim: int, vdim: int, seed=0, shard=None, residual=False):
        """shard = (index, count): this rank computes `heads` of the count*heads heads (its slice of
        the packed weight Wq[E][H][kdim] | Wk | Wv | Wo[H][vdim][E]); its output is a partial sum."""
        self.x = x
        dt, dev = x.data.dtype, x.data.device
        N, S, E = x.shape
        self.state = K.MultiHeadAttention.init_kernel(N, heads, E, E, E, kdim, kdim, vdim, E, S, S, dtype=dt, device=dev)
        key = (str(dev), self.state.workSpace.numel())
        self.state.workSpace = SelfAttention._shared_ws.setdefault(key, self.state.workSpace)
        nw = self.state.weightSize // x.data.element_size()
        g = torch.Generator(device=dev).manual_seed(977 + seed)
        idx, cnt = shard if shard else (0, 1)
        full = (torch.rand(nw * cnt, device=dev, generator=g) * 2 - 1) * (1.0 / E ** 0.5)
        if cnt > 1:
            H = heads * cnt
            sizes = [E * H * kdim, E * H * kdim, E * H * vdim, H * vdim * E]
            wq, wk, wv, wo = torch.split(full, sizes)
            h0, h1 = idx * heads, (idx + 1) * heads
            full = torch.cat([wq.view(E, H, kdim)[:, h0:h1].reshape(-1), wk.view(E, H, kdim)[:, h0:h1].reshape(-1),
                              wv.view(E, H, vdim)[:, h0:h1].reshape(-1), wo.view(H, vdim, E)[h0:h1].reshape(-1)])
        self.w = full.to(dt).contiguous()
        self.dw = torch.zeros_like(self.w)
        self.y = Tensor(x.shape, dt, dev)
        self.with_residual = bool(residual)
        self.fused = False        # Model(fuse=True): dq, dk, dv are accumulated straight into x.grad
        self.alias_x = False      # Model(fuse=True): x.grad IS y.grad (residual form, first writer)
        self._dqkv = None
        self.params, self.grads = [self.w], [self.dw]

    def forward(self):
        if self.with_residual:
            K.MultiHeadAttention.forward_residual_kernel(None, self.state, self.x.data, self.x.data, self.x.data,
                                                         self.w, self.y.data, self.x.data)
        else:
            K.MultiHeadAttention.forward_kernel(None, self.state, self.x.data, self.x.data, self.x.data, self.w,
                                                self.y.data)

    def backward(self):
        G, xg = self.y.grad, self.x.grad
        if self.fused:
            acc_q = True
            if self.with_residual:
                if not self.alias_x:   # x.grad (+)= y.grad, the residual branch
                    (K.copy if self.first(self.x) else K.accumulate)(None, xg, G)
            else:
                acc_q = not self.first(self.x)
            K.MultiHeadAttention.backward_accumulate_kernel(None, self.state, self.x.data, xg, self.x.data, xg,
                                                            self.x.data, xg, self.w, self.dw, G,
                                                            accumulate=(acc_q, True, True))
            return
        if self._dqkv is None:
            # dq | dk | dv stacked like the replicas of a Replicate op: their sum is one kernel
            self._dqkv = torch.zeros((3,) + tuple(self.x.data.shape), dtype=self.x.data.dtype, device=self.x.data.device)
        dq, dk, dv = self._dqkv[0], self._dqkv[1], self._dqkv[2]
        if self.with_residual:
            K.accumulate(None, xg, G)
        K.MultiHeadAttention.backward_kernel(None, self.state, self.x.data, dq, self.x.data, dk,
                                             self.x.data, dv, self.w, self.dw, G)
        K.Replicate.backward_kernel(None, xg, self._dqkv, 3)  # x.grad += dq + dk + dv


class ReduceSum(Layer):
    """Reduction over the tensor-parallel group (lib/op-attrs .../ops/reduction.cc): forward sums
    the ranks' partial tensors (all-reduce), backward hands the gradient through unchanged."""

    supports_overwrite = True

    def __init__(self, x: Tensor, group=None):
        self.x, self.group = x, group
        self.y = Tensor(x.shape, x.data.dtype, x.data.device)

    def forward(self):
        K.copy(None, self.y.data, self.x.data)
        dist.all_reduce(self.y.data, op=dist.ReduceOp.SUM, group=self.group)

    def backward(self):
        (K.copy if self.first(self.x) else K.accumulate)(None, self.x.grad, self.y.grad)


class ReplicateIn(Layer):
    """Replicate towards the tensor-parallel group (.../ops/replicate.cc): forward is the identity
    (every rank already holds x), backward sums the ranks' gradients of their replicas."""

    def __init__(self, x: Tensor, group=None):
        self.x, self.group = x, group
        self.y = Tensor.__new__(Tensor)
        self.y.data = x.data  # same storage: no copy in the forward pass
        self.y.grad = torch.zeros_like(x.data)

    def forward(self):
        pass

    def backward(self):
        dist.all_reduce(self.y.grad, op=dist.ReduceOp.SUM, group=self.group)
        K.accumulate(None, self.x.grad, self.y.grad)


class CombineGather(Layer):
    """Combine over the tensor-parallel group (lib/op-attrs .../ops/combine.cc; kernel
    combine_kernels.cu:50-61): forward concatenates the ranks' slabs along `dim` (all-gather,
    bit-exact), backward adds this rank's slab of the output gradient to the slab gradient."""

    def __init__(self, x: Tensor, dim: int, pops):
        self.x, self.dim, self.pops = x, dim % len(x.shape), pops
        shape = list(x.shape)
        shape[self.dim] *= pops.world
        self.y = Tensor(tuple(shape), x.data.dtype, x.data.device)

    def forward(self):
        self.pops.combine_forward(self.x.data, self.y.data, self.dim)

    def backward(self):
        self.pops.combine_backward(self.y.grad, self.x.grad, self.dim)


class RepartitionSlice(Layer):
    """Repartition over the tensor-parallel group (.../ops/repartition.cc; kernel
    partition_kernels.cu:53-72): forward keeps this rank's slab of the replicated input along `dim`
    (bit-exact), backward all-gathers the slab gradients and adds them to the input gradient."""

    def __init__(self, x: Tensor, dim: int, pops):
        self.x, self.dim, self.pops = x, dim % len(x.shape), pops
        self.y = Tensor(tuple(pops.slab_shape(x.shape, self.dim)), x.data.dtype, x.data.device)

    def forward(self):
        self.pops.repartition_forward(self.x.data, self.y.data, self.dim)

    def backward(self):
        self.pops.repartition_backward(self.x.grad, self.y.grad, self.dim)


class Softmax(Layer):
    supports_overwrite = True

    def __init__(self, x: Tensor):
        self.x = x
        self.y = Tensor(x.shape, x.data.dtype, x.data.device)
        self.overwrites = True   # legacy (unfused) mode: Model zero-fills x.grad unless this layer is x's sole reader
        self.alias = False       # fused mode: x.grad IS y.grad, nothing to do in backward

    def forward(self):
        C = self.x.shape[-1]
        K.Softmax.forward_kernel(None, self.x.data, self.y.data, self.x.data.numel() // C, C, 1)

    def backward(self):
        # the reference's softmax backward is a copy (it relies on the fused CE loss gradient)
        if self.alias:
            return
        if self.overwrites or self.first(self.x):
            K.Softmax.backward_kernel(None, self.x.grad, self.y.grad)
        else:
            K.accumulate(None, self.x.grad, self.y.grad)


@dataclass
class SGD:
    lr: float = 0.01
    momentum: float = 0.0
    nesterov: bool = False
    weight_decay: float = 1e-4  # FFConfig default (lib/local-execution/include/local-execution/config.h)
    velocity: Dict[int, torch.Tensor] = field(default_factory=dict)
    master: Dict[int, torch.Tensor] = field(default_factory=dict)

    def step(self, params, grads, num_replicas=1):
        for p, g in zip(params, grads):
            v = None
            if self.momentum > 0:
                v = self.velocity.get(id(p))
                if v is None:
                    v = self.velocity[id(p)] = torch.zeros(p.numel(), dtype=torch.float32, device=p.device)
            m = None
            if p.dtype != torch.float32:
                m = self.master.get(id(p))
                if m is None:  # fp32 master copy, made once (setdefault would rebuild it every step)
                    m = self.master[id(p)] = p.detach().float().clone()
            K.sgd_ps_update_task_gpu(None, self.lr, self.momentum, self.nesterov, self.weight_decay, g,
                                     p.numel(), num_replicas, p, v, m)


class Model:
    def __init__(self, inp: Tensor, layers: List[Layer], out: Tensor, fuse: bool = False):
        """fuse=False runs the reference's per-operator convention to the letter: every backward kernel
        ADDS to its input gradients and the driver zero-fills them first. fuse=True is fused-operator
        execution (the reference's `perform_fusion`, config.h:80): the gradient plan below decides,
        per tensor, which layer writes its gradient first (that layer overwrites: no fill, no read of
        the old value), lets the operands of a residual add SHARE the add's output-gradient buffer
        where the order of reads and writes allows it, and wires the ReLU(+dropout) gradient of a
        feed-forward layer into the data-gradient GEMM of the layer behind it."""
        self.input, self.layers, self.output, self.fuse = inp, layers, out, fuse
        readers: Dict[int, List[Layer]] = {}
        for l in layers:
            for t in l.input_tensors():
                readers.setdefault(id(t), []).append(l)
        self._overwritten = set()
        for l in layers:
            l._first = {}
        if fuse:
            self._plan_gradients(readers)
        else:
            # gradients that a backward kernel overwrites need no zero fill and no accumulate pass:
            # a Dropout whose input nobody else reads writes its result straight into x.grad
            for l in layers:
                sole = len(readers.get(id(l.x), [])) == 1 if isinstance(getattr(l, "x", None), Tensor) else False
                if isinstance(l, Dropout) and sole and l.x.grad is not None:
                    l.sole_reader = True
                    self._overwritten.add(id(l.x))
                # the reference's softmax backward is a copy: it overwrites its input gradient
                if isinstance(l, Softmax):
                    l.overwrites = sole
                    if sole and l.x.grad is not None:
                        self._overwritten.add(id(l.x))
        # every loss kernel writes the whole gradient of the model output
        if out.grad is not None and not readers.get(id(out)):
            self._overwritten.add(id(out))
        for tid in self._overwritten:
            assert len(readers.get(tid, [])) <= 1 or fuse, "an overwritten gradient must have a single writer"
        self._flatten_gradients()
        self.params = [p for l in layers for p in l.params]
        self.grads = [g for l in layers for g in l.grads]

    def _plan_gradients(self, readers: Dict[int, List[Layer]]):
        """Fused-operator execution: first writers, shared gradient buffers, gated data gradients."""
        order = list(reversed(self.layers))                   # backward order
        pos = {id(l): i for i, l in enumerate(order)}
        producer: Dict[int, Layer] = {}
        tensors: Dict[int, Tensor] = {}
        for l in self.layers:
            for t in l.output_tensors():
                producer[id(t)] = l
            for t in l.input_tensors() + l.output_tensors():
                tensors[id(t)] = t
        writers = {tid: sorted(ls, key=lambda l: pos[id(l)]) for tid, ls in readers.items()}
        for tid, ls in writers.items():
            t = tensors[tid]
            if t.grad is None:
                continue
            if ls[0].supports_overwrite:
                ls[0]._first[tid] = True
                self._overwritten.add(tid)                    # no zero fill: its first writer overwrites

        def later_writers_ok(z: Tensor, me: Layer, x_producer: Optional[Layer]) -> bool:
            """z.grad may share a buffer that x's producer still reads: every OTHER writer of z.grad must
            run after that read — or be that producer itself, if it reads before it writes."""
            if x_producer is None:
                return True
            for w in writers[id(z)]:
                if w is me:
                    continue
                if pos[id(w)] < pos[id(x_producer)]:
                    return False
                if pos[id(w)] == pos[id(x_producer)] and not w.safe_inplace:
                    return False
            return True

        def share(t: Tensor, buf_owner: Tensor):
            t.grad = buf_owner.grad

        for l in order:
            if isinstance(l, Add):
                ops = {"a": l.a, "b": l.b}
                first = {n: t.grad is not None and l._first.get(id(t), False) for n, t in ops.items()}
                sole = {n: len(readers[id(t)]) == 1 for n, t in ops.items()}
                for n in ops:
                    l.mode[n] = "copy" if first[n] else "acc"
                if l.a is l.b:
                    continue
                for n, m in (("a", "b"), ("b", "a")):
                    x, z = ops[n], ops[m]
                    if not (first[n] and sole[n]):
                        continue
                    # x is read only by this add: its gradient IS y.grad
                    l.mode[n] = "alias"
                    share(x, l.y)
                    if first[m] and later_writers_ok(z, l, producer.get(id(x))):
                        l.mode[m] = "alias"      # later writers add into the shared buffer after x's producer read it
                        share(z, l.y)
                    break
            elif isinstance(l, Linear):
                if l.residual is not None and l.residual.grad is not None:
                    # y.grad is only read by this layer's backward: the residual's gradient can be that buffer
                    if l._first.get(id(l.residual), False):
                        l.residual_mode = "alias"
                        share(l.residual, l.y)
                    else:
                        l.residual_mode = "acc"
                if l.preact is not None:
                    l.fused_preact = True
                src = producer.get(id(l.x))
                if (isinstance(src, Linear) and src.state.activation == K.ACT_RELU and src.residual is None
                        and len(readers[id(l.x)]) == 1 and l.x.grad is not None and l._first.get(id(l.x), False)
                        and l.x.data.dtype in (torch.bfloat16, torch.float16) and l.in_dim % 32 == 0):
                    l.gate_src = src                 # dX of this layer = pre-activation gradient of `src`
                    src.grad_is_preact = True
            elif isinstance(l, Softmax):
                # backward is a copy (softmax fused with the cross-entropy gradient): let the loss
                # kernel write the logits' gradient in place instead
                if l.x.grad is not None and l._first.get(id(l.x), False) and l.y.grad is not None:
                    l.alias = True
                    share(l.x, l.y)
            elif isinstance(l, SelfAttention):
                l.fused = True
                l._dqkv = None
                if l.with_residual and l.x.grad is not None and l._first.get(id(l.x), False):
                    l.alias_x = True
                    share(l.x, l.y)

    def rehome_gradients(self, flats: List[torch.Tensor]):
        """Move the flat gradient buckets into caller-provided storage of the same sizes (e.g. views
        of a peer-mapped region, so that the data-parallel sum can read them over NVLink)."""
        assert len(flats) == len(self.grad_buckets)
        for old, new in zip(self.grad_buckets, flats):
            assert old.dtype == new.dtype and old.numel() == new.numel()
        self._flatten_gradients(flats)
        self.grads = [g for l in self.layers for g in l.grads]

    def _flatten_gradients(self, flats: Optional[List[torch.Tensor]] = None):
        """Re-home every weight gradient in one flat bucket per dtype (16-byte aligned slots),
        so that zeroing is one fill and data-parallel sync is one all-reduce per bucket."""
        by_dtype: Dict[torch.dtype, List] = {}
        for l in self.layers:
            for g in l.grads:
                by_dtype.setdefault(g.dtype, []).append((l, g))
        self.grad_buckets: List[torch.Tensor] = []
        for bi, (dt, items) in enumerate(by_dtype.items()):
            align = 16 // torch.empty(0, dtype=dt).element_size()
            total = sum((g.numel() + align - 1) // align * align for _, g in items)
            flat = flats[bi] if flats is not None else torch.zeros(total, dtype=dt, device=items[0][1].device)
            assert flat.numel() == total
            off = 0
            for l, g in items:
                view = flat[off:off + g.numel()].view(g.shape)
                for name, val in list(vars(l).items()):
                    if val is g:
                        setattr(l, name, view)
                l.grads = [view if x is g else x for x in l.grads]
                off += (g.numel() + align - 1) // align * align
            self.grad_buckets.append(flat)

    def activations(self):
        seen, out = set(), []
        for l in self.layers:
            for t in l.input_tensors() + l.output_tensors():
                if id(t) not in seen:
                    seen.add(id(t))
                    out.append(t)
        return out

    def zero_gradients(self):
        for b in self.grad_buckets:
            _fill0(b)
        for t in self.activations():
            if t.grad is not None and id(t) not in self._overwritten:
                _fill0(t.grad)

    def forward(self):
        for l in self.layers:
            l.forward()

    def backward(self, after_layer=None):
        for l in reversed(self.layers):
            l.backward()
            if after_layer is not None and l.grads:
                after_layer(l)


def sparse_cce_step(model: Model, labels: torch.Tensor, opt: SGD, pops=None, global_batch=None):
    """One training step with softmax + sparse categorical cross-entropy:
    zero grads, forward, loss gradient (scale 1/global batch, lib/runtime/src/loss_functions.cc:68-72),
    backward, gradient sync (sum over ranks, no division), SGD."""
    model.zero_gradients()
    model.forward()
    probs = model.output
    rows = probs.data.numel() // probs.shape[-1]
    scale = 1.0 / float(global_batch or rows)
    K.sparse_categorical_crossentropy_loss_backward_kernel(None, probs.grad, probs.data, labels, rows,
                                                           probs.shape[-1], 1, scale)
    _backward_sync_update(model, opt, pops)


def _backward_sync_update(model: Model, opt: SGD, pops):
    """backward, data-parallel gradient sum (SUM over ranks, no division), optimizer. `pops` is a
    ParallelOps (NCCL all-reduce of the flat buckets after backward) or a PeerGradSync (peer-memory
    sums issued under the backward pass; the optimizer reads the summed buckets)."""
    if pops is not None and hasattr(pops, "after_layer"):
        model.backward(pops.after_layer)
        pops.finish()
        opt.step(model.params, pops.summed_grads)
        return
    model.backward()
    if pops is not None:
        pops.allreduce_gradients(model.grad_buckets)
    opt.step(model.params, model.grads)


def mse_step(model: Model, target: torch.Tensor, opt: SGD, pops=None, replicas=None):
    """One step with the mean-squared-error (avg) loss: scale = 2 / global volume (the volume of
    all `replicas` data-parallel batches; default: the size of `pops`' group)."""
    model.zero_gradients()
    model.forward()
    out = model.output
    per_rank = out.data.numel()
    world = replicas if replicas else (1 if pops is None else pops.world)
    scale = 2.0 / float(per_rank * world)
    K.mean_squared_error_avg_loss_backward_kernel(None, out.grad, out.data, target, scale)
    _backward_sync_update(model, opt, pops)


def build_mlp(batch=64, dims=(784, 512, 512, 10), dtype=torch.float32, device="cuda", fuse=False):
    """BASELINE configs[0]: MLP 784-512-512-10, ReLU, softmax head."""
    x = Tensor((batch, dims[0]), dtype, device, requires_grad=True)
    layers: List[Layer] = []
    cur = x
    for i, d in enumerate(dims[1:]):
        last = i == len(dims) - 2
        lin = Linear(cur, d, K.ACT_NONE if last else K.ACT_RELU, bias=True, seed=i)
        layers.append(lin)
        cur = lin.y
    sm = Softmax(cur)
    layers.append(sm)
    return Model(x, layers, sm.y, fuse=fuse)


def build_transformer_encoder(num_layers=12, d_model=1024, heads=16, seq=512, batch=64, dim_ff=4096,
                              head_dim: Optional[int] = None, dropout=0.1, dtype=torch.bfloat16,
                              device="cuda", seed=0, fuse=False):
    """BASELINE configs[2]: the encoder stack lib/models builds
    (lib/models/src/models/transformer/transformer.cc:32-65): MHA(x,x,x) -> add -> LayerNorm ->
    dense(dim_ff, ReLU) -> dropout -> dense(d_model) -> dropout -> add -> LayerNorm.
    head_dim None reproduces the reference's kdim = vdim = dim_ff / heads.

    fuse=True builds the same graph for fused-operator execution: per layer
    [MHA + residual] -> LayerNorm -> [dense + ReLU + dropout] -> [dense + dropout + residual] ->
    LayerNorm — five launches' worth of layers instead of nine, same seeds, same dropout masks, same
    results up to the rounding of the intermediates that are no longer stored."""
    hd = head_dim if head_dim is not None else dim_ff // heads
    x = Tensor((batch, seq, d_model), dtype, device)
    layers: List[Layer] = []
    cur = x
    if fuse:
        for i in range(num_layers):
            att = SelfAttention(cur, heads, hd, hd, seed=seed + i, residual=True)
            ln1 = LayerNorm(att.y)
            ff1 = Linear(ln1.y, dim_ff, K.ACT_RELU, True, seed=100 + i,
                         dropout=(dropout, 12345 + 2 * i) if dropout > 0 else None)
            ff2 = Linear(ff1.y, d_model, K.ACT_NONE, True, seed=200 + i,
                         dropout=(dropout, 12346 + 2 * i) if dropout > 0 else None, residual=ln1.y)
            ln2 = LayerNorm(ff2.y)
            layers += [att, ln1, ff1, ff2, ln2]
            cur = ln2.y
        return Model(x, layers, cur, fuse=True)
    for i in range(num_layers):
        att = SelfAttention(cur, heads, hd, hd, seed=seed + i)
        add1 = Add(cur, att.y)
        ln1 = LayerNorm(add1.y)
        ff1 = Linear(ln1.y, dim_ff, K.ACT_RELU, True, seed=100 + i)
        dp1 = Dropout(ff1.y, dropout, 12345 + 2 * i)
        ff2 = Linear(dp1.y, d_model, K.ACT_NONE, True, seed=200 + i)
        dp2 = Dropout(ff2.y, dropout, 12346 + 2 * i)
        add2 = Add(ln1.y, dp2.y)
        ln2 = LayerNorm(add2.y)
        layers += [att, add1, ln1, ff1, dp1, ff2, dp2, add2, ln2]
        cur = ln2.y
    return Model(x, layers, cur)


def build_bert(num_layers=12, hidden=768, heads=12, dim_ff=3072, seq=512, batch=64, vocab=30522,
               dropout=0.1, layer_norm_eps=1e-12, dtype=torch.bfloat16, device="cuda", seed=0,
               tp_rank=0, tp_world=1, tp_group=None, fuse=False):
    """BASELINE configs[3] (data-parallel form): the graph lib/models builds for BERT
    (lib/models/src/models/bert/bert.cc:52-155) — the input is the already embedded sequence
    [batch, seq, hidden]; each layer is MHA(x,x,x) with kdim = vdim = dim_ff / heads -> add ->
    LayerNorm -> dense(dim_ff, GELU) -> dropout -> dense(hidden) -> dropout -> add -> LayerNorm;
    the head is dense(vocab, GELU) -> softmax. The reference passes attention dropout and bias flags
    that its MHA kernel ignores (attention_kernels.cu:94,109); so does this."""
    hd = dim_ff // heads
    x = Tensor((batch, seq, hidden), dtype, device)
    layers: List[Layer] = []
    cur = x
    tp = tp_world > 1
    if fuse and not tp:
        # fused-operator execution: [MHA + residual] -> LayerNorm -> [dense + GELU + dropout, the
        # pre-activation stored by the same GEMM] -> [dense + dropout + residual] -> LayerNorm
        for i in range(num_layers):
            att = SelfAttention(cur, heads, hd, hd, seed=seed + i, residual=True)
            ln1 = LayerNorm(att.y, eps=layer_norm_eps)
            ff1 = Linear(ln1.y, dim_ff, K.ACT_GELU, True, seed=100 + i,
                         dropout=(dropout, 22345 + 2 * i) if dropout > 0 else None)
            ff2 = Linear(ff1.y, hidden, K.ACT_NONE, True, seed=200 + i,
                         dropout=(dropout, 22346 + 2 * i) if dropout > 0 else None, residual=ln1.y)
            ln2 = LayerNorm(ff2.y, eps=layer_norm_eps)
            layers += [att, ln1, ff1, ff2, ln2]
            cur = ln2.y
        head = Linear(cur, vocab, K.ACT_GELU, True, seed=300)
        sm = Softmax(head.y)
        layers += [head, sm]
        return Model(x, layers, sm.y, fuse=True)
    if tp:
        assert heads % tp_world == 0 and dim_ff % tp_world == 0
    sh = (tp_rank, tp_world) if tp else None
    for i in range(num_layers):
        # tensor-parallel form (the placement a searched strategy gives BERT): heads and the
        # feed-forward width are split over the group; Replicate in front of each split block,
        # Reduction behind it; LayerNorm, residuals and the head stay replicated
        a_in = ReplicateIn(cur, tp_group) if tp else None
        att = SelfAttention(a_in.y if tp else cur, heads // tp_world, hd, hd, seed=seed + i, shard=sh)
        a_out = ReduceSum(att.y, tp_group) if tp else None
        add1 = Add(a_out.y if tp else att.y, cur)
        ln1 = LayerNorm(add1.y, eps=layer_norm_eps)
        f_in = ReplicateIn(ln1.y, tp_group) if tp else None
        ff1 = Linear(f_in.y if tp else ln1.y, dim_ff // tp_world, K.ACT_GELU, True, seed=100 + i, shard_out=sh)
        dp1 = Dropout(ff1.y, dropout, 22345 + 2 * i + 7919 * tp_rank)
        ff2 = Linear(dp1.y, hidden, K.ACT_NONE, tp_rank == 0, seed=200 + i, shard_in=sh)
        f_out = ReduceSum(ff2.y, tp_group) if tp else None
        dp2 = Dropout(f_out.y if tp else ff2.y, dropout, 22346 + 2 * i)
        add2 = Add(ln1.y, dp2.y)
        ln2 = LayerNorm(add2.y, eps=layer_norm_eps)
        layers += [l for l in (a_in, att, a_out, add1, ln1, f_in, ff1, dp1, ff2, f_out, dp2, add2, ln2) if l is not None]
        cur = ln2.y
    head = Linear(cur, vocab, K.ACT_GELU, True, seed=300)
    sm = Softmax(head.y)
    layers += [head, sm]
    return Model(x, layers, sm.y)


# ------------------------------------------------------------------ models from a parallel computation graph
def build_from_strategy(strategy, batch: int, dtype=torch.bfloat16, device="cuda", tp_rank=0, tp_group=None,
                        seed=0, input_shape=None, fuse=False):
    """The layer list of a model described by a parallel computation graph in the reference's v1
    file format (strategy.load_strategy -> Strategy.placements): every compute layer of the graph
    becomes a step-driver layer, every parallel operator on an ACTIVATION becomes the matching
    collective layer of the tensor-parallel group —

        Replicate   -> ReplicateIn   (identity forward, all-reduce backward)
        Reduction   -> ReduceSum     (all-reduce forward, copy backward)
        Repartition -> RepartitionSlice along a feature dimension (local slab / all-gather + add)
        Combine     -> CombineGather along a feature dimension (all-gather / local slab)

    — and the Repartition operators found on a WEIGHT chain shard that weight: heads for
    multi_head_attention, in- or out-channels for linear (the reference lays a projection out as
    [in_channels][out_channels], op-attrs/ops/linear.cc:42-50, so ff dim 0 is the row-parallel and
    ff dim 1 the column-parallel split). Batch-dimension Repartition / Combine and weight Replicate
    are data parallelism: each rank simply holds its `batch` samples. Supported layers:
    multi_head_attention (self-attention), linear, layer_norm, dropout, element_binary(add), softmax."""
    from .parallel_ops import ParallelOps
    tp = strategy.tensor_parallel
    pops = ParallelOps(group=tp_group) if tp > 1 else None
    tensors: Dict[int, Tensor] = {}      # compute / input node id -> its output activation
    par_cache: Dict[int, Tensor] = {}    # parallel-op node id -> the activation behind it
    layers: List[Layer] = []
    model_input = None
    counters = {"linear": 0, "multi_head_attention": 0, "dropout": 0}

    def through(chain, t: Tensor) -> Tensor:
        for op in chain:
            if op.node in par_cache:
                t = par_cache[op.node]
                continue
            if op.op in ("repartition", "combine") and op.dim == 0:
                pass                                   # the batch split: this rank's samples are `t`
            elif tp == 1 or op.degree == 1:
                pass
            elif op.op == "replicate":
                l = ReplicateIn(t, tp_group)
                layers.append(l)
                t = l.y
            elif op.op == "reduction":
                l = ReduceSum(t, tp_group)
                layers.append(l)
                t = l.y
            elif op.op == "combine":
                l = CombineGather(t, op.dim, pops)
                layers.append(l)
                t = l.y
            else:
                l = RepartitionSlice(t, op.dim, pops)
                layers.append(l)
                t = l.y
            par_cache[op.node] = t
        return t

    def weight_shard(chain, dim):
        for op in chain:
            if op.op == "repartition" and op.dim == dim and op.degree > 1:
                assert op.degree == tp, f"weight shard degree {op.degree} != tensor-parallel degree {tp}"
                return (tp_rank, tp)
        return None

    # fused-operator execution on the imported graph: a dropout that is the only reader of a dense
    # layer's output (no parallel operator between them) runs in that layer's GEMM epilogue
    readers: Dict[int, List] = {}
    for pl in strategy.placements:
        for src in pl.producers:
            readers.setdefault(src, []).append(pl)
    by_node = {pl.node: pl for pl in strategy.placements}
    fused_dropout: Dict[int, Any] = {}     # linear node -> the dropout placement folded into it
    skip_nodes = set()
    if fuse:
        for pl in strategy.placements:
            if pl.kind == "dropout" and len(pl.producers) == 1 and not pl.inputs[0]:
                src = by_node.get(pl.producers[0])
                if src is not None and src.kind == "linear" and len(readers.get(src.node, [])) == 1 and not src.output:
                    fused_dropout[src.node] = pl
                    skip_nodes.add(pl.node)

    out = None
    for pl in strategy.placements:
        a = pl.attrs
        if pl.node in skip_nodes:
            continue
        if pl.kind == "input":
            shape = input_shape
            if shape is None:
                dims = (((a.get("tensor_shape") or {}).get("dims") or {}).get("ff_ordered")) or []
                shape = (batch,) + tuple(int(d) for d in dims[1:])
            model_input = Tensor(tuple(shape), dtype, device)
            tensors[pl.node] = model_input
            out = model_input
            continue
        ins = [through(chain, tensors[src]) for chain, src in zip(pl.inputs, pl.producers)]
        if pl.kind == "multi_head_attention":
            assert all(t is ins[0] for t in ins), "only self-attention (q = k = v) is supported"
            heads = int(a["num_heads"])
            sh = weight_shard(pl.weights[0], 1) if pl.weights else None
            i = counters["multi_head_attention"]
            counters["multi_head_attention"] += 1
            l = SelfAttention(ins[0], heads // (tp if sh else 1), int(a["kdim"]), int(a["vdim"]), seed=seed + i, shard=sh)
        elif pl.kind == "linear":
            act = a.get("activation")
            act = {None: K.ACT_NONE, "RELU": K.ACT_RELU, "SIGMOID": K.ACT_SIGMOID, "TANH": K.ACT_TANH,
                   "GELU": K.ACT_GELU}[act.upper() if isinstance(act, str) else act]
            w = pl.weights[0] if pl.weights else []
            col, row = weight_shard(w, 1), weight_shard(w, 0)
            i = counters["linear"]
            counters["linear"] += 1
            out_ch = int(a["out_channels"])
            use_bias = bool(a.get("use_bias", True))
            if row is not None and tp_rank != 0:
                use_bias = False                      # the bias of a row-parallel layer is added once
            drop = None
            dpl = fused_dropout.get(pl.node)
            if dpl is not None and float(dpl.attrs["rate"]) > 0:
                j = counters["dropout"]
                counters["dropout"] += 1
                drop = (float(dpl.attrs["rate"]), int(dpl.attrs.get("seed", 12345 + j)))
            l = Linear(ins[0], out_ch // (tp if col else 1), act, use_bias, seed=int(a.get("seed", 100 + i)),
                       shard_out=col, shard_in=row, dropout=drop)
        elif pl.kind == "layer_norm":
            l = LayerNorm(ins[0], eps=float(a.get("eps", 1e-5)))
        elif pl.kind == "dropout":
            i = counters["dropout"]
            counters["dropout"] += 1
            # dropout on a tensor every tensor-parallel rank holds in full must draw the same mask everywhere
            l = Dropout(ins[0], float(a["rate"]), int(a.get("seed", 12345 + i)))
        elif pl.kind == "element_binary":
            assert str(a.get("type", "EW_ADD")).upper().endswith("ADD"), "only element_binary add is supported"
            l = Add(ins[0], ins[1])
        elif pl.kind == "softmax":
            l = Softmax(ins[0])
        else:
            raise ValueError(f"parallel computation graph: layer kind {pl.kind!r} is not supported by the step driver")
        layers.append(l)
        tensors[pl.node] = l.y
        out = through(pl.output, l.y)
        dpl = fused_dropout.get(pl.node)
        if dpl is not None:                 # the folded dropout's readers see this layer's output
            tensors[dpl.node] = l.y
            out = through(dpl.output, l.y)
        # later consumers reach this output through its trailing parallel ops, which are cached by node id
    return Model(model_input, layers, out, fuse=fuse)


# ------------------------------------------------------------------ DLRM (BASELINE configs[4])
class Concat(Layer):
    """concat along the last dimension of 2-D activations (DLRM's "cat" interaction)."""

    def __init__(self, inputs: List[Tensor]):
        self.inputs = inputs
        rows = inputs[0].shape[0]
        width = sum(t.shape[1] for t in inputs)
        self.y = Tensor((rows, width), inputs[0].data.dtype, inputs[0].data.device)

    def forward(self):
        K.Concat.forward_kernel(None, self.y.data, [t.data for t in self.inputs], 1)

    def backward(self):
        K.Concat.backward_kernel(None, self.y.grad, [t.grad for t in self.inputs], 1)


class ShardedEmbeddings(Layer):
    """The embedding tables of a DLRM, table-sharded over the ranks (model parallel), feeding
    data-parallel MLPs: rank r owns tables [r*tl, (r+1)*tl), looks up the WHOLE global batch in
    them and the results are repartitioned along the batch dimension with one all-to-all, so that
    every rank ends with all tables' vectors for its own `batch` rows; backward runs the same
    exchange in reverse and scatters into the local tables. With one rank it is plain lookups.
    Tables and their gradients stay out of the data-parallel gradient buckets."""

    def __init__(self, indices: torch.Tensor, entries: int, dim: int, batch: int, tables: int, rank=0,
                 world=1, group=None, seed=0, device="cuda", sparse_update=True):
        """sparse_update: plain SGD touches only the looked-up rows (ffk_embedding_sgd_update scatters
        -lr * dOut straight into the table): no gradient table, no zero fill, no dense optimizer pass.
        With momentum or weight decay the dense path (gradient table + ffk_sgd_update) is used."""
        assert tables % world == 0
        self.sparse_update = sparse_update
        self.tl, self.T, self.B, self.W, self.rank, self.group = tables // world, tables, batch, world, rank, group
        self.dim, self.bag = dim, indices.shape[-1]
        assert tuple(indices.shape[:2]) == (self.tl, world * batch)
        self.idx = indices.contiguous()
        self.tables, self.dtables = [], []
        for j in range(self.tl):
            g = torch.Generator(device=device).manual_seed(5000 + seed + rank * self.tl + j)
            bound = (1.0 / entries) ** 0.5
            self.tables.append((torch.rand(entries, dim, device=device, generator=g) * 2 - 1) * bound)
            self.dtables.append(None if sparse_update else torch.zeros(entries, dim, device=device))
        self._grad_src = None
        self.y = Tensor((tables, batch, dim), torch.float32, device)          # [table][row][dim] for MY rows
        self.send = torch.zeros(world, self.tl, batch, dim, device=device)    # [dest rank][local table][row][dim]
        self.gsend = torch.zeros_like(self.send) if world > 1 else None
        self.params, self.grads = [], []

    def views(self) -> List[Tensor]:
        """One activation per table (rows of this rank), sharing storage with y."""
        out = []
        for t in range(self.T):
            v = Tensor.__new__(Tensor)
            v.data, v.grad = self.y.data[t], self.y.grad[t]
            out.append(v)
        return out

    def forward(self):
        dst = self.send if self.W > 1 else self.y.data.view(1, self.tl, self.B, self.dim)
        for j in range(self.tl):
            for q in range(self.W):
                K.Embedding.forward_kernel(None, self.idx[j, q * self.B:(q + 1) * self.B], dst[q, j],
                                           self.tables[j], K.AGGR_SUM, self.bag, self.dim, self.B)
        if self.W > 1:
            dist.all_to_all_single(self.y.data.view(-1), self.send.view(-1), group=self.group)

    def backward(self):
        if self.W > 1:
            dist.all_to_all_single(self.gsend.view(-1), self.y.grad.view(-1), group=self.group)
            src = self.gsend
        else:
            src = self.y.grad.view(1, self.tl, self.B, self.dim)
        self._grad_src = src
        if self.sparse_update:
            return                      # the rows are updated straight from dOut in update()
        for j in range(self.tl):
            _fill0(self.dtables[j])
            for q in range(self.W):
                K.Embedding.backward_kernel(None, self.idx[j, q * self.B:(q + 1) * self.B], src[q, j],
                                            self.dtables[j], K.AGGR_SUM, self.bag, self.dim, self.B)

    def update(self, opt: "SGD"):
        if self.sparse_update:
            assert opt.momentum == 0 and opt.weight_decay == 0, "row-sparse update is plain SGD only"
            src = self._grad_src
            for j in range(self.tl):
                for q in range(self.W):
                    K.Embedding.sgd_update_kernel(None, self.idx[j, q * self.B:(q + 1) * self.B], src[q, j],
                                                  self.tables[j], opt.lr, K.AGGR_SUM, self.bag, self.dim, self.B)
            return
        for w, g in zip(self.tables, self.dtables):
            K.sgd_ps_update_task_gpu(None, opt.lr, 0.0, False, opt.weight_decay, g, w.numel(), 1, w, None, None)


class DLRM:
    """examples/cpp/DLRM/dlrm.cc:139-166: dense features -> bottom MLP; one embedding lookup per
    sparse feature; "cat" interaction; top MLP (sigmoid on its last layer); MSE loss.
    Defaults are those of the example's run script (examples/cpp/DLRM/run_random.sh:9): 8 tables of
    1 M x 64, bag size 1, bottom MLP 64-512-512-64, top MLP 576-1024-1024-1024-1 (576 = 64 + 8 x 64
    is the width of the concatenation, so mlp_top lists the layer widths behind it)."""

    def __init__(self, batch=8192, embedding_size=(1000000,) * 8, sparse_feature_size=64, bag=1,
                 mlp_bot=(64, 512, 512, 64), mlp_top=(1024, 1024, 1024, 1), rank=0, world=1, group=None,
                 device="cuda", seed=0, sparse_update=True):
        self.batch, self.world, self.rank = batch, world, rank
        T = len(embedding_size)
        assert len(set(embedding_size)) == 1, "one table size (the example's default)"
        tl = T // world
        # synthetic indices for the GLOBAL batch, the same on every rank (common seed); a rank
        # keeps the rows of its own tables only
        g = torch.Generator(device=device).manual_seed(99 + seed)
        all_idx = torch.randint(0, embedding_size[0], (T, world * batch, bag), device=device, generator=g,
                                dtype=torch.int64)
        self.dense = Tensor((batch, mlp_bot[0]), torch.float32, device)
        layers: List[Layer] = []
        cur = self.dense
        for i, d in enumerate(mlp_bot[1:]):
            lin = Linear(cur, d, K.ACT_RELU, True, seed=400 + i)
            layers.append(lin)
            cur = lin.y
        self.emb = ShardedEmbeddings(all_idx[rank * tl:(rank + 1) * tl], embedding_size[0], sparse_feature_size,
                                     batch, T, rank, world, group, seed, device, sparse_update=sparse_update)
        layers.append(self.emb)
        cat = Concat([cur] + self.emb.views())
        layers.append(cat)
        cur = cat.y
        for i, d in enumerate(mlp_top):
            last = i == len(mlp_top) - 1
            lin = Linear(cur, d, K.ACT_SIGMOID if last else K.ACT_RELU, True, seed=500 + i)
            layers.append(lin)
            cur = lin.y
        self.model = Model(self.dense, layers, cur)


def dlrm_step(net: DLRM, target: torch.Tensor, opt: SGD, pops=None):
    """zero grads, forward (with the table exchange), MSE gradient, backward, data-parallel sum
    of the MLP gradients, SGD on the MLPs and on this rank's tables."""
    m = net.model
    m.zero_gradients()
    m.forward()
    out = m.output
    scale = 2.0 / float(out.data.numel() * net.world)
    K.mean_squared_error_avg_loss_backward_kernel(None, out.grad, out.data, target, scale)
    m.backward()
    if pops is not None:
        pops.allreduce_gradients(m.grad_buckets)
    opt.step(m.params, m.grads)
    net.emb.update(opt)
