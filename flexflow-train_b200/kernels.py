"""Host-side mirror of the reference's `FlexFlow::Kernels::<Op>` namespaces over torch CUDA
tensors. Each class corresponds to one `lib/kernels/include/kernels/*_kernels.h` header and
keeps its function names (`init_kernel`, `forward_kernel`, `backward_kernel`); the bodies
only unpack device pointers and call the C ABI in libffk.so. torch is used for device
memory and streams, nothing else: there is no torch arithmetic on these paths and no CPU
fallback (a missing libffk.so raises at import of the first op).
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional, Sequence

import torch

from ._lib import lib

# ffk_dtype_t
BOOL, INT32, INT64, HALF, FLOAT, DOUBLE, BFLOAT16 = range(7)
_DT = {
    torch.bool: BOOL,
    torch.int32: INT32,
    torch.int64: INT64,
    torch.float16: HALF,
    torch.float32: FLOAT,
    torch.float64: DOUBLE,
    torch.bfloat16: BFLOAT16,
}

ACT_NONE, ACT_RELU, ACT_SIGMOID, ACT_TANH, ACT_GELU = -1, 0, 1, 2, 3
GEMM_AUTO, GEMM_TENSOR, GEMM_SIMT, GEMM_SPLIT_BF16 = 0, 1, 2, 3


class OperatorType:
    """Values of the reference's OperatorType enum that the kernels accept."""

    SCALAR_MULTIPLY, SCALAR_ADD, SCALAR_TRUE_DIV, SCALAR_SUB = 8, 9, 11, 12
    RELU, IDENTITY, SIGMOID, TANH, ELU = 13, 14, 15, 16, 17
    EW_ADD, EW_MUL, EW_SUB, EW_DIV, EW_MAX, EW_MIN = 28, 29, 35, 36, 40, 41
    REDUCE_MEAN, REDUCE_SUM = 45, 48
    EXP, SIN, COS, GELU, RSQRT, POW = 56, 61, 62, 67, 70, 71


def dtype_code(t: torch.Tensor) -> int:
    return _DT[t.dtype]


def _ptr(t: Optional[torch.Tensor]):
    if t is None:
        return None
    if not t.is_cuda:
        raise ValueError("kernel operands must live on the GPU (no CPU fallback exists)")
    if not t.is_contiguous():
        raise ValueError("kernel operands must be dense row-major")
    return t.data_ptr()


def _stream(stream=None):
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


def _dims(shape: Sequence[int]):
    return (ctypes.c_longlong * len(shape))(*shape)


def _call(name, *args):
    return lib().call(name, *args)


# ------------------------------------------------------------------ dense
_gemm_ws = {}  # device index -> registered scratch tensor


def set_gemm_workspace(nbytes_or_tensor, fp32_tensor_math=False):
    """Registers scratch for the GEMM-based entry points on the CURRENT device
    (ffk_set_gemm_workspace): unaligned 16-bit operands are repacked into it. fp32_tensor_math
    (ffk_set_gemm_fp32_tensor_math, the reference handle's allowTensorOpMathConversion) lets
    GEMM_AUTO run large fp32 products on the tensor cores through the bf16 split.
    Pass None to unregister (and switch the fp32 policy off)."""
    dev = torch.cuda.current_device()
    if nbytes_or_tensor is None:
        _gemm_ws.pop(dev, None)
        _call("ffk_set_gemm_workspace", None, 0)
        _call("ffk_set_gemm_fp32_tensor_math", 0)
        return None
    t = nbytes_or_tensor
    if not isinstance(t, torch.Tensor):
        t = torch.empty(int(t), dtype=torch.uint8, device="cuda")
    _gemm_ws[dev] = t  # keep it alive
    _call("ffk_set_gemm_workspace", t.data_ptr(), t.numel() * t.element_size())
    _call("ffk_set_gemm_fp32_tensor_math", 1 if fp32_tensor_math else 0)
    return t


def set_gemm_kcut(enable: bool):
    """ffk_set_gemm_kcut: the K-cut schedule of 16-bit products on the CURRENT device (default on)."""
    _call("ffk_set_gemm_kcut", 1 if enable else 0)


class Gemm:
    @staticmethod
    def run(a, b, d, M, N, K, batch=1, a_mn=False, b_mn=False, alpha=1.0, accumulate=False,
            bias=None, act=ACT_NONE, path=GEMM_AUTO, stream=None):
        lda, ldb = a.shape[-1], b.shape[-1]
        _call("ffk_gemm", _stream(stream), _ptr(a), _ptr(b), _ptr(d), M, N, K, batch, lda, ldb, N,
              a.shape[-2] * lda, b.shape[-2] * ldb, M * N, dtype_code(a), dtype_code(d), int(a_mn),
              int(b_mn), alpha, int(accumulate), _ptr(bias), act, path)
        return d


@dataclass
class LinearPerDeviceState:
    activation: int = ACT_NONE
    l2_lambda: float = 0.0
    use_bias: bool = True
    path: int = GEMM_AUTO
    workspace: Optional[torch.Tensor] = None  # PerDeviceFFHandle::workSpace stand-in


class Linear:
    """linear_kernels.h"""

    @staticmethod
    def init_kernel(activation=ACT_NONE, regularizer_lambda=0.0, use_bias=True, path=GEMM_AUTO,
                    out_dim=0, device="cuda"):
        ws = None
        if out_dim and use_bias:
            nbytes = lib().cdll.ffk_linear_backward_workspace_bytes(out_dim)
            ws = torch.empty(nbytes // 4, dtype=torch.float32, device=device)
        return LinearPerDeviceState(activation, regularizer_lambda, use_bias, path, ws)

    @staticmethod
    def forward_kernel(stream, m, input, output, weight, bias, in_dim, out_dim, batch_size):
        _call("ffk_linear_forward", _stream(stream), _ptr(input), _ptr(output), _ptr(weight),
              _ptr(bias), in_dim, out_dim, batch_size, dtype_code(input), m.activation, m.path)

    @staticmethod
    def backward_kernel(stream, m, input, input_grad, output, output_grad, weight, weight_grad,
                        bias_grad, in_dim, out_dim, batch_size, preact=None):
        _call("ffk_linear_backward", _stream(stream), _ptr(input), _ptr(input_grad), _ptr(output),
              _ptr(output_grad), _ptr(weight), _ptr(weight_grad), _ptr(bias_grad), in_dim, out_dim,
              batch_size, dtype_code(input), m.activation, m.l2_lambda, _ptr(preact), m.path,
              _ptr(m.workspace), m.workspace.numel() * 4 if m.workspace is not None else 0)


class _LinearForwardFusion(ctypes.Structure):      # ffk_linear_forward_fusion_t
    _fields_ = [("preact_out", ctypes.c_void_p), ("residual", ctypes.c_void_p), ("dropout_mask", ctypes.c_void_p),
                ("dropout_rate", ctypes.c_float), ("dropout_seed", ctypes.c_ulonglong),
                ("dropout_offset", ctypes.c_ulonglong)]


class _LinearBackwardFusion(ctypes.Structure):     # ffk_linear_backward_fusion_t
    _fields_ = [("accumulate_input_grad", ctypes.c_int), ("output_grad_is_preactivation", ctypes.c_int),
                ("dropout_mask", ctypes.c_void_p), ("dropout_rate", ctypes.c_float),
                ("input_gate", ctypes.c_void_p), ("input_gate_scale", ctypes.c_float)]


class FusedLinear:
    """Fused-operator forms of Linear (ffk_linear_forward_fused / ffk_linear_backward_fused):
    Linear -> activation -> Dropout -> + residual as one GEMM; backward with the dropout / activation
    gradient in one pass, a first-writer (overwrite) data gradient, and a data gradient gated by the
    stored ReLU(+dropout) output of the layer before."""

    @staticmethod
    def forward_kernel(stream, m, input, output, weight, bias, in_dim, out_dim, batch_size, preact_out=None,
                       residual=None, dropout=None):
        """dropout: a DropoutPerDeviceState (rate, seed, mask, offset); its offset advances per call."""
        fu = _LinearForwardFusion(_ptr(preact_out), _ptr(residual), _ptr(dropout.mask) if dropout else None,
                                  dropout.rate if dropout else 0.0, dropout.seed if dropout else 0,
                                  dropout.offset if dropout else 0)
        _call("ffk_linear_forward_fused", _stream(stream), _ptr(input), _ptr(output), _ptr(weight), _ptr(bias),
              in_dim, out_dim, batch_size, dtype_code(input), m.activation, m.path, ctypes.byref(fu))
        if dropout:
            dropout.offset += 1

    @staticmethod
    def backward_kernel(stream, m, input, input_grad, output, output_grad, weight, weight_grad, bias_grad, in_dim,
                        out_dim, batch_size, preact=None, accumulate_input_grad=True,
                        output_grad_is_preactivation=False, dropout=None, input_gate=None, input_gate_scale=1.0):
        fu = _LinearBackwardFusion(int(accumulate_input_grad), int(output_grad_is_preactivation),
                                   _ptr(dropout.mask) if dropout else None, dropout.rate if dropout else 0.0,
                                   _ptr(input_gate), float(input_gate_scale))
        _call("ffk_linear_backward_fused", _stream(stream), _ptr(input), _ptr(input_grad), _ptr(output),
              _ptr(output_grad), _ptr(weight), _ptr(weight_grad), _ptr(bias_grad), in_dim, out_dim, batch_size,
              dtype_code(input), m.activation, m.l2_lambda, _ptr(preact), m.path, _ptr(m.workspace),
              m.workspace.numel() * 4 if m.workspace is not None else 0, ctypes.byref(fu))


class BatchMatmul:
    """batch_matmul_kernels.h"""

    @staticmethod
    def forward_kernel(stream, output, a_input, b_input, m, n, k, batch, a_seq_length_dim=-1,
                       b_seq_length_dim=-1, seq_length=-1, path=GEMM_AUTO):
        _call("ffk_batch_matmul_forward", _stream(stream), _ptr(output), _ptr(a_input),
              _ptr(b_input), m, n, k, batch, a_seq_length_dim, b_seq_length_dim, seq_length,
              dtype_code(a_input), path)

    @staticmethod
    def backward_kernel(stream, o, o_grad, a, a_grad, b, b_grad, m, n, k, batch, path=GEMM_AUTO):
        _call("ffk_batch_matmul_backward", _stream(stream), _ptr(o), _ptr(o_grad), _ptr(a),
              _ptr(a_grad), _ptr(b), _ptr(b_grad), m, n, k, batch, dtype_code(a), path)


# ------------------------------------------------------------------ elementwise
class ElementUnary:
    """element_unary_kernels.h"""

    @staticmethod
    def forward_kernel(stream, op_type, scalar, input, output):
        _call("ffk_element_unary_forward", _stream(stream), op_type, float(scalar or 0.0),
              _ptr(input), _ptr(output), input.numel(), dtype_code(input))

    @staticmethod
    def backward_kernel(stream, op_type, scalar, input, input_grad, output, output_grad):
        _call("ffk_element_unary_backward", _stream(stream), op_type, float(scalar or 0.0),
              _ptr(input), _ptr(input_grad), _ptr(output), _ptr(output_grad), input.numel(),
              dtype_code(input))


class ElementBinary:
    """element_binary_kernels.h"""

    @staticmethod
    def forward_kernel(stream, op_type, lhs, rhs, out):
        nd = out.dim()
        _call("ffk_element_binary_forward", _stream(stream), op_type, _ptr(lhs), _ptr(rhs),
              _ptr(out), nd, _dims(lhs.shape), _dims(rhs.shape), _dims(out.shape), dtype_code(out))

    @staticmethod
    def backward_kernel(stream, op_type, out_grad, lhs, rhs, lhs_grad, rhs_grad):
        nd = out_grad.dim()
        lshape = lhs_grad.shape if lhs_grad is not None else lhs.shape
        rshape = rhs_grad.shape if rhs_grad is not None else rhs.shape
        _call("ffk_element_binary_backward", _stream(stream), op_type, _ptr(out_grad), _ptr(lhs),
              _ptr(rhs), _ptr(lhs_grad), _ptr(rhs_grad), nd, _dims(lshape), _dims(rshape),
              _dims(out_grad.shape), dtype_code(out_grad))


class Cast:
    """cast_kernels.h"""

    @staticmethod
    def forward_kernel(stream, input, output):
        _call("ffk_cast_forward", _stream(stream), _ptr(input), _ptr(output), input.numel(),
              dtype_code(input), dtype_code(output))

    @staticmethod
    def backward_kernel(stream, src, dst):
        _call("ffk_cast_backward", _stream(stream), _ptr(src), _ptr(dst), src.numel(),
              dtype_code(src), dtype_code(dst))


def copy(stream, dst, src):
    _call("ffk_copy", _stream(stream), _ptr(dst), _ptr(src), src.numel(), dtype_code(src))


def accumulate(stream, dst, src):
    _call("ffk_accumulate", _stream(stream), _ptr(dst), _ptr(src), src.numel(), dtype_code(src))


class Flat:
    """flat_kernels.h"""
    forward_kernel = staticmethod(lambda stream, input, output: copy(stream, output, input))
    backward_kernel = staticmethod(
        lambda stream, input, input_grad, output_grad: accumulate(stream, input_grad, output_grad))


class Reshape:
    """reshape_kernels.h"""
    forward_kernel = staticmethod(lambda stream, input, output: copy(stream, output, input))
    backward_kernel = staticmethod(
        lambda stream, input_grad, output_grad: accumulate(stream, input_grad, output_grad))


class Combine:
    """combine_kernels.h (single-device form; the multi-GPU form lives in parallel_ops.py)"""
    forward_kernel = staticmethod(lambda stream, input, output: copy(stream, output, input))
    backward_kernel = staticmethod(
        lambda stream, output_grad, input_grad: accumulate(stream, input_grad, output_grad))


class Repartition:
    """partition_kernels.h (single-device form)"""
    forward_kernel = staticmethod(lambda stream, input, output: copy(stream, output, input))
    # kernel contract pinned by the reference test: first (W) accessor += second (R) accessor
    backward_kernel = staticmethod(
        lambda stream, output_grad, input_grad: accumulate(stream, output_grad, input_grad))


class Replicate:
    """replicate_kernels.h (single-device form)"""
    forward_kernel = staticmethod(lambda stream, input, output: copy(stream, output, input))

    @staticmethod
    def backward_kernel(stream, input_grad, output_grad, num_replicas):
        _call("ffk_sum_replicas", _stream(stream), _ptr(input_grad), _ptr(output_grad),
              input_grad.numel(), num_replicas, 1, dtype_code(input_grad))


class Reduction:
    """reduction_kernels.h (single-device form)"""

    @staticmethod
    def forward_kernel(stream, input, output, num_replicas):
        _call("ffk_sum_replicas", _stream(stream), _ptr(output), _ptr(input), output.numel(),
              num_replicas, 0, dtype_code(output))

    backward_kernel = staticmethod(
        lambda stream, input_grad, output_grad: copy(stream, input_grad, output_grad))


@dataclass
class DropoutPerDeviceState:
    rate: float
    seed: int
    mask: torch.Tensor
    offset: int = 0


class Dropout:
    """dropout_kernels.h"""

    @staticmethod
    def init_kernel(rate, seed, numel, device="cuda"):
        nbytes = lib().cdll.ffk_dropout_mask_bytes(numel)
        mask = torch.zeros(max(nbytes // 4, 1), dtype=torch.int32, device=device)
        return DropoutPerDeviceState(rate, seed, mask)

    @staticmethod
    def forward_kernel(stream, m, input, output):
        _call("ffk_dropout_forward", _stream(stream), _ptr(input), _ptr(output), _ptr(m.mask),
              input.numel(), m.rate, m.seed, m.offset, dtype_code(input))
        m.offset += 1

    @staticmethod
    def backward_kernel(stream, m, output_grad, input_grad):
        _call("ffk_dropout_backward", _stream(stream), _ptr(output_grad), _ptr(input_grad),
              _ptr(m.mask), output_grad.numel(), m.rate, dtype_code(output_grad))


# ------------------------------------------------------------------ normalisations
class Softmax:
    """softmax_kernels.h"""

    @staticmethod
    def forward_kernel(stream, input, output, outer, dim, inner=1):
        _call("ffk_softmax_forward", _stream(stream), _ptr(input), _ptr(output), outer, dim, inner,
              dtype_code(input))

    @staticmethod
    def backward_kernel(stream, input_grad, output_grad, output=None, outer=0, dim=0, inner=1,
                        mode=0):
        if mode == 0:
            outer, dim, inner = output_grad.numel(), 1, 1
        _call("ffk_softmax_backward", _stream(stream), _ptr(input_grad), _ptr(output_grad),
              _ptr(output), outer, dim, inner, dtype_code(output_grad), mode)


@dataclass
class LayerNormPerDeviceState:
    elementwise_affine: bool
    effective_batch_size: int
    effective_num_elements: int
    eps: float
    mean: torch.Tensor
    rstd: torch.Tensor
    workspace: Optional[torch.Tensor]


class LayerNorm:
    """layer_norm_kernels.h"""

    @staticmethod
    def init_kernel(elementwise_affine, effective_batch_size, effective_num_elements, eps,
                    device="cuda", workspace=True):
        M, N = effective_batch_size, effective_num_elements
        ws = None
        if workspace:
            nbytes = lib().cdll.ffk_layer_norm_backward_workspace_bytes(N)
            ws = torch.empty(max(nbytes // 4, 1), dtype=torch.float32, device=device)
        return LayerNormPerDeviceState(
            elementwise_affine, M, N, eps,
            torch.empty(M, dtype=torch.float32, device=device),
            torch.empty(M, dtype=torch.float32, device=device), ws)

    @staticmethod
    def forward_kernel(stream, m, input, output, gamma, beta):
        _call("ffk_layer_norm_forward", _stream(stream), _ptr(input), _ptr(output), _ptr(gamma),
              _ptr(beta), _ptr(m.mean), _ptr(m.rstd), m.effective_batch_size,
              m.effective_num_elements, m.eps, dtype_code(input))

    @staticmethod
    def backward_kernel(stream, m, output_grad, input, input_grad, gamma, gamma_grad, beta_grad,
                        accumulate_param_grads=False, accumulate_input_grad=True):
        """accumulate_input_grad=False (ffk_layer_norm_backward_ex) WRITES input_grad: for the first
        writer of that gradient under fused-operator execution."""
        ws = m.workspace
        if accumulate_input_grad:
            _call("ffk_layer_norm_backward", _stream(stream), _ptr(output_grad), _ptr(input),
                  _ptr(input_grad), _ptr(gamma), _ptr(gamma_grad), _ptr(beta_grad), _ptr(m.mean),
                  _ptr(m.rstd), m.effective_batch_size, m.effective_num_elements,
                  dtype_code(input), int(accumulate_param_grads), _ptr(ws),
                  ws.numel() * 4 if ws is not None else 0)
        else:
            _call("ffk_layer_norm_backward_ex", _stream(stream), _ptr(output_grad), _ptr(input),
                  _ptr(input_grad), _ptr(gamma), _ptr(gamma_grad), _ptr(beta_grad), _ptr(m.mean),
                  _ptr(m.rstd), m.effective_batch_size, m.effective_num_elements,
                  dtype_code(input), int(accumulate_param_grads), 0, _ptr(ws),
                  ws.numel() * 4 if ws is not None else 0)


@dataclass
class BatchNormPerDeviceState:
    n: int
    c: int
    h: int
    w: int
    relu: bool
    running_mean: torch.Tensor
    running_var: torch.Tensor
    save_mean: torch.Tensor
    save_invstd: torch.Tensor
    eps: float = 0.0  # CUDNN_BN_MIN_EPSILON of cuDNN 9, which the reference passes


class BatchNorm:
    """batch_norm_kernels.h"""

    @staticmethod
    def init_kernel(n, c, h, w, relu, device="cuda", eps=0.0):
        z = lambda: torch.zeros(c, dtype=torch.float32, device=device)
        return BatchNormPerDeviceState(n, c, h, w, relu, z(), z(), z(), z(), eps)

    @staticmethod
    def forward_kernel(stream, m, input, output, scale, bias):
        _call("ffk_batch_norm_forward", _stream(stream), _ptr(input), _ptr(output), _ptr(scale),
              _ptr(bias), _ptr(m.running_mean), _ptr(m.running_var), _ptr(m.save_mean),
              _ptr(m.save_invstd), m.n, m.c, m.h, m.w, m.eps, 1.0)

    @staticmethod
    def backward_kernel(stream, m, input, output_grad, output, input_grad, scale, scale_grad,
                        bias_grad):
        _call("ffk_batch_norm_backward", _stream(stream), _ptr(input), _ptr(output_grad),
              _ptr(output), _ptr(input_grad), _ptr(scale), _ptr(scale_grad), _ptr(bias_grad),
              _ptr(m.save_mean), _ptr(m.save_invstd), m.n, m.c, m.h, m.w, int(m.relu))


# ------------------------------------------------------------------ index-driven
AGGR_NONE, AGGR_SUM, AGGR_AVG = -1, 0, 1


class Embedding:
    """embedding_kernels.h"""

    @staticmethod
    def forward_kernel(stream, input, output, weight, aggr, in_dim, out_dim, batch_size):
        _call("ffk_embedding_forward", _stream(stream), _ptr(input), _ptr(output), _ptr(weight),
              dtype_code(input), dtype_code(weight), aggr, in_dim, out_dim, batch_size)

    @staticmethod
    def backward_kernel(stream, input, output_grad, weight_grad, aggr, in_dim, out_dim, batch_size):
        _call("ffk_embedding_backward", _stream(stream), _ptr(input), _ptr(output_grad),
              _ptr(weight_grad), dtype_code(input), dtype_code(weight_grad), aggr, in_dim, out_dim,
              batch_size)

    @staticmethod
    def sgd_update_kernel(stream, input, output_grad, weight, lr, aggr, in_dim, out_dim, batch_size):
        """Row-sparse SGD on the table (ffk_embedding_sgd_update): W[idx] -= lr * dOut."""
        _call("ffk_embedding_sgd_update", _stream(stream), _ptr(input), _ptr(output_grad), _ptr(weight),
              float(lr), dtype_code(input), dtype_code(weight), aggr, in_dim, out_dim, batch_size)

    @staticmethod
    def rand_generate_int(stream, t, p):
        _call("ffk_rand_generate_int", _stream(stream), _ptr(t), t.numel(), p, dtype_code(t))


class Gather:
    """gather_kernels.h — `dim` is the FF (torch) dimension gathered along."""

    @staticmethod
    def _geom(input, output, dim):
        stride = 1
        for s in output.shape[dim + 1:]:
            stride *= s
        return output.numel(), stride, input.shape[dim], output.shape[dim]

    @staticmethod
    def forward_kernel(stream, dim, input, index, output):
        n, stride, idim, odim = Gather._geom(input, output, dim)
        _call("ffk_gather_forward", _stream(stream), _ptr(input), _ptr(index), _ptr(output), n,
              stride, idim, odim, dtype_code(index), dtype_code(input))

    @staticmethod
    def backward_kernel(stream, dim, output_grad, index, input_grad):
        n, stride, idim, odim = Gather._geom(input_grad, output_grad, dim)
        _call("ffk_gather_backward", _stream(stream), _ptr(output_grad), _ptr(index),
              _ptr(input_grad), n, stride, idim, odim, dtype_code(index), dtype_code(input_grad))


class TopK:
    """topk_kernels.h"""

    @staticmethod
    def forward_kernel(stream, input, output, indices, batch_size, length, k, sorted=True):
        _call("ffk_topk_forward", _stream(stream), _ptr(input), _ptr(output), _ptr(indices),
              batch_size, length, k, int(sorted))

    @staticmethod
    def backward_kernel(stream, out_grad, indices, in_grad, batch_size, length, k):
        _call("ffk_topk_backward", _stream(stream), _ptr(out_grad), _ptr(indices), _ptr(in_grad),
              batch_size, length, k)


# ------------------------------------------------------------------ loss / optimizer / metrics
def sparse_categorical_crossentropy_loss_backward_kernel(stream, logit_grad, logit, label,
                                                         num_samples, num_classes, k, scale_factor):
    _call("ffk_sparse_categorical_crossentropy_loss_backward", _stream(stream), _ptr(logit_grad),
          _ptr(logit), _ptr(label), num_samples, num_classes, k, scale_factor, dtype_code(logit))


def categorical_crossentropy_loss_backward_kernel(stream, logit_grad, logit, label, scale_factor):
    _call("ffk_categorical_crossentropy_loss_backward", _stream(stream), _ptr(logit_grad),
          _ptr(logit), _ptr(label), logit.numel(), scale_factor, dtype_code(logit))


def mean_squared_error_avg_loss_backward_kernel(stream, logit_grad, logit, label, scale_factor):
    _call("ffk_mean_squared_error_avg_loss_backward", _stream(stream), _ptr(logit_grad),
          _ptr(logit), _ptr(label), logit.numel(), scale_factor, dtype_code(logit))


def identity_loss_backward_kernel(stream, loss_grad, loss, scale_factor):
    _call("ffk_identity_loss_backward", _stream(stream), _ptr(loss_grad), _ptr(loss),
          loss_grad.numel(), scale_factor, dtype_code(loss_grad))


def sgd_ps_update_task_gpu(stream, lr, momentum, nesterov, weight_decay, weight_grad, size,
                           num_replicas, weight, sgd_v, master_weight=None):
    _call("ffk_sgd_update", _stream(stream), lr, momentum, int(nesterov), weight_decay,
          _ptr(weight_grad), size, num_replicas, _ptr(weight), _ptr(sgd_v), dtype_code(weight),
          _ptr(master_weight))


def adam_ps_update_task_gpu(stream, alpha_t, beta1, beta2, weight_decay, epsilon, weight_grad,
                            size, num_replicas, adam_m, adam_v, weight, master_weight=None):
    _call("ffk_adam_update", _stream(stream), alpha_t, beta1, beta2, weight_decay, epsilon,
          _ptr(weight_grad), size, num_replicas, _ptr(adam_m), _ptr(adam_v), _ptr(weight),
          dtype_code(weight), _ptr(master_weight))


METRIC_ACCURACY, METRIC_CCE, METRIC_SPARSE_CCE, METRIC_MSE, METRIC_RMSE, METRIC_MAE = 1, 2, 4, 8, 16, 32


class PerfMetrics:
    """Device-side accumulator (ffk_perf_metrics_t) with the reference's accessors."""

    def __init__(self, device="cuda"):
        self.buf = torch.zeros(7, dtype=torch.int32, device=device)

    def read(self):
        raw = self.buf.cpu()
        f = raw.view(torch.float32)
        return dict(train_all=int(raw[0]), train_correct=int(raw[1]), cce_loss=float(f[2]),
                    sparse_cce_loss=float(f[3]), mse_loss=float(f[4]), rmse_loss=float(f[5]),
                    mae_loss=float(f[6]))


def update_metrics_sparse_label_kernel(stream, metrics, logit, label, num_samples, num_classes, perf):
    _call("ffk_update_metrics_sparse_label", _stream(stream), metrics, _ptr(logit), _ptr(label),
          num_samples, num_classes, _ptr(perf.buf))


def update_metrics_label_kernel(stream, metrics, logit, label, num_samples, num_classes, perf):
    _call("ffk_update_metrics_label", _stream(stream), metrics, _ptr(logit), _ptr(label),
          num_samples, num_classes, _ptr(perf.buf))


# ------------------------------------------------------------------ layout ops
def _ints(vals):
    return (ctypes.c_int * len(vals))(*vals)


class Concat:
    """concat_kernels.h — `axis` is the FF (torch) dimension."""

    @staticmethod
    def _blocks(shape, axis):
        num_blocks = 1
        for s in shape[:axis]:
            num_blocks *= s
        blk = 1
        for s in shape[axis:]:
            blk *= s
        return num_blocks, blk

    @staticmethod
    def forward_kernel(stream, output, inputs, axis):
        nb, oblk = Concat._blocks(output.shape, axis)
        off = 0
        for t in inputs:
            _, iblk = Concat._blocks(t.shape, axis)
            _call("ffk_copy_with_stride", _stream(stream), output.data_ptr() + off * output.element_size(),
                  _ptr(t), nb, oblk, iblk, dtype_code(t))
            off += iblk

    @staticmethod
    def backward_kernel(stream, output_grad, input_grads, axis):
        nb, oblk = Concat._blocks(output_grad.shape, axis)
        off = 0
        for g in input_grads:
            _, iblk = Concat._blocks(g.shape, axis)
            _call("ffk_add_with_stride", _stream(stream), _ptr(g),
                  output_grad.data_ptr() + off * output_grad.element_size(), nb, iblk, oblk,
                  dtype_code(g))
            off += iblk


class Split:
    """split_kernels.h"""

    @staticmethod
    def forward_kernel(stream, outputs, input, out_blk_sizes, in_blk_size, num_blks):
        off = 0
        for t, blk in zip(outputs, out_blk_sizes):
            _call("ffk_copy_with_stride", _stream(stream), _ptr(t),
                  input.data_ptr() + off * input.element_size(), num_blks, blk, in_blk_size,
                  dtype_code(t))
            off += blk

    @staticmethod
    def backward_kernel(stream, input_grad, output_grads, out_blk_sizes, in_blk_size, num_blks):
        off = 0
        for g, blk in zip(output_grads, out_blk_sizes):
            _call("ffk_add_with_stride", _stream(stream),
                  input_grad.data_ptr() + off * input_grad.element_size(), _ptr(g), num_blks,
                  in_blk_size, blk, dtype_code(g))
            off += blk


class Transpose:
    """transpose_kernels.h — perm in FF order, torch.permute semantics."""

    @staticmethod
    def init_kernel(num_dim, perm):
        assert len(perm) == num_dim
        return list(perm)

    @staticmethod
    def forward_kernel(stream, m, input, output):
        _call("ffk_transpose", _stream(stream), _ptr(input), _ptr(output), input.dim(),
              _dims(input.shape), _ints(m), 0, dtype_code(input))

    @staticmethod
    def backward_kernel(stream, m, in_grad, out_grad):
        inv = [0] * len(m)
        for i, p in enumerate(m):
            inv[p] = i
        _call("ffk_transpose", _stream(stream), _ptr(out_grad), _ptr(in_grad), out_grad.dim(),
              _dims(out_grad.shape), _ints(inv), 1, dtype_code(out_grad))


class Reverse:
    """reverse_kernels.h"""

    @staticmethod
    def forward_kernel(stream, in_t, out_t, num_out_blks, reverse_dim_size, in_blk_size):
        _call("ffk_reverse", _stream(stream), _ptr(in_t), _ptr(out_t), num_out_blks,
              reverse_dim_size, in_blk_size, dtype_code(in_t))

    backward_kernel = forward_kernel


class Reduce:
    """reduce_kernels.h"""

    @staticmethod
    def forward_kernel(stream, op_type, input, output):
        _call("ffk_reduce_forward", _stream(stream), _ptr(input), _ptr(output), input.dim(),
              _dims(input.shape), _dims(output.shape), op_type, dtype_code(input))

    @staticmethod
    def backward_kernel(stream, op_type, output_grad, input_grad):
        _call("ffk_reduce_backward", _stream(stream), _ptr(output_grad), _ptr(input_grad),
              input_grad.dim(), _dims(input_grad.shape), _dims(output_grad.shape), op_type,
              dtype_code(input_grad))


POOL_MAX, POOL_AVG = 0, 1


@dataclass
class Pool2DPerDeviceState:
    geom: tuple  # (n, c, ih, iw, oh, ow, kh, kw, sh, sw, ph, pw)
    pool_type: int


class Pool2D:
    """pool_2d_kernels.h"""

    @staticmethod
    def init_kernel(input_w, input_h, input_c, input_n, output_w, output_h, output_c, output_n,
                    pad_h, pad_w, kernel_h, kernel_w, stride_h, stride_w, pool_type):
        assert output_h == (input_h + 2 * pad_h - kernel_h) // stride_h + 1
        assert output_w == (input_w + 2 * pad_w - kernel_w) // stride_w + 1
        assert output_c == input_c and output_n == input_n
        return Pool2DPerDeviceState((input_n, input_c, input_h, input_w, output_h, output_w,
                                     kernel_h, kernel_w, stride_h, stride_w, pad_h, pad_w), pool_type)

    @staticmethod
    def forward_kernel(stream, m, input, output):
        _call("ffk_pool2d_forward", _stream(stream), _ptr(input), _ptr(output), *m.geom, m.pool_type)

    @staticmethod
    def backward_kernel(stream, m, input, input_grad, output, output_grad):
        _call("ffk_pool2d_backward", _stream(stream), _ptr(input), _ptr(input_grad), _ptr(output),
              _ptr(output_grad), *m.geom, m.pool_type)


# ------------------------------------------------------------------ attention
class _MhaConfig(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int) for n in (
        "num_samples", "num_heads", "qSize", "kSize", "vSize", "qProjSize", "kProjSize",
        "vProjSize", "oProjSize", "qoSeqLength", "kvSeqLength")]


@dataclass
class MHAPerDeviceState:
    cfg: _MhaConfig
    weightSize: int          # bytes, as the reference's field
    reserveSpaceSize: int
    reserveSpace: torch.Tensor
    workSpace: torch.Tensor
    dtype: torch.dtype
    path: int = GEMM_AUTO


class MultiHeadAttention:
    """attention_kernels.h"""

    @staticmethod
    def init_kernel(num_samples, num_heads, qSize, kSize, vSize, qProjSize, kProjSize, vProjSize,
                    oProjSize, qoSeqLength, kvSeqLength, add_bias_kv=False, dtype=torch.float32,
                    device="cuda", path=GEMM_AUTO):
        assert not add_bias_kv, "the reference asserts !add_bias_kv (attention_kernels.cu:93)"
        cfg = _MhaConfig(num_samples, num_heads, qSize, kSize, vSize, qProjSize, kProjSize,
                         vProjSize, oProjSize, qoSeqLength, kvSeqLength)
        dt = _DT[dtype]
        c = lib().cdll
        ref = ctypes.byref(cfg)
        wbytes = c.ffk_mha_weight_bytes(ref, dt)
        rbytes = c.ffk_mha_reserve_bytes(ref, dt)
        sbytes = c.ffk_mha_workspace_bytes(ref, dt)
        if wbytes == 0:
            from ._lib import FFKError
            raise FFKError("ffk_mha_weight_bytes", -1, lib().last_error())
        mk = lambda b: torch.empty(max(b, 16), dtype=torch.uint8, device=device)
        return MHAPerDeviceState(cfg, wbytes, rbytes, mk(rbytes), mk(sbytes), dtype, path)

    @staticmethod
    def forward_kernel(stream, m, query, key, value, weight, output):
        _call("ffk_mha_forward", _stream(stream), ctypes.byref(m.cfg), _ptr(query), _ptr(key),
              _ptr(value), _ptr(weight), _ptr(output), _ptr(m.reserveSpace), _DT[m.dtype], m.path)

    @staticmethod
    def forward_residual_kernel(stream, m, query, key, value, weight, output, residual):
        """output = MHA(q, k, v) + residual (ffk_mha_forward_residual)."""
        _call("ffk_mha_forward_residual", _stream(stream), ctypes.byref(m.cfg), _ptr(query), _ptr(key),
              _ptr(value), _ptr(weight), _ptr(output), _ptr(m.reserveSpace), _DT[m.dtype], m.path,
              _ptr(residual))

    @staticmethod
    def backward_accumulate_kernel(stream, m, query, query_grad, key, key_grad, value, value_grad, weight,
                                   weight_grad, output_grad, accumulate=(True, True, True)):
        """ffk_mha_backward_accumulate: data gradients are added where `accumulate` says so; the three
        gradient pointers may be one buffer (self-attention)."""
        _call("ffk_mha_backward_accumulate", _stream(stream), ctypes.byref(m.cfg), _ptr(query),
              _ptr(query_grad), _ptr(key), _ptr(key_grad), _ptr(value), _ptr(value_grad),
              _ptr(weight), _ptr(weight_grad), _ptr(output_grad), _ptr(m.reserveSpace),
              _ptr(m.workSpace), m.workSpace.numel(), _DT[m.dtype], m.path, int(accumulate[0]),
              int(accumulate[1]), int(accumulate[2]))

    @staticmethod
    def backward_kernel(stream, m, query, query_grad, key, key_grad, value, value_grad, weight,
                        weight_grad, output_grad):
        _call("ffk_mha_backward", _stream(stream), ctypes.byref(m.cfg), _ptr(query),
              _ptr(query_grad), _ptr(key), _ptr(key_grad), _ptr(value), _ptr(value_grad),
              _ptr(weight), _ptr(weight_grad), _ptr(output_grad), _ptr(m.reserveSpace),
              _ptr(m.workSpace), m.workSpace.numel(), _DT[m.dtype], m.path)


# ------------------------------------------------------------------ conv
class _Conv2DConfig(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int) for n in (
        "n", "in_c", "in_h", "in_w", "out_c", "out_h", "out_w", "kernel_h", "kernel_w", "pad_h",
        "pad_w", "stride_h", "stride_w", "groups")]


@dataclass
class Conv2DPerDeviceState:
    cfg: _Conv2DConfig
    workspace: torch.Tensor
    relu: bool


class Conv2D:
    """conv_2d_kernels.h"""

    @staticmethod
    def init_kernel(activation, kernel_h, kernel_w, groups, padding_h, padding_w, stride_h, stride_w,
                    input_shape, output_shape, device="cuda"):
        n, ci, ih, iw = input_shape
        _, co, oh, ow = output_shape
        cfg = _Conv2DConfig(n, ci, ih, iw, co, oh, ow, kernel_h, kernel_w, padding_h, padding_w,
                            stride_h, stride_w, groups)
        nbytes = lib().cdll.ffk_conv2d_workspace_bytes(ctypes.byref(cfg))
        if nbytes == 0:
            from ._lib import FFKError
            raise FFKError("ffk_conv2d_workspace_bytes", -1, lib().last_error())
        ws = torch.empty(nbytes // 4, dtype=torch.float32, device=device)
        return Conv2DPerDeviceState(cfg, ws, activation is not None and activation != ACT_NONE)

    @staticmethod
    def forward_kernel(stream, m, input, output, filter, bias):
        _call("ffk_conv2d_forward", _stream(stream), ctypes.byref(m.cfg), _ptr(input), _ptr(output),
              _ptr(filter), _ptr(bias), int(m.relu), _ptr(m.workspace), m.workspace.numel() * 4)

    @staticmethod
    def backward_kernel(stream, m, input, input_grad, output, output_grad, filter, filter_grad, bias_grad):
        _call("ffk_conv2d_backward", _stream(stream), ctypes.byref(m.cfg), _ptr(input), _ptr(input_grad),
              _ptr(output), _ptr(output_grad), _ptr(filter), _ptr(filter_grad), _ptr(bias_grad),
              int(m.relu), _ptr(m.workspace), m.workspace.numel() * 4)
