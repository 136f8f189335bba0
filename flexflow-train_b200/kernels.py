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
GEMM_AUTO, GEMM_TENSOR, GEMM_SIMT = 0, 1, 2


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


class Linear:
    """linear_kernels.h"""

    @staticmethod
    def init_kernel(activation=ACT_NONE, regularizer_lambda=0.0, use_bias=True, path=GEMM_AUTO):
        return LinearPerDeviceState(activation, regularizer_lambda, use_bias, path)

    @staticmethod
    def forward_kernel(stream, m, input, output, weight, bias, in_dim, out_dim, batch_size):
        _call("ffk_linear_forward", _stream(stream), _ptr(input), _ptr(output), _ptr(weight),
              _ptr(bias), in_dim, out_dim, batch_size, dtype_code(input), m.activation, m.path)

    @staticmethod
    def backward_kernel(stream, m, input, input_grad, output, output_grad, weight, weight_grad,
                        bias_grad, in_dim, out_dim, batch_size, preact=None):
        _call("ffk_linear_backward", _stream(stream), _ptr(input), _ptr(input_grad), _ptr(output),
              _ptr(output_grad), _ptr(weight), _ptr(weight_grad), _ptr(bias_grad), in_dim, out_dim,
              batch_size, dtype_code(input), m.activation, m.l2_lambda, _ptr(preact), m.path)


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
                        accumulate_param_grads=False):
        ws = m.workspace
        _call("ffk_layer_norm_backward", _stream(stream), _ptr(output_grad), _ptr(input),
              _ptr(input_grad), _ptr(gamma), _ptr(gamma_grad), _ptr(beta_grad), _ptr(m.mean),
              _ptr(m.rstd), m.effective_batch_size, m.effective_num_elements,
              dtype_code(input), int(accumulate_param_grads), _ptr(ws),
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
