"""The allocation-free LayerNorm backward (ffk_layer_norm_backward_static).

Some users of this library run where device memory cannot be requested at call time: a task
body under a runtime that reserved the GPU's memory for its own pool when the process started
(the reference's legacy Legion/Realm runtime works like that — a cudaMalloc from inside a task
can fail outright and always stalls the device), or a caller that simply has no workspace
buffer to lend. For them the scratch of the backward pass lives in an array inside libffk.so.

This file checks, through the C ABI, that the variant gives the same gradients as the ordinary
entry point (which is handed a torch-allocated workspace), twice in a row with different shapes
so that whatever the first call left in the shared array cannot leak into the second, and
that oversized rows are refused. The file name sorts last on purpose: under `pytest -x` every
other test has already run when this one starts."""
import numpy as np
import pytest
import torch

from helpers import SEED_ACT, SEED_W, assert_close, to_dev, uniform

pytestmark = pytest.mark.gpu


def _backward_pair(K, ffk, M, N, dtype, accumulate, oracle=None):
    x, xn = to_dev(uniform((M, N), SEED_ACT) + 0.25, dtype)
    gamma, gn = to_dev(uniform((N,), SEED_W) + 1.0, dtype)
    beta, _ = to_dev(uniform((N,), SEED_W + 1), dtype)
    dy, dyn = to_dev(uniform((M, N), SEED_ACT + 1), dtype)
    y = torch.empty_like(x)
    state = K.LayerNorm.init_kernel(True, M, N, 1e-5)      # owns a torch-allocated workspace
    K.LayerNorm.forward_kernel(None, state, x, y, gamma, beta)

    def fresh():
        dx, _ = to_dev(uniform((M, N), SEED_ACT + 2), dtype)
        dg = torch.full((N,), 3.0, dtype=dtype, device="cuda")
        db = torch.full((N,), -2.0, dtype=dtype, device="cuda")
        return dx, dg, db

    dx_ref, dg_ref, db_ref = fresh()
    K.LayerNorm.backward_kernel(None, state, dy, x, dx_ref, gamma, dg_ref, db_ref,
                                accumulate_param_grads=accumulate)
    dx, dg, db = fresh()
    stream = torch.cuda.current_stream().cuda_stream
    ffk.lib().call("ffk_layer_norm_backward_static", stream, dy.data_ptr(), x.data_ptr(),
                   dx.data_ptr(), gamma.data_ptr(), dg.data_ptr(), db.data_ptr(),
                   state.mean.data_ptr(), state.rstd.data_ptr(), M, N, K.dtype_code(x),
                   int(accumulate))
    torch.cuda.synchronize()
    # same kernels, possibly a different number of partial slabs => fp32 summation order may
    # differ: compare within the suite's tolerances (1e-4 fp32, 2e-2 bf16)
    assert_close(dx, dx_ref.float().cpu().numpy(), dtype, what=f"static dX {M}x{N}")
    assert_close(dg, dg_ref.float().cpu().numpy(), dtype, what=f"static dgamma {M}x{N}")
    assert_close(db, db_ref.float().cpu().numpy(), dtype, what=f"static dbeta {M}x{N}")
    if oracle is not None:
        # second referee: the CPU oracle on the same inputs (it starts from the same dX / dgamma /
        # dbeta contents and follows the same accumulate rule)
        dxo, _, _ = fresh()
        dxn = dxo.float().cpu().numpy()
        dgn = np.full(N, 3.0, np.float32)
        dbn = np.full(N, -2.0, np.float32)
        oracle.layer_norm_backward(dyn, xn, dxn, gn, dgn, dbn, state.mean[:M].cpu().numpy(),
                                   state.rstd[:M].cpu().numpy(), M, N, int(accumulate))
        assert_close(dx, dxn, dtype, what=f"static dX vs oracle {M}x{N}")
        if dtype == torch.float32:   # 16-bit parameter sums over hundreds of rows: judged against the workspace variant above
            assert_close(dg, dgn, dtype, what=f"static dgamma vs oracle {M}x{N}")
            assert_close(db, dbn, dtype, what=f"static dbeta vs oracle {M}x{N}")


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_static_scratch_backward_matches_workspace_backward_twice_in_a_row(K, ffk, oracle, dtype):
    # first call: wide rows, many slabs; second call: fewer, narrower slabs over the same array
    _backward_pair(K, ffk, 600, 4096, dtype, accumulate=False, oracle=oracle)
    _backward_pair(K, ffk, 37, 1024, dtype, accumulate=True, oracle=oracle)
    # and once more at the documented limit
    _backward_pair(K, ffk, 5, 16384, dtype, accumulate=False)


def test_static_scratch_backward_refuses_rows_beyond_its_limit(K, ffk):
    M, N = 2, 16384 + 8
    x = torch.zeros(M, N, device="cuda")
    st = torch.zeros(M, device="cuda")
    with pytest.raises(ffk.FFKError) as e:
        ffk.lib().call("ffk_layer_norm_backward_static", None, x.data_ptr(), x.data_ptr(),
                       x.data_ptr(), None, None, None, st.data_ptr(), st.data_ptr(), M, N, K.FLOAT, 0)
    assert e.value.code == -5 and "16384" in str(e.value)
