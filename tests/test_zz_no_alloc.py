"""pantr_b200_eval_psi_grad_psi_noalloc_device (include/pantr_b200.h): the allocation-free variant of the batched
psi / grad psi evaluation, scratch in a static device array of the library.  No counterpart in the reference.
Checked through the C ABI against the existing entry point
pantr_b200_eval_psi_grad_psi_batch: bit-exact (the tolerance of every parity test here), on calls in a row that
share the scratch array.  Collected last (test_zz_*): a failure here must not hide other tests under pytest -x."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
PROBLEM_IDS = {"hanging_chain": 0, "quadcopter": 1, "realistic_quadcopter": 2}  # include/pantr_b200.h:33


def _inputs(prob, name, B, rng):
    x0 = np.tile(prob.initial_state(), (B, 1)) + 0.02 * rng.standard_normal((B, prob.nx))
    if name == "hanging_chain":
        x = rng.uniform(-0.8, 0.8, (B, prob.n))
    else:
        x = np.tile(prob.initial_guess(), (B, 1)) + 0.05 * rng.standard_normal((B, prob.n))
    y = rng.standard_normal((B, prob.m)) * 3
    S = 10.0 ** rng.uniform(0, 3, (B, prob.m))
    return x0, x, y, S


def _noalloc(name, N, x0, x, y, S, n, m):
    import torch
    from pantr_cdc2023_experiments_b200 import _lib
    lib = _lib.load()
    B = x.shape[0]
    dev = torch.device("cuda:0")
    with torch.cuda.device(dev):
        t = [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (x0, x)]
        ty, tS = (torch.from_numpy(y).to(dev), torch.from_numpy(S).to(dev)) if m else (None, None)
        psi = torch.full((B,), float("nan"), dtype=torch.float64, device=dev)
        grad = torch.full((B, n), float("nan"), dtype=torch.float64, device=dev)
        torch.cuda.synchronize()
        stream = torch.cuda.current_stream().cuda_stream
        rc = lib.pantr_b200_eval_psi_grad_psi_noalloc_device(
            PROBLEM_IDS[name], N, B, t[0].data_ptr(), t[1].data_ptr(), ty.data_ptr() if m else None,
            tS.data_ptr() if m else None, psi.data_ptr(), grad.data_ptr(), stream or None)
        assert rc == 0, lib.pantr_b200_last_error().decode()
        torch.cuda.synchronize()
        return psi.cpu().numpy(), grad.cpu().numpy()


def test_noalloc_matches_existing_entry_point_on_calls_in_a_row():
    import pantr_cdc2023_experiments_b200 as pb
    rng = np.random.default_rng(2024)
    # calls in a row share the static scratch: different models, horizons and batch sizes follow each other, so
    # each call finds the previous one's (differently laid out) contents in the array; the first pair repeats
    cases = [("quadcopter", 20, 300), ("quadcopter", 20, 300), ("hanging_chain", 40, 700),
             ("realistic_quadcopter", 60, 2000), ("hanging_chain", 8, 3), ("quadcopter", 120, 5)]
    for name, N, B in cases:
        prob = pb.BatchedProblem(name, N)
        x0, x, y, S = _inputs(prob, name, B, rng)
        psi_ref, grad_ref = prob.eval_psi_grad_psi(x, y, S, x0=x0)  # pantr_b200_eval_psi_grad_psi_batch
        for _ in range(2):  # two calls in a row on the same inputs
            psi, grad = _noalloc(name, N, x0, x, y, S, prob.n, prob.m)
            np.testing.assert_array_equal(psi, psi_ref)
            np.testing.assert_array_equal(grad, grad_ref)
        assert np.isfinite(psi_ref).all() and np.isfinite(grad_ref).all()


def test_noalloc_rejects_sizes_beyond_the_maximum():
    import torch
    from pantr_cdc2023_experiments_b200 import _lib
    lib = _lib.load()
    buf = torch.zeros(64, dtype=torch.float64, device="cuda:0")
    p = buf.data_ptr()
    OUT_OF_RANGE, INVALID = -4, -1
    assert lib.pantr_b200_eval_psi_grad_psi_noalloc_device(1, 121, 1, p, p, p, p, p, p, None) == OUT_OF_RANGE
    assert b"horizon" in lib.pantr_b200_last_error()
    assert lib.pantr_b200_eval_psi_grad_psi_noalloc_device(1, 0, 1, p, p, p, p, p, p, None) == OUT_OF_RANGE
    assert lib.pantr_b200_eval_psi_grad_psi_noalloc_device(1, 10, (1 << 22) + 1, p, p, p, p, p, p, None) == OUT_OF_RANGE
    assert b"batch" in lib.pantr_b200_last_error()
    assert lib.pantr_b200_eval_psi_grad_psi_noalloc_device(1, 10, -1, p, p, p, p, p, p, None) == OUT_OF_RANGE
    assert lib.pantr_b200_eval_psi_grad_psi_noalloc_device(7, 10, 1, p, p, p, p, p, p, None) == INVALID
    assert lib.pantr_b200_eval_psi_grad_psi_noalloc_device(1, 10, 1, p, p, None, p, p, p, None) == INVALID  # m > 0 needs y
    assert lib.pantr_b200_eval_psi_grad_psi_noalloc_device(1, 10, 1, None, p, p, p, p, p, None) == INVALID
    assert lib.pantr_b200_eval_psi_grad_psi_noalloc_device(0, 10, 0, p, p, None, None, p, p, None) == 0  # empty batch
    torch.cuda.synchronize()
