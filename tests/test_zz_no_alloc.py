"""GPU (-m gpu), collected LAST on purpose: `psob_layer_run_dot_static` -- the fused discriminator head with its partial sums in
a static device array inside libpsob.so (no allocation by the caller or the library,
no counterpart in the reference) -- against the existing pair psob_layer_run_dot + psob_dot_finish, through the C ABI, on two
calls in a row that share the static array.  Passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest
import torch

from oracle import port

pytestmark = pytest.mark.gpu


def _existing(ffi, engine, dnet, hidden, P):
    """psob_layer_run_dot + psob_dot_finish with a caller-allocated partial-sum buffer -> score fp32 [P]."""
    lib, hd, layer = ffi.lib(), dnet.head, dnet.layers[-2]
    part = torch.full((hd["slots"], P), float("nan"), device="cuda")
    out = torch.empty(P, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    ffi.check(lib.psob_layer_run_dot(layer.handle, hidden.data_ptr(), hidden.stride(0), P, hd["w"].data_ptr(), part.data_ptr(), s), "run_dot")
    ffi.check(lib.psob_dot_finish(part.data_ptr(), hd["slots"], P, hd["bias"], hd["act"], out.data_ptr(), s), "dot_finish")
    return out


def _static(ffi, dnet, hidden, P, out):
    lib, hd, layer = ffi.lib(), dnet.head, dnet.layers[-2]
    x, pitch = (hidden.data_ptr(), hidden.stride(0)) if hidden is not None else (None, 0)
    return lib.psob_layer_run_dot_static(layer.handle, x, pitch, P, hd["w"].data_ptr(), hd["bias"], hd["act"], out.data_ptr(),
                                         torch.cuda.current_stream().cuda_stream)


def test_static_scratch_head_equals_existing_entry_points_twice_in_a_row():
    from psob import engine, _ffi as ffi
    engine.require_cuda()
    torch.manual_seed(11)
    disc = port.Discriminator(1, 64).eval()
    with torch.no_grad():
        for prm in disc.parameters():
            if prm.dim() == 4:
                prm.mul_(3.0)                      # spread the scores (default init keeps them near 0.5)
    results = []
    # two calls in a row on the shared array: different inputs AND a smaller, ragged P the second time, so stale partial sums of
    # the first call (another P, another layout) would be caught
    for P, seed in ((1000, 1), (333, 2)):
        dnet = engine.DiscriminatorNet(disc.state_dict(), P)
        assert dnet.head is not None
        img = torch.zeros(P, 31, 32)
        img[:, 1:29, 1:29] = torch.tanh(torch.randn(P, 28, 28, generator=torch.Generator().manual_seed(seed)))
        x = img.reshape(P, -1).cuda().bfloat16().contiguous()
        plan = dnet.launch_plan(x, finish=False)
        for name, launch in plan[:-1]:             # bias columns + first conv: the hidden activation the head layer reads
            launch()
        hidden = dnet.acts[0]
        ref = _existing(ffi, engine, dnet, hidden, P)
        out = torch.full((P,), float("nan"), device="cuda")
        ffi.check(_static(ffi, dnet, hidden, P, out), "psob_layer_run_dot_static")
        torch.cuda.synchronize()
        got, want = out.cpu().numpy(), ref.cpu().numpy()
        assert np.isfinite(got).all() and want.max() - want.min() > 1e-3
        # same kernels, same summation order: identical bits; the gate is the fused-head tests' own tolerance (1e-5)
        np.testing.assert_allclose(got, want, atol=1e-5, rtol=0)
        assert np.array_equal(got, want)
        results.append(got)
        # and once more on the same inputs: the array now holds THIS call's partial sums
        out2 = torch.empty(P, device="cuda")
        ffi.check(_static(ffi, dnet, hidden, P, out2), "psob_layer_run_dot_static")
        torch.cuda.synchronize()
        assert np.array_equal(out2.cpu().numpy(), want)
    # the score itself against the fp32 oracle on the second batch (bf16 activations: the discriminator gate 5e-3)
    with torch.no_grad():
        oracle = disc(img[:, 1:29, 1:29].reshape(P, 1, 28, 28).bfloat16().float()).reshape(-1).numpy()
    np.testing.assert_allclose(results[-1], oracle, atol=5e-3, rtol=0)
    # beyond the static array's size: rejected with a clear error before anything runs (the pointers are never touched)
    too_many = ffi.STATIC_SCRATCH_BYTES // 4 // dnet.head["slots"] + 1
    rc = _static(ffi, dnet, hidden, too_many, out)
    assert rc == -2
    msg = ffi.lib().psob_last_error().decode()
    assert "exceed the static scratch" in msg and str(ffi.STATIC_SCRATCH_BYTES) in msg
    torch.cuda.synchronize()                       # no sticky error: nothing was launched
    assert _static(ffi, dnet, None, P, out) < 0    # null input
