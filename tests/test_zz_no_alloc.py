"""ts_eval_step_static -- the evaluate entry point for callers that cannot obtain device memory (a process whose
GPU memory belongs to another framework's pool, a loop that must not take the device-wide synchronisation of
cudaMalloc/cudaFree, a plug-in without an allocator). Its intermediates live in a 16 MiB array inside
libtsb200.so, so the handle needs no workspace at all.

The test drives the C ABI directly with ctypes: one handle WITHOUT a workspace calls the new entry point, a
second handle WITH a workspace calls the existing ts_eval_step on the same weights and data; predictions must be
identical and the loss equal within the fp32 tolerance of the other tests (1e-5 relative). It makes two calls in
a row with different batches, because the built-in array still holds the first call's data during the second.
The file name keeps it last in collection order; it passes on its own in a fresh process as well."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

DIMS = [784, 256, 10]


def _handle(lib, check, device_index):
    h = C.c_void_p()
    dims = (C.c_int32 * len(DIMS))(*DIMS)
    check(lib.ts_mlp_create(C.byref(h), dims, len(DIMS), 0, 0.1, 0, device_index), 'ts_mlp_create')   # ReLU, fp32
    n = 2 * (len(DIMS) - 1)
    offs, total = (C.c_int64 * n)(), C.c_int64()
    check(lib.ts_mlp_param_layout(h, offs, C.byref(total)), 'ts_mlp_param_layout')
    return h, list(offs), total.value


def test_static_evaluate_matches_workspace_evaluate_twice_in_a_row():
    from torchsystem_b200 import _lib
    lib, check = _lib.lib, _lib.check
    dev = torch.device('cuda', 0)
    stream = torch.cuda.current_stream(dev).cuda_stream

    h_static, offs, total = _handle(lib, check, 0)
    h_ws, _, _ = _handle(lib, check, 0)
    try:
        g = torch.Generator().manual_seed(2024)
        params = torch.zeros(total, dtype=torch.float32)
        shapes = [(256, 784), (256,), (10, 256), (10,)]
        for o, s in zip(offs, shapes):
            n = int(np.prod(s))
            params[o:o + n] = torch.randn(n, generator=g) * 0.05
        params = params.to(dev)
        grads = torch.zeros_like(params)
        for h in (h_static, h_ws):
            check(lib.ts_mlp_bind_params(h, params.data_ptr(), grads.data_ptr()), 'ts_mlp_bind_params')
        need = lib.ts_mlp_workspace_bytes(h_ws, 4096)
        ws = torch.empty(need, dtype=torch.uint8, device=dev)
        check(lib.ts_mlp_set_workspace(h_ws, ws.data_ptr(), need, 4096), 'ts_mlp_set_workspace')

        for B in (4096, 1000):                                    # second call: smaller, ragged batch over stale scratch
            x = torch.randn(B, 784, generator=g).to(dev)
            y = torch.randint(0, 10, (B,), generator=g).to(dev)
            out = {}
            for name, fn, h in (('static', lib.ts_eval_step_static, h_static), ('workspace', lib.ts_eval_step, h_ws)):
                pred = torch.full((B,), -1, dtype=torch.int64, device=dev)
                rec = torch.zeros(4, dtype=torch.float32, device=dev)
                check(fn(h, x.data_ptr(), _lib.TS_DT_F32, y.data_ptr(), B, pred.data_ptr(), rec.data_ptr(), stream), name)
                torch.cuda.synchronize()
                out[name] = (pred.cpu().numpy(), rec[0].item(), rec.view(torch.int32)[1].item(), rec.view(torch.int32)[2].item())
            (ps, ls, cs, ns), (pw, lw, cw, nw) = out['static'], out['workspace']
            assert np.array_equal(ps, pw), 'predictions differ between the static and the workspace entry point'
            assert ls == pytest.approx(lw, rel=1e-5)
            assert (cs, ns) == (cw, nw) == (int((pw == y.cpu().numpy()).sum()), B)
            # and against the CPU oracle, so this file stands on its own: same weights, fp32 arithmetic
            from oracle import mlp_numpy as onp
            host = params.cpu().numpy()
            tensors = [host[o:o + int(np.prod(sh))].reshape(sh) for o, sh in zip(offs, shapes)]
            oracle = onp.OracleMLP(weights=tensors[0::2], biases=tensors[1::2])
            opred, oloss = oracle.evaluate(x.cpu().numpy(), y.cpu().numpy())
            assert np.array_equal(ps, opred), 'static entry point: predictions differ from the oracle'
            assert ls == pytest.approx(float(oloss), rel=1e-5)

        # beyond the stated maximum: refused with a clear message, nothing launched
        B = 20000                                                 # 20000 * 256 * 4 = 20.5 MB > 16 MiB
        x = torch.zeros(B, 784, device=dev)
        y = torch.zeros(B, dtype=torch.int64, device=dev)
        pred = torch.zeros(B, dtype=torch.int64, device=dev)
        rec = torch.zeros(4, dtype=torch.float32, device=dev)
        before = lib.ts_launch_count()
        rc = lib.ts_eval_step_static(h_static, x.data_ptr(), _lib.TS_DT_F32, y.data_ptr(), B, pred.data_ptr(), rec.data_ptr(), stream)
        assert rc == -5 and b'16 MiB' in lib.ts_last_error() and lib.ts_launch_count() == before
        rc = lib.ts_eval_step_static(h_static, x.data_ptr(), _lib.TS_DT_BF16, y.data_ptr(), 16, pred.data_ptr(), rec.data_ptr(), stream)
        assert rc == -4
    finally:
        torch.cuda.synchronize()
        lib.ts_mlp_destroy(h_static)
        lib.ts_mlp_destroy(h_ws)
