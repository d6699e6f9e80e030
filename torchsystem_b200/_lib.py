"""ctypes binding of libtsb200.so (the C ABI declared in include/tsb200.h).

There is deliberately no fallback: if the shared library is missing the import raises, and every
non-zero status from the library becomes a RuntimeError carrying ts_last_error()."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libtsb200.so')

TS_ACT_RELU, TS_ACT_GELU = 0, 1
TS_PREC_F32, TS_PREC_BF16 = 0, 1
TS_DT_F32, TS_DT_BF16, TS_DT_U8 = 0, 1, 2
TS_OPT_SGD, TS_OPT_ADAM = 0, 1
RECORD_BYTES = 16
STATE_BYTES = 32


class OptHyper(C.Structure):
    _fields_ = [('lr', C.c_double), ('beta1', C.c_double), ('beta2', C.c_double), ('eps', C.c_double),
                ('weight_decay', C.c_double), ('momentum', C.c_double), ('dampening', C.c_double),
                ('nesterov', C.c_int32), ('amsgrad', C.c_int32), ('decoupled_weight_decay', C.c_int32), ('reserved', C.c_int32)]


class MetricsState(C.Structure):
    _fields_ = [('loss_sum', C.c_double), ('n_batches', C.c_int64), ('n_correct', C.c_int64), ('n_total', C.c_int64)]


class TsError(RuntimeError):
    pass


# name -> (restype, argtypes); every symbol include/tsb200.h declares must be here (tests check both ways)
PROTOTYPES = {
    'ts_abi_version': (C.c_int32, []),
    'ts_last_error': (C.c_char_p, []),
    'ts_mlp_create': (C.c_int32, [C.POINTER(C.c_void_p), C.POINTER(C.c_int32), C.c_int32, C.c_int32, C.c_float,
                                  C.c_int32, C.c_int32]),
    'ts_mlp_destroy': (C.c_int32, [C.c_void_p]),
    'ts_mlp_status': (C.c_int32, [C.c_void_p]),
    'ts_mlp_clear_status': (C.c_int32, [C.c_void_p]),
    'ts_mlp_param_layout': (C.c_int32, [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    'ts_mlp_set_input_lut': (C.c_int32, [C.c_void_p, C.POINTER(C.c_float)]),
    'ts_mlp_bind_params': (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p]),
    'ts_opt_bind': (C.c_int32, [C.c_void_p, C.c_int32, C.POINTER(OptHyper), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]),
    'ts_opt_step_count': (C.c_int64, [C.c_void_p]),
    'ts_dp_bind': (C.c_int32, [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]),
    'ts_dp_bind_ll': (C.c_int32, [C.c_void_p, C.POINTER(C.c_void_p), C.c_int64]),
    'ts_dp_set_timeout': (C.c_int32, [C.c_void_p, C.c_double]),
    'ts_dp_status': (C.c_int32, [C.c_void_p]),
    'ts_mlp_shadow_elems': (C.c_int64, [C.c_void_p]),
    'ts_dp_bind_nvls': (C.c_int32, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.POINTER(C.c_void_p)]),
    'ts_dp_nvls_second_shadow': (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p]),
    'ts_mlp_workspace_bytes': (C.c_int64, [C.c_void_p, C.c_int64]),
    'ts_mlp_set_workspace': (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]),
    'ts_mlp_sync_weights': (C.c_int32, [C.c_void_p, C.c_void_p]),
    'ts_fit_step': (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.POINTER(C.c_void_p),
                                C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]),
    'ts_fit_backward': (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.POINTER(C.c_void_p),
                                    C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]),
    'ts_fit_backward_overlap': (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.POINTER(C.c_void_p),
                                            C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p),
                                            C.POINTER(C.c_int32)]),
    'ts_dp_reserve_sms': (C.c_int32, [C.c_void_p, C.c_int32]),
    'ts_opt_apply': (C.c_int32, [C.c_void_p, C.c_float, C.c_void_p]),
    'ts_opt_apply_range': (C.c_int32, [C.c_void_p, C.c_float, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    'ts_eval_step': (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p,
                                 C.c_void_p]),
    'ts_eval_step_static': (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p,
                                        C.c_void_p]),
    'ts_metrics_accumulate': (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]),
    'ts_metrics_update_raw': (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    'ts_metrics_reset': (C.c_int32, [C.c_void_p, C.c_void_p]),
    'ts_metrics_read': (C.c_int32, [C.c_void_p, C.POINTER(MetricsState), C.c_void_p]),
    'ts_test_gemm_bf16': (C.c_int32, [C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_int32, C.c_int64, C.c_int32, C.c_int32,
                                      C.c_int32, C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p]),
    'ts_test_fused_trace': (C.c_int32, [C.POINTER(C.c_uint64)]),
    'ts_test_step_trace': (C.c_int32, [C.POINTER(C.c_uint64)]),
    'ts_test_wgrad_trace': (C.c_int32, [C.POINTER(C.c_uint64)]),
    'ts_test_eval_trace': (C.c_int32, [C.POINTER(C.c_uint64)]),
    'ts_launch_count': (C.c_int64, []),
    'ts_profile_begin': (C.c_int32, []),
    'ts_profile_end': (C.c_int32, [C.POINTER(C.c_double), C.POINTER(C.c_int64), C.c_int32]),
    'ts_profile_kind_count': (C.c_int32, []),
    'ts_profile_kind_name': (C.c_char_p, [C.c_int32]),
}


ABI_VERSION = 2      # TS_ABI_VERSION in include/tsb200.h


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f'{LIB_PATH} not found: build it with `python -c "import __graft_entry__ as g; g.build()"` '
            '(nvcc, sm_100a). torchsystem_b200 has no CPU or PyTorch fallback.')
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    if lib.ts_abi_version() != ABI_VERSION:
        raise ImportError(f'{LIB_PATH} has ABI version {lib.ts_abi_version()}, this package binds version {ABI_VERSION}: '
                          'rebuild it (python -c "import __graft_entry__ as g; g.build()")')
    return lib


lib = _load()


def check(rc: int, what: str = '') -> None:
    if rc != 0:
        msg = lib.ts_last_error().decode('utf-8', 'replace')
        raise TsError(f'{what or "tsb200"} failed (status {rc}): {msg}')


def profile_begin() -> None:
    check(lib.ts_profile_begin(), 'ts_profile_begin')


def profile_end() -> dict:
    """{kernel kind: (total ms, launches)} measured with CUDA events on the launching stream."""
    n = lib.ts_profile_kind_count()
    ms, cnt = (C.c_double * n)(), (C.c_int64 * n)()
    check(lib.ts_profile_end(ms, cnt, n), 'ts_profile_end')
    return {lib.ts_profile_kind_name(i).decode(): (ms[i], cnt[i]) for i in range(n) if cnt[i]}
