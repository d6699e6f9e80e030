"""CPU-side checks of the drop-in boundary: the shared library loads without a GPU and exports exactly the
symbols include/tsb200.h declares, the ctypes table covers them all, and calls fail loudly (no fallback)."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, 'include', 'tsb200.h')


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return set(re.findall(r'\b(ts_[a-z0-9_]+)\s*\(', text))


@pytest.fixture(scope='module')
def built():
    import __graft_entry__ as g
    g.build()
    from torchsystem_b200 import _lib
    return _lib


def test_header_symbols_are_exported_and_bound(built):
    names = declared_symbols()
    assert len(names) >= 20
    out = subprocess.run(['nm', '-D', '--defined-only', built.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r'\bT (ts_[a-z0-9_]+)', out))
    assert names <= exported, f'declared but not exported: {sorted(names - exported)}'
    assert exported <= names, f'exported but not declared in include/tsb200.h: {sorted(exported - names)}'
    assert names == set(built.PROTOTYPES), f'ctypes table out of sync: {sorted(names ^ set(built.PROTOTYPES))}'


def test_abi_version_and_error_string(built):
    assert built.lib.ts_abi_version() == built.ABI_VERSION == 2
    h = ctypes.c_void_p()
    dims = (ctypes.c_int32 * 2)(784, 10)
    rc = built.lib.ts_mlp_create(ctypes.byref(h), dims, 2, 0, 0.0, 0, 0)      # no hidden layer: rejected before any CUDA call
    assert rc == -1 and b'hidden' in built.lib.ts_last_error()
    with pytest.raises(built.TsError):
        built.check(rc, 'ts_mlp_create')


def test_inbox_size_agrees_between_header_kernels_and_python():
    """TS_DP_LL_INBOX_BYTES (include/tsb200.h) = 2 parities x kLlParityBytes (csrc/kernels.h) = parallel.LL_INBOX_BYTES: the inbox
    of the low-latency exchange is allocated in Python, checked in the C ABI and indexed in the kernel."""
    header = int(re.search(r'#define\s+TS_DP_LL_INBOX_BYTES\s+(\d+)', open(HEADER).read()).group(1))
    kh = open(os.path.join(ROOT, 'torchsystem_b200', 'csrc', 'kernels.h')).read()
    ranks = int(re.search(r'constexpr int kMaxRanks = (\d+);', kh).group(1))
    ctas, quads = map(int, re.search(r'constexpr int kLlCtas = (\d+), kLlQuads = (\d+);', kh).groups())
    assert '(kMaxRanks + 1) * kLlCtas * kLlQuads * 256 * 16' in kh
    assert header == 2 * (ranks + 1) * ctas * quads * 256 * 16
    from torchsystem_b200 import parallel
    assert parallel.LL_INBOX_BYTES == header and header % 16 == 0


def test_no_torch_types_in_the_abi():
    text = open(HEADER).read()
    assert 'torch' not in re.sub(r'/\*.*?\*/', '', text, flags=re.S).lower()
    assert 'at::' not in text and 'c10::' not in text


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, 'torchsystem_b200')
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r'^\s*(from|import)\s+oracle\b', src, flags=re.M), f'{f} imports the oracle'
                assert '/root/reference' not in src.replace('/root/reference/', '').replace('(/root/reference', ''), f


def test_missing_library_fails_loudly(tmp_path, built):
    """Import of the binding with no .so next to it must raise, not fall back."""
    import shutil, sys, textwrap
    pkg = tmp_path / 'torchsystem_b200'
    shutil.copytree(os.path.join(ROOT, 'torchsystem_b200'), pkg, ignore=shutil.ignore_patterns('*.so', '__pycache__', 'csrc'))
    code = textwrap.dedent(f'''
        import sys; sys.path.insert(0, {str(tmp_path)!r})
        try:
            import torchsystem_b200
        except ImportError as e:
            assert 'no CPU' in str(e) or 'not found' in str(e); print('ok')
    ''')
    out = subprocess.run([sys.executable, '-c', code], capture_output=True, text=True)
    assert out.stdout.strip() == 'ok', out.stderr
