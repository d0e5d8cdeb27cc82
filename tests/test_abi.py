"""CPU checks of the drop-in boundary: the C-ABI library loads, exports every
symbol include/salsa_b200.h declares, and the ctypes binding agrees with the
header on every argument count.  No compute calls (no GPU here)."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_prototypes():
    src = open(os.path.join(ROOT, 'include', 'salsa_b200.h')).read()
    src = re.sub(r'/\*.*?\*/', ' ', src, flags=re.S)
    src = re.sub(r'^\s*#.*$', '', src, flags=re.M)
    protos = {}
    for m in re.finditer(r'([A-Za-z_][\w\s\*]*?)\b(salsa_\w+)\s*\(([^;{}]*?)\)\s*;', src, flags=re.S):
        name, args = m.group(2), m.group(3).strip()
        n = 0 if args in ('', 'void') else len(args.split(','))
        protos[name] = n
    return protos


@pytest.fixture(scope='module')
def built():
    r = subprocess.run(['make', '-C', os.path.join(ROOT, 'salsa_b200', 'csrc'), '-j4'], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    from salsa_b200 import _lib
    return _lib


def test_library_exports_every_declared_symbol(built):
    protos = header_prototypes()
    assert len(protos) >= 20
    L = C.CDLL(built.LIB_PATH)
    for name in protos:
        assert hasattr(L, name), 'libsalsa_b200.so does not export %s' % name


def test_binding_matches_header(built):
    protos = header_prototypes()
    for name, n in protos.items():
        assert name in built.SIGNATURES, 'no ctypes signature for %s' % name
        assert len(built.SIGNATURES[name][1]) == n, '%s: header has %d parameters, binding %d' % (
            name, n, len(built.SIGNATURES[name][1]))
    for name in built.SIGNATURES:
        assert name in protos, '%s bound but not declared in include/salsa_b200.h' % name
    built.lib()


def test_no_cpu_fallback(built):
    """Without a CUDA device the product refuses to run instead of falling back."""
    import torch
    if torch.cuda.is_available():
        pytest.skip('a GPU is present')
    L = built.lib()
    h = C.c_void_p()
    rc = L.salsa_ctx_create(0, C.byref(h))
    assert rc != 0 and b'no CPU fallback' in L.salsa_last_error()
    from salsa_b200.engine import Engine
    with pytest.raises(RuntimeError):
        Engine(0)


def test_product_never_imports_oracle():
    """oracle/ is test infrastructure: nothing under salsa_b200/ may reference it."""
    bad = []
    for dp, dn, fn in os.walk(os.path.join(ROOT, 'salsa_b200')):
        for f in fn:
            if f.endswith(('.py', '.cu', '.cuh', '.h', 'Makefile')):
                txt = open(os.path.join(dp, f), errors='replace').read()
                if re.search(r'^\s*(import|from)\s+oracle\b', txt, flags=re.M) or 'liboracle' in txt or 'salsa_oracle' in txt:
                    bad.append(os.path.join(dp, f))
    assert not bad, bad
