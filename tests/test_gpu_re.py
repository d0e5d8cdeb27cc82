"""GPU parity tests of restriction-site counting (RE_sites.py) against the oracle."""
import gzip
import io
import json
import os
import sys
import tempfile

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, 'salsa_b200', 'bin')
GOLD = os.path.join(ROOT, 'tests', 'golden')


@pytest.fixture(scope='module')
def eng():
    sys.path.insert(0, BIN)
    from _common import engine
    return engine()


def gpu_lines(eng, fasta, enzyme):
    from salsa_b200.re_sites import count_sites
    n, l, r, s = count_sites(eng, np.frombuffer(fasta, dtype=np.uint8), enzyme)
    return ['%s %d %d' % (a.decode(), b, c) for a, b, c in zip(n, l, r)], s


def oracle_lines(o, fasta, enzyme):
    n, l, r, s = o.re_sites(fasta, enzyme)
    return ['%s %d %d' % (a.decode(), b, c) for a, b, c in zip(n, l, r)], s


ENZ = ['GATC', 'GATC,GANTC', 'DNASE', 'AAGCTT', 'GATC, GATC', 'CTNAG,TTAA', 'GANNTC', 'A', 'AAAA', 'GAGAG,ACACA', 'ATAT,GATC,TT',
       'ACGTACGTACGTACGTACGTACGTACGTACGTACGTACGTACGTACGTACGTACGTACGTACG']


@pytest.mark.parametrize('enz', ENZ)
def test_golden_fasta(eng, oracle_lib, enz):
    fa = gzip.open(os.path.join(GOLD, 'links_case.fa.gz'), 'rb').read()
    want, ws = oracle_lines(oracle_lib, fa, enz)
    got, gs = gpu_lines(eng, fa, enz)
    assert got == want and (ws == gs).all()
    g = json.load(open(os.path.join(GOLD, 'links_twin.json')))
    if 're_' + enz in g:
        assert got == g['re_' + enz]          # the py3 twin's committed vectors


def _messy_fasta(seed):
    rng = np.random.default_rng(seed)
    out = []
    alpha = b'ACGT'
    for i in range(25):
        L = int(rng.choice([0, 1, 2, 3, 5, 63, 64, 65, 127, 128, 129, 1000, 4095, 4096, 4097, 9000, 70000]))
        # low-complexity stretches make self-overlapping patterns meaningful
        s = bytearray(alpha[int(x)] for x in rng.integers(0, 4, L))
        for _ in range(L // 300):
            p = int(rng.integers(0, max(L - 40, 1)))
            unit = [b'A', b'AT', b'GA', b'AC', b'T'][int(rng.integers(0, 5))]
            rep = (unit * 40)[:int(rng.integers(5, 40))]
            s[p:p + len(rep)] = rep[:max(0, min(len(rep), L - p))]
        for _ in range(L // 2000 + 1):
            if L > 10:
                p = int(rng.integers(0, L - 5))
                s[p:p + 3] = [b'NNN', b'acg', b'n-*', b' \t '][int(rng.integers(0, 4))][:3]
        s = bytes(s[:L])
        out.append(b'>rec%d some description %d\n' % (i % 20, i))      # names repeat: later record wins
        w = int(rng.choice([1, 7, 60, 61, 80, 100000]))
        eol = [b'\n', b'\r\n', b' \n', b'\t \r\n'][int(rng.integers(0, 4))]
        for j in range(0, len(s), w):
            out.append(s[j:j + w] + eol)
        if rng.random() < 0.3:
            out.append(b'\n')                 # blank line inside a record
    txt = b''.join(out)
    if seed % 2:
        txt = txt.rstrip(b'\n')               # no trailing newline
    return txt


@pytest.mark.parametrize('seed', [0, 1, 2, 3])
def test_messy_fasta(eng, oracle_lib, seed):
    fa = _messy_fasta(seed)
    for enz in ('GATC,GANTC', 'AAAA', 'ATAT,TATA,AA', 'GAGAG', 'DNASE', 'T'):
        want, ws = oracle_lines(oracle_lib, fa, enz)
        got, gs = gpu_lines(eng, fa, enz)
        assert (ws == gs).all(), enz
        assert got == want, enz


def test_kats(eng):
    assert gpu_lines(eng, b'>c1 x\nGATCGATC\n', 'GATC')[0] == ['c1 1 1']
    assert gpu_lines(eng, b'>a\nGAGGTCGAAATCGATTTCGACCTCGAGATC\n', 'GANNTC')[0] == ['a 3 1']   # starts 0,6,12 | 18
    assert gpu_lines(eng, b'>a\nAAAAA\nAAAA\n', 'AAAA')[0] == ['a 1 1']          # non-overlapping finditer
    assert gpu_lines(eng, b'>a\nGATCgatcGANC\n', 'GATC,GATC')[0] == ['a 2 0']
    assert gpu_lines(eng, b'>a d\nACGTA \nCC\n>b\nA\n', 'DNASE')[0] == ['a 3 3', 'b 0 0']
    assert gpu_lines(eng, b'>a\nGA\nTC\n', 'GATC')[0] == ['a 1 0']               # a site across a line break


def test_script_and_errors(eng):
    import RE_sites
    d = tempfile.mkdtemp()
    fa = gzip.open(os.path.join(GOLD, 'links_case.fa.gz'), 'rb').read()
    open(d + '/a.fa', 'wb').write(fa)
    buf = io.StringIO()
    assert RE_sites.main(['-a', d + '/a.fa', '-e', 'GATC,GANTC'], out=buf) == 0
    g = json.load(open(os.path.join(GOLD, 'links_twin.json')))
    assert buf.getvalue().splitlines() == g['re_GATC,GANTC']
    open(d + '/bad.fa', 'wb').write(b'>a:b\nACGT\n')
    assert RE_sites.main(['-a', d + '/bad.fa', '-e', 'GATC'], out=io.StringIO()) == 1
    assert RE_sites.main(['-a', d + '/a.fa', '-e', 'MboI'], out=io.StringIO()) == 1
    assert RE_sites.main(['-a', d + '/a.fa', '-e', 'GATC,'], out=io.StringIO()) == 1
