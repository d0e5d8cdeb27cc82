"""correct_links (GFA-guided link re-orientation, correct_links.cpp:102-347): the boost-free C++ host program and the
Python twin against the REFERENCE: tests/golden/correct_links_ref.json holds the stdout of the unmodified
correct_links.cpp, compiled against the test-only boost stand-in oracle/_shim (tests/golden/make_golden.py), on the
cases of tests/cases.py -- and, when oracle/_ref/correct_links travelled with the repo, the binary is run live too.
CPU only (the program does no GPU work)."""
import json
import os
import subprocess
import tempfile

import pytest

import cases

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, 'salsa_b200', 'bin', 'correct_links')
REF = os.path.join(ROOT, 'oracle', '_ref', 'correct_links')
GOLD = json.load(open(os.path.join(ROOT, 'tests', 'golden', 'correct_links_ref.json')))


@pytest.fixture(scope='module')
def prog():
    r = subprocess.run(['make', '-C', os.path.join(ROOT, 'salsa_b200', 'csrc'), os.path.join(ROOT, 'salsa_b200', 'bin', 'correct_links')],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    return BIN


def run(binary, d):
    r = subprocess.run([binary, '-g', d + '/g.gfa', '-l', d + '/l.txt'], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return r.stdout.split('\n')[:-1]


@pytest.mark.parametrize('name', sorted(cases.CORRECT_LINKS_CASES))
def test_against_reference(prog, name):
    from oracle import py_twin
    gfa, links = cases.CORRECT_LINKS_CASES[name]()
    d = tempfile.mkdtemp()
    open(d + '/g.gfa', 'w').write(gfa)
    open(d + '/l.txt', 'w').write(links)
    want = GOLD[name]
    assert 0 < len(want) <= links.count('\n')
    assert run(prog, d) == want                                              # the drop-in
    assert py_twin.correct_links(d + '/g.gfa', d + '/l.txt') == want         # the oracle's twin
    if os.path.exists(REF):
        assert run(REF, d) == want                                           # the reference itself, live


def test_kat_lines():
    """What the fixture says about the hand-made case: direct edge, two-hop path, isolated contig (nothing printed),
    unknown contigs (echoed), ratio < 1 (echoed), 6 significant digits, the operator[] vertex-0 quirk."""
    got = GOLD['kat']
    assert got[0] == 'a:E\tb:B\t5.5\t1\t5.5\t30\tFalse\tFalse'
    assert got[1].startswith('a:E\tc:E\t4\t2\t2\t25')
    assert 'x:B\ty:E\t3.0\t2.0\t1.5\t22\tFalse\tFalse' in got
    assert any(ln.startswith('a:B\td:E\t0.5') for ln in got)
    assert not any(ln.startswith('c:') for ln in got)
    assert any('123457\t3\t41152.3' in ln for ln in got)
