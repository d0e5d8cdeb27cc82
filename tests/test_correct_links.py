"""correct_links (GFA-guided link re-orientation, correct_links.cpp:70-354): the boost-free C++
host program against the independent Python twin.  CPU only (the program does no GPU work).
The reference itself needs boost::graph and cannot be built here: parity unpinned."""
import os
import subprocess
import tempfile

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, 'salsa_b200', 'bin', 'correct_links')


@pytest.fixture(scope='module')
def prog():
    r = subprocess.run(['make', '-C', os.path.join(ROOT, 'salsa_b200', 'csrc'), os.path.join(ROOT, 'salsa_b200', 'bin', 'correct_links')],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    return BIN


def run_both(prog, gfa_text, links_text):
    from oracle import py_twin
    d = tempfile.mkdtemp()
    open(d + '/g.gfa', 'w').write(gfa_text)
    open(d + '/l.txt', 'w').write(links_text)
    r = subprocess.run([prog, '-g', d + '/g.gfa', '-l', d + '/l.txt'], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    got = r.stdout.split('\n')[:-1]
    want = py_twin.correct_links(d + '/g.gfa', d + '/l.txt')
    return got, want


def test_kats(prog):
    gfa = 'S\ta\t*\tLN:i:10\nS\tb\t*\tLN:i:10\nS\tc\t*\tLN:i:10\nS\td\t*\tLN:i:10\nL\ta\t+\tb\t+\t0M\nL\tb\t+\tc\t-\t0M\n'
    links = ('a:B\tb:B\t5.5\t1.0\t5.5\t30\tFalse\tFalse\n'        # direct edge a:FOW -> b:FOW  == orientation EB
             'a:E\tc:E\t4.0\t2.0\t2.0\t25\tFalse\tFalse\n'        # both ends "added" as bare names only: still processed; path a+ -> b+ -> c-
             'c:B\td:B\t3.0\t2.0\t1.5\t22\tFalse\tFalse\n'        # d is isolated: no finite path -> nothing printed
             'x:B\ty:E\t3.0\t2.0\t1.5\t22\tFalse\tFalse\n'        # unknown contigs: echoed
             'a:B\td:E\t0.5\t2.0\t0.25\t9\tFalse\tFalse\n')       # ratio < 1: echoed
    got, want = run_both(prog, gfa, links)
    assert got == want
    assert got[0] == 'a:E\tb:B\t5.5\t1\t5.5\t30\tFalse\tFalse'
    assert got[1].startswith('a:E\tc:E\t4\t2\t2\t25')
    assert 'x:B\ty:E\t3.0\t2.0\t1.5\t22\tFalse\tFalse' in got and got[-1].startswith('a:B\td:E\t0.5')
    assert not any(ln.startswith('c:') for ln in got)


@pytest.mark.parametrize('seed', [0, 1, 2])
def test_random_graphs(prog, seed):
    rng = np.random.default_rng(seed)
    n = 60
    names = ['tig%05d' % i for i in range(n)]
    gfa = ['S\t%s\t*\tLN:i:%d' % (nm, rng.integers(1000, 9000)) for nm in names[:-5]]       # the last 5 have no S line
    for _ in range(90):
        a, b = rng.integers(0, n, 2)
        gfa.append('L\t%s\t%s\t%s\t%s\t0M' % (names[a], '+-'[rng.integers(0, 2)], names[b], '+-'[rng.integers(0, 2)]))
    links = []
    for _ in range(400):
        a, b = rng.choice(n, 2, replace=False)
        w = float(rng.integers(1, 400)) / 7
        alt = float(rng.integers(1, 400)) / 7
        links.append('%s:%s\t%s:%s\t%r\t%r\t%r\t%d\tFalse\tFalse' % (names[a], 'BE'[rng.integers(0, 2)], names[b],
                                                                   'BE'[rng.integers(0, 2)], w, alt, w / alt, rng.integers(1, 999)))
    links.sort(key=lambda s: -float(s.split('\t')[4]))
    got, want = run_both(prog, '\n'.join(gfa) + '\n', '\n'.join(links) + '\n')
    assert got == want
    assert 0 < len(got) <= len(links)
