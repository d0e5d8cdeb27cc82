"""Host logic of the drop-in step scripts that needs no GPU: the vectorised reader of
contig_links_iteration_N must give what the line loop gives, or decline."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'salsa_b200', 'bin'))


def same(a, b):
    na, ua, va, wa, la = a
    nb, ub, vb, wb, lb = b
    ea = [(na[x], na[y]) for x, y in zip(ua, va)]
    eb = [(nb[x], nb[y]) for x, y in zip(ub, vb)]
    return ea == eb and (wa == wb).all() and (la == lb).all() and sorted(na) == sorted(nb)


def test_canonical_reader_equals_line_loop():
    import fast_scaled_scores as F
    rng = np.random.default_rng(4)
    lines, seen = [], set()
    while len(lines) < 5000:
        a, b = (int(x) for x in rng.integers(0, 300, 2))
        if a == b or (min(a, b), max(a, b)) in seen:
            continue
        seen.add((min(a, b), max(a, b)))
        w = ['0', '1.0', '0.285714285714', '1e-05', '12345.678', '103.138095238'][int(rng.integers(6))]
        lines.append(b'ctg%d:%s\tctg%d:%s\t%s\t%d\n' % (a, b'BE'[a % 2:a % 2 + 1], b, b'BE'[b % 2:b % 2 + 1], w.encode(), rng.integers(1, 99999)))
    data = b''.join(lines)
    assert same(F.parse_links_canonical(data), F.parse_links(lines))
    assert same(F.parse_links_canonical(data[:-1]), F.parse_links(lines))          # no final newline


@pytest.mark.parametrize('data', [
    b'a:B b:E 1.0 3\n',                                  # spaces
    b'a:B\tb:E\t1.0\t3\r\n',                             # CR
    b'a:B\tb:E\t1.0\t3\t9\n',                            # a fifth field
    b'a:B\tb:E\t1.0\n',                                  # three fields
    b'a:B\tb:E\t1.0\t3\n\nc:B\td:E\t1.0\t3\n',           # empty line
    b'a:B\tb:E\t1.0\t3\nb:E\ta:B\t2.0\t1\n',             # the same edge twice
    b'a:B\ta:B\t1.0\t3\n',                               # self edge
    b'a:B\tb:E\t-1.0\t3\n',                              # negative weight
    b'a:B\tb:E\tx\t3\n',                                 # not a number
    b'a:B\tb:E\t\t3\n',                                  # empty field
    b'',
])
def test_canonical_reader_declines(data):
    import fast_scaled_scores as F
    assert F.parse_links_canonical(data) is None
