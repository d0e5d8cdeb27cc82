#!/usr/bin/env python3
"""Drop-in for the reference's refactor_breaks.py:   refactor_breaks.py -d DIR -i N
Splits scaffolds at the joins misasm_iteration_N.report flags, writes links_avoid_iteration_N and
rewrites scaffolds_iteration_{N-1}.p on the host; re-projects the BED on the GPU."""
import argparse
import sys

import _common  # noqa: F401
from salsa_b200.host import breaks


def main(argv=None, bed_pass=None):
    ap = argparse.ArgumentParser()
    ap.add_argument('-i', '--iteration')
    ap.add_argument('-d', '--directory')
    a = ap.parse_args(argv)
    breaks.refactor(a.directory, int(a.iteration), bed_pass or _common.bed_pass(), re_pass=None if bed_pass else _common.re_pass())
    return 0


if __name__ == '__main__':
    sys.exit(main())
