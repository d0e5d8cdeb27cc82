#!/usr/bin/env python3
"""Drop-in for the reference's correct.py:   correct.py ASSEMBLY INPUT_BREAKS BED OUTDIR
Host-side step of -m yes: breaks contigs at the positions break_contigs_start reported."""
import sys

import _common  # noqa: F401
from salsa_b200.host import clean

if __name__ == '__main__':
    clean.clean(sys.argv[1], sys.argv[2], sys.argv[3], sys.argv[4])
