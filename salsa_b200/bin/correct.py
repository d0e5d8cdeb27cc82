#!/usr/bin/env python3
"""Drop-in for the reference's correct.py:   correct.py ASSEMBLY INPUT_BREAKS BED OUTDIR
Host-side step of -m yes: breaks contigs at the positions break_contigs_start reported."""
import sys

import _common  # noqa: F401
from salsa_b200.host import clean



def main(argv=None):
    a = sys.argv[1:] if argv is None else argv
    clean.clean(a[0], a[1], a[2], a[3])
    return 0


if __name__ == '__main__':
    sys.exit(main())
