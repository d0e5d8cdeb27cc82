#!/usr/bin/env python3
"""Drop-in for the reference's correct.py:   correct.py ASSEMBLY INPUT_BREAKS BED OUTDIR
Step of -m yes: breaks contigs at the positions break_contigs_start reported.  The FASTA split and the new lengths are
host work; the rewrite of the BED (correct.py:62-92) runs on the GPU."""
import sys

import _common  # noqa: F401
from salsa_b200.host import clean



def main(argv=None, bed_pass=None):
    a = sys.argv[1:] if argv is None else argv
    clean.clean(a[0], a[1], a[2], a[3], bed_pass=bed_pass or _common.clean_pass())
    return 0


if __name__ == '__main__':
    sys.exit(main())
