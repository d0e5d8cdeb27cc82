#!/usr/bin/env python3
"""Drop-in for the reference's get_seq.py:   get_seq.py -a CLEANED -f FASTA_OUT -g AGP_OUT -p MAP [-b BREAKS -G AGP_ORIG]
Host-side step: final scaffold FASTA and AGP."""
import argparse
import sys

import _common  # noqa: F401
from salsa_b200.host import agp


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument('-a', '--cleaned')
    ap.add_argument('-b', '--breaks')
    ap.add_argument('-f', '--scaffold')
    ap.add_argument('-g', '--agp')
    ap.add_argument('-G', '--agp-orig-coords', dest='agp_oc')
    ap.add_argument('-p', '--map')
    a = ap.parse_args(argv)
    agp.write_outputs(a.cleaned, a.map, a.scaffold, a.agp, a.breaks, a.agp_oc)
    return 0


if __name__ == '__main__':
    sys.exit(main())
