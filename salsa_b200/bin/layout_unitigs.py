#!/usr/bin/env python3
"""Drop-in for the reference's layout_unitigs.py as run_pipeline.py calls it:

    layout_unitigs.py -x GFA|abc -l SORTED_LINKS -c CUTOFF -i N -d DIR

The greedy layout stays on the host (it consumes the byte-identical sorted links the GPU path
produces); the re-projection of the iteration-1 BED that follows it runs on the GPU.  Writes scaffolds_iteration_N.p, breakpoints_iteration_{N+1}.txt,
alignment_iteration_{N+1}.bed, scaffold_length_iteration_{N+1}, re_counts_iteration_{N+1}.
`-x`, `-u`, `-t` are accepted and ignored: the reference never loads them on this code path
(layout_unitigs.py:98,206-211)."""
import argparse
import pickle
import sys

import _common  # noqa: F401  (puts the repository root on sys.path)
from salsa_b200.host import reproject, scaffold_graph


def main(argv=None, bed_pass=None):
    ap = argparse.ArgumentParser()
    ap.add_argument('-x', '--graph', default='abc')
    ap.add_argument('-l', '--links', required=True)
    ap.add_argument('-c', '--cutoff', default=1000)
    ap.add_argument('-i', '--iteration', default=1)
    ap.add_argument('-u', '--unitigs', default='abc')
    ap.add_argument('-t', '--tenx', default='abc')
    ap.add_argument('-d', '--directory', default='out')
    a = ap.parse_args(argv)
    it = int(a.iteration)
    previous = {}
    if it > 1:
        try:
            with open('%s/scaffolds_iteration_%d.p' % (a.directory, it - 1), 'rb') as f:
                previous = pickle.load(f)
        except Exception:
            previous = {}
    expanded, contig_length = scaffold_graph.layout(a.directory, it, int(a.cutoff), previous)
    with open('%s/scaffolds_iteration_%d.p' % (a.directory, it), 'wb') as f:
        pickle.dump(expanded, f)
    # an injected bed_pass (the tests' CPU path) also keeps the host RE loop; the product runs both on the GPU
    reproject.update_bed(a.directory, expanded, contig_length, it + 1, refactor=False, bed_pass=bed_pass or _common.bed_pass(),
                         re_pass=None if bed_pass else _common.re_pass())
    return 0


if __name__ == '__main__':
    sys.exit(main())
