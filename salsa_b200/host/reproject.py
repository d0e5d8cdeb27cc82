"""Contig -> scaffold re-projection of the iteration-1 BED, scaffold lengths and per-scaffold
RE counts: what layout_unitigs.py:708-886 (after every layout) and refactor_breaks.py:35-203
(after every misjoin check) do with their `update_bed`.  The placement tables, scaffold lengths
and RE proration are host-side; the pass over the BED itself is injected (`bed_pass`): the product
uses the CUDA one (salsa_b200/reproject.py, csrc/reproject.cu), the tests' CPU runs the oracle's.

Differences between the two reference variants, both kept:
  * layout variant  : writes alignment_iteration_{N+1}.bed only if it does not exist yet
                      (layout_unitigs.py:821); keeps read pairs /1 then /2 whose contigs are both
                      placed, wherever they land;
  * refactor variant: always rewrites alignment_iteration_N.bed; same-contig adjacent lines are
                      skipped with a `continue` that does NOT advance prev_line
                      (refactor_breaks.py:150-151), and only pairs that land on two DIFFERENT
                      scaffolds are kept (:159).
"""
import os


def placement(scaffolds, contig_length, step_to):
    """contig -> scaffold and (offset, end, 'FOW'|'REV'); scaffold lengths.
    step_to: len(path) for the layout variant, len(path)-1 for the refactor variant (identical
    for the even-length paths both produce)."""
    c2s, info, slen = {}, {}, {}
    for key, path in scaffolds.items():
        slen[key] = 0
        off = 0
        for i in range(0, step_to(path), 2):
            contig = path[i].split(':')[0]
            c2s[contig] = key
            ori = path[i].split(':')[1] + path[i + 1].split(':')[1]
            info[contig] = (off, off + contig_length[contig], 'FOW' if ori == 'BE' else 'REV')
            off += contig_length[contig]
            slen[key] += contig_length[contig]
    return c2s, info, slen


def scaffold_re_counts(scaffolds, info, slen, contig_length, re_counts):
    """Midpoint-split proration of contig RE counts (layout_unitigs.py:736-814, refactor_breaks.py:63-131):
    Python-2 integer `length/2`, float proration, int() truncation."""
    out = {}
    for key, path in scaffolds.items():
        if len(path) == 2:
            contig, ori = path[0].split(':')
            left, right = re_counts[contig]
            out[key] = (left, right) if ori == 'B' else (right, left)
            continue
        mid = slen[key] // 2
        s_left = s_right = 0
        for i in range(0, len(path), 2):
            contig = path[i].split(':')[0]
            left, right = re_counts[contig]
            start, end = info[contig][0], info[contig][1]
            if end < mid:
                s_left += left + right
            if start > mid:
                s_right += left + right
            if start <= mid and end >= mid:
                s_left += float(left + right) / contig_length[contig] * (mid - start)
                s_right += float(left + right) / contig_length[contig] * (end - mid)
        out[key] = (int(s_left), int(s_right))
    return out


def update_bed(directory, scaffolds, contig_length, out_iteration, refactor, bed_pass, re_pass=None):
    """Writes alignment_iteration_{out}.bed, scaffold_length_iteration_{out}, re_counts_iteration_{out}.
    bed_pass(src, dst, c2s, info, contig_length, refactor) does the BED pass itself: the step
    scripts pass the GPU one (salsa_b200.reproject.bed_pass).  re_pass(scaffolds, contig_length, re_counts) ->
    {scaffold: (left, right)} does the RE proration: the step scripts pass the GPU one
    (salsa_b200.reproject.re_pass); without it the host loop above runs (the tests' CPU path)."""
    re_counts = {}
    with open(directory + '/re_counts_iteration_1', 'r', encoding='latin-1') as f:
        for line in f:
            a = line.split()
            re_counts[a[0]] = (int(a[1]), int(a[2]))
    c2s, info, slen = placement(scaffolds, contig_length, (lambda p: len(p) - 1) if refactor else len)
    sre = re_pass(scaffolds, contig_length, re_counts) if re_pass else scaffold_re_counts(scaffolds, info, slen, contig_length, re_counts)
    dst = '%s/alignment_iteration_%d.bed' % (directory, out_iteration)
    if refactor or not os.path.isfile(dst):
        bed_pass(directory + '/alignment_iteration_1.bed', dst, c2s, info, contig_length, refactor)
    with open('%s/scaffold_length_iteration_%d' % (directory, out_iteration), 'w') as f:
        for key in slen:
            f.write('%s\t%d\n' % (key, slen[key]))
    with open('%s/re_counts_iteration_%d' % (directory, out_iteration), 'w') as f:
        for key in sre:
            f.write('%s\t%d\t%d\n' % (key, sre[key][0], sre[key][1]))
