"""Contig -> scaffold re-projection of the iteration-1 BED, scaffold lengths and per-scaffold
RE counts: what layout_unitigs.py:708-886 (after every layout) and refactor_breaks.py:35-203
(after every misjoin check) do with their `update_bed`.  Host-side; SURVEY section 8(f) ranks a
GPU version of the BED pass as the next widening of the hot path.

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


def _map(line_attrs, contig, info, contig_length):
    off, _, ori = info[contig]
    s, e = int(line_attrs[1]), int(line_attrs[2])
    if ori == 'FOW':
        return off + s, off + e
    return off + contig_length[contig] - e, off + contig_length[contig] - s


def reproject_bed(src, dst, c2s, info, contig_length, refactor):
    """The BED pass of update_bed.  Sliding adjacent-line window over the iteration-1 BED."""
    with open(src, 'r', encoding='latin-1', newline='\n') as f, open(dst, 'w', encoding='latin-1', newline='\n') as out:
        prev_line = ''
        buf = []
        for line in f:
            if prev_line == '':
                prev_line = line
                continue
            pa, ca = prev_line.split(), line.split()
            pc, cc = pa[0], ca[0]
            if refactor and pc == cc:
                continue                                    # prev_line stays (refactor_breaks.py:150-151)
            pr, cr = pa[3].split('/'), ca[3].split('/')
            first, second = pr[1], cr[1]                    # IndexError without '/', like the reference
            if pc in c2s and cc in c2s:
                ps, cs = c2s[pc], c2s[cc]
                if pr[0] == cr[0] and first == '1' and second == '2' and (not refactor or ps != cs):
                    a, b = _map(pa, pc, info, contig_length)
                    buf.append('%s\t%d\t%d\t%s\n' % (ps, a, b, pa[3]))
                    a, b = _map(ca, cc, info, contig_length)
                    buf.append('%s\t%d\t%d\t%s\n' % (cs, a, b, ca[3]))
                    if len(buf) >= 1000000:
                        out.writelines(buf)
                        buf = []
            prev_line = line
        out.writelines(buf)


def update_bed(directory, scaffolds, contig_length, out_iteration, refactor):
    """Writes alignment_iteration_{out}.bed, scaffold_length_iteration_{out}, re_counts_iteration_{out}."""
    re_counts = {}
    with open(directory + '/re_counts_iteration_1', 'r', encoding='latin-1') as f:
        for line in f:
            a = line.split()
            re_counts[a[0]] = (int(a[1]), int(a[2]))
    c2s, info, slen = placement(scaffolds, contig_length, (lambda p: len(p) - 1) if refactor else len)
    sre = scaffold_re_counts(scaffolds, info, slen, contig_length, re_counts)
    dst = '%s/alignment_iteration_%d.bed' % (directory, out_iteration)
    if refactor or not os.path.isfile(dst):
        reproject_bed(directory + '/alignment_iteration_1.bed', dst, c2s, info, contig_length, refactor)
    with open('%s/scaffold_length_iteration_%d' % (directory, out_iteration), 'w') as f:
        for key in slen:
            f.write('%s\t%d\n' % (key, slen[key]))
    with open('%s/re_counts_iteration_%d' % (directory, out_iteration), 'w') as f:
        for key in sre:
            f.write('%s\t%d\t%d\n' % (key, sre[key][0], sre[key][1]))
