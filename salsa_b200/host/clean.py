"""Breaking contigs at the positions break_contigs_start found (-m yes): host-side consumer of
`input_breaks` -- what correct.py does.  FASTA split, BED coordinate/name rewrite, new lengths.

A quirk kept for parity: the BED rewrite buffers output lines and flushes only when 1,000,000
have accumulated; the final partial buffer is never written (correct.py:87-92 closes the file
without a last write), so the cleaned BED holds floor(n / 1e6) * 1e6 lines.
"""


def read_fasta(path):
    """name -> sequence (lines rstripped and joined), later duplicate wins (correct.py:4-19)."""
    fa, cur = {}, None
    with open(path, 'r', encoding='latin-1', newline='\n') as fh:
        for ln in fh:
            if ln[0] == '>':
                cur = ln[1:].rstrip().split()[0]
                fa[cur] = []
            else:
                fa[cur].append(ln.rstrip())
    return {k: ''.join(v) for k, v in fa.items()}


def clean(assembly, breaks_path, bed, outdir, flush_every=1000000):
    seqs = read_fasta(assembly)
    cut = {}
    with open(breaks_path, 'r', encoding='latin-1') as f:
        for line in f:
            a = line.split()
            cut[a[0]] = int(a[1])
    new_seq = {}
    for name, s in seqs.items():
        if name in cut:
            new_seq[name + '_1'] = s[:cut[name]]
            new_seq[name + '_2'] = s[cut[name]:]
        else:
            new_seq[name] = s
    with open(outdir + '/alignment_iteration_1.tmp.bed', 'w', encoding='latin-1', newline='\n') as out, \
            open(bed, 'r', encoding='latin-1', newline='\n') as f:
        buf = []
        for line in f:
            a = line.split()
            if a[0] not in cut:
                buf.append('%s\t%s\t%s\t%s\n' % (a[0], a[1], a[2], a[3]))
            else:
                p1, p2, bp = int(a[1]), int(a[2]), cut[a[0]]
                if p1 < bp and p2 > bp:
                    continue                                   # straddles the break: dropped
                if p2 < bp:
                    buf.append('%s_1\t%s\t%s\t%s\n' % (a[0], a[1], a[2], a[3]))
                else:
                    buf.append('%s_2\t%d\t%d\t%s\n' % (a[0], p1 - bp, p2 - bp, a[3]))
            if len(buf) >= flush_every:
                out.writelines(buf)
                buf = []
        # (no final flush: see module docstring)
    with open(outdir + '/asm.cleaned.fasta', 'w') as fo:
        for name, s in new_seq.items():
            fo.write('>' + name + '\n')
            for i in range(0, len(s), 80):
                fo.write(s[i:i + 80] + '\n')
    with open(outdir + '/scaffold_length_iteration_1', 'w') as fo:
        for name, s in new_seq.items():
            fo.write('%s\t%d\n' % (name, len(s)))
