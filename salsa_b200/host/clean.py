"""Breaking contigs at the positions break_contigs_start found (-m yes): host-side consumer of
`input_breaks` -- what correct.py does.  FASTA split, BED coordinate/name rewrite, new lengths.

A quirk kept for parity: the BED rewrite buffers output lines and flushes only when 1,000,000
have accumulated; the final partial buffer is never written (correct.py:87-92 closes the file
without a last write), so the cleaned BED holds floor(n / 1e6) * 1e6 lines.
"""


import numpy as np


def _read_fasta_lines(path):
    fa, cur = {}, None
    with open(path, 'r', encoding='latin-1', newline='\n') as fh:
        for ln in fh:
            if ln[0] == '>':
                cur = ln[1:].rstrip().split()[0]
                fa[cur] = []
            else:
                fa[cur].append(ln.rstrip())
    return {k: ''.join(v) for k, v in fa.items()}


def read_fasta(path, as_bytes=False):
    """name -> sequence (lines rstripped and joined; latin-1 bytes when as_bytes), later duplicate wins (correct.py:4-19).
    Whole records at a time (a 3 Gb assembly has 50 million lines); anything unusual -- text before the first
    header, whitespace other than the newlines inside a record -- goes through the line loop, which is the
    reference's own and raises what it raises."""
    with open(path, 'rb') as fh:
        data = fh.read()
    if not data:
        return {}
    slow = (lambda: {k: v.encode('latin-1') for k, v in _read_fasta_lines(path).items()}) if as_bytes else (lambda: _read_fasta_lines(path))
    if data[:1] != b'>':
        return slow()
    starts, q = [0], data.find(b'\n>')
    while q >= 0:
        starts.append(q + 1)
        q = data.find(b'\n>', q + 1)
    starts.append(len(data))
    fa = {}
    for b, e in zip(starts[:-1], starts[1:]):
        nl = data.find(b'\n', b, e)
        head = data[b + 1:nl if nl >= 0 else e]
        body = data[nl + 1:e] if nl >= 0 else b''
        if not head.split() or any(c in body for c in (b' ', b'\t', b'\r', b'\x0b', b'\x0c')) or b'\r' in head:
            return slow()
        seq = _unwrap(body)
        fa[head.decode('latin-1').rstrip().split()[0]] = seq if as_bytes else seq.decode('latin-1')
    return fa


def _unwrap(body):
    """body without its newlines.  Evenly wrapped records (every line but the last as wide as the first) are a
    strided copy; anything else goes through bytes.translate."""
    w = body.find(b'\n')
    if w <= 0:
        return body.translate(None, b'\n')
    k = len(body) // (w + 1)
    tail = body[k * (w + 1):]
    tn = tail.count(b'\n')
    if tn > 1 or (tn == 1 and tail[-1:] != b'\n'):
        return body.translate(None, b'\n')
    a = np.frombuffer(body, dtype=np.uint8, count=k * (w + 1)).reshape(k, w + 1)
    if np.count_nonzero(a == 10) != k or not (a[:, w] == 10).all():       # exactly one newline per row, at its end
        return body.translate(None, b'\n')
    return a[:, :w].copy().tobytes() + tail.rstrip(b"\n")


def _wrapped(s, width):
    b = s if isinstance(s, (bytes, bytearray)) else s.encode('latin-1')
    k = len(b) // width
    out = np.empty((k, width + 1), np.uint8)
    out[:, :width] = np.frombuffer(b, dtype=np.uint8, count=k * width).reshape(k, width)
    out[:, width] = 10
    tail = b[k * width:]
    return out, (tail + b'\n' if tail else b'')


def wrap_fasta(s, width=80):
    """The sequence (str or latin-1 bytes) in lines of `width` characters, every line (the last, shorter one too) ending
    in a newline: what `for i in range(0, len(s), width): write(s[i:i + width] + '\\n')` writes, as one bytes object."""
    out, tail = _wrapped(s, width)
    return out.tobytes() + tail


def write_wrapped(fh, s, width=80):
    """fh.write(wrap_fasta(s)) without the intermediate copy (fh: binary file)."""
    out, tail = _wrapped(s, width)
    if out.size:
        fh.write(out)
    if tail:
        fh.write(tail)


def clean(assembly, breaks_path, bed, outdir, flush_every=1000000, bed_pass=None):
    """bed_pass(src, dst, names, cut, flush_every) does the BED rewrite: the step script passes the CUDA one
    (salsa_b200.reproject.clean_pass); without it the line loop below runs (the tests' CPU path)."""
    seqs = read_fasta(assembly, as_bytes=True)
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
    if bed_pass is not None:
        bed_pass(bed, outdir + '/alignment_iteration_1.tmp.bed', list(seqs.keys()), cut, flush_every)
    else:
        _clean_bed_lines(bed, outdir + '/alignment_iteration_1.tmp.bed', seqs, cut, flush_every)
    with open(outdir + '/asm.cleaned.fasta', 'wb') as fo:
        for name, s in new_seq.items():
            fo.write(('>' + name + '\n').encode('latin-1'))
            write_wrapped(fo, s)
    with open(outdir + '/scaffold_length_iteration_1', 'w') as fo:
        for name, s in new_seq.items():
            fo.write('%s\t%d\n' % (name, len(s)))


def _clean_bed_lines(bed, dst, seqs, cut, flush_every):
    with open(dst, 'w', encoding='latin-1', newline='\n') as out, open(bed, 'r', encoding='latin-1', newline='\n') as f:
        buf = []
        for line in f:
            a = line.split()
            if a[0] not in cut:
                if a[0] not in seqs:
                    raise KeyError(a[0])                       # contig2new[attrs[0]], correct.py:69
                buf.append('%s\t%s\t%s\t%s\n' % (a[0], a[1], a[2], a[3]))
            else:
                p1, p2, bp = int(a[1]), int(a[2]), cut[a[0]]
                if p1 < bp and p2 > bp:
                    continue                                   # straddles the break: dropped
                if a[0] not in seqs:
                    raise KeyError(a[0])                       # contig2new[attrs[0]][k], correct.py:81,86
                if p2 < bp:
                    buf.append('%s_1\t%s\t%s\t%s\n' % (a[0], a[1], a[2], a[3]))
                else:
                    buf.append('%s_2\t%d\t%d\t%s\n' % (a[0], p1 - bp, p2 - bp, a[3]))
            if len(buf) >= flush_every:
                out.writelines(buf)
                buf = []
        # (no final flush: see module docstring)
