"""Final scaffold FASTA + AGP writer: host-side consumer of scaffolds_iteration_N.p -- what
get_seq.py does.  Scaffolds are numbered by decreasing length (get_seq.py:104-114), contigs are
joined with 100-N gaps and `U 100 scaffold yes proximity_ligation` AGP lines (:160-170); the
optional second AGP maps `_1`/`_2` pieces back to the pre-cleaning names and coordinates
(:49-75,147-153)."""
import pickle
import re

from .clean import read_fasta, write_wrapped

_COMP = {'A': 'T', 'B': 'N', 'C': 'G', 'G': 'C', 'T': 'A', 'N': 'N', 'R': 'N', 'M': 'N', 'Y': 'N', 'S': 'N', 'W': 'N', 'K': 'N',
         'V': 'N', 'H': 'N', 'D': 'N', 'a': 't', 'c': 'g', 'g': 'c', 't': 'a', 'n': 'n', ' ': ''}


_KNOWN = ''.join(_COMP.keys()).encode('latin-1')
# complement table over all 256 byte values: anything outside the reference's table becomes NUL (itself outside it)
_TABLE = bytes(ord(COMP_) if COMP_ else 0 for COMP_ in (_COMP.get(chr(i), '') for i in range(256)))
_DROP = ''.join(k for k, v in _COMP.items() if not v).encode('latin-1')       # ' ' maps to nothing


def revcomp(s):
    """Reverse complement by the reference's table (get_seq.py:8-16) of a sequence given as latin-1 bytes (or str:
    returned as str).  A character outside the table raises KeyError, as the reference's dict lookup does."""
    b = s if isinstance(s, (bytes, bytearray)) else s.encode('latin-1', 'replace')
    r = b.translate(_TABLE, _DROP)
    if b'\0' in r:
        bad = b.translate(None, _KNOWN)                                # what is left is outside the table
        raise KeyError(min(set(s) - _COMP.keys()) if isinstance(s, str) else chr(min(bad)))
    r = r[::-1]
    return r if isinstance(s, (bytes, bytearray)) else r.decode('latin-1')


def original_coords(name, position, breaks):
    if re.search(r"_[12]$", name) and name[:-2] in breaks:
        orig = name[:-2]
        return orig, position + (breaks[orig] if name.endswith('_2') else 0)
    return name, position


def write_outputs(cleaned, scaffold_map_path, fasta_out, agp_out, breaks_path=None, agp_orig_out=None):
    if bool(breaks_path) ^ bool(agp_orig_out):
        raise SystemExit("ERROR: Both -b, --breaks and -G, --agp-orig-coords must be supplied together. "
                         "None or both are fine, but not one only.")
    with open(scaffold_map_path, 'rb') as f:
        scaff_map = pickle.load(f)
    breaks = None
    if breaks_path:
        breaks = {}
        with open(breaks_path, 'r') as fh:
            for line in fh:
                t = line.rstrip('\n').split('\t')
                assert t[0] not in breaks
                breaks[t[0]] = int(t[1])
    seq = read_fasta(cleaned, as_bytes=True)
    clen = {k: len(v) for k, v in seq.items()}
    total = {s: sum(clen[p[i].split(':')[0]] for i in range(0, len(p) - 1, 2)) for s, p in scaff_map.items()}
    ranked = sorted(total.items(), key=lambda kv: kv[1], reverse=True)        # stable: ties keep map order
    orig = open(agp_orig_out, 'w') if agp_orig_out else None
    with open(agp_out, 'w') as agp, open(fasta_out, 'wb') as fa:
        for cid, (key, _) in enumerate(ranked, 1):
            path = scaff_map[key]
            start, comp, parts = 1, 1, []
            for i in range(0, len(path) - 1, 2):
                cur, nxt = path[i].split(':'), path[i + 1].split(':')
                n = clen[cur[0]]
                end = n + start - 1
                direction = ''
                if cur[1] == 'B' and nxt[1] == 'E':
                    parts.append(seq[cur[0]])
                    direction = '+'
                if cur[1] == 'E' and nxt[1] == 'B':
                    parts.append(revcomp(seq[cur[0]]))
                    direction = '-'
                cols = ['scaffold_%d\t%d\t%d\t%d\tW' % (cid, start, end, comp), cur[0], '1', str(n), direction + '\n']
                agp.write('\t'.join(cols))
                if orig:
                    oname, ostart = original_coords(cur[0], 1, breaks)
                    cols[1], cols[2], cols[3] = oname, str(ostart), str(n + ostart - 1)
                    orig.write('\t'.join(cols))
                start, comp = end + 1, comp + 1
                if i != len(path) - 2:
                    parts.append(b'N' * 100)
                    end = 100 + start - 1
                    gap = 'scaffold_%d\t%d\t%d\t%d\tU\t100\tscaffold\tyes\tproximity_ligation\n' % (cid, start, end, comp)
                    agp.write(gap)
                    if orig:
                        orig.write(gap)
                    start, comp = end + 1, comp + 1
            fa.write(b'>scaffold_%d\n' % cid)
            write_wrapped(fa, b''.join(parts))
    if orig:
        orig.close()
