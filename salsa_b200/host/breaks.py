"""Splitting scaffolds at suspicious joins: host-side consumer of misasm_iteration_N.report
(what refactor_breaks.py:15-28,205-282 does), followed by the re-projection of
salsa_b200.host.reproject (refactor variant)."""
import pickle

from . import reproject


def read_report(path):
    """Only lines that start with 's' and carry at least one position count (refactor_breaks.py:15-28)."""
    bp = {}
    with open(path, 'r', encoding='latin-1') as f:
        for line in f:
            a = line.strip().split()
            if not line or line[0] != 's' or len(a) < 2:
                continue
            try:
                vals = [int(x) for x in a[1:]]
            except ValueError:
                # the reference assigns the empty list first, then stops at the bad token
                vals = []
                for x in a[1:]:
                    try:
                        vals.append(int(x))
                    except ValueError:
                        break
            bp[a[0]] = vals
    return bp


def split_scaffolds(current, breakpoints, unitig_length, avoid_out):
    """refactor_breaks.py:212-278.  A join (odd index i of the path) is broken when its cumulative
    length lies within 10 kb of a reported position.  Every break records the two new neighbours in
    links_avoid so make_links will not re-join them."""
    new, sid = {}, 1
    for key, path in current.items():
        if key not in breakpoints:
            new['scaffold_%d' % sid] = path
            sid += 1
            continue
        cum, cuts = 0, []
        for i in range(1, len(path), 2):
            cum += unitig_length[path[i].split(':')[0]]
            for pos in breakpoints[key]:
                if pos - 10000 < cum < pos + 10000:
                    cuts.append(i)
                    break
        if not cuts:
            new['scaffold_%d' % sid] = path
            sid += 1
        elif len(cuts) == 1:
            a, b = 'scaffold_%d' % sid, 'scaffold_%d' % (sid + 1)
            sid += 2
            avoid_out.write('%s\t%s\n' % (a, b))
            new[a] = path[:cuts[0] + 1]
            new[b] = path[cuts[0] + 1:]
        else:
            prev, prev_name = 0, ''
            for k, c in enumerate(cuts):
                name = 'scaffold_%d' % sid
                new[name] = path[prev:c + 1]
                sid += 1
                if k == len(cuts) - 1:
                    tail = 'scaffold_%d' % sid
                    new[tail] = path[c + 1:]
                    # the reference pairs the tail with the piece BEFORE the last one (refactor_breaks.py:264)
                    avoid_out.write('%s\t%s\n' % (prev_name, tail))
                    sid += 1
                    break
                prev = c + 1
                if prev_name != '':
                    avoid_out.write('%s\t%s\n' % (prev_name, name))
                prev_name = name
    return new


def refactor(directory, iteration, bed_pass, re_pass=None):
    """refactor_breaks.py -d directory -i iteration"""
    with open('%s/scaffolds_iteration_%d.p' % (directory, iteration - 1), 'rb') as f:
        current = pickle.load(f)
    breakpoints = read_report('%s/misasm_iteration_%d.report' % (directory, iteration))
    unitig_length = {}
    with open(directory + '/scaffold_length_iteration_1', 'r', encoding='latin-1') as f:
        for line in f:
            a = line.split()
            unitig_length[a[0]] = int(a[1])
    with open('%s/links_avoid_iteration_%d' % (directory, iteration), 'w', 1) as avoid_out:
        new = split_scaffolds(current, breakpoints, unitig_length, avoid_out)
    with open('%s/scaffolds_iteration_%d.p' % (directory, iteration - 1), 'wb') as f:
        pickle.dump(new, f)
    reproject.update_bed(directory, new, unitig_length, iteration, refactor=True, bed_pass=bed_pass, re_pass=re_pass)
    # DONE vouches for a complete alignment_iteration_N.bed (refactor_breaks.py writes the BED, then the marker): a pass
    # that writes the BED behind the pipeline's back creates the marker itself, behind the finished file
    marker = '%s/misasm_%d.DONE' % (directory, iteration)
    after = getattr(bed_pass, 'after_write', None)
    if after is not None:
        after('%s/alignment_iteration_%d.bed' % (directory, iteration), marker)
    else:
        open(marker, 'w').close()
    return new
