"""py_twin.py -- line-by-line Python-3 twin of the reference's Python-2 hot-path scripts
(make_links, fast_scaled_scores, RE_sites, correct_links, the BED pass of update_bed).

TEST INFRASTRUCTURE ONLY (see oracle/salsa_oracle.c header): an independent
second restatement used to cross-check the C oracle on small inputs.  Pure
Python loops, so only for small cases.

The reference has no tests of its own and its scripts are Python-2 only; parity of make_links /
fast_scaled_scores / RE_sites is PINNED by tests/golden/links_ref.json, which the reference's own
scripts produced under tests/golden/ref_py2emu.py (the twin and the C oracle both reproduce it), and
update_bed's BED pass by tests/golden/pipeline_ref/ (a stage-by-stage run of the reference's programs);
correct_links by tests/golden/correct_links_ref.json (the unmodified correct_links.cpp built against the
boost stand-in of oracle/_shim).  Deviations from a
literal transliteration are exactly the Python-2 semantics that Python 3
changed, each emulated explicitly:
  * ``str(float)``  -> ``py2_str`` ('%.12g' + '.0'),  evidence
    fast_scaled_scores.py:39-40
  * ``int / int``   -> ``//``                          RE_sites.py:42,71
  * ``long(text)``  -> ``int(text)``                   make_links.py:81-82
  * dict iteration order: Python-2 hash order is not reproducible; callers
    compare outputs as sorted line lists.
Files are read as latin-1 so ``str`` comparison equals byte comparison
(make_links.py:90).
"""
import re


def py2_str(x):
    """Python-2 ``str()`` of an int or float."""
    if isinstance(x, float):
        if x != x:
            return 'nan'
        if x in (float('inf'), float('-inf')):
            return 'inf' if x > 0 else '-inf'
        s = '%.12g' % x
        if '.' not in s and 'e' not in s:
            s += '.0'
        return s
    return str(x)


def make_links(mapping, directory, iteration):
    """make_links.py:11-172.  Returns the lines written to contig_links_iteration_N
    (in insertion order) and also writes the file."""
    contig_lengths = {}
    RF_counts_left = {}
    RF_counts_right = {}
    iteration = str(iteration)
    # dup_links: dead code in the reference (str == int), make_links.py:20-27
    avoid_links = {}
    if int(iteration) > 1:
        with open(directory + '/links_avoid_iteration_' + iteration, 'r', encoding='latin-1') as f:
            for line in f:
                attrs = line.split()
                avoid_links[attrs[0] + '$' + attrs[1]] = 1
                avoid_links[attrs[1] + '$' + attrs[0]] = 1
    with open(directory + '/scaffold_length_iteration_' + iteration, 'r', encoding='latin-1') as cfile:
        for line in cfile.readlines():
            attrs = line.split()
            contig_lengths[attrs[0]] = float(attrs[1])
    with open(directory + '/re_counts_iteration_' + iteration, 'r', encoding='latin-1') as f:
        for line in f:
            attrs = line.split()
            RF_counts_left[attrs[0]] = int(attrs[1])
            RF_counts_right[attrs[0]] = int(attrs[2])

    contig_links = {}
    norm_score = {}
    prev_line = ''
    with open(mapping, 'r', encoding='latin-1', newline='\n') as f:
        for line in f:
            attrs = line.split()
            if prev_line == '':
                prev_line = line
                continue
            prev_attrs = prev_line.split()
            if prev_attrs[0] != attrs[0]:
                ks = prev_attrs[0] + '$' + attrs[0]
                if ks in avoid_links:
                    continue
                prev_read = prev_attrs[3]
                curr_read = attrs[3]
                if prev_read.split('/')[0] == curr_read.split('/')[0]:
                    pos1 = (int(prev_attrs[1]) + int(prev_attrs[2])) / 2.0
                    pos2 = (int(attrs[1]) + int(attrs[2])) / 2.0
                    key = ''
                    if prev_attrs[0] < attrs[0]:
                        first = prev_attrs[0]
                        second = attrs[0]
                    else:
                        first = attrs[0]
                        second = prev_attrs[0]
                        pos1, pos2 = pos2, pos1
                    len1 = contig_lengths[first]
                    len2 = contig_lengths[second]
                    l1 = len1 / 2
                    l2 = len2 / 2
                    r1 = l1 / len1
                    r2 = l2 / len2
                    if pos1 <= l1 and pos2 <= l2:
                        key = first + ":B$" + second + ':B'
                        norm_score[key] = RF_counts_left[first] * r1 + RF_counts_left[second] * r2
                    if pos1 <= l1 and pos2 > len2 - l2:
                        key = first + ":B$" + second + ':E'
                        norm_score[key] = RF_counts_left[first] * r1 + RF_counts_right[second] * r2
                    if pos1 > len1 - l1 and pos2 <= l2:
                        key = first + ":E$" + second + ':B'
                        norm_score[key] = RF_counts_right[first] * r1 + RF_counts_left[second] * r2
                    if pos1 > len1 - l1 and pos2 > len2 - l2:
                        key = first + ":E$" + second + ':E'
                        norm_score[key] = RF_counts_right[first] * r1 + RF_counts_right[second] * r2
                    if key not in contig_links and key != '':
                        contig_links[key] = 0
                    if key != '':
                        contig_links[key] += 1
            prev_line = line

    out = []
    for key in contig_links:
        if key != '':
            edge = key.split('$')
            if norm_score[key] == 0:
                score = 0
            else:
                score = contig_links[key] * 1.0 / norm_score[key]
            out.append(edge[0] + '\t' + edge[1] + '\t' + py2_str(score) + '\t' + py2_str(contig_links[key]) + '\n')
    with open(directory + '/contig_links_iteration_' + iteration, 'w', encoding='latin-1') as ofile:
        ofile.writelines(out)
    return out


def fast_scaled_scores(directory, iteration):
    """fast_scaled_scores.py:1-90, with the nx.Graph replaced by the adjacency
    dict it is used as (G.successors raises AttributeError on an undirected
    graph, so the ``except`` branch with G.neighbors is the live one, :22-25)."""
    G = {}        # node -> {nbr: attrs}

    def get_max_incident(start, end):
        mx = 0
        for e in G[start]:
            if e != end and G[start][e]['weight'] > mx:
                mx = G[start][e]['weight']
        return mx

    def get_max_weight(start, end):
        return max(get_max_incident(start, end), get_max_incident(end, start))

    node2weight = {}
    path = directory + '/contig_links_iteration_' + str(iteration)
    with open(path, 'r', encoding='latin-1') as f:
        for line in f:
            line = line.strip().split()
            G.setdefault(line[0], {})
            G.setdefault(line[1], {})
            for k in (0, 1):
                if line[k] not in node2weight:
                    node2weight[line[k]] = float(line[2])
                elif float(line[2]) >= node2weight[line[k]]:
                    node2weight[line[k]] = float(line[2])
            attrs = dict(weight=float(line[2]), isNeighbor="False", isGood="False", links=int(line[3]))
            G[line[0]][line[1]] = attrs
            G[line[1]][line[0]] = attrs
    out = []
    with open(path, 'r', encoding='latin-1') as f1:
        for line in f1:
            attrs = line.strip().split()
            u = attrs[0]
            v = attrs[1]
            w = float(attrs[2])
            max_u = node2weight.get(u, 0)
            max_v = node2weight.get(v, 0)
            bestAlt = max(max_u, max_v)
            if bestAlt == w:
                bestAlt = get_max_weight(u, v)
            if bestAlt == 0:
                bestAlt = 1
            out.append(u + '\t' + v + '\t' + py2_str(w) + '\t' + py2_str(bestAlt) + '\t' + py2_str(w / bestAlt)
                       + '\t' + str(G[u][v]['links']) + '\t' + G[u][v]['isNeighbor'] + '\t' + G[u][v]['isGood'] + '\n')
    with open(directory + '/contig_links_scaled_iteration_' + str(iteration), 'w', encoding='latin-1') as ofile:
        ofile.writelines(out)
    return out


class Exit1(Exception):
    """stands for sys.exit(1) in RE_sites.py:13-15,48-50"""


def parse_fasta(fh):
    """RE_sites.py:6-28"""
    fa = {}
    current_short_name = None
    for ln in fh:
        if ln[0] == '>':
            if ":" in ln:
                raise Exit1("fasta records contain ':'")
            long_name = ln[1:].rstrip()
            current_short_name = long_name.split()[0]
            fa[current_short_name] = []
        else:
            fa[current_short_name].append(ln.rstrip())
    for short_name, nuc_list in fa.items():
        fa[short_name] = ''.join(nuc_list)
    return fa


def re_sites(assembly, enzyme):
    """RE_sites.py:30-88.  Returns the stdout lines ('id left right')."""
    with open(assembly, 'r', encoding='latin-1', newline='\n') as fh:
        f = parse_fasta(fh)
    out = []
    if enzyme == "DNASE":
        for key in f:
            out.append('%s %d %d' % (key, len(f[key]) // 2, len(f[key]) // 2))
        return out
    enzymes_input = enzyme.replace(' ', '').split(',')
    final_enzymes = []
    for each in enzymes_input:
        if not re.match("^[ACGTN]+$", each):
            raise Exit1("bad enzyme %s" % each)
        if 'N' in each:
            final_enzymes.append(each.replace('N', 'G'))
            final_enzymes.append(each.replace('N', 'A'))
            final_enzymes.append(each.replace('N', 'T'))
            final_enzymes.append(each.replace('N', 'C'))
        else:
            final_enzymes.append(each)
    for key in f:
        seq = f[key]
        left_count = 0
        rigt_count = 0
        for enzyme_ in final_enzymes:
            pos = [m.start(0) for m in re.finditer(enzyme_, seq)]
            length = len(seq)
            for each in pos:
                if each < length // 2:
                    left_count += 1
                else:
                    rigt_count += 1
        out.append('%s %d %d' % (key, left_count, rigt_count))
    return out


def correct_links(gfa, hic_links):
    """correct_links.cpp:70-354 as a second, independent restatement (pinned by
    tests/golden/correct_links_ref.json, the compiled reference's stdout).  Dijkstra on unit weights is a
    breadth-first search.  Emulated C++ behaviours: unordered_map::operator[] inserting id 0 for
    unknown names, the vecS graph growing to the largest edge endpoint, (long)FLT_MAX failing the
    INT_MIN/INT_MAX test, ostream's '%g' formatting of doubles, the inconsistent `added` keys.
    Returns the stdout lines."""
    from collections import deque

    INT_MAX, INT_MIN, UNREACH = 2 ** 31 - 1, -2 ** 31, -2 ** 63

    class IdMap(dict):
        def __missing__(self, k):
            self[k] = 0
            return 0

    unitig2id = IdMap()
    out_edges = []
    nid = 0

    def add_edge(a, b):
        while len(out_edges) <= max(a, b):
            out_edges.append([])
        out_edges[a].append(b)

    with open(gfa, 'r', encoding='latin-1') as f:
        for line in f:
            line = line.rstrip('\n')
            if not line:
                continue
            t = line.split()
            if line[0] == 'S':
                v = t[1] if len(t) > 1 else ''
                for suf in (':FOW', ':REV'):
                    if (v + suf) not in unitig2id:
                        unitig2id[v + suf] = nid
                        nid += 1
            if line[0] == 'L':
                t = t + [''] * 6
                c1, o1, c2, o2 = t[1], t[2], t[3], t[4]
                v1 = v2 = v3 = v4 = ''
                if o1 == '+' and o2 == '+':
                    v1, v2, v3, v4 = c1 + ':FOW', c2 + ':FOW', c2 + ':REV', c1 + ':REV'
                if o1 == '+' and o2 == '-':
                    v1, v2, v3, v4 = c1 + ':FOW', c2 + ':REV', c2 + ':FOW', c1 + ':REV'
                if o1 == '-' and o2 == '+':
                    v1, v2, v3, v4 = c1 + ':REV', c2 + ':FOW', c2 + ':REV', c1 + ':FOW'
                if o1 == '-' and o2 == '-':
                    v1, v2, v3, v4 = c1 + ':REV', c2 + ':REV', c2 + ':FOW', c1 + ':FOW'
                a, b = unitig2id[v1], unitig2id[v2]
                add_edge(a, b)
                a, b = unitig2id[v3], unitig2id[v4]
                add_edge(a, b)

    def bfs(src):
        d = [UNREACH] * len(out_edges)
        if 0 <= src < len(out_edges):
            d[src] = 0
            q = deque([src])
            while q:
                x = q.popleft()
                for y in out_edges[x]:
                    if d[y] == UNREACH:
                        d[y] = d[x] + 1
                        q.append(y)
        return d

    def g6(x):
        return '%g' % x

    out = []
    added = {}
    with open(hic_links, 'r', encoding='latin-1') as f:
        for line in f:
            line = line.rstrip('\n')
            t = line.split()
            a = t[0] if len(t) > 0 else ''
            b = t[1] if len(t) > 1 else ''
            try:
                c, d, e, ff = (float(t[i]) for i in (2, 3, 4, 5))
            except (IndexError, ValueError):
                c = d = e = ff = 0.0
            x = t[6] if len(t) > 6 else ''
            h = t[7] if len(t) > 7 else ''
            if e < 1:
                out.append(line)
                continue
            u, v = a[:-2] if len(a) >= 2 else '', b[:-2] if len(b) >= 2 else ''
            if all((k not in unitig2id) for k in (u + ':FOW', u + ':REV', v + ':FOW', v + ':REV')):
                out.append(line)
                continue
            dist_fow = bfs(unitig2id[u + ':FOW'])
            dist_rev = bfs(unitig2id[u + ':REV'])
            minpath, ori, ratio = -1, '', 0.1
            done = printed = inlimits = False
            for orientation in ('BB', 'BE', 'EB', 'EE'):
                printed = False
                v1 = u + (':REV' if orientation[0] == 'B' else ':FOW')
                v2 = v + (':FOW' if orientation[1] == 'B' else ':REV')
                if a not in added and b not in added:
                    i1, i2 = unitig2id[v1], unitig2id[v2]
                    if 0 <= i1 < len(out_edges) and i2 in out_edges[i1]:
                        added[u] = True
                        added[v] = True
                        out.append('\t'.join([u + ':' + orientation[0], v + ':' + orientation[1], g6(c), g6(d), g6(e), g6(ff), x, h]))
                        done = printed = True
                        break
                    dv = dist_rev if orientation[0] == 'B' else dist_fow
                    path_len = dv[i2] if 0 <= i2 < len(dv) else UNREACH
                    if INT_MIN < path_len < INT_MAX:
                        inlimits = True
                        if minpath == -1:
                            minpath, ori = path_len, orientation
                        elif path_len < minpath:
                            ratio = 1.0 * path_len / minpath
                            minpath, ori = path_len, orientation
            if not inlimits:
                continue
            if 0 < ratio < 1 and INT_MIN < minpath < INT_MAX and not done:
                out.append('\t'.join([u + ':' + ori[0], v + ':' + ori[1], g6(c), g6(d), g6(e), g6(ff), x, h]))
                added[a] = True
                printed = True
                added[b] = True
            elif not printed:
                out.append(line)
                added[a] = True
                added[b] = True
    return out


def reproject_bed(src, dst, c2s, info, contig_length, refactor):
    """The BED pass of update_bed: layout_unitigs.py:816-886 (refactor=False) and
    refactor_breaks.py:133-203 (refactor=True).  Sliding adjacent-line window over the
    iteration-1 BED; byte strings, so ``split()`` has the Python-2 whitespace set.
    c2s: contig -> scaffold key; info: contig -> (offset, end, 'FOW'|'REV'); str keys."""
    def mapped(attrs, contig):
        off, _, ori = info[contig]
        s, e = int(attrs[1]), int(attrs[2])
        if ori == 'FOW':
            return off + s, off + e
        return off + contig_length[contig] - e, off + contig_length[contig] - s

    with open(src, 'rb') as f, open(dst, 'wb') as out:
        prev_line = b''
        for line in f:
            if prev_line == b'':
                prev_line = line
                continue
            pa, ca = prev_line.split(), line.split()
            pc, cc = pa[0].decode('latin-1'), ca[0].decode('latin-1')
            if refactor and pc == cc:
                continue                                    # prev_line stays (refactor_breaks.py:150-151)
            pr, cr = pa[3].split(b'/'), ca[3].split(b'/')
            first, second = pr[1], cr[1]                    # IndexError without '/', like the reference
            if pc in c2s and cc in c2s:
                ps, cs = c2s[pc], c2s[cc]
                if pr[0] == cr[0] and first == b'1' and second == b'2' and (not refactor or ps != cs):
                    a, b = mapped(pa, pc)
                    out.write(b'%s\t%d\t%d\t%s\n' % (ps.encode('latin-1'), a, b, pa[3]))
                    a, b = mapped(ca, cc)
                    out.write(b'%s\t%d\t%d\t%s\n' % (cs.encode('latin-1'), a, b, ca[3]))
            prev_line = line
