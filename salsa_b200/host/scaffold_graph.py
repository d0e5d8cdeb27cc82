"""Greedy scaffold layout over the sorted best-buddy links: host-side consumer of
`contig_links_scaled_sorted_iteration_N` (stays on the host by design; it must see
byte-identical edges, which the GPU path guarantees).

Restates what the reference's layout_unitigs.py does when driven by run_pipeline.py
(the unitig / 10x / GFA branches of that script are unreachable from the pipeline:
`-u`/`-t` are never passed and load_GFA is never called, layout_unitigs.py:98,206-211):

  * generate_scaffold_graph (layout_unitigs.py:372-583): stream the links in file order;
    every line read joins the "all links" graph; iteration 1 skips contigs <= cutoff;
    a line with ratio < 1 and >= 100 links ends the scan once one of its ends is taken;
    otherwise, if both ends are free, the edge is kept iff links >= 20 and ratio >= 1.12,
    and both contigs are marked seen either way; every seen contig gets its B-E edge;
  * get_seed_scaffold (:589-613): components with exactly two degree-1 nodes are paths;
  * assign_small_to_seed / insert (:621-705): leftover contigs go where their links sum highest;
  * expansion through the previous iteration's scaffolds + breakpoints file (:975-1022).

Python-2 dict/set iteration order is not reproducible; every container here is
insertion ordered, which makes the layout a deterministic function of its input files.
"""


class Graph:
    """Undirected simple graph with insertion-ordered nodes and edge attributes."""

    def __init__(self):
        self.adj = {}

    def add_edge(self, u, v, **attrs):
        self.adj.setdefault(u, {})
        self.adj.setdefault(v, {})
        d = self.adj[u].get(v)
        if d is None:
            d = {}
            self.adj[u][v] = d
            self.adj[v][u] = d
        d.update(attrs)

    def __contains__(self, n):
        return n in self.adj

    def has_edge(self, u, v):
        return u in self.adj and v in self.adj[u]

    def edge(self, u, v):
        return self.adj[u][v]

    def degree(self, n):
        return len(self.adj[n])

    def components(self):
        seen = set()
        order = {n: i for i, n in enumerate(self.adj)}
        for s in self.adj:
            if s in seen:
                continue
            comp, stack = [], [s]
            seen.add(s)
            while stack:
                x = stack.pop()
                comp.append(x)
                for y in self.adj[x]:
                    if y not in seen:
                        seen.add(y)
                        stack.append(y)
            comp.sort(key=order.__getitem__)
            yield comp

    def walk(self, a, b):
        """The path from a to b in a component whose nodes all have degree <= 2."""
        path, prev, cur = [a], None, a
        while cur != b:
            nxt = [y for y in self.adj[cur] if y != prev]
            prev, cur = cur, nxt[0]
            path.append(cur)
        return path


def split_end(node):
    c, o = node.rsplit(':', 1) if node.count(':') > 1 else node.split(':')
    return c, o


def build_link_graphs(sorted_links_path, iteration, contig_length, cutoff):
    """Returns (G, all_G, seen contigs) -- layout_unitigs.py:372-583."""
    G, all_G, seen = Graph(), Graph(), {}
    with open(sorted_links_path, 'r', encoding='latin-1') as f:
        for line in f:
            a = line.split()
            if len(a) < 5:
                break
            v1, v2 = a[0], a[1]
            c1, c2 = v1.split(':')[0], v2.split(':')[0]
            ratio = float(a[4])
            all_G.add_edge(v1, v2, score=ratio)
            if iteration == 1 and (contig_length[c1] <= cutoff or contig_length[c2] <= cutoff):
                continue
            both_free = v1 not in G and v2 not in G
            if ratio < 1 and float(a[5]) >= 100:
                if both_free:
                    continue            # nothing is added for such a link (its add_edge sits behind -u, never given)
                break                   # an end is already taken: the scan stops for good
            if both_free:
                if int(a[5]) >= 20 and ratio >= 1.12:
                    G.add_edge(v1, v2, score=ratio, linktype='hic')
                seen[c1] = True
                seen[c2] = True
    for ctg in list(seen):
        G.add_edge(ctg + ':B', ctg + ':E', score=0, linktype='implicit')
    return G, all_G, seen


def seed_scaffolds(G):
    """layout_unitigs.py:589-613"""
    seeds, leftovers, idx = {}, {}, 1
    for comp in G.components():
        ends = [n for n in comp if G.degree(n) == 1]
        if len(ends) == 2:
            seeds[idx] = G.walk(ends[0], ends[1])
            idx += 1
        else:
            for n in comp:
                leftovers[n.split(':')[0]] = True
    return seeds, leftovers


def assign_leftovers(leftovers, seeds, all_G):
    """layout_unitigs.py:621-655"""
    assignment = {}
    for contig in leftovers:
        best_sum, best_key = -1, -1
        b, e = contig + ':B', contig + ':E'
        for key, path in seeds.items():
            total, cnt = 0, 0
            for node in path:
                if all_G.has_edge(b, node):
                    total += all_G.edge(b, node)['score']
                    cnt += 1
                if all_G.has_edge(e, node):
                    total += all_G.edge(e, node)['score']
                    cnt += 1
            if cnt != 0 and total > best_sum:
                best_sum, best_key = total, key
        if best_sum != -1:
            assignment[contig] = best_key
    return assignment


def insert_leftovers(assignment, seeds, all_G):
    """layout_unitigs.py:661-705"""
    later = {}
    for contig, key in assignment.items():
        path = seeds[key]
        b, e = contig + ':B', contig + ':E'
        best, orient, pos = -1, '', -1
        for i in range(1, len(path) - 1, 2):
            fow = rev = -1
            if all_G.has_edge(b, path[i]) and all_G.has_edge(e, path[i + 1]):
                fow = all_G.edge(b, path[i])['score'] + all_G.edge(e, path[i + 1])['score']
            if all_G.has_edge(e, path[i]) and all_G.has_edge(b, path[i + 1]):
                rev = all_G.edge(e, path[i])['score'] + all_G.edge(b, path[i + 1])['score']
            if fow >= rev:
                if fow > best:
                    best, orient, pos = fow, 'fow', i
            elif rev > best:
                best, orient, pos = rev, 'rev', i
        if best != -1:
            pair = [b, e] if orient == 'fow' else [e, b]
            path[pos + 1:pos + 1] = pair
            seeds[key] = path
        else:
            later[contig] = True
    return seeds, later


def layout(directory, iteration, cutoff, previous_scaffolds):
    """One layout_unitigs.py run minus update_bed.  Returns the expanded scaffold map
    (name -> [ctg:B|E, ...]) and writes breakpoints_iteration_{N+1}.txt."""
    contig_length = {}
    with open(directory + '/scaffold_length_iteration_1', 'r', encoding='latin-1') as f:
        for line in f:
            a = line.split()
            contig_length[a[0]] = int(a[1])
    G, all_G, _ = build_link_graphs('%s/contig_links_scaled_sorted_iteration_%d' % (directory, iteration), iteration,
                                    contig_length, cutoff)
    seeds, leftovers = seed_scaffolds(G)
    seeds, later = insert_leftovers(assign_leftovers(leftovers, seeds, all_G), seeds, all_G)
    final, sid = {}, 1
    for key in seeds:
        final[sid] = seeds[key]
        sid += 1
    for key in later:
        final[sid] = [key + ':B', key + ':E']
        sid += 1
    prev_len = {}
    with open('%s/scaffold_length_iteration_%d' % (directory, iteration), 'r', encoding='latin-1') as f:
        for line in f:
            a = line.split()
            prev_len[a[0]] = int(a[1])
    placed, expanded = {}, {}
    with open('%s/breakpoints_iteration_%d.txt' % (directory, iteration + 1), 'w') as bp:       # :975-1022
        for key, path in final.items():
            new_path = []
            for i in range(0, len(path) - 1, 2):
                contig = path[i].split(':')[0]
                placed[contig] = True
                if contig[0] == 's' and iteration > 1:
                    inner = previous_scaffolds[contig]
                    if path[i].split(':')[1] + path[i + 1].split(':')[1] != 'BE':
                        inner = inner[::-1]
                    new_path.extend(inner)
                else:
                    new_path.extend([path[i], path[i + 1]])
            expanded['scaffold_%d' % key] = new_path
            cum, cells = 0, []
            for i in range(0, len(path) - 1, 2):
                cum += prev_len[path[i].split(':')[0]]
                cells.append(str(cum))
            bp.write('scaffold_%d\t%s\t\n' % (key, '\t'.join(cells)))
    if iteration > 1:
        for key in previous_scaffolds:
            if key not in placed:
                expanded['scaffold_%d' % sid] = previous_scaffolds[key]
                sid += 1
    else:
        for contig in contig_length:
            if contig not in placed:
                expanded['scaffold_%d' % sid] = [contig + ':B', contig + ':E']
                sid += 1
    return expanded, contig_length
