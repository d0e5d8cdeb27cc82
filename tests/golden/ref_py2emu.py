"""Runs the reference's OWN Python-2 scripts (read from /root/reference at call time, never
copied) under Python 3 with Python-2 semantics restored, to generate golden vectors.

TEST INFRASTRUCTURE, generation time only: /root/reference does not exist on the GPU box; the
vectors this produces are committed (links_ref.json) together with make_golden.py.

The three hot-path scripts (make_links.py, fast_scaled_scores.py, RE_sites.py) use exactly these
Python-2-only constructs, each restored mechanically -- the program logic that runs is the
reference's source text:
  * ``print x, y`` / ``print >> sys.stderr, x``  -> print(...) calls        (line rewrite)
  * ``d.iteritems()``                            -> d.items()                (text rewrite)
  * ``long(x)``                                  -> int                      (injected builtin)
  * ``a / b`` on two ints floors                 -> every ast.Div node calls _div(a, b), which
                                                    floors iff both operands are ints (AST)
  * ``str(float)`` is '%.12g' (+ '.0')           -> injected ``str`` that formats floats the
                                                    Python-2 way (oracle.py_twin.py2_str)
Not restorable: Python-2 dict iteration order (callers compare sorted lines) -- the same caveat
the reference has between two of its own runs on different builds.  networkx: the image has 3.x;
fast_scaled_scores.py only uses Graph/add_node/add_edge/neighbors/G[u][v], and its
``G.successors`` call raises AttributeError on an undirected Graph in 1.x and 3.x alike, which
its own ``except:`` turns into the neighbours loop.
"""
import ast
import builtins
import contextlib
import io
import os
import re
import sys

REF = '/root/reference'


def _py2_str(x=''):
    from oracle.py_twin import py2_str
    if isinstance(x, float):
        return py2_str(x)
    return builtins.str(x)


def _div(a, b):
    if isinstance(a, int) and isinstance(b, int) and not isinstance(a, bool) and not isinstance(b, bool):
        return a // b
    return a / b


class _Div(ast.NodeTransformer):
    def visit_BinOp(self, node):
        self.generic_visit(node)
        if isinstance(node.op, ast.Div):
            return ast.copy_location(ast.Call(func=ast.Name(id='_div', ctx=ast.Load()), args=[node.left, node.right], keywords=[]), node)
        return node

    def visit_AugAssign(self, node):
        self.generic_visit(node)
        if isinstance(node.op, ast.Div):
            load = ast.parse(ast.unparse(node.target), mode='eval').body
            return ast.copy_location(ast.Assign(targets=[node.target], value=ast.Call(
                func=ast.Name(id='_div', ctx=ast.Load()), args=[load, node.value], keywords=[])), node)
        return node


def to_py3_text(src):
    out = []
    for line in src.split('\n'):
        m = re.match(r'^(\s*)print\s*>>\s*sys\.stderr\s*,\s*(.*)$', line)
        if m:
            line = '%sprint(%s, file=sys.stderr)' % (m.group(1), m.group(2))
        else:
            m = re.match(r'^(\s*)print\s+(?!\()(.*)$', line) or re.match(r'^(\s*)print\s+(\(.*\)\s*%.*)$', line)
            if m:
                line = '%sprint(%s)' % (m.group(1), m.group(2))
            elif re.match(r'^\s*print\s*$', line):
                line = line.replace('print', 'print()')
        out.append(line.replace('.iteritems()', '.items()'))
    return '\n'.join(out)


def run(script, argv, stdin=None, as_function=False):
    """Executes REF/script as __main__ with sys.argv = [script] + argv.  -> (stdout text, exit code).
    as_function: run the module body as a function body (same statements, names become locals).  Only for
    speed: CPython 3 appends to a str in place for local names but copies it for module-level names, which
    makes the reference's `oline += ...` accumulation quadratic on million-line inputs."""
    src = open(os.path.join(REF, script)).read()
    text = to_py3_text(src)
    if as_function:
        text = 'def _module_body():\n' + ''.join('    ' + ln + '\n' for ln in text.split('\n')) + '_module_body()\n'
    tree = _Div().visit(ast.parse(text, filename=script))
    ast.fix_missing_locations(tree)
    code = compile(tree, os.path.join(REF, script), 'exec')
    g = {'__name__': '__main__', '__file__': os.path.join(REF, script), 'long': int, 'str': _py2_str, '_div': _div,
         'xrange': range, 'unicode': builtins.str}
    buf, err = io.StringIO(), io.StringIO()
    if 'networkx' in src:
        import networkx
        if not hasattr(networkx, 'connected_component_subgraphs'):       # networkx 1.x API, removed in 2.4
            networkx.connected_component_subgraphs = lambda G: [G.subgraph(c).copy() for c in networkx.connected_components(G)]
    old_argv = sys.argv
    sys.argv = [script] + [builtins.str(a) for a in argv]
    rc = 0
    # Python-2 files opened 'w' / 'r' carry bytes: pickle.dump / load on them go to the binary layer
    import pickle as real_pickle
    import types
    shim = types.ModuleType('pickle')
    shim.__dict__.update(real_pickle.__dict__)
    shim.dump = lambda obj, f, *a, **k: real_pickle.dump(obj, getattr(f, 'buffer', f), *a, **k)
    shim.load = lambda f, *a, **k: real_pickle.load(getattr(f, 'buffer', f), *a, **k)
    sys.modules['pickle'] = shim
    try:
        with contextlib.redirect_stdout(buf), contextlib.redirect_stderr(err):
            try:
                exec(code, g)
            except SystemExit as e:
                rc = e.code if isinstance(e.code, int) else (0 if e.code is None else 1)
    finally:
        sys.argv = old_argv
        sys.modules['pickle'] = real_pickle
    return buf.getvalue(), rc
