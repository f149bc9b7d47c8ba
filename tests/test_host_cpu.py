"""Host-side logic that runs without a GPU: the flat input arena of the step graph and the prefix fork's gradient rule."""
import torch


def test_tree_arena_layout_and_single_copy():
    """learner.tree_arena: every tensor of the tree is a view of ONE flat buffer (256-byte aligned), values preserved; two
    arenas built from trees of the same shapes have the same layout, so one flat copy moves a whole batch."""
    from applestar_b200.learner import tree_arena
    g = torch.Generator().manual_seed(0)
    tree = {'a': torch.randn(3, 5, generator=g), 'b': {'c': torch.randint(0, 9, (7,), generator=g, dtype=torch.int64),
                                                        'd': torch.rand(2, 2, generator=g) > 0.5},
            'h': [(torch.randn(4, generator=g), torch.randn(4, generator=g))], 'n': 3}
    flat, view = tree_arena(tree, torch.device('cpu'))
    assert view['n'] == 3 and torch.equal(view['a'], tree['a']) and torch.equal(view['b']['c'], tree['b']['c'])
    assert torch.equal(view['b']['d'], tree['b']['d']) and torch.equal(view['h'][0][1], tree['h'][0][1])
    base = flat.data_ptr()
    for t in (view['a'], view['b']['c'], view['b']['d'], view['h'][0][0], view['h'][0][1]):
        off = t.data_ptr() - base
        assert 0 <= off < flat.numel() and off % 256 == 0
    other = {'a': torch.randn(3, 5, generator=g), 'b': {'c': torch.randint(0, 9, (7,), generator=g, dtype=torch.int64),
                                                         'd': torch.rand(2, 2, generator=g) > 0.5},
             'h': [(torch.randn(4, generator=g), torch.randn(4, generator=g))], 'n': 3}
    flat2, view2 = tree_arena(other, torch.device('cpu'))
    assert flat2.numel() == flat.numel()
    flat.copy_(flat2)                                   # the whole batch with one copy
    assert torch.equal(view['a'], other['a']) and torch.equal(view['b']['d'], other['b']['d'])
    assert torch.equal(view['h'][0][0], other['h'][0][0])


def test_prefix_fork_gradient_equals_slicing():
    """ops.prefix_fork(x, n) -> (x, x[:n]) with the sum of both consumers' gradients formed in one pass: same values as
    autograd's own slice + accumulate (reference: model.py:128-137 feeds the heads `[:-batch_size]`)."""
    from applestar_b200 import ops
    g = torch.Generator().manual_seed(1)
    x0 = torch.randn(6, 4, 3, generator=g)
    wa, wb = torch.randn(6, 4, 3, generator=g), torch.randn(4, 4, 3, generator=g)
    x = x0.clone().requires_grad_(True)
    full, head = ops.prefix_fork(x, 4)
    ((full * wa).sum() + (head * wb).sum()).backward()
    y = x0.clone().requires_grad_(True)
    ((y * wa).sum() + (y[:4] * wb).sum()).backward()
    assert torch.allclose(x.grad, y.grad)
    # only the heads use it / only the full-size consumers use it
    x = x0.clone().requires_grad_(True)
    _, head = ops.prefix_fork(x, 4)
    (head * wb).sum().backward()
    assert torch.allclose(x.grad[:4], wb) and float(x.grad[4:].abs().sum()) == 0
    x = x0.clone().requires_grad_(True)
    full, _ = ops.prefix_fork(x, 4)
    (full * wa).sum().backward()
    assert torch.allclose(x.grad, wa)
    # no gradient wanted: plain slices
    a, b = ops.prefix_fork(x0, 4)
    assert a is x0 and torch.equal(b, x0[:4])


def test_product_never_touches_the_oracle():
    """The oracle is test infrastructure: nothing under applestar_b200/ (nor the shim under integration/) may import it,
    bench.py only inside its CPU-baseline / library-baseline / reference-arm functions, __graft_entry__ only in smoke()."""
    import ast
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

    def oracle_imports(path):
        tree = ast.parse(open(path).read())
        hits = []
        for node in ast.walk(tree):
            names = []
            if isinstance(node, ast.Import):
                names = [a.name for a in node.names]
            elif isinstance(node, ast.ImportFrom):
                names = [node.module or '']
            if any(n == 'oracle' or n.startswith('oracle.') for n in names):
                hits.append(node.lineno)
        return tree, hits

    for base in ('applestar_b200', 'integration'):
        for dirpath, _, files in os.walk(os.path.join(root, base)):
            for f in files:
                if f.endswith('.py'):
                    _, hits = oracle_imports(os.path.join(dirpath, f))
                    assert not hits, (os.path.join(dirpath, f), hits)
    allowed = {'bench.py': {'cpu_learner_throughput', 'cpu_inference_throughput', 'gpu_library_baseline', 'reference_arm'},
               '__graft_entry__.py': {'smoke'}}
    for fname, funcs in allowed.items():
        tree, hits = oracle_imports(os.path.join(root, fname))
        inside = set()
        for node in ast.walk(tree):
            if isinstance(node, ast.FunctionDef):
                for sub in ast.walk(node):
                    if isinstance(sub, (ast.Import, ast.ImportFrom)) and sub.lineno in hits:
                        inside.add((sub.lineno, node.name))
        assert {ln for ln, _ in inside} == set(hits), (fname, 'module-level oracle import')
        assert {fn for _, fn in inside} <= funcs, (fname, sorted({fn for _, fn in inside}))
