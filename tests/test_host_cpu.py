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


def test_actor_copy_helpers_slot_semantics():
    """actor_batch.copy_input_data / copy_output_data (agent.py:37-89) without the reference tree: only the addressed row of
    the shared buffers changes, ragged parts (entities beyond entity_num, selected units beyond their count, map area
    beyond the game's h x w) keep their padding, outputs land only in the rows whose flag was raised."""
    from applestar_b200 import actor_batch
    E, N = 3, 8
    shared = {'entity_num': torch.zeros(E, dtype=torch.int64), 'selected_units_num': torch.zeros(E, dtype=torch.int64),
              'entity_info': {'x': torch.full((E, N), -1, dtype=torch.int64)},
              'spatial_info': {'height_map': torch.full((E, 6, 6), 9, dtype=torch.uint8), 'effect_a': torch.zeros(E, 4, dtype=torch.int16)},
              'scalar_info': {'time': torch.zeros(E)},
              'action_info': {'selected_units': torch.full((E, 5), -1, dtype=torch.int64), 'delay': torch.zeros(E, dtype=torch.int64)},
              'hidden_state': [(torch.zeros(E, 4), torch.zeros(E, 4))]}
    step = {'entity_num': torch.tensor(3), 'selected_units_num': torch.tensor(2),
            'entity_info': {'x': torch.tensor([5, 6, 7])},
            'spatial_info': {'height_map': torch.ones(4, 5, dtype=torch.uint8), 'effect_a': torch.tensor([1, 2, 3, 4], dtype=torch.int16)},
            'scalar_info': {'time': torch.tensor(12.0)},
            'action_info': {'selected_units': torch.tensor([4, 3]), 'delay': torch.tensor(7)},
            'hidden_state': [(torch.ones(4), 2 * torch.ones(4))], 'value_feature': {'ignored': torch.zeros(1)}}
    actor_batch.copy_input_data(shared, step, 1)
    assert shared['entity_info']['x'][1].tolist() == [5, 6, 7, -1, -1, -1, -1, -1]
    assert shared['entity_info']['x'][0].tolist() == [-1] * N and shared['entity_info']['x'][2].tolist() == [-1] * N
    hm = shared['spatial_info']['height_map']
    assert int(hm[1, :4, :5].sum()) == 20 and int(hm[1].sum()) == 20            # rest of the row's map zeroed
    assert int(hm[0].sum()) == 9 * 36                                           # other rows untouched
    assert shared['action_info']['selected_units'][1].tolist() == [4, 3, -1, -1, -1]
    assert float(shared['scalar_info']['time'][1]) == 12.0 and int(shared['action_info']['delay'][1]) == 7
    assert shared['hidden_state'][0][1][1].tolist() == [2.0] * 4 and int(shared['entity_num'][1]) == 3
    out = {'logit': {'selected_units': torch.arange(E * 2 * 3, dtype=torch.float32).view(E, 2, 3)},
           'selected_units_num': torch.tensor([1, 2, 3]), 'not_in_buffers': torch.zeros(E)}
    buf = {'logit': {'selected_units': torch.full((E, 4, 5), -7.0)}, 'selected_units_num': torch.zeros(E, dtype=torch.int64)}
    actor_batch.copy_output_data(buf, out, torch.tensor([False, True, False]))
    assert torch.equal(buf['logit']['selected_units'][1, :2, :3], out['logit']['selected_units'][1])
    assert float(buf['logit']['selected_units'][1, 2:].min()) == -7.0 and float(buf['logit']['selected_units'][0].max()) == -7.0
    assert buf['selected_units_num'].tolist() == [0, 2, 0]


def test_staging_pack_random_sparsity_patterns():
    """staging.pack on arbitrary row patterns (not only the prefix pattern the data loader produces): a dense row travels
    iff it holds anything other than the fill value, row_of is a bijection onto the kept rows, the dense tensor is
    recovered exactly; all-padding and all-data batches included."""
    from applestar_b200 import staging
    g = torch.Generator().manual_seed(9)
    for density in (0.0, 0.03, 0.5, 1.0):
        T, B, S, K = 3, 2, 64, 17
        dense = torch.full((T, B, S, K), staging.FILL)
        keep = torch.rand(T, B, S, generator=g) < density
        vals = torch.randn(T, B, S, K, generator=g)
        dense[keep] = vals[keep]
        batch = {'teacher_logit': {'selected_units': dense, 'delay': torch.randn(T, B, 5, generator=g)}, 'step': torch.arange(T * B)}
        packed = staging.pack(batch, pin=False)
        c = packed['teacher_logit']['selected_units']
        assert c['rows'].shape == (int(keep.sum()), K) and c['row_of'].shape == (T * B * S,)
        sel = c['row_of'] >= 0
        assert torch.equal(sel.view(T, B, S), keep)
        assert sorted(c['row_of'][sel].tolist()) == list(range(int(keep.sum())))
        rebuilt = torch.full((T * B * S, K), staging.FILL)
        rebuilt[sel] = c['rows'][c['row_of'][sel].long()]
        assert torch.equal(rebuilt.view(T, B, S, K), dense)
        assert packed['teacher_logit']['delay'] is batch['teacher_logit']['delay'] and packed['step'] is batch['step']
        assert batch['teacher_logit']['selected_units'] is dense                  # the caller's batch is not modified
