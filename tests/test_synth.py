import torch

from applestar_b200 import config as C, synth
from util import load_spec


def test_synth_is_deterministic_and_in_range():
    a = synth.gen_trajectories(2, 2, 32, 7)
    b = synth.gen_trajectories(2, 2, 32, 7)
    assert torch.equal(a['entity_info']['unit_type'], b['entity_info']['unit_type'])
    assert torch.equal(a['teacher_logit']['selected_units'], b['teacher_logit']['selected_units'])
    sun = a['selected_units_num']
    assert ((sun == 0) | (sun >= 2)).all()
    en = a['entity_num'][:4].view(2, 2)
    su = a['action_info']['selected_units']
    for t in range(2):
        for bb in range(2):
            n = int(sun[t, bb])
            if n:
                assert su[t, bb, n - 1] == en[t, bb]
                assert len(set(su[t, bb, :n].tolist())) == n
                assert (su[t, bb, :n - 1] < en[t, bb]).all()
    for name, dt, arc, num in C.ENTITY_FIELDS:
        v = a['entity_info'][name]
        assert v.shape == (6, 32)
        if arc == 'one_hot':
            assert 0 <= int(v.min()) and int(v.max()) < num


def test_state_dict_spec_counts():
    spec = load_spec('winloss')
    assert len(spec) == 432
    assert sum(1 for _, _, t in spec if t) == 394
    n = 0
    for _, s, t in spec:
        if t:
            k = 1
            for d in s:
                k *= d
            n += k
    assert n == 31051023
    sd = synth.synth_state_dict([(k, s) for k, s, t in spec if t][:5], 0)
    sd2 = synth.synth_state_dict([(k, s) for k, s, t in spec if t][:5], 0)
    for k in sd:
        assert torch.equal(sd[k], sd2[k])


def test_trajectory_pack_is_lossless_and_compact():
    """staging.pack (collation side of the compact trajectory layout): only the teacher's emitted selected-units rows
    travel; the dense [T,B,64,N+1] tensor is recovered exactly from rows + row_of."""
    import torch
    from applestar_b200 import staging, synth
    data = synth.gen_trajectories(4, 3, 64, 11)
    packed = staging.pack(data, pin=False)
    c = packed['teacher_logit']['selected_units']
    dense = data['teacher_logit']['selected_units']
    assert c['shape'] == tuple(dense.shape)
    rebuilt = torch.full((c['row_of'].numel(), dense.shape[-1]), staging.FILL)
    sel = c['row_of'] >= 0
    rebuilt[sel] = c['rows'][c['row_of'][sel].long()]
    assert torch.equal(rebuilt.view(dense.shape), dense)
    emitted = int(torch.clamp(data['selected_units_num'], min=1).sum())
    assert c['rows'].shape[0] == emitted                       # exactly the rows the teacher emitted
    saved = staging.packed_bytes(data) - staging.packed_bytes(packed)
    assert saved == (dense.numel() - c['rows'].numel()) * 4 - c['row_of'].numel() * 4 and saved > 0
