"""TEST INFRASTRUCTURE (container only): regenerates applestar_b200/tables.py from the
reference's literal action tables. Run: python oracle/make_tables.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.ref_loader import load_reference  # noqa: E402


def bits(seq):
    return ''.join('1' if b else '0' for b in seq)


def chunks(s, indent):
    return ''.join("%s'%s'\n" % (indent, s[i:i + 100]) for i in range(0, len(s), 100))


def main():
    load_reference()
    from distar.agent.default.lib.actions import ACTIONS, SELECTED_UNITS_MASK
    from distar.agent.default.lib.stat import ACTION_RACE_MASK
    assert bits(SELECTED_UNITS_MASK.tolist()) == bits(a['selected_units'] for a in ACTIONS)
    out = ['"""Action-argument tables for the 327 raw actions.\n\n'
           'Derived data (not code) from the reference literals: `ACTIONS` flags\n'
           '(distar/agent/default/lib/actions.py:5-333), `SELECTED_UNITS_MASK` (:403-406) and\n'
           '`ACTION_RACE_MASK` (distar/agent/default/lib/stat.py:533). One character per action,\n'
           'index = action_type id. Regenerate with oracle/make_tables.py (container only).\n"""\n',
           'NUM_ACTIONS = %d\n' % len(ACTIONS)]
    for k in ['queued', 'selected_units', 'target_unit', 'target_location']:
        out.append('ACTION_HAS_%s = (\n%s)\n' % (k.upper(), chunks(bits(a[k] for a in ACTIONS), '    ')))
    out.append('ACTION_RACE_MASK_BITS = {\n')
    for r, m in ACTION_RACE_MASK.items():
        out.append("    '%s': (\n%s    ),\n" % (r, chunks(bits(m.tolist()), '        ')))
    out.append('}\n\n\ndef flags(bits):\n    """bit-string -> list[bool]"""\n    return [c == \'1\' for c in bits]\n')
    with open(os.path.join(ROOT, 'applestar_b200', 'tables.py'), 'w') as f:
        f.write(''.join(out))


if __name__ == '__main__':
    main()
