"""TEST INFRASTRUCTURE — loads the *unmodified* reference (DI-star / Applestar) from
/root/reference so golden vectors can be generated in the build container.

Nothing here is shipped or measured as product; /root/reference does not exist on the
GPU box, so only `oracle/make_golden.py` (run here, outputs committed under
tests/golden/) and the container-only tests call `load_reference()`.

Recipe follows SURVEY.md Appendix E: five third-party modules that contribute no
arithmetic are stubbed (oracle/stubs), and the numpy aliases the vendored pysc2
expects are restored.
"""
import os
import sys

REFERENCE_ROOT = os.environ.get('APPLESTAR_REFERENCE', '/root/reference')
_STUBS = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'stubs')


def reference_available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, 'distar'))


def load_reference():
    """Import `distar` from the read-only reference tree. Returns the `distar` module."""
    if not reference_available():
        raise RuntimeError('reference tree not present at %s' % REFERENCE_ROOT)
    import numpy as np
    for name, typ in (('int', int), ('float', float), ('bool', bool), ('object', object)):
        if name not in np.__dict__:
            setattr(np, name, typ)
    for p in (_STUBS, REFERENCE_ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    import distar  # noqa: F401
    return distar
