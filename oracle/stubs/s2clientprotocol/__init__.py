"""Permissive stand-in for s2clientprotocol: the reference imports protobuf enums at
module import time for tables the model never reads. Test infrastructure only."""
import sys
import types


class _Dummy:
    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Dummy()

    def __getattr__(self, name):
        if name.startswith('__') and name.endswith('__'):
            raise AttributeError(name)
        return _Dummy()

    def __iter__(self):
        return iter(())

    def __int__(self):
        return 0

    def __index__(self):
        return 0

    def __hash__(self):
        return 0

    def __eq__(self, o):
        return isinstance(o, _Dummy)

    def __len__(self):
        return 0

    def __getitem__(self, k):
        return _Dummy()

    def __or__(self, o):
        return self

    __ror__ = __or__


def _mk(name):
    m = types.ModuleType(name)
    m.__getattr__ = lambda attr: _Dummy()
    return m


for _n in ['sc2api_pb2', 'raw_pb2', 'spatial_pb2', 'ui_pb2', 'common_pb2', 'debug_pb2', 'error_pb2',
           'data_pb2', 'query_pb2', 'score_pb2']:
    _m = _mk(__name__ + '.' + _n)
    sys.modules[__name__ + '.' + _n] = _m
    globals()[_n] = _m
