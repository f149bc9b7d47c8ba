"""Minimal EasyDict stand-in (attribute-access dict, recursive) so the reference's
config_helper imports. Test infrastructure only."""


class EasyDict(dict):
    def __init__(self, d=None, **kwargs):
        super().__init__()
        if d is None:
            d = {}
        if kwargs:
            d = dict(d, **kwargs)
        for k, v in d.items():
            setattr(self, k, v)

    @classmethod
    def _wrap(cls, v):
        if isinstance(v, dict) and not isinstance(v, EasyDict):
            return cls(v)
        if isinstance(v, (list, tuple)):
            return type(v)(cls._wrap(x) for x in v)
        return v

    def __setattr__(self, k, v):
        v = self._wrap(v)
        super().__setitem__(k, v)
        super().__setattr__(k, v)

    __setitem__ = __setattr__

    def update(self, e=None, **f):
        d = dict(e or {})
        d.update(f)
        for k, v in d.items():
            setattr(self, k, v)

    def pop(self, k, *args):
        if hasattr(self, k) and k in self.__dict__:
            super().__delattr__(k)
        return super().pop(k, *args)

    def __deepcopy__(self, memo):
        import copy
        return EasyDict({k: copy.deepcopy(v, memo) for k, v in self.items()})
