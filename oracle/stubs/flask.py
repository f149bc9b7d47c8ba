"""Stub of `flask` for importing the reference learner in the build container (TEST INFRASTRUCTURE: contributes no
arithmetic; the reference uses Flask only for its debug / control routes)."""


class Flask:
    def __init__(self, *a, **k):
        self.routes = {}

    def route(self, rule, **k):
        def deco(fn):
            self.routes[rule] = fn
            return fn
        return deco

    def run(self, *a, **k):
        raise RuntimeError('flask stub: no server in the test container')


class _Request:
    json = {}
    args = {}


request = _Request()
