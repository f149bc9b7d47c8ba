"""No-op SummaryWriter so the reference's log_helper imports. Test infrastructure only."""


class SummaryWriter:
    def __init__(self, *a, **k):
        pass

    def __getattr__(self, name):
        return lambda *a, **k: None
