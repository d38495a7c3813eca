"""Monte-Carlo fantasy-tree posterior engine (host driver) - placeholder until the kernels land."""


class MCTree:
    def __init__(self, *a, **k):
        raise NotImplementedError("HES-MC kernels: in progress")


def baseline_forward(*a, **k):
    raise NotImplementedError("baseline kernels: in progress")
