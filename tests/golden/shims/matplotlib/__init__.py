from . import pyplot  # noqa: F401
rcParams = {}
def rc(*a, **k):
    pass
