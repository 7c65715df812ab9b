class MCSamples:
    pass
def loadMCSamples(*a, **k):
    raise NotImplementedError
