rcParams = {}
def __getattr__(name):
    def _noop(*a, **k):
        raise NotImplementedError('matplotlib is not available in this image')
    return _noop
