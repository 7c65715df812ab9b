"""Names only; new_metals / cosmo-dependent paths are out of scope."""
ABSORBER_IGM = {}


class Cosmo:
    def __init__(self, *a, **k):
        raise NotImplementedError('picca.Cosmo is not available in this image')
