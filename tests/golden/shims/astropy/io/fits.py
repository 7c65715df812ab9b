"""astropy.io.fits surface the reference touches, backed by vega_b200.fitsio."""
from vega_b200.fitsio import open, HDUList, HDU  # noqa: F401,A004


class _Unsupported:
    def __init__(self, *a, **k):
        raise NotImplementedError('FITS writing through the shim is not supported')


PrimaryHDU = Column = BinTableHDU = Header = _Unsupported
