"""picca.constants stand-in for the golden generator: the restated cosmology and rest
wavelengths of vega_b200/cosmo.py (picca itself is not in the image; see that module)."""
from vega_b200.cosmo import ABSORBER_IGM, SPEED_LIGHT, Cosmo  # noqa: F401
