"""vega_b200: B200-native model-evaluation and likelihood path of Vega."""
from .vega_interface import VegaInterface  # noqa: F401

__version__ = '0.1.0'
