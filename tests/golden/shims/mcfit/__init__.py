"""mcfit stand-in: the restated P2xi of oracle/hankel.py (mcfit is not installed)."""
from oracle.hankel import P2xi  # noqa: F401
