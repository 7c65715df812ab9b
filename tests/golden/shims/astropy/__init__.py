"""Import-time stand-in so the unmodified reference package can be imported here
(astropy is not installed).  Only used by tests/golden/make_golden.py."""
