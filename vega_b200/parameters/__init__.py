"""Default prior boxes and minimiser step sizes (parameters/default_values.txt, the
table the reference reads at vega/parameters/param_utils.py)."""
from pathlib import Path


def get_default_values():
    table = {}
    with open(Path(__file__).resolve().parent / 'default_values.txt') as f:
        for line in f:
            line = line.strip()
            if not line or line.startswith('#'):
                continue
            words = line.split()
            table[words[0]] = {'limits': (float(words[1]), float(words[2])),
                               'error': float(words[3])}
    return table
