"""One-off probe (SURVEY 8c iii): is the real `mcfit` importable on the GPU
box?  If it is, diff the restated P2xi (oracle/hankel.py, the algorithm both the oracle and
vega_b200/hankel.py follow) against it for ell = 0, 2, 4, 6 on the template's 814-point k grid and
write the maximum relative difference; if not, record that outcome.  Output:
gpurun_out/mcfit_probe.json (copied to profiles/ by hand)."""
import json
import sys
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(REPO))
out = {'probe': 'import mcfit'}
try:
    import mcfit
    out['mcfit'] = getattr(mcfit, '__version__', 'unknown version')
except Exception as exc:      # noqa: BLE001
    out['mcfit'] = None
    out['error'] = f'{type(exc).__name__}: {exc}'
if out['mcfit'] is not None:
    from oracle.hankel import P2xi
    from vega_b200 import VegaInterface
    vega = VegaInterface('configs/configs/main.ini')
    k, pk = vega.fiducial['k'], vega.fiducial['pk_full']
    out['diff'] = {}
    for lowring in (True, False):
        for ell in (0, 2, 4, 6):
            r_ref, xi_ref = mcfit.P2xi(k, l=ell, lowring=lowring)(pk, extrap=False)
            r_own, xi_own = P2xi(k, l=ell, lowring=lowring)(pk, extrap=False)
            out['diff'][f'lowring={lowring},ell={ell}'] = {
                'max_rel_r': float(np.max(np.abs(r_own / r_ref - 1))),
                'max_rel_xi_over_max': float(np.max(np.abs(xi_own - xi_ref)) / np.max(np.abs(xi_ref)))}
Path(REPO / 'gpurun_out').mkdir(exist_ok=True)
(REPO / 'gpurun_out' / 'mcfit_probe.json').write_text(json.dumps(out, indent=1))
print(json.dumps(out))
