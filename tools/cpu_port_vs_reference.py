"""CPU timing of the oracle (the numpy port bench.py times as `cpu_baseline`, kind "port") next to
the UNMODIFIED reference, in the build container where both run (the reference needs the I/O shims
of tests/golden/shims and cannot travel to the GPU box).  One process, one BLAS thread, the
reference's own configurations: C1 (tests/configs/main.ini) and its 4-correlation configuration
(tests/full_configs/main.ini); warm caches on both sides (first call excluded), parameter values
changed on every call so that neither side answers from a cache of the previous call.
Output: profiles/r02_port_vs_reference_cpu.json."""
import json
import os
import sys
import time
from pathlib import Path

os.environ.setdefault('OMP_NUM_THREADS', '1')
REPO = Path(__file__).resolve().parents[1]
REF = Path('/root/reference')
sys.path[:0] = [str(REPO / 'tests' / 'golden' / 'shims'), str(REF), str(REPO), str(REPO / 'tests')]
import numpy as np  # noqa: E402


def time_calls(fn, n):
    fn(0)
    t0 = time.perf_counter()
    for i in range(1, n + 1):
        fn(i)
    return (time.perf_counter() - t0) / n


def main():
    out = {'threads': 1, 'note': 's per chi2 call, mean over n calls after one warm-up call'}
    cwd = os.getcwd()
    for key, ref_ini, own_ini, n in (('c1', 'tests/configs/main.ini', 'configs/configs/main.ini', 6),
                                     ('four_correlations', 'tests/full_configs/main.ini',
                                      'configs/full_configs/main.ini', 3)):
        pars = [{'bias_eta_LYA': -0.2 - 0.001 * i, 'beta_LYA': 1.6 + 0.01 * i} for i in range(n + 1)]
        os.chdir(REF)
        from vega import VegaInterface as RefVega
        ref = RefVega(ref_ini)
        t_ref = time_calls(lambda i: ref.chi2(pars[i]), n)
        values_ref = [ref.chi2(p) for p in pars[:2]]
        os.chdir(cwd)
        from helpers import make_oracle
        vega, oracle = make_oracle(own_ini)
        t_port = time_calls(lambda i: oracle.chi2(pars[i]), n)
        values_port = [oracle.chi2(p) for p in pars[:2]]
        out[key] = {'reference_s_per_chi2': t_ref, 'port_s_per_chi2': t_port, 'port_over_reference': t_port / t_ref,
                    'calls': n, 'max_abs_dchi2': float(np.max(np.abs(np.array(values_ref) - np.array(values_port))))}
        print(key, out[key], flush=True)
    (REPO / 'profiles' / 'r02_port_vs_reference_cpu.json').write_text(json.dumps(out, indent=1))


if __name__ == '__main__':
    main()
