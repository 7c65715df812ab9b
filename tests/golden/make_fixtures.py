"""Stage the INPUT fixtures the parity tests need (run HERE only, once).

/root/reference does not exist on the GPU box, so the data tables its own tests
read (tests/data/*.fits*, the P(k) template, the Fvoigt table, the DESI
instrumental-systematics table) are staged into this repo as data fixtures, and
its benchmark ini files are re-emitted (comments stripped) under tests/configs/.
No reference source code is copied; these are numeric tables and option files.

    python tests/golden/make_fixtures.py
"""
import configparser
import gzip
import shutil
from pathlib import Path

REF = Path('/root/reference')
REPO = Path(__file__).resolve().parents[2]

DATA_FILES = [
    'dr16_simple_auto.fits', 'dr16_simple_cross.fits', 'picca_bench_data.fits',
    'cf_lya-exp.fits.gz', 'cf_lyb-exp.fits.gz', 'xcf_lya-exp.fits.gz', 'xcf_lyb-exp.fits.gz',
    'metal_dmat_lya.fits.gz', 'metal_dmat_lyb.fits.gz',
    'metal_xdmat_lya.fits.gz', 'metal_xdmat_lyb.fits.gz',
]

MODEL_FILES = [
    'PlanckDR16/PlanckDR16.fits',
    'PlanckDR16/PlanckDR16_z_2.334.fits',
    'Planck18/DESI-2024_z_2.33.fits',
    'instrumental_systematics/desi-instrument-syst-for-forest-auto-correlation.csv',
]


def rewrite_ini(src, dst, path_map=None):
    """Re-emit an ini file through configparser (drops comments), remapping paths."""
    cfg = configparser.ConfigParser()
    cfg.optionxform = lambda option: option
    cfg.read(src)
    for section in cfg.sections():
        for key, value in cfg.items(section):
            if path_map is not None:
                for old, new in path_map.items():
                    value = value.replace(old, new)
            cfg[section][key] = value
    dst.parent.mkdir(parents=True, exist_ok=True)
    with open(dst, 'w') as f:
        f.write(f'# re-emitted from the reference\'s {src.relative_to(REF)} (options only)\n')
        cfg.write(f)


def main():
    data_dir = REPO / 'tests' / 'data'
    data_dir.mkdir(parents=True, exist_ok=True)
    for name in DATA_FILES:
        shutil.copyfile(REF / 'tests' / 'data' / name, data_dir / name)

    models = REPO / 'vega_b200' / 'models'
    for name in MODEL_FILES:
        (models / name).parent.mkdir(parents=True, exist_ok=True)
        shutil.copyfile(REF / 'vega' / 'models' / name, models / name)
    fv = models / 'fvoigt_models'
    fv.mkdir(parents=True, exist_ok=True)
    with open(REF / 'vega' / 'models' / 'fvoigt_models' / 'Fvoigt_exp.txt', 'rb') as fin, \
            gzip.open(fv / 'Fvoigt_exp.txt.gz', 'wb') as fout:
        shutil.copyfileobj(fin, fout)
    shutil.copyfile(REF / 'vega' / 'parameters' / 'default_values.txt',
                    REPO / 'vega_b200' / 'parameters' / 'default_values.txt')

    cfg_dir = REPO / 'tests' / 'configs'
    # reference tests/configs and tests/full_configs
    for sub in ('configs', 'full_configs'):
        for src in sorted((REF / 'tests' / sub).glob('*.ini')):
            rewrite_ini(src, cfg_dir / sub / src.name)
    # picca benchmark configs (the golden-vector path)
    bench = REF / 'examples' / 'picca_benchmarks' / 'configs' / 'vega'
    for src in sorted(bench.glob('*.ini')):
        rewrite_ini(src, cfg_dir / 'picca_bench' / src.name,
                    {'examples/picca_benchmarks/configs/vega/': 'configs/picca_bench/'})


if __name__ == '__main__':
    main()
