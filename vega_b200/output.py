"""Result files in the reference's FITS layout (vega/output.py:13-350): one `MODEL_<name>` table
per correlation (model, masks, data, variance, coordinates; every parameter and the per-
correlation fit statistics as header cards), `BESTFIT` (names, values, errors, covariance;
FVAL / VALID / ACCURATE) and `SCAN` (one column per sampled parameter + fval, grid description
in the header).  Written with the package's own FITS writer (astropy is not in the image).

`write_pk` / `write_cf` add the reference's `PK_<name>` / `Xi_<name>` component tables
(vega/output.py:100-110,351-440) from `Model.save_components`.  `write_monte_carlo` writes the
`Bestfit` / `FitInfo` / `Mocks` tables of a Monte Carlo run (vega/output.py:442-520).  Not written:
the HDF variant (h5py is not in the image).
"""
import os
from pathlib import Path

import numpy as np

from . import fitsio


def _pad(array, size, value=np.nan):
    array = np.asarray(array)
    return np.pad(array, (0, size - len(array)), constant_values=value)


class Output:
    def __init__(self, config, data, corr_items, analysis=None):
        self.data = data
        self.analysis = analysis
        self.corr_items = corr_items
        self.type = config.get('type', 'fits')
        self.overwrite = config.getboolean('overwrite', False) if hasattr(config, 'getboolean') \
            else bool(config.get('overwrite', False))
        self.outfile = os.path.expandvars(config['filename'])
        self.output_cf = config.getboolean('write_cf', False) if hasattr(config, 'getboolean') else False
        self.output_pk = config.getboolean('write_pk', False) if hasattr(config, 'getboolean') else False
        self.mc_output = config.get('mc_output', None)

    # vega/output.py:37-123
    def write_results(self, corr_funcs, params, minimizer=None, bestfit_corr_stats=None,
                      scan_results=None, models=None):
        if self.type != 'fits':
            raise NotImplementedError('Only type = fits is written by this package.')
        if self.data is None:
            raise ValueError('Output object was initialized with an invalid data object.'
                             ' Reinitialize with a valid vega.data object.')
        tables = self._model_tables(corr_funcs, params, bestfit_corr_stats)
        if minimizer is not None:
            tables.append(self._bestfit_table(minimizer))
        if self.output_pk or self.output_cf:        # vega/output.py:100-110
            assert models is not None
            for model in models.values():
                model.save_components(params)
            if self.output_pk:
                tables += [('PK_' + key, {}, self._get_components(model.pk)) for key, model in models.items()]
            if self.output_cf:
                for key, model in models.items():
                    columns = {**self._get_components(model.xi, 'raw_'),
                               **self._get_components(model.xi_distorted, 'distorted_')}
                    size = max(v.shape[0] for v in columns.values())     # model grid > distorted grid: NaN tail
                    tables.append(('Xi_' + key, {}, {k: _pad(v, size) for k, v in columns.items()}))
        if scan_results is not None:
            assert minimizer is not None
            tables.append(self._scan_table(scan_results))
        if self.outfile[-5:] != '.fits':
            self.outfile += '.fits'
        path = Path(self.outfile)
        if path.exists() and not self.overwrite:
            raise OSError(f"File {path} already exists. Set overwrite = True in [output] to replace it.")
        fitsio.write_tables(path, tables)
        return path

    # vega/output.py:401-440
    @staticmethod
    def _get_components(model_components, name_prefix=''):
        """Columns `<prefix><part>_core` (and `<prefix><part>_<m1>_<m2>` for saved metal terms) of
        the peak / smooth parts that hold something."""
        columns = {}
        for part, data in model_components.items():
            if not data:
                continue
            for key, item in data.items():
                name = name_prefix + part + ('_core' if key == 'core' else f'_{key[0]}_{key[1]}')
                columns[name] = np.asarray(item, dtype=float)
        return columns

    # vega/output.py:144-235
    def _model_tables(self, corr_funcs, params, bestfit_corr_stats=None):
        tables = []
        for name, cf in corr_funcs.items():
            data, item = self.data[name], self.corr_items[name]
            n = len(cf)
            if len(data.data_vec) > n:
                raise ValueError('Data coordinate grid is larger than the model grid.')
            columns = {
                name + '_MODEL': _pad(cf, n),
                name + '_MODEL_MASK': _pad(data.model_mask, n, False),
                name + '_MASK': _pad(data.data_mask, n, False),
                name + '_DATA': _pad(data.data_vec, n),
                name + '_VAR': _pad(data.variance, n),
                name + '_RP': _pad(item.dist_model_coordinates.rp_grid, n),
                name + '_RT': _pad(item.dist_model_coordinates.rt_grid, n),
            }
            z = np.broadcast_to(np.asarray(item.model_coordinates.z_grid, dtype=float),
                                item.model_coordinates.rp_grid.shape)
            columns[name + '_Z'] = np.zeros(n) if n < z.size else _pad(z, n)
            if data.nb is not None:
                columns[name + '_NB'] = _pad(np.asarray(data.nb, dtype=np.int64), n, 0)
            header = {}
            for par, val in params.items():
                header['hierarch ' + par] = val
            if bestfit_corr_stats is not None:
                for par, val in bestfit_corr_stats[name].items():
                    if par == 'bestfit_marg_coeff':
                        for i, v in enumerate(val if val is not None else []):
                            header[f'hierarch marg_coeff_{i}'] = float(v)
                    else:
                        header['hierarch ' + par] = val
            tables.append(('MODEL_' + name, header, columns))
        return tables

    # vega/output.py:237-289
    def _bestfit_table(self, minimizer):
        names = np.array(list(minimizer.values.keys()))
        columns = {
            'names': names,
            'values': np.array([minimizer.values[n] for n in names], dtype=float),
            'errors': np.array([minimizer.errors[n] for n in names], dtype=float),
            'covariance': np.array(minimizer.covariance, dtype=float).reshape(len(names), len(names)),
        }
        valid = bool(minimizer.fmin.is_valid)
        header = {'FVAL': float(minimizer.fmin.fval), 'VALID': valid,
                  'ACCURATE': bool(getattr(minimizer.fmin, 'has_accurate_covar', valid))}
        return ('BESTFIT', header, columns)

    # vega/output.py:291-349
    def _scan_table(self, scan_results):
        names = list(scan_results[0].keys())
        results = np.array([[res[par] for par in names] for res in scan_results], dtype=float)
        # columns of different lengths: astropy's BinTableHDU.from_columns (vega/output.py:326) makes
        # the table as long as the longest and fills the rest with blanks / zeros -- same here
        rows = max(len(names), results.shape[0])
        columns = {'names': np.array(names + [''] * (rows - len(names)))}
        results = np.vstack([results, np.zeros((rows - results.shape[0], results.shape[1]))])
        for col, name in zip(results.T, names):
            columns[name] = col
        header = {f'hierarch name_{i}': n for i, n in enumerate(names)}
        if self.analysis is not None and getattr(self.analysis, 'grids', None):
            for par, grid in self.analysis.grids.items():
                header[f'hierarch {par}_min'] = float(grid[0])
                header[f'hierarch {par}_max'] = float(grid[-1])
                header[f'hierarch {par}_num_bins'] = int(len(grid))
        return ('SCAN', header, columns)

    # vega/output.py:442-520
    def write_monte_carlo(self, cpu_id=None):
        """Best fits and mock data vectors of Analysis.run_monte_carlo as a FITS file:
        `Bestfit` (one row per parameter: names, values[n_mocks], errors[n_mocks],
        covariance[n_mocks * P]), `FitInfo` (one row per mock: chisq, valid_minima, valid_hesse,
        failed_mask) and `Mocks` (one row per mock, one vector column per correlation, or `global`).
        File: <output dir>/monte_carlo/monte_carlo[_<cpu_id>].fits, or under [output] mc_output."""
        assert self.analysis is not None, 'Output.write_monte_carlo requires an Analysis object'
        assert self.analysis.has_monte_carlo, (
            'No Monte Carlo results found. Run Analysis.run_monte_carlo() first.')
        tables = []
        bestfits = self.analysis.mc_bestfits
        if not bestfits:
            print('No MC bestfit data to write.')
        else:
            names = np.array(list(bestfits.keys()))
            values = np.array([bestfits[name][:, 0] for name in names])
            errors = np.array([bestfits[name][:, 1] for name in names])
            covariances = np.array(self.analysis.mc_covariances)
            covariances = covariances.reshape(values.shape[1] * len(names), len(names)).T
            tables.append(('Bestfit', {}, {'names': names, 'values': values, 'errors': errors,
                                           'covariance': np.ascontiguousarray(covariances)}))
            tables.append(('FitInfo', {}, {
                'chisq': np.asarray(self.analysis.mc_chisq, dtype=float),
                'valid_minima': np.asarray(self.analysis.mc_valid_minima, dtype=bool),
                'valid_hesse': np.asarray(self.analysis.mc_valid_hesse, dtype=bool),
                'failed_mask': np.asarray(self.analysis.mc_failed_mask, dtype=bool)}))
        mocks = self.analysis.mc_mocks
        tables.append(('Mocks', {}, {name: np.array(mocks[name], dtype=float) for name in mocks}))
        dir_path = Path(self.outfile).parent / 'monte_carlo' if self.mc_output is None else Path(self.mc_output)
        dir_path.mkdir(parents=True, exist_ok=True)
        filepath = dir_path / ('monte_carlo.fits' if cpu_id is None else f'monte_carlo_{cpu_id}.fits')
        if filepath.exists() and not self.overwrite:
            raise OSError(f'File {filepath} already exists. Set overwrite = True in [output] to replace it.')
        fitsio.write_tables(filepath, tables)
        return filepath
