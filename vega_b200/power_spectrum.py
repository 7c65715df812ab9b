"""`PowerSpectrum.compute` with the reference's call shape (vega/power_spectrum.py:87-196):

    pk = PowerSpectrum.compute(pk_lin, params, fast_metals=False)   -> ndarray [n_mu, n_k]

`pk_lin` is any linear P(k) on the template's k grid, `params` a parameter dictionary that
carries the 'peak' flag (vega/model.py:176,179 set it before the call).  An instance stands for
one PowerSpectrum object of the reference -- the core one of a correlation (Model.Pk_core) or the
one of a metal correlation (Model.metals.Pk_metal[pair]) -- and evaluates on the device through
vg_pk_batch: every factor of the fused kernel, materialised on the (mu, k) grid.
"""
import ctypes as C

import numpy as np

from . import _lib, utils


class PowerSpectrum:
    def __init__(self, vega, corr_name, metal_slot=None):
        self._vega = vega
        self._name = corr_name
        self._index = list(vega.corr_items).index(corr_name)
        self._metal_slot = metal_slot
        fid = vega.fiducial
        self.k_grid = fid['k']
        item = vega.corr_items[corr_name]
        section = item.config['model'] if metal_slot is None else item.config['metals']
        nmu = section.getint('num_bins_muk', 1000)
        self.muk_grid = ((np.arange(nmu) + 0.5) / nmu)[:, None]       # vega/power_spectrum.py:76-78
        self.dmuk = 1 / nmu
        self.k_par_grid = self.k_grid * self.muk_grid
        self.k_trans_grid = self.k_grid * np.sqrt(1 - self.muk_grid**2)

    def _metal_eval(self, params):
        """Evaluation of the plan this metal object stands for: one per pair, or -- with
        no-metal-decomp = False (vega/model.py:119-131) -- the peak / smooth one of the pair."""
        item = self._vega.corr_items[self._name]
        if item.config['model'].getboolean('no-metal-decomp', True):
            return 2 + self._metal_slot
        return 2 + 2 * self._metal_slot + (0 if params.get('peak') else 1)

    def _eval_index(self, params):
        if self._metal_slot is not None:
            return self._metal_eval(params)
        if 'peak' not in params:
            raise KeyError("params['peak'] selects the component (vega/model.py:176,179).")
        return 0 if params['peak'] else 1

    def compute(self, pk_lin, params, fast_metals=False):
        import torch
        vega = self._vega
        changed = {k: v for k, v in params.items() if k != 'peak' and not (k in vega.params and vega.params[k] == v)}
        engine = vega._engine_for(tuple(changed), changed)
        theta = engine.theta_tensor(vega._theta_row(engine, changed))
        nmu, nk = self.muk_grid.size, self.k_grid.size
        out = torch.empty(1, nmu, nk, dtype=torch.float64, device=engine.device)
        status = torch.zeros(1, dtype=torch.int32, device=engine.device)
        pk_lin = np.ascontiguousarray(pk_lin, dtype=np.float64)
        if pk_lin.size != nk:
            raise ValueError('pk_lin must be sampled on the k grid of the fiducial template.')
        _lib.check(engine.lib.vg_pk_batch(engine.handle, self._index, self._eval_index(params),
                                          pk_lin.ctypes.data_as(C.c_void_p), int(bool(fast_metals)),
                                          C.c_void_p(theta.data_ptr()), 1, C.c_void_p(out.data_ptr()),
                                          C.c_void_p(status.data_ptr()), engine._stream()))
        if (int(status.cpu()[0]) | engine.static_status) & _lib.STATUS_ARINYO:
            raise utils.VegaArinyoError
        return out.cpu().numpy()[0]
