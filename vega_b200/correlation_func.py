"""`CorrelationFunction.compute` with the reference's call shape (vega/correlation_func.py:117-198):

    xi = CorrelationFunction.compute(pk, pk_lin, PktoXi_obj, params)   -> ndarray [n_model]

`pk` is a caller-supplied P(k, mu) array [n_mu, n_k] (what PowerSpectrum.compute returns),
`pk_lin` the isotropic P(k) the relativistic / asymmetry terms are built from, `params` a
parameter dictionary with the 'peak' flag.  The Alcock-Paczynski rescaling, the spline read-out of
the multipoles, bias evolution, growth and the additive terms run on the device
(vg_xi_from_grid_batch) on the coordinates of the object the instance stands for: the core
CorrelationFunction of a correlation (Model.Xi_core) or one of a metal correlation
(Model.metals.Xi_metal[pair]).
"""
import ctypes as C

import numpy as np

from . import _lib, utils


class CorrelationFunction:
    def __init__(self, vega, corr_name, metal_slot=None, coordinates=None):
        self._vega = vega
        self._name = corr_name
        self._index = list(vega.corr_items).index(corr_name)
        self._metal_slot = metal_slot
        item = vega.corr_items[corr_name]
        coords = item.model_coordinates if coordinates is None else coordinates
        self._r, self._mu, self._z = coords.r_grid, coords.mu_grid, coords.z_grid
        cfg = item.config['model'] if metal_slot is None else item.config['metals']
        self._needs_pk_lin = any(key in cfg and cfg.getboolean(key)
                                 for key in ('relativistic correction', 'standard asymmetry'))

    def _pk_lin_sel(self, pk_lin):
        """Which template the relativistic / asymmetry tables of the plan belong to."""
        fid = self._vega.fiducial
        pk_lin = np.asarray(pk_lin, dtype=float)
        for sel, ref in enumerate((fid['pk_full'] - fid['pk_smooth'], fid['pk_smooth'], fid['pk_full'])):
            if pk_lin.shape == ref.shape and np.array_equal(pk_lin, ref):
                return sel
        if self._needs_pk_lin:
            raise NotImplementedError('relativistic / asymmetry terms are tabulated for the peak, smooth and full '
                                      'templates of the plan; pass one of them as pk_lin.')
        return -1

    def _metal_eval(self, params):
        """Evaluation of the plan this metal object stands for: one per pair, or -- with
        no-metal-decomp = False (vega/model.py:119-131) -- the peak / smooth one of the pair."""
        item = self._vega.corr_items[self._name]
        if item.config['model'].getboolean('no-metal-decomp', True):
            return 2 + self._metal_slot
        return 2 + 2 * self._metal_slot + (0 if params.get('peak') else 1)

    def compute(self, pk, pk_lin, PktoXi_obj, params):
        import torch
        vega = self._vega
        if self._metal_slot is not None:
            eval_index = self._metal_eval(params)
        else:
            if 'peak' not in params:
                raise KeyError("params['peak'] selects the component (vega/model.py:176,179).")
            eval_index = 0 if params['peak'] else 1
        changed = {k: v for k, v in params.items() if k != 'peak' and not (k in vega.params and vega.params[k] == v)}
        engine = vega._engine_for(tuple(changed), changed)
        theta = engine.theta_tensor(vega._theta_row(engine, changed))
        nk = vega.fiducial['k'].size
        pk = np.ascontiguousarray(pk, dtype=np.float64)
        if pk.ndim != 2 or pk.shape[1] != nk:
            pk = np.ascontiguousarray(np.broadcast_to(pk, (PktoXi_obj.muk_grid.size, nk)))
        grid = torch.from_numpy(pk[None]).to(engine.device)
        n_model = self._r.size
        out = torch.empty(1, n_model, dtype=torch.float64, device=engine.device)
        status = torch.zeros(1, dtype=torch.int32, device=engine.device)
        _lib.check(engine.lib.vg_xi_from_grid_batch(
            engine.handle, self._index, eval_index, C.c_void_p(grid.data_ptr()), self._pk_lin_sel(pk_lin),
            C.c_void_p(theta.data_ptr()), 1, C.c_void_p(out.data_ptr()), C.c_void_p(status.data_ptr()),
            engine._stream()))
        if int(status.cpu()[0]) & _lib.STATUS_BOUNDS:
            raise utils.VegaBoundsError
        return out.cpu().numpy()[0]
