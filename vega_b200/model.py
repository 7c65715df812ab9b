"""`Model` call shapes of the reference (vega/model.py:8-208) on top of the device engine.

`VegaInterface.models[name]` exposes, per correlation, what the reference's scripts reach for:
`compute(pars, pk_full, pk_smooth)`, `compute_direct(pars, pk_full)` and the per-stage
components (`Pk_core.compute`, `PktoXi.compute_pk_ells`, `Xi_core` outputs) through the
library's stage taps.  Everything is evaluated on the device for a batch of one; `pars` may be a
complete parameter dictionary (as `_get_lcl_prms` builds) or only the entries to change.
"""
import numpy as np

from . import utils


def _shared_pktoxi(vega, name, pk_obj):
    from .pktoxi import PktoXi
    engine = vega._engine_for()
    ci = list(vega.corr_items).index(name)
    cfg = vega.corr_items[name].config['model']
    return PktoXi(pk_obj.k_grid, pk_obj.muk_grid, pk_obj.dmuk, cfg,
                  _shared=(engine.handle, engine.infos[ci]['hankel_id'], engine))


class _Metals:
    """The per-pair objects of the reference's Metals (vega/metals.py:95-142): Pk_metal, PktoXi,
    Xi_metal, keyed by the metal correlation."""

    def __init__(self, vega, name):
        from .correlation_func import CorrelationFunction
        from .power_spectrum import PowerSpectrum
        item, data = vega.corr_items[name], vega.data[name]
        self.Pk_metal, self.Xi_metal, self._vega, self._name = {}, {}, vega, name
        for slot, pair in enumerate(item.metal_correlations):
            coords = data.get_metal(pair)[0] if data is not None else None
            self.Pk_metal[pair] = PowerSpectrum(vega, name, metal_slot=slot)
            self.Xi_metal[pair] = CorrelationFunction(vega, name, metal_slot=slot, coordinates=coords)
        self._pktoxi = None

    @property
    def PktoXi(self):
        if self._pktoxi is None:
            any_pk = next(iter(self.Pk_metal.values()))
            shared = _shared_pktoxi(self._vega, self._name, any_pk)
            self._pktoxi = {pair: shared for pair in self.Pk_metal}
        return self._pktoxi


class Model:
    def __init__(self, vega, name):
        self._vega = vega
        self._name = name
        self._index = list(vega.corr_items).index(name)
        self._corr_item = vega.corr_items[name]
        # the per-stage objects of the reference (vega/model.py:60-75), each with its call shape
        from .correlation_func import CorrelationFunction
        from .pktoxi import PktoXi
        from .power_spectrum import PowerSpectrum
        self.Pk_core = PowerSpectrum(vega, name)
        self.Xi_core = CorrelationFunction(vega, name)
        self._pktoxi = None
        self.metals = _Metals(vega, name) if self._corr_item.has_metals else None
        # the save-components dictionaries (vega/model.py:41-45), filled by save_components()
        self.pk = {'peak': {}, 'smooth': {}, 'full': {}}
        self.xi = {'peak': {}, 'smooth': {}, 'full': {}}
        self.xi_distorted = {'peak': {}, 'smooth': {}, 'full': {}}

    # -- helpers ------------------------------------------------------------------------------
    def _changed(self, pars):
        """Entries of `pars` that differ from the interface's [parameters] values."""
        pars = pars or {}
        base = self._vega.params
        return {k: v for k, v in pars.items() if k in base and base[k] != v}

    @property
    def PktoXi(self):
        """The transform object of this correlation (vega/model.py:63): shares the Hankel + spline
        operator of the device plan."""
        if self._pktoxi is None:
            self._pktoxi = _shared_pktoxi(self._vega, self._name, self.Pk_core)
        return self._pktoxi

    # -- vega/model.py:157-208 ------------------------------------------------------------------
    def compute(self, pars, pk_full, pk_smooth):
        """bao_amp * peak + smooth (+ metals, systematics, broadband, distortion) of this
        correlation (vega/model.py:157-187)."""
        with self._vega.templates(pk_full, pk_smooth):
            return self._vega.compute_model(self._changed(pars), run_init=False)[self._name]

    def compute_direct(self, pars, pk_full):
        """The model directly from a full P(k) (vega/model.py:189-208)."""
        return self._vega.compute_model(self._changed(pars), run_init=False, direct_pk=pk_full)[self._name]

    # -- per-stage components (the save-components dictionaries, vega/model.py:113-116) -----------
    def _stage(self, pars, stage, component):
        changed = self._changed(pars)
        engine = self._vega._engine_for(tuple(changed))
        theta = self._vega._theta_row(engine, changed)
        eval_index = {'peak': 0, 'smooth': 1}[component] if isinstance(component, str) else int(component)
        out, status = engine.stage(self._index, stage, theta, eval_index=eval_index)
        from . import _lib
        if status[0] & _lib.STATUS_ARINYO:
            raise utils.VegaArinyoError
        if status[0] & _lib.STATUS_BOUNDS:
            raise utils.VegaBoundsError
        return out[0], engine

    def pk_grid(self, pars=None, component='smooth'):
        """P(k, mu) of one component on the (mu, k) grid -- PowerSpectrum.compute
        (vega/power_spectrum.py:87-196) with that component's linear P(k)."""
        out, engine = self._stage(pars, 0, component)
        nk = self._vega.fiducial['k'].size
        return out.reshape(-1, nk)

    def pk_ells(self, pars=None, component='smooth'):
        """Multipoles P_ell(k), [n_ell, nk] -- PktoXi.compute_pk_ells (vega/pktoxi.py:79-97)."""
        out, engine = self._stage(pars, 1, component)
        nk = self._vega.fiducial['k'].size
        nk_pad = (nk + 15) // 16 * 16
        n_ell = len(engine.infos[self._index]['ells'])
        eval_index = {'peak': 0, 'smooth': 1}[component] if isinstance(component, str) else int(component)
        return out.reshape(-1, n_ell, nk_pad)[eval_index][:, :nk]

    def xi_of(self, pars=None, component='smooth'):
        """xi of the peak or smooth component on the model grid, before metals / broadband /
        distortion -- CorrelationFunction.compute (vega/correlation_func.py:117-161)."""
        out, engine = self._stage(pars, 3, component)
        n_model = self._corr_item.model_coordinates.rp_grid.size
        ld = (n_model + 15) // 16 * 16
        return out.reshape(-1, ld)[{'peak': 0, 'smooth': 1}[component]][:n_model]

    def save_components(self, pars=None):
        """Fill `pk`, `xi` and `xi_distorted` as the reference does with `save-components` on
        (vega/model.py:113-116,151-153; write_pk / write_cf of vega/output.py:100-110).  With the
        default `no-metal-decomp = True` the reference stores the `core` entry of the peak and
        the smooth component only (the metal dictionaries are merged in the other branch,
        :125-131): P(k, mu) and xi before metals, and the component after metals, broadband and
        distortion.  The distorted components come from one batch of two vectors -- the model is
        bao_amp * peak + smooth with every later stage linear (:186), so they are M(bao_amp = 0)
        and M(1) - M(0)."""
        changed = self._changed(pars)
        changed.pop('bao_amp', None)
        names = list(changed) + ['bao_amp']
        theta = np.array([[changed[n] for n in names[:-1]] + [a] for a in (0.0, 1.0)], dtype=np.float64)
        models, status = self._vega.compute_model_batch(theta, names)
        from . import _lib
        if np.any(status & _lib.STATUS_ARINYO):
            raise utils.VegaArinyoError
        if np.any(status & _lib.STATUS_BOUNDS):
            raise utils.VegaBoundsError
        m = models[self._name]
        self.xi_distorted['smooth']['core'] = m[0].copy()
        self.xi_distorted['peak']['core'] = m[1] - m[0]
        for component in ('peak', 'smooth'):
            self.pk[component]['core'] = self.pk_grid(changed, component).copy()
            self.xi[component]['core'] = self.xi_of(changed, component).copy()
        return self
