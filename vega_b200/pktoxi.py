"""`PktoXi` with the reference's call shapes (vega/pktoxi.py:10-163) on the device.

    PktoXi(k_grid, muk_grid, dmuk, config)            constructor            :17-59
    PktoXi.init_from_Pk(pk_obj, config)               from a PowerSpectrum   :61-77
    compute_pk_ells(pk) -> [n_ell, nk]                mu projection          :79-97
    compute(r_grid, mu_grid, pk, single_ell=-1)       P(k,mu) -> xi(r,mu)    :99-163

`pk` is a caller-supplied P(k, mu) array [n_mu, n_k]; the projection, the Hankel + spline
operator GEMM and the read-out run in the CUDA library (vg_pktoxi_batch).  An instance either
shares the transform of a VegaInterface plan (Model.PktoXi) or, built from its constructor,
owns a small plan of its own -- the stand-alone use of bin/make_template-style scripts.
"""
import ctypes as C

import numpy as np

from . import _lib, hankel, utils


def _legendre(ell, x):
    return np.polynomial.legendre.legval(x, [0] * ell + [1])


class PktoXi:
    cache = {}            # the reference's class attribute (vega/pktoxi.py:15); unused here

    def __init__(self, k_grid, muk_grid, dmuk, config, _shared=None):
        self.k_grid = np.ascontiguousarray(k_grid, dtype=np.float64).ravel()
        self.muk_grid = np.asarray(muk_grid, dtype=np.float64)
        self.dmuk = float(dmuk)
        self._config = config
        get = config.get if config is not None else (lambda key, default=None: default)
        getb = config.getboolean if config is not None else (lambda key, default=False: default)
        geti = config.getint if config is not None else (lambda key, default=None: default)
        self._old_fftlog = getb('old_fftlog', False)
        self._extrap = getb('fht_extrap', False)
        if self._extrap:
            raise NotImplementedError('fht_extrap = True is not built.')
        self.ell_max = geti('ell_max', 6)
        self.ells = tuple(int(x) for x in np.arange(0, self.ell_max + 1, 2))
        self.cache_pars = None
        self._lib = _lib.load()
        if _shared is not None:                  # the transform of an existing plan
            self._handle, self._hankel_id, self._owner = _shared
            self._own = False
            return
        mu = self.muk_grid.ravel()
        wleg = np.ascontiguousarray(np.stack([self.dmuk * (2 * ell + 1) * _legendre(ell, mu) for ell in self.ells]))
        ones = np.ones(self.k_grid.size)
        base, cols = np.zeros(1), np.zeros(1, dtype=np.int32)
        handle = C.c_void_p()
        ptr = lambda a: a.ctypes.data_as(C.c_void_p)    # noqa: E731
        mu_c = np.ascontiguousarray(mu)
        _lib.check(self._lib.vg_plan_create(1, ptr(base), 0, ptr(cols), self.k_grid.size, ptr(self.k_grid), ptr(ones),
                                            ptr(ones), ptr(ones), mu_c.size, ptr(mu_c), len(self.ells), ptr(wleg),
                                            -1, C.byref(handle)))
        xk, op = hankel.build_operator_set(self.k_grid, self.ells, self._old_fftlog, getb('fht_lowring', True))
        ells_arr = np.ascontiguousarray(self.ells, dtype=np.int32)
        hid = C.c_int32()
        _lib.check(self._lib.vg_plan_add_hankel(handle, len(self.ells), ptr(ells_arr), xk.shape[1], ptr(xk), ptr(op),
                                                0 if self._old_fftlog else 1, C.byref(hid)))
        self._handle, self._hankel_id, self._owner, self._own = handle, hid.value, None, True

    def __del__(self):
        try:
            if getattr(self, '_own', False) and self._handle:
                self._lib.vg_plan_destroy(self._handle)
                self._handle = None
        except Exception:
            pass

    @classmethod
    def init_from_Pk(cls, pk_obj, config):
        """vega/pktoxi.py:61-77."""
        return cls(pk_obj.k_grid, pk_obj.muk_grid, pk_obj.dmuk, config)

    # -- device calls ---------------------------------------------------------------------
    def _run(self, pk, r=None, mu=None, single_ell=-1, want_ells=False):
        import torch
        pk = np.ascontiguousarray(pk, dtype=np.float64)
        nmu, nk = self.muk_grid.size, self.k_grid.size
        if pk.shape != (nmu, nk):
            pk = np.ascontiguousarray(np.broadcast_to(pk, (nmu, nk)))
        dev = torch.device('cuda', torch.cuda.current_device())
        grid = torch.from_numpy(pk[None]).to(dev)
        status = torch.zeros(1, dtype=torch.int32, device=dev)
        nk_pad = (nk + 15) // 16 * 16
        ells = torch.zeros(1, len(self.ells), nk_pad, dtype=torch.float64, device=dev) if want_ells else None
        n = 0 if r is None else int(np.size(r))
        xi = torch.zeros(1, max(n, 1), dtype=torch.float64, device=dev)
        r_c = np.ascontiguousarray(r, dtype=np.float64).ravel() if n else np.zeros(1)
        mu_c = np.ascontiguousarray(mu, dtype=np.float64).ravel() if n else np.zeros(1)
        _lib.check(self._lib.vg_pktoxi_batch(
            self._handle, self._hankel_id, C.c_void_p(grid.data_ptr()), 1, r_c.ctypes.data_as(C.c_void_p),
            mu_c.ctypes.data_as(C.c_void_p), n, int(single_ell), C.c_void_p(xi.data_ptr()),
            C.c_void_p(ells.data_ptr()) if want_ells else None, C.c_void_p(status.data_ptr()),
            C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        if int(status.cpu()[0]) & _lib.STATUS_BOUNDS:
            raise utils.VegaBoundsError
        return (xi.cpu().numpy()[0, :n] if n else None), (ells.cpu().numpy()[0, :, :nk] if want_ells else None)

    def compute_pk_ells(self, pk):
        """Multipoles of P(k, mu): [n_ell, nk] (vega/pktoxi.py:79-97)."""
        return self._run(pk, want_ells=True)[1]

    def compute(self, r_grid, mu_grid, pk, single_ell=-1):
        """xi(r, mu) of a P(k, mu) array (vega/pktoxi.py:99-163); `single_ell` >= 0 returns that
        multipole alone.  r outside the transform's range raises VegaBoundsError."""
        if single_ell >= 0 and single_ell not in self.ells:
            raise ValueError(f'single_ell = {single_ell} is not one of the multipoles {self.ells}.')
        return self._run(pk, r_grid, mu_grid, single_ell)[0]
