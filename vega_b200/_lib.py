"""ctypes binding of the CUDA library (include/vega_b200.h).

The library is the product: when it is missing, or there is no CUDA device, every entry
point raises VegaDeviceError -- there is no CPU fallback.
"""
import ctypes as C
from pathlib import Path

from .utils import VegaDeviceError

LIB_PATH = Path(__file__).resolve().parent / 'libvega_b200.so'

VG_MAX_BB_TERMS = 8
F_ARINYO, F_MCDONALD, F_GK, F_PEAKNL, F_FSGAUSS, F_FSEXP, F_VELO, F_KDAMP, F_MOCKGK = (1 << i for i in range(9))
STATUS_BOUNDS, STATUS_ARINYO, STATUS_NONFINITE, STATUS_DEVICE = 1, 2, 4, 8
OPT_MOMENTS, OPT_DIRECT_PK, OPT_FOLD_METALS, OPT_OZAKI, OPT_OZAKI_DM, OPT_VELO_TABLES, OPT_FUSE_DM_METALS, OPT_WHITEN = 0, 1, 2, 3, 4, 5, 6, 7
OPT_OZAKI_MIN_BATCH, OPT_LANES = 8, 9
ERR_CUDA, ERR_ARG, ERR_STATE = 1, 2, 3
STATIC_SCRATCH_DOUBLES = 1 << 21      # VG_STATIC_SCRATCH_DOUBLES of include/vega_b200.h

i32, u32, f64 = C.c_int32, C.c_uint32, C.c_double


class EvalDesc(C.Structure):
    """Mirror of vg_eval_desc (field order and types must match the header)."""
    _fields_ = [
        ('pk_lin_sel', i32), ('is_peak', i32), ('fast_metals', i32), ('same_tracer', i32),
        ('idx_bias', i32 * 2), ('idx_beta', i32 * 2), ('idx_bias_eta', i32 * 2),
        ('idx_growth_rate', i32),
        ('uv_on', i32 * 2), ('uvb', i32), ('heii', i32),
        ('idx_lambda_uv', i32), ('idx_bias_gamma', i32), ('idx_bias_prim', i32),
        ('idx_lambda_heii', i32), ('idx_bias_gamma_e', i32),
        ('hcd_model', i32), ('hcd_on', i32 * 2),
        ('idx_bias_hcd', i32), ('idx_beta_hcd', i32), ('idx_L0', i32),
        ('hcd_grid', i32), ('static_grid', i32),
        ('dyn_mask', u32), ('static_mask', u32),
        ('arinyo_mode', i32), ('idx_arinyo', i32 * 6), ('idx_sigmaNL', i32 * 2),
        ('idx_fs_gauss', i32 * 4), ('idx_fs_exp', i32 * 4),
        ('velo_kind', i32), ('idx_velo', i32 * 2), ('kdamp_power', i32),
        ('gk_drp', f64), ('gk_drt', f64), ('kdamp_scale', f64),
        ('coord_set', i32), ('scale_form', i32), ('idx_scale', i32 * 2),
        ('scale_form_sys', i32), ('idx_scale_sys', i32 * 2),
        ('idx_drp', i32), ('single_ell', i32),
        ('zevol_kind', i32 * 2), ('idx_alpha', i32 * 2), ('idx_croom', i32 * 2),
        ('radiation', i32), ('idx_qso_rad', i32 * 4), ('rescale_sys', i32),
        ('rel', i32), ('asy', i32), ('idx_Arel', i32 * 2), ('idx_Aasy', i32 * 3),
        ('uv_shot', i32), ('idx_uvshot', i32 * 3),
        ('out_slot', i32), ('weight_bias', i32), ('idx_weight_extra', i32),
        ('weight_const', f64),
        ('table_owner', i32), ('mock_los_mode', i32), ('idx_los_smooth_amp', i32),
        ('mock_bin_size', f64), ('direct', i32), ('frozen', i32), ('static_xi', i32), ('pk_twin', i32),
        ('static_evol', i32), ('idx_weight_extra2', i32), ('uv_static', i32),
    ]

    def __init__(self):
        super().__init__()
        # every index defaults to "absent"
        for name, typ in self._fields_:
            if name.startswith('idx_'):
                if typ is i32:
                    setattr(self, name, -1)
                else:
                    setattr(self, name, typ(*([-1] * typ._length_)))
        self.hcd_grid = -1
        self.static_grid = -1
        self.single_ell = -1
        self.weight_const = 1.0
        self.table_owner = -1
        self.pk_twin = -1


class BBTerm(C.Structure):
    """Mirror of vg_bb_term."""
    _fields_ = [('kind', i32), ('position', i32), ('n_coef', i32), ('design_offset', i32),
                ('idx_coef_first', i32), ('idx_sky', i32 * 2), ('idx_coef', i32 * 32)]


_PTR = C.c_void_p
_SIGNATURES = {
    'vg_last_error': (C.c_char_p, []),
    'vg_version': (C.c_int, []),
    'vg_sizeof_eval_desc': (C.c_int, []),
    'vg_sizeof_bb_term': (C.c_int, []),
    'vg_launch_count': (C.c_int64, [C.c_int]),
    'vg_plan_create': (C.c_int, [i32, _PTR, i32, _PTR, i32, _PTR, _PTR, _PTR, _PTR, i32, _PTR, i32,
                                 _PTR, i32, C.POINTER(_PTR)]),
    'vg_plan_destroy': (C.c_int, [_PTR]),
    'vg_plan_set_chunk': (C.c_int, [_PTR, C.c_int64]),
    'vg_plan_set_option': (C.c_int, [_PTR, i32, i32]),
    'vg_corr_n_moment': (C.c_int, [_PTR, i32]),
    'vg_corr_n_static_xi': (C.c_int, [_PTR, i32]),
    'vg_corr_velo_tables': (C.c_int, [_PTR, i32, C.POINTER(i32), C.POINTER(f64)]),
    'vg_plan_set_theta_limits': (C.c_int, [_PTR, _PTR, _PTR]),
    'vg_plan_velo_tables_active': (C.c_int, [_PTR]),
    'vg_corr_fused_metal_operator': (C.c_int, [_PTR, i32, C.POINTER(i32), _PTR]),
    'vg_corr_ozaki_slices': (C.c_int, [_PTR, i32, _PTR]),
    'vg_corr_ozaki_dm_slices': (C.c_int, [_PTR, i32, _PTR]),
    'vg_corr_n_folded': (C.c_int, [_PTR, i32]),
    'vg_plan_set_direct_pk': (C.c_int, [_PTR, _PTR]),
    'vg_plan_set_theta_transform': (C.c_int, [_PTR, _PTR, _PTR]),
    'vg_plan_set_fvoigt': (C.c_int, [_PTR, i32, _PTR, _PTR]),
    'vg_plan_set_uvshot_table': (C.c_int, [_PTR, i32, _PTR, _PTR]),
    'vg_plan_set_priors': (C.c_int, [_PTR, i32, _PTR, _PTR, _PTR]),
    'vg_plan_add_hankel': (C.c_int, [_PTR, i32, _PTR, i32, _PTR, _PTR, i32, C.POINTER(i32)]),
    'vg_plan_add_corr': (C.c_int, [_PTR, i32, i32, i32, C.POINTER(i32)]),
    'vg_corr_add_coords': (C.c_int, [_PTR, i32, _PTR, f64, C.POINTER(i32)]),
    'vg_corr_add_eval': (C.c_int, [_PTR, i32, C.POINTER(EvalDesc), C.POINTER(i32)]),
    'vg_corr_set_static_xi': (C.c_int, [_PTR, i32, i32, i32, i32, _PTR]),
    'vg_corr_set_metal_matrix': (C.c_int, [_PTR, i32, i32, _PTR]),
    'vg_corr_set_metal_matrix_rp': (C.c_int, [_PTR, i32, i32, _PTR, i32, i32]),
    'vg_corr_set_distortion': (C.c_int, [_PTR, i32, _PTR]),
    'vg_corr_set_inst_sys': (C.c_int, [_PTR, i32, _PTR, i32, f64]),
    'vg_corr_set_broadband': (C.c_int, [_PTR, i32, i32, _PTR, i32, _PTR, i32, _PTR, _PTR, _PTR]),
    'vg_corr_set_data': (C.c_int, [_PTR, i32, i32, _PTR, _PTR, _PTR]),
    'vg_corr_set_whitening': (C.c_int, [_PTR, i32, i32, _PTR]),
    'vg_corr_set_data_vector': (C.c_int, [_PTR, i32, i32, _PTR]),
    'vg_corr_whitened': (C.c_int, [_PTR, i32]),
    'vg_plan_set_global_invcov': (C.c_int, [_PTR, i32, _PTR]),
    'vg_plan_finalize': (C.c_int, [_PTR, _PTR]),
    'vg_plan_device_bytes': (C.c_int64, [_PTR]),
    'vg_plan_graph_replays': (C.c_int64, [_PTR]),
    'vg_plan_static_status': (C.c_int, [_PTR]),
    'vg_profile_enable': (C.c_int, [_PTR, C.c_int]),
    'vg_profile_read': (C.c_int, [_PTR, _PTR, _PTR]),
    'vg_model_batch': (C.c_int, [_PTR, _PTR, C.c_int64, _PTR, _PTR, _PTR]),
    'vg_chi2_batch': (C.c_int, [_PTR, _PTR, C.c_int64, _PTR, _PTR, _PTR]),
    'vg_corr_set_mocks': (C.c_int, [_PTR, i32, i32, _PTR]),
    'vg_chi2_batch_mc': (C.c_int, [_PTR, _PTR, _PTR, C.c_int64, _PTR, _PTR, _PTR]),
    'vg_chi2_batch_host': (C.c_int, [_PTR, _PTR, C.c_int64, _PTR, _PTR, _PTR]),
    'vg_plan_set_templates': (C.c_int, [_PTR, _PTR, _PTR, _PTR]),
    'vg_pk_batch': (C.c_int, [_PTR, i32, i32, _PTR, i32, _PTR, C.c_int64, _PTR, _PTR, _PTR]),
    'vg_pktoxi_batch': (C.c_int, [_PTR, i32, _PTR, C.c_int64, _PTR, _PTR, i32, i32, _PTR, _PTR, _PTR, _PTR]),
    'vg_xi_from_grid_batch': (C.c_int, [_PTR, i32, i32, _PTR, i32, _PTR, C.c_int64, _PTR, _PTR, _PTR]),
    'vg_stage_batch': (C.c_int, [_PTR, i32, i32, i32, _PTR, C.c_int64, _PTR, _PTR, _PTR]),
    'vg_stage_size': (C.c_int64, [_PTR, i32, i32]),
    'vg_probe_dfma': (C.c_int, [C.c_int, C.c_int, _PTR, C.POINTER(f64), _PTR]),
    'vg_metal_pair_ratio_range': (C.c_int, [_PTR, i32, _PTR, i32, _PTR]),
    'vg_metal_pair_histograms': (C.c_int, [_PTR, i32, _PTR, i32, i32, f64, f64, _PTR, i32, _PTR, i32,
                                           _PTR, _PTR, _PTR, _PTR]),
    'vg_ozaki_gemm': (C.c_int, [_PTR, _PTR, _PTR, C.c_int, C.c_int, C.c_int, C.c_longlong, C.c_longlong,
                                C.c_longlong, C.c_int, _PTR, _PTR]),
    'vg_dgemm_tn': (C.c_int, [_PTR, _PTR, _PTR, C.c_int, C.c_int, C.c_int, C.c_longlong,
                              C.c_longlong, C.c_longlong, f64, f64, C.c_int, _PTR, _PTR]),
    'vg_dgemm_tn_static': (C.c_int, [_PTR, _PTR, _PTR, C.c_int, C.c_int, C.c_int, C.c_longlong,
                                     C.c_longlong, C.c_longlong, f64, C.c_int, _PTR]),
}

_lib = None


def exported_symbols():
    return sorted(_SIGNATURES)


def load():
    """Load libvega_b200.so (built in-tree by `make` / __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise VegaDeviceError(
                f'{LIB_PATH} is missing: build it with `make` (nvcc, sm_100a). '
                'vega_b200 has no CPU fallback.')
        lib = C.CDLL(str(LIB_PATH))
        for name, (restype, argtypes) in _SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = restype
            fn.argtypes = argtypes
        if lib.vg_sizeof_eval_desc() != C.sizeof(EvalDesc) or lib.vg_sizeof_bb_term() != C.sizeof(BBTerm):
            raise VegaDeviceError('vega_b200: descriptor layout mismatch between _lib.py and the header')
        _lib = lib
    return _lib


def check(rc):
    if rc != 0:
        raise VegaDeviceError(f'vega_b200 CUDA library error {rc}: '
                              f'{load().vg_last_error().decode(errors="replace")}')
