"""ctypes binding of ``libviabel_b200.so`` (the C ABI declared in ``include/viabel_b200.h``).

There is no CPU fallback: if the library is missing or a call fails, this raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libviabel_b200.so')

F32, F64 = 0, 1
FAMILY_MF_GAUSS, FAMILY_MF_T, FAMILY_FR_T = 0, 1, 2
OBJ_KLVI, OBJ_CHIVI, OBJ_KLVI_PD = 0, 1, 2
OPT_ADAGRAD, OPT_RMSPROP, OPT_ADAM = 0, 1, 2
MODEL_FUNNEL, MODEL_MIXTURE, MODEL_EIGHT_SCHOOLS, MODEL_REGRESSION, MODEL_GAUSS = range(5)

_p, _i, _ll, _d, _sz = C.c_void_p, C.c_int, C.c_longlong, C.c_double, C.c_size_t
_ull, _u = C.c_ulonglong, C.c_uint

# name -> (restype, argtypes); mirrors include/viabel_b200.h one to one
SIGNATURES = {
    'vb_abi_version': (_i, []),
    'vb_compiled_sm': (_i, []),
    'vb_error_string': (C.c_char_p, [_i]),
    'vb_measure_dmma_peak_workspace_bytes': (_sz, []),
    'vb_measure_dmma_peak': (_i, [C.POINTER(_d), _p, _sz, _p]),
    'vb_model_create_funnel': (_i, [_d, C.POINTER(_p)]),
    'vb_model_create_mixture': (_i, [_d, C.POINTER(_p)]),
    'vb_model_create_eight_schools': (_i, [C.POINTER(_d), C.POINTER(_d), _i, C.POINTER(_p)]),
    'vb_model_create_regression': (_i, [_p, _p, _ll, _i, _d, _d, _i, C.POINTER(_p)]),
    'vb_model_regression_attach_split': (_i, [_p, _p]),
    'vb_split_f16_bytes': (_sz, [_ll, _ll]),
    'vb_split_f16': (_i, [_p, _ll, _ll, _ll, _p, _p]),
    'vb_model_regression_attach_split_f16': (_i, [_p, _p]),
    'vb_model_create_gauss': (_i, [_p, _p, _i, C.POINTER(_p)]),
    'vb_model_gauss_cache_bytes': (_sz, [_i]),
    'vb_model_gauss_attach_cache': (_i, [_p, _p, _p]),
    'vb_model_kind': (_i, [_p]),
    'vb_model_dim': (_i, [_p]),
    'vb_model_destroy': (None, [_p]),
    'vb_model_logp_grad_workspace_bytes': (_sz, [_p, _ll, _i]),
    'vb_model_logp_grad': (_i, [_p, _i, _p, _ll, _p, _p, _p, _sz, _p]),
    'vb_base_draws': (_i, [_i, _d, _i, _ll, _i, _ull, _u, _ull, _p, _p, _p]),
    'vb_base_draws_fr_split': (_i, [_d, _ll, _i, _ull, _u, _ull, _p, _p, _p, _p]),
    'vb_mf_step_workspace_bytes': (_sz, [_i]),
    'vb_mf_step': (_i, [_i, _d, _p, _i, _d, _i, _p, _ll, _p, _ull, _u, _ull, _p, _p, _p, _p, _sz, _p]),
    'vb_mf_adagrad_state_bytes': (_sz, [_i, _i]),
    'vb_mf_adagrad_create': (_i, [_i, _d, _p, _i, _d, _i, _p, _ll, _p, _i, _i, _d, _p, _p, _p, _p, _p, _p, _sz,
                                  _p, _sz, _i, _i, _p, C.POINTER(_p)]),
    'vb_mf_optimizer_create': (_i, [_i, _i, _d, _p, _i, _d, _i, _p, _ll, _p, _i, _d, _p, _p, _p, _p, _p, _p, _p, _p, _sz,
                                    _p, _sz, _i, _i, _p, C.POINTER(_p)]),
    'vb_mf_adagrad_run': (_i, [_p, _i, _p]),
    'vb_mf_adagrad_destroy': (_i, [_p]),
    'vb_mf_finalize': (_i, [_i, _i, _d, _p, _i, _p, _i, _p, _p, _p]),
    'vb_mf_reduce_workspace_bytes': (_sz, [_ll, _i]),
    'vb_mf_reduce': (_i, [_i, _d, _i, _d, _i, _p, _ll, _i, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    'vb_mf_samples_logweights': (_i, [_i, _d, _p, _i, _p, _ll, _p, _ull, _u, _ull, _p, _p, _p, _p]),
    'vb_weighted_moments_workspace_bytes': (_sz, [_ll, _i]),
    'vb_weighted_moments': (_i, [_i, _p, _p, _ll, _i, _p, _p, _p, _sz, _p]),
    'vb_mf_sample': (_i, [_i, _p, _ll, _i, _p, _p, _p]),
    'vb_mf_logdensity': (_i, [_i, _d, _i, _p, _ll, _i, _p, _p, _p]),
    'vb_fr_sample_workspace_bytes': (_sz, [_i, _ll, _i]),
    'vb_fr_sample': (_i, [_i, _p, _p, _d, _ll, _i, _p, _p, _p, _p, _sz, _p]),
    'vb_fr_logdensity_workspace_bytes': (_sz, [_ll, _i]),
    'vb_fr_logdensity': (_i, [_i, _p, _p, _d, _d, _ll, _i, _p, _p, _p, _sz, _p]),
    'vb_fr_chain_workspace_bytes': (_sz, [_i]),
    'vb_fr_chain_forward': (_i, [_p, _i, _i, _p, _p, _p, _p, _p, _sz, _p]),
    'vb_fr_chain_backward': (_i, [_i, _i, _i, _d, _d, _p, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    'vb_fr_chain_info': (_i, [_p, _i, C.POINTER(_i), _p]),
    'vb_fr_rows_workspace_bytes': (_sz, [_ll]),
    'vb_fr_rows': (_i, [_i, _d, _d, _i, _ll, _i, _p, _p, _p, _d, _p, _p, _p, _p, _sz, _p]),
    'vb_fr_weighted_workspace_bytes': (_sz, [_ll, _i]),
    'vb_fr_weighted': (_i, [_i, _d, _ll, _i, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    'vb_fr_weighted_split': (_i, [_d, _ll, _i, _p, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    'vb_gemm_workspace_bytes': (_sz, [_i, _ll, _ll, _ll]),
    'vb_gemm': (_i, [_i, _i, _i, _ll, _ll, _ll, _d, _p, _ll, _p, _ll, _d, _p, _ll, _p, _sz, _p]),
    'vb_split_bytes': (_sz, [_ll, _ll]),
    'vb_split_f32': (_i, [_p, _ll, _ll, _ll, _p, _p]),
    'vb_gemm_tf32x3_split': (_i, [_i, _i, _ll, _ll, _ll, _p, _p, _p, _ll, _p]),
    'vb_bounds_divergence_workspace_bytes': (_sz, [_ll]),
    'vb_bounds_divergence': (_i, [_p, _ll, _d, _p, _p, _sz, _p]),
    'vb_bounds_divergence_static_max_n': (_ll, []),
    'vb_bounds_divergence_static': (_i, [_p, _ll, _d, _p, _p]),
    'vb_bounds_divergence_partials': (_i, [_p, _ll, _d, _p, _p, _sz, _p]),
    'vb_bounds_divergence_merge': (_i, [_p, _i, _d, _p, _p]),
    'vb_bounds_moments_workspace_bytes': (_sz, [_ll, _i]),
    'vb_bounds_colsum': (_i, [_i, _p, _ll, _i, _p, _p, _sz, _p]),
    'vb_bounds_central': (_i, [_i, _p, _ll, _i, _p, _p, _p, _p, _sz, _p]),
    'vb_psis_tail_capacity': (_ll, [_ll, _d]),
    'vb_psis_smooth_workspace_bytes': (_sz, [_ll, _d]),
    'vb_psis_smooth': (_i, [_p, _p, _ll, _d, _p, _p, _p, _sz, _p]),
    'vb_psis_sample': (_i, [_p, _ll, _ll, _p, _p]),
    'vb_psis_pivot': (_i, [_p, _ll, _ll, _d, _p, _p]),
    'vb_psis_filter_workspace_bytes': (_sz, [_ll]),
    'vb_psis_filter': (_i, [_p, _ll, _ll, _p, _p, _ll, _p, _p, _p, _sz, _p]),
    'vb_psis_prepare': (_i, [_p, _ll, _p, _p, _p, _p, _p]),
    'vb_psis_refine_workspace_bytes': (_sz, []),
    'vb_psis_refine': (_i, [_p, _ll, _p, _p, _ll, _p, _p, _p, _p, _sz, _p]),
    'vb_psis_hist_bins': (_i, []),
    'vb_psis_refine_hist': (_i, [_p, _ll, _p, _p, _p, _p, _sz, _p]),
    'vb_psis_refine_select': (_i, [_p, _ll, _p, _p, _p, _ll, _p, _p, _p, _p, _sz, _p]),
    'vb_psis_tail_workspace_bytes': (_sz, [_ll, _ll]),
    'vb_psis_tail': (_i, [_p, _ll, _ll, _p, _p, _p, _i, _p, _p, _p, _sz, _p]),
    'vb_psis_write': (_i, [_p, _p, _ll, _ll, _p, _ll, _ll, _p, _p]),
    'vb_psis_tail_dev': (_i, [_p, _ll, _p, _ll, _p, _p, _p, _i, _p, _ll, _p, _p, _p, _p, _sz, _p]),
    'vb_psis_prepare_hist_dev': (_i, [_p, _ll, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    'vb_psis_refine_select_dev': (_i, [_p, _ll, _p, _p, _p, _p, _ll, _p, _ll, _p, _ll, _p, _sz, _p]),
    'vb_psis_gather_compact': (_i, [_p, _i, _ll, _p, _p, _p, _p, _p, _p]),
    'vb_psis_div_workspace_bytes': (_sz, []),
    'vb_psis_write_div': (_i, [_p, _p, _ll, _ll, _p, _ll, _ll, _p, _p, _p, _sz, _p]),
    'vb_psis_smooth_div': (_i, [_p, _p, _ll, _d, _p, _p, _p, _p, _sz, _p]),
    'vb_psis_smooth_host': (_i, [_p, _p, _ll, _d, _p, _p, _p, _p, _p, _sz, _p]),
    'vb_psis_columns_max_n': (_ll, []),
    'vb_psis_smooth_columns': (_i, [_p, _p, _ll, _ll, _d, _p, _p, _p, _p]),
    'vb_gpd_fit_workspace_bytes': (_sz, [_ll]),
    'vb_gpd_fit': (_i, [_p, _ll, _p, _p, _sz, _p]),
    'vb_sort_f64_workspace_bytes': (_sz, [_ll]),
    'vb_sort_f64': (_i, [_p, _ll, _p, _p, _sz, _p]),
    'vb_gpd_fit_quadrature': (_i, [_p, _ll, _p, _p, _p, _p, _sz, _p]),
    'vb_gpd_inv': (_i, [_p, _ll, _d, _d, _p, _p]),
    'vb_transpose_scale_f64': (_i, [_p, _ll, _ll, _d, _p, _p]),
    'vb_logsumexp_workspace_bytes': (_sz, [_ll]),
    'vb_logsumexp': (_i, [_p, _ll, _p, _p, _sz, _p]),
    'vb_psis_normalise': (_i, [_p, _ll, _d, _p]),
    'vb_max_workspace_bytes': (_sz, [_ll]),
    'vb_max': (_i, [_p, _ll, _p, _p, _sz, _p]),
    'vb_rhat_workspace_bytes': (_sz, [_i, _ll, _i]),
    'vb_rhat_windows': (_i, [_p, _i, _ll, _ll, _p, _i, _p, _p, _p, _sz, _p]),
    'vb_iterate_average': (_i, [_p, _ll, _ll, _ll, _ll, _p, _p]),
}


class ViabelB200Error(RuntimeError):
    pass


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            'viabel_b200: %s not found. Build it with `python viabel_b200/build.py` '
            '(nvcc, sm_100a). There is no CPU fallback.' % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is missing: fail loudly
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


def check(rc, what=''):
    if rc != 0:
        msg = lib.vb_error_string(rc).decode()
        raise ViabelB200Error('viabel_b200 %s failed: %s (status %d)' % (what, msg, rc))
