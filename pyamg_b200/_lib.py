"""ctypes binding of libb200amg.so (include/b200amg.h).  There is NO fallback: if the library is
missing, or a call returns non-zero, this raises."""
import ctypes
import os
from ctypes import POINTER, Structure, c_double, c_int, c_longlong, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('B200AMG_LIB') or os.path.join(_HERE, 'libb200amg.so')   # override: kernel experiments

p = c_void_p          # every device pointer / stream travels as an integer address


class LevelDesc(Structure):
    """mirror of struct b200amg_level_desc"""
    _fields_ = [
        ('n', c_int), ('bs', c_int), ('flavor', c_int),
        ('Ap', p), ('Aj', p), ('Ax', p),
        ('nc', c_int), ('p_br', c_int), ('p_bc', c_int),
        ('Pp', p), ('Pj', p), ('Px', p),
        ('Rp', p), ('Rj', p), ('Rx', p),
        ('pre_kind', c_int), ('post_kind', c_int),
        ('pre_omega', c_double), ('post_omega', c_double),
        ('pre_iters', c_int), ('post_iters', c_int),
        ('pre_sweep', c_int), ('post_sweep', c_int),
        ('Id', p), ('color_ptr_host', POINTER(c_int)), ('K', c_int),
        ('pre_ncoef', c_int), ('post_ncoef', c_int),
        ('pre_coef', c_double * 8), ('post_coef', c_double * 8),
        ('Dinv', p),
        ('gs_csr', c_int),
    ]


_SIGS = {
    'b200amg_last_error': (ctypes.c_char_p, []),
    'b200amg_version': (c_int, []),
    'b200amg_launch_count': (c_longlong, []),
    'b200amg_csr_matvec': (c_int, [c_int, c_int, p, p, p, p, p, p]),
    'b200amg_csr_spmv': (c_int, [c_int, c_int, p, p, p, p, p, c_double, c_double, p]),
    'b200amg_csr_residual': (c_int, [c_int, p, p, p, p, p, p, p]),
    'b200amg_csr_matvec_add': (c_int, [c_int, p, p, p, p, p, p]),
    'b200amg_bsr_matvec': (c_int, [c_int, c_int, c_int, c_int, p, p, p, p, p, p, c_int, p]),
    'b200amg_jacobi': (c_int, [p, p, p, p, p, p, c_int, c_int, c_int, c_double, p]),
    'b200amg_bsr_jacobi': (c_int, [p, p, p, p, p, p, c_int, c_int, c_int, c_int, c_double, p]),
    'b200amg_gauss_seidel_indexed': (c_int, [p, p, p, p, p, p, c_int, c_int, c_int, POINTER(c_int), c_int, p]),
    'b200amg_gauss_seidel': (c_int, [p, p, p, p, p, c_int, c_int, c_int, p]),
    'b200amg_bsr_gauss_seidel': (c_int, [p, p, p, p, p, c_int, c_int, c_int, c_int, p, POINTER(c_int), c_int, p]),
    'b200amg_op_create': (c_int, [c_int, p, p, p, c_int, POINTER(c_void_p), p]),
    'b200amg_op_apply': (c_int, [p, c_int, p, p, p, c_double, p]),
    'b200amg_op_layout': (c_int, [p]),
    'b200amg_bsr_op_create': (c_int, [c_int, c_int, c_int, p, p, p, POINTER(c_void_p), p]),
    'b200amg_vcycle_layout': (c_int, [p, c_int, c_int]),
    'b200amg_op_destroy': (c_int, [p]),
    'b200amg_halo_push': (c_int, [p, c_int, p, p, p, p, p, c_int, p, p]),
    'b200amg_halo_wait': (c_int, [p, c_int, p, c_longlong, p]),
    'b200amg_halo_exchange': (c_int, [p, c_int, p, p, p, p, p, p, c_int, p, p, c_longlong, p]),
    'b200amg_dot': (c_int, [c_longlong, p, p, p, p, p]),
    'b200amg_dot_noalloc': (c_int, [c_longlong, p, p, p, p]),
    'b200amg_pcg_scalars': (c_int, [p, c_int, p]),
    'b200amg_pcg_update_xr': (c_int, [c_longlong, p, p, p, p, p, c_int, p]),
    'b200amg_pcg_update_p': (c_int, [c_longlong, p, p, p, p]),
    'b200amg_axpy': (c_int, [c_longlong, c_double, p, p, p]),
    'b200amg_scal': (c_int, [c_longlong, c_double, p, p, p]),
    'b200amg_scale_diag': (c_int, [c_longlong, c_double, p, p, p, c_int, p]),
    'b200amg_extract_diag': (c_int, [c_int, c_int, p, p, p, p, p]),
    'b200amg_dense_matvec': (c_int, [c_int, p, p, p, p]),
    'b200amg_spgemm_count': (c_int, [c_int, p, p, p, p, p]),
    'b200amg_spgemm_numeric': (c_int, [c_int, p, p, p, p, p, p, p, p, p, p, p, p, p, c_int, p]),
    'b200amg_spgemm_compact': (c_int, [c_int, p, p, p, p, p, p, p]),
    'b200amg_exclusive_scan_i64': (c_int, [c_longlong, p, p, p, p]),
    'b200amg_exclusive_scan_i32': (c_int, [c_longlong, p, p, p, p]),
    'b200amg_scan_scratch_elems': (c_longlong, [c_longlong]),
    'b200amg_csr_transpose_work_bytes': (c_longlong, [c_int, c_int, c_longlong]),
    'b200amg_csr_transpose': (c_int, [c_int, c_int, c_longlong, p, p, p, p, p, p, p, p]),
    'b200amg_csr_scale_rows': (c_int, [c_int, p, p, p, c_double, c_int, p, p]),
    'b200amg_tentative_scalar': (c_int, [c_int, p, p, p, c_double, p, p, p]),
    'b200amg_flag_nonneg': (c_int, [c_int, p, p, p]),
    'b200amg_tentative_fill': (c_int, [c_int, p, p, p, p, p, p]),
    'b200amg_csr_sub_sorted': (c_int, [c_int, p, p, p, p, p, p, c_int, p, p, p, p, p]),
    'b200amg_csr_sub_general': (c_int, [c_int, p, p, p, p, p, p, c_int, p, p, p, p, p]),
    'b200amg_block_sqnorm': (c_int, [c_longlong, c_int, c_int, p, p, p]),
    'b200amg_abs_diag': (c_int, [c_int, p, p, p, p, p]),
    'b200amg_symmetric_strength': (c_int, [c_int, c_double, p, p, p, p, c_int, p, p, p, p, p]),
    'b200amg_bsr_scale_rows': (c_int, [c_int, c_int, c_int, p, p, p, c_double, c_int, p, p]),
    'b200amg_bsr_gather_blocks': (c_int, [c_longlong, c_int, c_int, p, p, p, c_int, p]),
    'b200amg_fit_candidates': (c_int, [c_int, c_int, c_int, p, p, p, c_double, p, p, p]),
    'b200amg_tentative_scatter': (c_int, [c_longlong, c_int, c_int, p, p, p, p, p]),
    'b200amg_bsr_sub': (c_int, [c_int, c_int, c_int, p, p, p, p, p, p, c_int, c_int, p, p, p, p, p]),
    'b200amg_bsr_slot_map': (c_int, [c_int, p, p, p, p, p, p]),
    'b200amg_bsr_matmat_numeric': (c_int, [c_int, c_int, c_int, c_int, p, p, p, p, p, p, p, p, p, c_longlong, p, p]),
    'b200amg_block_jacobi': (c_int, [p, p, p, p, p, p, p, c_int, c_int, c_int, c_double, c_int, p]),
    'b200amg_block_gauss_seidel': (c_int, [p, p, p, p, p, p, c_int, c_int, c_int, c_int, p, POINTER(c_int), c_int, p]),
    'b200amg_block_diag_inverse': (c_int, [c_int, c_int, p, p, p, p, p]),
    'b200amg_block_diag_inverse_gj': (c_int, [c_int, c_int, p, p, p, p, p]),
    'b200amg_pinv_array': (c_int, [p, c_int, c_int, c_int, p]),
    'b200amg_rs_strength': (c_int, [c_int, c_double, p, p, p, c_int, c_int, p, p, p, p, p]),
    'b200amg_csr_not_canonical': (c_int, [c_int, p, p, p, p]),
    'b200amg_rs_direct_interpolation_pass1': (c_int, [c_int, p, p, p, p, p]),
    'b200amg_rs_direct_interpolation_pass2': (c_int, [c_int, p, p, p, p, p, p, p, p, p, p, p, p]),
    'b200amg_rs_cf_splitting_host': (c_int, [c_int, p, p, p, p, p]),
    'b200amg_vertex_coloring_mis_host': (c_int, [c_int, p, p, p]),
    'b200amg_vertex_coloring_jones_plassmann_host': (c_int, [c_int, p, p, p, p]),
    'b200amg_vertex_coloring_ldf_host': (c_int, [c_int, p, p, p, p]),
    'b200amg_naive_aggregation_host': (c_int, [c_int, p, p, p, p]),
    'b200amg_gs_wavefronts_host': (c_int, [c_int, p, p, p]),
    'b200amg_standard_aggregation_host': (c_int, [c_int, p, p, p, p]),
    'b200amg_level_desc_size': (c_int, []),
    'b200amg_vcycle_work_bytes': (c_longlong, [c_int, POINTER(LevelDesc)]),
    'b200amg_vcycle_create': (c_int, [c_int, POINTER(LevelDesc), p, p, c_longlong, POINTER(c_void_p)]),
    'b200amg_vcycle_run': (c_int, [p, p, p, c_int, c_int, p]),
    'b200amg_vcycle_run_zero': (c_int, [p, p, p, c_int, c_int, p]),
    'b200amg_vcycle_apply_op': (c_int, [p, c_int, c_int, p, p, p, p]),
    'b200amg_vcycle_profile': (c_int, [p, p, p, c_int, c_int, POINTER(c_int), POINTER(c_int),
                                       POINTER(ctypes.c_float), POINTER(c_int), p]),
    'b200amg_vcycle_residual_norm2': (c_int, [p, p, p, p, p]),
    'b200amg_vcycle_destroy': (c_int, [p]),
}

EXPORTS = tuple(_SIGS)


class B200AmgError(RuntimeError):
    pass


_lib = None


def load():
    """Load libb200amg.so (built by pyamg_b200/csrc/Makefile or __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200AmgError('libb200amg.so not built: run `make -C pyamg_b200/csrc` '
                               '(there is no CPU fallback)')
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            f = getattr(lib, name)
            f.restype, f.argtypes = res, args
        _lib = lib
    return _lib


def call(name, *args):
    """Call an int-returning entry point and raise on a non-zero status."""
    lib = load()
    rc = getattr(lib, name)(*args)
    if rc != 0:
        raise B200AmgError('%s failed (%d): %s' % (name, rc, lib.b200amg_last_error().decode()))


def launch_count():
    return int(load().b200amg_launch_count())
