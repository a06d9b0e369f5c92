"""ctypes binding of libsevennet_b200.so (the C-ABI declared in include/sevennet_b200.h).

The library is the only compute path of this package: if it is missing or a call fails, an
exception is raised - there is no PyTorch/CPU fallback.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('B200_LIB_PATH') or os.path.join(HERE, 'libsevennet_b200.so')   # override: kernel experiments

B200_CUTOFF_POLY, B200_CUTOFF_XPLOR = 0, 1


class EdgeParams(C.Structure):
    _fields_ = [('num_basis', C.c_int), ('cutoff_kind', C.c_int), ('normalize_sph', C.c_int),
                ('r_cut', C.c_float), ('r_on', C.c_float), ('poly_p', C.c_float)]


class MlpDesc(C.Structure):
    _fields_ = [('nb', C.c_int), ('h1', C.c_int), ('h2', C.c_int), ('wn', C.c_int)]


class LinBlock(C.Structure):
    _fields_ = [('d', C.c_int), ('mul_in', C.c_int), ('mul_out', C.c_int), ('in_off', C.c_int),
                ('out_off', C.c_int), ('w_off', C.c_int), ('scale', C.c_float)]


class LinDesc(C.Structure):
    _fields_ = [('nblk', C.c_int), ('blk', LinBlock * 4), ('in_dim', C.c_int), ('out_dim', C.c_int)]


class ModelConfig(C.Structure):
    _fields_ = [('n_layers', C.c_int), ('n_species', C.c_int), ('lmax', C.c_int),
                ('mul', (C.c_int * 3) * 8), ('hidden', C.c_int * 2), ('readout_hidden', C.c_int),
                ('edge', EdgeParams), ('species_wise_rescale', C.c_int)]


_vp, _i, _i64, _f, _sz = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t
_P = C.POINTER
HALO_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int)
B200_HALO_FWD_X, B200_HALO_BWD_X, B200_HALO_FORCE = 0, 1, 2
B200_TIMING_CATEGORIES = 9

# name -> (restype, argtypes); must list every symbol include/sevennet_b200.h declares
SIGNATURES = {
    'b200_last_error': (C.c_char_p, []),
    'b200_version': (_i, []),
    'b200_launch_count': (C.c_longlong, []),
    'b200_edge_embed_fwd': (_i, [_P(EdgeParams), _i64, _vp, _vp, _vp, _vp, _vp, _vp]),
    'b200_edge_embed_bwd': (_i, [_P(EdgeParams), _i64, _vp, _vp, _vp, _vp, _vp, _vp]),
    'b200_edge_embed_bwd2': (_i, [_P(EdgeParams), _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    'b200_edge_embed_dk': (_i, [_P(EdgeParams), _i64, _vp, _vp, _vp, _vp, _vp]),
    'b200_radial_mlp_fwd': (_i, [_P(MlpDesc), _i64, _vp, _vp, _vp, _vp, _vp, _vp]),
    'b200_radial_mlp_bwd': (_i, [_P(MlpDesc), _i64, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp]),
    'b200_edge_embed_bwd_r': (_i, [_P(EdgeParams), _i64, _vp, _vp, _vp, _vp, _vp]),
    'b200_radial_table_floats': (_sz, [_P(EdgeParams), _P(MlpDesc)]),
    'b200_radial_table_build': (_i, [_P(EdgeParams), _P(MlpDesc), _vp, _vp, _vp, _vp, _vp, _vp]),
    'b200_radial_table_fwd': (_i, [_P(EdgeParams), _P(MlpDesc), _i64, _vp, _vp, _vp, _vp, _vp]),
    'b200_radial_table_bwd': (_i, [_P(EdgeParams), _P(MlpDesc), _i64, _vp, _vp, _vp, _vp, _vp, _i, _vp]),
    'b200_conv_dims': (_i, [_i, _i, _i, _i, _P(_i), _P(_i)]),
    'b200_conv_fwd': (_i, [_i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _f, _vp, _vp]),
    'b200_conv_bwd_edge': (_i, [_i, _i, _i, _i, _i, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _f, _vp, _vp, _i, _vp]),
    'b200_conv_bwd': (_i, [_i, _i, _i, _i, _i, _i, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _f, _vp, _vp, _i, _vp, _vp]),
    'b200_conv_bwd_scaled': (_i, [_i, _i, _i, _i, _i, _i, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _f, _vp, _vp, _i, _vp, _vp, _i, _vp]),
    'b200_conv_fwd_scaled': (_i, [_i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _f, _vp, _vp, _vp]),
    'b200_conv_bwd_x': (_i, [_i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _f, _vp, _vp]),
    'b200_irreps_linear': (_i, [_P(LinDesc), _i, _vp, _vp, _vp, _i, _i, _vp]),
    'b200_irreps_linear_wgrad': (_i, [_P(LinDesc), _i, _vp, _vp, _vp, _i, _vp]),
    'b200_gate_fwd': (_i, [_i, _i, _i, _i, _vp, _vp, _vp]),
    'b200_gate_bwd': (_i, [_i, _i, _i, _i, _vp, _vp, _vp, _vp]),
    'b200_gate_bwd2': (_i, [_i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    'b200_table_scatter_dot': (_i, [_i, _i, _i, _i64, _f, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    'b200_table_hermite': (_i, [_i, _i, C.c_double, _vp, _vp, _vp, _vp]),
    'b200_table_hermite_adj': (_i, [_i, _i, C.c_double, _vp, _vp, _vp, _vp]),
    'b200_table_interp': (_i, [_i, _i, _i64, _f, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    'b200_table_dot': (_i, [_i, _i, _i64, _f, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    'b200_table_scatter': (_i, [_i, _i, _i64, _f, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    'b200_embed_fwd': (_i, [_i, _i, _i, _vp, _vp, _vp, _vp]),
    'b200_rescale_fwd': (_i, [_i, _vp, _vp, _vp, _vp, _i, _vp, _vp]),
    'b200_colsum': (_i, [_i, _i, _vp, _vp, _vp]),
    'b200_partition_keys': (_i, [_i64, _vp, _P(C.c_double), _P(C.c_int), _P(C.c_int), _P(C.c_double), _P(C.c_int), _vp, _vp]),
    'b200_halo_pack': (_i, [_i64, _i, _vp, _vp, _vp, _vp]),
    'b200_halo_unpack_add': (_i, [_i64, _i, _vp, _vp, _vp, _vp]),
    'b200_force_virial': (_i, [_i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    'b200_graph_workspace_bytes': (_sz, [_i]),
    'b200_graph_count': (_i, [_i, _vp, _P(C.c_double), _P(_i), C.c_double, _vp, _vp, _P(_i64), _vp]),
    'b200_graph_fill': (_i, [_i, _vp, _P(C.c_double), _P(_i), C.c_double, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    'b200_transpose_workspace_bytes': (_sz, [_i, _i64]),
    'b200_csr_transpose': (_i, [_i, _i64, _vp, _vp, _vp, _vp, _vp]),
    'b200_csr_transpose_static_max_nodes': (_i, []),
    'b200_csr_transpose_static': (_i, [_i, _i64, _vp, _vp, _vp, _vp]),
    'b200_model_create': (_i, [_P(ModelConfig), _P(_vp)]),
    'b200_model_set_param': (_i, [_vp, C.c_char_p, _vp, _i64]),
    'b200_model_destroy': (None, [_vp]),
    'b200_model_workspace_bytes': (_sz, [_vp, _i, _i64]),
    'b200_model_set_timing': (_i, [_vp, _i]),
    'b200_model_get_timing': (_i, [_vp, _vp, _i]),
    'b200_model_workspace_bytes_halo': (_sz, [_vp, _i, _i, _i64]),
    'b200_model_compute_halo': (_i, [_vp, _i, _i, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp, _vp, _vp, _vp, _vp, _vp,
                                     _vp]),
    'b200_model_compute': (_i, [_vp, _i, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp, _vp, _vp, _vp, _vp]),
}

_lib = None


def lib():
    """Loads the shared library (once).  Raises if it has not been built: no fallback exists."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f'{LIB_PATH} is missing: build it with `python -m sevennet_finetuning_b200.build` '
                '(nvcc, sm_100a). This package has no CPU or PyTorch fallback.')
        h = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(h, name)
            fn.restype = res
            fn.argtypes = args
        _lib = h
    return _lib


def check(rc, what=''):
    if rc != 0:
        msg = lib().b200_last_error().decode()
        raise RuntimeError(f'sevennet_b200 {what} failed (status {rc}): {msg}')
