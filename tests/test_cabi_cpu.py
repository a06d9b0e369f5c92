"""CPU-side checks: the C-ABI library builds, loads, and exports every symbol the header declares;
the product's coupling tables match the reference constants; host-side descriptors are right."""
import ctypes
import os
import re

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from sevennet_finetuning_b200 import _lib, build
    build.build()
    h = ctypes.CDLL(_lib.LIB_PATH)
    header = open(os.path.join(ROOT, 'include', 'sevennet_b200.h')).read()
    declared = set(re.findall(r'\b(b200_[a-z0-9_]+)\s*\(', header))
    assert declared, 'no declarations parsed'
    for name in declared:
        assert hasattr(h, name), f'{name} declared in include/sevennet_b200.h but not exported'
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    assert h.b200_version() >= 1
    # the stand-alone C++ caller (csrc/pair_driver.cpp) is built and linked against the same library
    assert os.path.exists(build.DRIVER) and os.access(build.DRIVER, os.X_OK)


def test_struct_layouts_match_header():
    from sevennet_finetuning_b200 import _lib
    assert ctypes.sizeof(_lib.EdgeParams) == 24
    assert ctypes.sizeof(_lib.MlpDesc) == 16
    assert ctypes.sizeof(_lib.LinBlock) == 28
    assert ctypes.sizeof(_lib.LinDesc) == 4 + 4 * 28 + 8
    assert ctypes.sizeof(_lib.ModelConfig) == 12 + 96 + 8 + 4 + 24 + 4


def test_coupling_tables_match_reference_constants(gold):
    from sevennet_finetuning_b200 import cg
    g = gold('cg_tables.npz')
    for key in ['111', '112', '121', '122', '211', '212', '221', '222']:
        l1, l2, l3 = (int(c) for c in key)
        assert np.abs(cg.coupling(l1, l2, l3) - g[key]).max() < 1e-6, key
    # the committed generated header is in sync with the derivation
    cur = open(os.path.join(ROOT, 'sevennet_finetuning_b200', 'csrc', 'cg_generated.cuh')).read()
    assert cur == cg.generate_header()


def test_conv_dims_no_gpu_needed():
    from sevennet_finetuning_b200 import cg, ops
    for muls, lout in [((128, 0, 0), 2), ((128, 64, 32), 2), ((128, 64, 32), 0), ((5, 0, 3), 2)]:
        _, wn, od = cg.tp_paths(*muls, lout)
        assert ops.conv_dims(*muls, lout) == (wn, od)
    assert ops.conv_dims(128, 64, 32, 2) == (960, 3136)
    assert ops.conv_dims(128, 0, 0, 2) == (384, 1152)
    assert ops.conv_dims(128, 64, 32, 0) == (224, 224)


def test_linear_descriptor():
    from sevennet_finetuning_b200 import ops
    d, n = ops.make_lin_desc('224x0e+384x1e+352x2e', '224x0e+64x1e+32x2e')
    assert n == 86016 and d.nblk == 3 and d.in_dim == 3136 and d.out_dim == 576
    assert abs(d.blk[1].scale - 0.0510310) < 1e-6 and d.blk[2].in_off == 224 + 3 * 384
    d, n = ops.make_lin_desc('128x0e', '224x0e+64x1e+32x2e')
    assert d.nblk == 1 and n == 128 * 224 and d.out_dim == 576


def test_cpu_tensor_is_an_error_not_a_fallback():
    import pytest
    import torch
    from sevennet_finetuning_b200 import ops
    with pytest.raises(RuntimeError):
        ops.gate_fwd(4, 0, 0, torch.zeros(3, 4))


def test_torch_operator_library_registers_the_deployment_op():
    """csrc/torch_ops.cpp: the C++-registered dispatcher operator a deployed TorchScript file calls; scripting the deployment
    module needs no GPU (the operator is resolved, not run) and the saved file carries the reference's extra files."""
    import tempfile
    import numpy as np
    import torch
    from sevennet_finetuning_b200 import build, deploy
    build.build_torch_ops()
    deploy.ops_library()
    schema = str(torch.ops.b200.sevennet_efs.default._schema)
    assert 'Tensor x, Tensor pos, Tensor edge_index, Tensor pbc_shift, Tensor cell, Tensor cell_volume, Tensor flat' in schema
    assert 'CUDA' in torch._C._dispatch_dump('b200::sevennet_efs')
    w = dict(np.load(os.path.join(ROOT, 'tests', 'golden', 'sevennet0_weights.npz')))
    with tempfile.TemporaryDirectory() as d:
        pt = deploy.deploy(w, None, os.path.join(d, 'm'))
        extra = {'chemical_symbols_to_index': '', 'cutoff': '', 'num_species': '', 'dtype': '', 'model_type': ''}
        m = torch.jit.load(pt, _extra_files=extra)
        assert 'b200.sevennet_efs' in m.code or 'b200::sevennet_efs' in str(m.graph)
        assert extra['cutoff'] in (b'5.0', '5.0') and len((extra['chemical_symbols_to_index'].decode()
                                                            if isinstance(extra['chemical_symbols_to_index'], bytes)
                                                            else extra['chemical_symbols_to_index']).split()) == 89
