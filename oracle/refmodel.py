"""TEST INFRASTRUCTURE - restated node-side leaves, force/stress outputs and model assembly.

Follows ``sevenn/nn/{scale,node_embedding,linear,force_output,sequential}.py`` and the layer
ordering of ``sevenn/model_build.py:186-445`` / ``sevenn/nn/interaction_blocks.py:21-84``.
Not part of the product (see ``oracle/__init__.py``).
"""
from collections import OrderedDict

import torch
import torch.nn as nn

from .refnn import (BesselBasis, EdgeEmbedding, EquivariantGate, IrrepsConvolution,
                    IrrepsLinear, PolynomialCutoff, SelfConnectionIntro, SelfConnectionLinearIntro,
                    SelfConnectionOutro, SphericalEncoding, XPLORCutoff, irreps_str,
                    parse_irreps, sh_irreps)


class OnehotEmbedding(nn.Module):
    """``sevenn/nn/node_embedding.py:14-48``."""

    def __init__(self, num_classes, data_key_x='x', data_key_save='atom_type',
                 data_key_additional='node_attr'):
        super().__init__()
        self.num_classes = num_classes
        self.key_x, self.key_save, self.key_add = data_key_x, data_key_save, data_key_additional

    def forward(self, data):
        inp = data[self.key_x]
        embd = torch.nn.functional.one_hot(inp, self.num_classes).to(torch.get_default_dtype())
        data[self.key_x] = embd
        if self.key_add is not None:
            data[self.key_add] = embd
        if self.key_save is not None:
            data[self.key_save] = inp
        return data


class Rescale(nn.Module):
    """``sevenn/nn/scale.py:12-38``."""

    def __init__(self, shift, scale, train_shift_scale=False, **kw):
        super().__init__()
        self.shift = nn.Parameter(torch.FloatTensor([shift]), requires_grad=train_shift_scale)
        self.scale = nn.Parameter(torch.FloatTensor([scale]), requires_grad=train_shift_scale)

    def forward(self, data):
        data['atomic_energy'] = data['scaled_atomic_energy'] * self.scale + self.shift
        return data


class SpeciesWiseRescale(nn.Module):
    """``sevenn/nn/scale.py:42-73``."""

    def __init__(self, shift, scale, train_shift_scale=False, **kw):
        super().__init__()
        self.shift = nn.Parameter(torch.FloatTensor(shift), requires_grad=train_shift_scale)
        self.scale = nn.Parameter(torch.FloatTensor(scale), requires_grad=train_shift_scale)

    def forward(self, data):
        idx = data['atom_type']
        data['atomic_energy'] = (data['scaled_atomic_energy'] * self.scale[idx].view(-1, 1)
                                 + self.shift[idx].view(-1, 1))
        return data


class AtomReduce(nn.Module):
    """``sevenn/nn/linear.py:53-90``."""

    def __init__(self, data_key_in, data_key_out, constant=1.0):
        super().__init__()
        self.key_input, self.key_output, self.constant = data_key_in, data_key_out, constant
        self._is_batch_data = True

    def forward(self, data):
        if self._is_batch_data:
            src = data[self.key_input].squeeze(1)
            size = int(data['batch'].max()) + 1
            out = torch.zeros(size, dtype=src.dtype, device=src.device)
            out.index_add_(0, data['batch'], src)
            data[self.key_output] = out * self.constant
        else:
            data[self.key_output] = torch.sum(data[self.key_input]) * self.constant
        return data


class ForceOutput(nn.Module):
    """``sevenn/nn/force_output.py:12-48``."""

    def forward(self, data):
        g = torch.autograd.grad([data['inferred_total_energy'].sum()], [data['pos']],
                                create_graph=self.training, allow_unused=True)[0]
        if g is not None:
            data['inferred_force'] = -g
        return data


class ForceStressOutput(nn.Module):
    """``sevenn/nn/force_output.py:52-130`` (deployed serial model)."""

    def __init__(self):
        super().__init__()
        self._is_batch_data = True

    def forward(self, data):
        g = torch.autograd.grad([data['inferred_total_energy'].sum()],
                                [data['pos'], data['_strain']],
                                create_graph=self.training, allow_unused=True)
        if g[0] is not None:
            data['inferred_force'] = -g[0]
        volume = data['cell_volume']
        vlim = 1e-3
        if self._is_batch_data:
            volume = volume.clone()
            volume[volume < vlim] = vlim
        elif volume < vlim:
            volume = torch.tensor(vlim)
        if g[1] is not None:
            if self._is_batch_data:
                s = -(g[1] / volume.view(-1, 1, 1))
                data['inferred_stress'] = torch.vstack(
                    (s[:, 0, 0], s[:, 1, 1], s[:, 2, 2], s[:, 0, 1], s[:, 1, 2], s[:, 0, 2])
                ).transpose(0, 1)
            else:
                s = -(g[1] / volume)
                data['inferred_stress'] = torch.stack(
                    (s[0, 0], s[1, 1], s[2, 2], s[0, 1], s[1, 2], s[0, 2]))
        return data


class ForceStressOutputFromEdge(nn.Module):
    """``sevenn/nn/force_output.py:134-215`` (training / checkpoint calculator / parallel)."""

    def __init__(self):
        super().__init__()
        self._is_batch_data = True

    def forward(self, data):
        tot_num = int(torch.sum(data['num_atoms']))
        rij, edge_idx = data['edge_vec'], data['edge_index']
        fij = torch.autograd.grad([data['inferred_total_energy'].sum()], [rij],
                                  create_graph=self.training, allow_unused=True)[0]
        if fij is None:
            return data
        pf = torch.zeros(tot_num, 3, dtype=fij.dtype, device=fij.device).index_add_(0, edge_idx[0], fij)
        nf = torch.zeros(tot_num, 3, dtype=fij.dtype, device=fij.device).index_add_(0, edge_idx[1], fij)
        data['inferred_force'] = pf - nf
        voigt = torch.cat([rij * fij, (rij[:, 0] * fij[:, 1]).unsqueeze(-1),
                           (rij[:, 1] * fij[:, 2]).unsqueeze(-1),
                           (rij[:, 2] * fij[:, 0]).unsqueeze(-1)], -1)
        s = torch.zeros(tot_num, 6, dtype=fij.dtype, device=fij.device).index_add_(0, edge_idx[1], voigt)
        if self._is_batch_data:
            nb = int(data['batch'].max().item()) + 1
            sout = torch.zeros(nb, 6, dtype=s.dtype, device=s.device).index_add_(0, data['batch'], s)
        else:
            sout = s.sum(0)
        data['inferred_stress'] = -sout / data['cell_volume'].unsqueeze(-1)
        return data


class AtomGraphSequential(nn.Sequential):
    """``sevenn/nn/sequential.py:14-90``."""

    def __init__(self, modules, cutoff=None, type_map=None):
        super().__init__(OrderedDict(modules))
        self.cutoff, self.type_map = cutoff, type_map

    def set_is_batch_data(self, flag: bool):
        for m in self:
            if hasattr(m, '_is_batch_data'):
                m._is_batch_data = flag

    def replace_module(self, key, module):
        self._modules.update({key: module})

    def prepand_module(self, key, module):
        # (torch >= 2.x keeps _modules in a plain dict: rebuild it with the new key first instead of move_to_end)
        rest = [(k, v) for k, v in self._modules.items() if k != key]
        self._modules.clear()
        self._modules[key] = module
        for k, v in rest:
            self._modules[k] = v

    def delete_module_by_key(self, key):
        if key in self._modules:
            del self._modules[key]

    def forward(self, data):
        for m in self:
            data = m(data)
        return data


SEVENNET0_CONFIG = dict(
    cutoff=5.0, num_species=89, channel=128, lmax=2, num_convolution_layer=5, is_parity=False,
    irreps_manual=['128x0e', '128x0e+64x1e+32x2e', '128x0e+64x1e+32x2e', '128x0e+64x1e+32x2e',
                   '128x0e+64x1e+32x2e', '128x0e'],
    radial_basis=dict(bessel_basis_num=8), cutoff_function=dict(name='XPLOR', cutoff_on=4.5),
    weight_nn_hidden_neurons=[64, 64], act_radial='silu', act_scalar='silu', act_gate='silu',
    conv_denominator=35.989574, shift=[0.0] * 89, scale=[1.0] * 89,
    use_species_wise_shift_scale=True, self_connection_type='linear', normalize_sph=True,
)


def build_model(config=None, force_from='edge'):
    """Restated ``build_E3_equivariant_model`` (``sevenn/model_build.py:186-445``) for
    'nequip' interaction + 'linear' self connection with ``irreps_manual`` (SevenNet-0).
    force_from: 'edge' (checkpoint/training model), 'pos' (deployed serial: EdgePreprocess +
    ForceStressOutput, ``sevenn/scripts/deploy.py:22-23``) or None."""
    cfg = dict(SEVENNET0_CONFIG)
    cfg.update(config or {})
    cutoff, nsp = cfg['cutoff'], cfg['num_species']
    parity = -1 if cfg['is_parity'] else 1
    irreps_filter = sh_irreps(cfg['lmax'], parity)
    man = [parse_irreps(s) for s in cfg['irreps_manual']]
    nconv = cfg['num_convolution_layer']
    assert len(man) == nconv + 1
    cf = dict(cfg['cutoff_function'])
    env = (XPLORCutoff(cutoff, cf['cutoff_on']) if cf['name'] == 'XPLOR'
           else PolynomialCutoff(cutoff, cf.get('poly_cut_p_value', 6)))
    L = OrderedDict()
    if force_from == 'pos':
        L['edge_preprocess'] = EdgePreprocess_(True)
    L['edge_embedding'] = EdgeEmbedding(
        BesselBasis(cutoff, **cfg['radial_basis']), env,
        SphericalEncoding(cfg['lmax'], parity, normalize=cfg['normalize_sph']))
    L['onehot_idx_to_onehot'] = OnehotEmbedding(nsp)
    irreps_x = man[0]
    L['onehot_to_feature_x'] = IrrepsLinear(f'{nsp}x0e', irreps_x, 'x', is_embed=True)
    den = cfg['conv_denominator']
    den = den if isinstance(den, list) else [den] * nconv
    for t in range(nconv):
        irreps_out = man[t + 1]
        lmax_node = 0 if t == nconv - 1 else cfg['lmax']
        # util.infer_irreps_out(..., drop_l=lmax_node, parity_mode 'full'/'even')
        tp_target = sorted({(l3, p1 * p2) for _, l1, p1 in irreps_x for _, l2, p2 in irreps_filter
                            for l3 in range(abs(l1 - l2), l1 + l2 + 1)
                            if l3 <= lmax_node and (t < nconv - 1 or p1 * p2 == 1)})
        irreps_out_tp = [(1, l, p) for l, p in tp_target]
        gate = EquivariantGate(irreps_out, cfg['act_scalar'], cfg['act_gate'])
        gin = gate.get_gate_irreps_in()
        if cfg['self_connection_type'] == 'nequip':   # model_build.py:45-56
            L[f'{t}_self_connection_intro'] = SelfConnectionIntro(irreps_x, f'{nsp}x0e', gin)
        else:
            L[f'{t}_self_connection_intro'] = SelfConnectionLinearIntro(irreps_x, gin)
        L[f'{t}_self_interaction_1'] = IrrepsLinear(irreps_x, irreps_x, 'x')
        conv = IrrepsConvolution(irreps_x, irreps_filter, irreps_out_tp,
                                 [cfg['radial_basis']['bessel_basis_num']] + cfg['weight_nn_hidden_neurons'],
                                 cfg['act_radial'], den[t])
        L[f'{t}_convolution'] = conv
        L[f'{t}_self_interaction_2'] = IrrepsLinear(conv.convolution.irreps_out, gin, 'x')
        L[f'{t}_self_connection_outro'] = SelfConnectionOutro()
        L[f'{t}_equivariant_gate'] = gate
        irreps_x = irreps_out
    mid = sum(m for m, _, _ in man[-1])
    L['reduce_input_to_hidden'] = IrrepsLinear(irreps_x, f'{mid // 2}x0e', 'x')
    L['reduce_hidden_to_energy'] = IrrepsLinear(f'{mid // 2}x0e', '1x0e', 'x', 'scaled_atomic_energy')
    if cfg['use_species_wise_shift_scale']:
        L['rescale_atomic_energy'] = SpeciesWiseRescale(cfg['shift'], cfg['scale'])
    else:
        L['rescale_atomic_energy'] = Rescale(cfg['shift'], cfg['scale'])
    L['reduce_total_enegy'] = AtomReduce('atomic_energy', 'inferred_total_energy')
    if force_from == 'edge':
        L['force_output'] = ForceStressOutputFromEdge()
    elif force_from == 'pos':
        L['force_output'] = ForceStressOutput()
    return AtomGraphSequential(L, cutoff)


def EdgePreprocess_(is_stress):
    from .refnn import EdgePreprocess
    return EdgePreprocess(is_stress)
