"""OUTCAR -> CUDA neighbour list -> collated batch -> one evaluation through the module chain (dataio.to_graphs)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _write_outcar(path, g, energy):
    """a one-step OUTCAR of the golden HfO2 cell (the fields dataio.read_outcar reads)"""
    order = np.argsort(g['types'], kind='stable')
    types = g['types'][order]
    uniq = sorted(set(types.tolist()))
    sym = {uniq[0]: 'Hf', uniq[1]: 'O'} if (types == uniq[0]).sum() < (types == uniq[1]).sum() else {uniq[0]: 'O', uniq[1]: 'Hf'}
    with open(path, 'w') as f:
        for t in uniq:
            f.write(f'   VRHFIN ={sym[t]}: xx\n')
        f.write('   ions per type = ' + ' '.join(str(int((types == t).sum())) for t in uniq) + '\n')
        f.write('  in kB     1.00000   2.00000   3.00000     0.10000    0.20000     0.30000\n')
        f.write('      direct lattice vectors                 reciprocal lattice vectors\n')
        for r in g['cell']:
            f.write('    ' + ' '.join(f'{v:13.9f}' for v in r) + '     0.0 0.0 0.0\n')
        f.write(' POSITION                                       TOTAL-FORCE (eV/Angst)\n ---\n')
        for p in g['pos'][order]:
            f.write('  ' + ' '.join(f'{v:12.5f}' for v in p) + '   0.000000 0.000000 0.000000\n')
        f.write(f' ---\n  free  energy   TOTEN  =       {energy:.8f} eV\n')
    return [sym[t] for t in uniq], order


def test_outcar_to_batch_to_model(gold, weights, tmp_path):
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    from oracle.graph import radius_graph
    from sevennet_finetuning_b200 import dataio, finetune as ft, nn as b200nn, stats
    g = gold('hfo2_96.npz')
    syms, order = _write_outcar(tmp_path / 'OUTCAR', g, -1000.0)
    (tmp_path / 'structure_list').write_text('[hfo2]\nOUTCAR :\nOUTCAR 0\n')
    structs = dataio.structure_list_reader(str(tmp_path / 'structure_list'))['hfo2']
    assert len(structs) == 2
    species = ['X'] * 89                      # the model's species list: put Hf and O where the golden type indices are
    for t, s in zip(sorted(set(g['types'].tolist())), syms):
        species[t] = s
    graphs = dataio.to_graphs(structs, 5.0, species, 'cuda:0')
    ei, S, D = radius_graph(np.round(g['pos'], 5), g['cell'], [True] * 3, 5.0)   # positions as printed (5 decimals)
    assert graphs[0]['edge_index'].shape[1] == ei.shape[1]
    assert abs(stats.DatasetStatistics(graphs, 89).get_avg_num_neigh() - ei.shape[1] / len(g['pos'])) < 1e-9
    batch = ft.collate(graphs, device='cuda:0')
    assert batch['num_atoms'].tolist() == [96, 96] and torch.allclose(batch['stress'][0].cpu() * 1602.1766208,
                                                                       torch.tensor([1, 2, 3, 0.1, 0.2, 0.3]), atol=1e-5)
    m = b200nn.build_E3_equivariant_model(dict(shift=weights['rescale_atomic_energy.shift'].tolist(),
                                               scale=weights['rescale_atomic_energy.scale'].tolist(),
                                               conv_denominator=float(weights['0_convolution.denominator'][0])))
    m.load_state_dict({k: torch.tensor(v) for k, v in weights.items()})
    m = m.cuda().eval()
    m.set_is_batch_data(True)
    out = m(batch)
    e = out['inferred_total_energy'].detach().cpu().numpy()
    # same structure twice; energy of the golden structure up to the 5-decimal rounding of the printed positions
    assert abs(e[0] - e[1]) < 1e-3 and abs(e[0] - float(g['o64_energy'])) / 96 < 5e-4


def test_checkpoint_dict_round_trip(weights, gold, tmp_path):
    """checkpoint_{epoch}.pth as the reference writes it (trainer.py:98-107 + processing_epoch.py:34-40) -> model built from the
    twins by `model_from_checkpoint` (util.py:186-231): same E/F/S as the golden vectors; optimizer state restored."""
    import torch
    from sevennet_finetuning_b200 import dataio, nn as b200nn
    ref_cfg = {'chemical_species': 'auto', '_number_of_species': 89, 'cutoff': 5.0, 'channel': 128, 'is_parity': False, 'lmax': 2,
               'num_convolution_layer': 5, 'irreps_manual': ['128x0e'] + ['128x0e+64x1e+32x2e'] * 4 + ['128x0e'],
               'weight_nn_hidden_neurons': [64, 64], 'radial_basis': {'radial_basis_name': 'bessel', 'bessel_basis_num': 8},
               'cutoff_function': {'cutoff_function_name': 'XPLOR', 'cutoff_on': 4.5},
               'conv_denominator': float(weights['0_convolution.denominator'][0]), 'train_shift_scale': False,
               'train_denominator': False, 'self_connection_type': 'linear', 'shift': weights['rescale_atomic_energy.shift'].tolist(),
               'scale': weights['rescale_atomic_energy.scale'].tolist(), 'use_species_wise_shift_scale': True}
    m = b200nn.build_E3_equivariant_model(dataio.config_from_reference(ref_cfg))
    m.load_state_dict({k: torch.tensor(v) for k, v in weights.items()})
    opt = torch.optim.Adam([p for p in m.parameters() if p.requires_grad], lr=1e-4)
    path = str(tmp_path / 'checkpoint_3.pth')
    torch.save(dataio.checkpoint_dict(m, opt, None, ref_cfg, epoch=3), path)
    ck = torch.load(path, map_location='cpu', weights_only=False)
    assert set(ck) == {'model_state_dict', 'optimizer_state_dict', 'config', 'epoch'} and ck['epoch'] == 3
    m2, cfg = dataio.model_from_checkpoint(path, device='cuda:0')
    assert cfg['cutoff'] == 5.0
    opt2 = torch.optim.Adam([p for p in m2.parameters() if p.requires_grad], lr=1.0)
    opt2.load_state_dict(ck['optimizer_state_dict'])
    assert opt2.param_groups[0]['lr'] == 1e-4
    import numpy as np
    from oracle.graph import radius_graph
    g = gold('si216.npz')
    ei, S, D = radius_graph(g['pos'], g['cell'], [True] * 3, 5.0)
    dev = torch.device('cuda:0')
    m2.eval()
    m2.set_is_batch_data(False)
    out = m2({'x': torch.tensor(g['types'], dtype=torch.long, device=dev), 'edge_index': torch.tensor(ei, device=dev),
              'edge_vec': torch.tensor(D, dtype=torch.float32, device=dev).requires_grad_(True),
              'num_atoms': torch.tensor(len(g['pos']), device=dev),
              'cell_volume': torch.tensor(abs(np.linalg.det(g['cell'])), dtype=torch.float32, device=dev),
              'batch': torch.zeros(len(g['pos']), dtype=torch.long, device=dev)})
    assert abs(float(out['inferred_total_energy'].detach()) - float(g['o64_energy'])) / len(g['pos']) <= 1e-6
    assert np.abs(out['inferred_force'].detach().cpu().numpy() - g['o64_force']).max() <= 1e-5
