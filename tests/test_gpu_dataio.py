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
