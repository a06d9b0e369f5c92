"""sevennet_finetuning_b200/dataio.py: structure_list + OUTCAR reading without ASE (sevenn/train/dataload.py:223-300)."""
import os

import numpy as np
import pytest

from sevennet_finetuning_b200 import dataio

_HEAD = '''   VRHFIN =Hf: 6s5d
   TITEL  = PAW_PBE Hf 20Jan2003
   VRHFIN =O: s2p4
   TITEL  = PAW_PBE O 08Apr2002
   number of dos      NEDOS =    301   number of ions     NIONS =      3
   ions per type =               1   2
      direct lattice vectors                 reciprocal lattice vectors
     9.000000000  0.000000000  0.000000000     0.111111111  0.000000000  0.000000000
     0.000000000  9.000000000  0.000000000     0.000000000  0.111111111  0.000000000
     0.000000000  0.000000000  9.000000000     0.000000000  0.000000000  0.111111111
'''


def _step(k, with_stress=True):
    s = ''
    if with_stress:
        s += '  FORCE on cell =-STRESS in cart. coord.  units (eV):\n'
        s += f'  in kB     {10.0 + k:.5f}   20.00000   30.00000     1.00000    -2.00000     3.00000\n'
    s += '''      direct lattice vectors                 reciprocal lattice vectors
     5.000000000  0.000000000  0.000000000     0.200000000  0.000000000  0.000000000
     0.500000000  6.000000000  0.000000000     0.000000000  0.166666667  0.000000000
     0.000000000-0.250000000  7.000000000     0.000000000  0.000000000  0.142857143
 POSITION                                       TOTAL-FORCE (eV/Angst)
 -----------------------------------------------------------------------------------
'''
    for a in range(3):
        s += f'      {0.1 * a + k:.5f}      {1.0 + a:.5f}      {2.0 * a:.5f}         {0.01 * a:.6f}     {-0.02 * a:.6f}      {0.5 - k:.6f}\n'
    s += ' -----------------------------------------------------------------------------------\n'
    s += f'  free  energy   TOTEN  =       {-20.5 - k:.8f} eV\n'
    s += f'  energy  without entropy=      {-20.4 - k:.8f}  energy(sigma->0) =      {-20.45 - k:.8f}\n'
    return s


def test_outcar_steps_and_conventions(tmp_path):
    p = tmp_path / 'OUTCAR_1'
    p.write_text(_HEAD + _step(0) + _step(1, with_stress=False) + _step(2))
    st = dataio.read_outcar(str(p))
    assert len(st) == 3
    assert st[0]['numbers'].tolist() == [72, 8, 8]
    assert st[1]['energy'] == -21.5                                    # the free energy (TOTEN), not energy(sigma->0)
    assert np.allclose(st[0]['cell'], [[5, 0, 0], [0.5, 6, 0], [0, -0.25, 7]])   # the step's cell, glued minus sign handled
    assert np.allclose(st[2]['pos'][1], [2.1, 2.0, 2.0]) and np.allclose(st[2]['forces'][2], [0.02, -0.04, -1.5])
    # VASP's "in kB" numbers (xx yy zz xy yz zx), in eV/A^3 (dataload.py:99-106 undoes ASE's sign and Voigt order)
    assert np.allclose(st[0]['stress'], np.array([10, 20, 30, 1, -2, 3]) / 1602.1766208)
    assert np.isnan(st[1]['stress']).all() and st[2]['stress'][0] == pytest.approx(12 / 1602.1766208)
    assert [s['info']['ionic_step'] for s in dataio.read_outcar(str(p), slice(1, None))] == [1, 2]
    assert len(dataio.read_outcar(str(p), 2)) == 1


def test_structure_list_reader(tmp_path):
    os.makedirs(tmp_path / 'data')
    for k in (1, 2, 3):
        (tmp_path / 'data' / f'OUTCAR_{k}').write_text(_HEAD + _step(k) + _step(k + 10))
    (tmp_path / 'run').mkdir()
    sl = tmp_path / 'run' / 'structure_list'
    sl.write_text('[bulk]\n../data/OUTCAR_{1..2} :\n../data/OUTCAR_3 1:\n\n[other]\n../data/OUTCAR_{1,3} 0\n')
    d = dataio.structure_list_reader(str(sl))
    assert list(d) == ['bulk', 'other'] and len(d['bulk']) == 5 and len(d['other']) == 2
    assert [s['energy'] for s in d['bulk']] == [-21.5, -31.5, -22.5, -32.5, -33.5]
    assert [s['energy'] for s in d['other']] == [-21.5, -23.5]
    sl.write_text('[bulk\n../data/OUTCAR_1 :\n')
    with pytest.raises(ValueError):
        dataio.structure_list_reader(str(sl))
    sl.write_text('[bulk]\n../data/OUTCAR_1 -2:\n')
    with pytest.raises(ValueError):
        dataio.structure_list_reader(str(sl))
    assert dataio.braceexpand('a{x,y}b{08..10}') == ['axb08', 'axb09', 'axb10', 'ayb08', 'ayb09', 'ayb10']
    assert dataio.string2index('3') == 3 and dataio.string2index('::2') == slice(None, None, 2)


@pytest.mark.skipif(not os.path.exists('/root/reference/example_inputs/training/structure_list'),
                    reason='the reference tree exists in the build container only')
def test_shipped_example_data_parses():
    d = dataio.structure_list_reader('/root/reference/example_inputs/training/structure_list')
    assert {k: len(v) for k, v in d.items()} == {'label_1': 10, 'label_2': 10}
    for v in d.values():
        for s in v:
            n = len(s['numbers'])
            assert s['pos'].shape == (n, 3) and s['forces'].shape == (n, 3) and set(s['numbers'].tolist()) <= {72, 8}
            assert np.abs(s['forces'].sum(0)).max() < 1e-3 and np.isfinite(s['energy']) and np.isfinite(s['stress']).all()
            assert abs(np.linalg.det(s['cell'])) / n > 5.0             # A^3 per atom


def test_extxyz_round_trip_and_conventions(tmp_path):
    """ASE-free extended-XYZ reader (the reference's default data format, dataload.py:192-196 through ase.io.read): frames,
    index strings, and the label conventions of dataload.py:94-106 (free energy preferred; stress = minus ASE's, xx yy zz xy yz zx)."""
    import numpy as np
    from sevennet_finetuning_b200 import dataio
    rng = np.random.default_rng(0)
    frames = [dict(numbers=np.array([72, 8, 8, 72]), pos=rng.uniform(0, 5, (4, 3)), cell=np.eye(3) * 5 + rng.normal(0, 0.1, (3, 3)),
                   energy=-30.0 - k, forces=rng.normal(0, 1, (4, 3)), stress=rng.normal(0, 0.01, 6), pbc=[True, True, True], info={})
              for k in range(3)]
    p = str(tmp_path / 'a.extxyz')
    dataio.write_extxyz(p, frames)
    back = dataio.read_extxyz(p)
    assert len(back) == 3
    for a, b in zip(frames, back):
        assert (a['numbers'] == b['numbers']).all() and b['pbc'] == [True] * 3
        for k in ('pos', 'cell', 'forces', 'stress'):
            assert np.array_equal(np.asarray(a[k], float), b[k]), k
        assert a['energy'] == b['energy']
    assert dataio.read_extxyz(p, '-1')[0]['energy'] == -32.0 and [f['energy'] for f in dataio.read_extxyz(p, '1:')] == [-31.0, -32.0]
    # a hand-written frame as ASE writes it: free_energy wins over energy, 3x3 stress with ASE's sign, Z column, no forces
    q = str(tmp_path / 'b.xyz')
    with open(q, 'w') as f:
        f.write('2\nLattice="4.0 0.0 0.0 0.0 4.0 0.0 0.0 0.0 4.0" Properties=species:S:1:pos:R:3:Z:I:1 energy=-1.0 free_energy=-1.25 '
                'stress="0.1 0.4 0.6 0.4 0.2 0.5 0.6 0.5 0.3" pbc="T T F"\nSi 0.0 0.0 0.0 14\nO 1.0 1.0 1.0 8\n')
    (fr,) = dataio.read_extxyz(q)
    assert fr['energy'] == -1.25 and fr['pbc'] == [True, True, False] and list(fr['numbers']) == [14, 8]
    assert np.allclose(fr['stress'], [-0.1, -0.2, -0.3, -0.4, -0.5, -0.6]) and np.isnan(fr['forces']).all()


def test_checkpoint_config_mapping():
    """The reference's checkpoint config (keys of sevenn/_keys.py, values of pretrained_potentials/.../pre_train.yaml) maps to the
    twin builder's keywords; options the CUDA path does not implement raise."""
    import pytest
    from sevennet_finetuning_b200 import dataio
    ref_cfg = {'chemical_species': ['Hf', 'O'], '_number_of_species': 2, 'cutoff': 5.0, 'channel': 128, 'is_parity': False, 'lmax': 2,
               'num_convolution_layer': 5,
               'irreps_manual': ['128x0e'] + ['128x0e+64x1e+32x2e'] * 4 + ['128x0e'], 'weight_nn_hidden_neurons': [64, 64],
               'radial_basis': {'radial_basis_name': 'bessel', 'bessel_basis_num': 8},
               'cutoff_function': {'cutoff_function_name': 'XPLOR', 'cutoff_on': 4.5}, 'conv_denominator': 35.989574,
               'train_shift_scale': True, 'train_denominator': False, 'self_connection_type': 'linear', 'shift': [0.1, 0.2],
               'scale': [1.5, 1.6], 'use_species_wise_shift_scale': True}
    c = dataio.config_from_reference(ref_cfg)
    assert c['num_species'] == 2 and c['cutoff_function'] == {'name': 'XPLOR', 'cutoff_on': 4.5} and c['train_shift_scale'] is True
    assert c['irreps_manual'][1] == '128x0e+64x1e+32x2e' and c['bessel_basis_num'] == 8 and c['shift'] == [0.1, 0.2]
    d = dict(ref_cfg)
    d.pop('irreps_manual')
    assert dataio.config_from_reference(d)['irreps_manual'] == ['128x0e'] + ['128x0e+128x1e+128x2e'] * 4 + ['128x0e']
    with pytest.raises(ValueError):
        dataio.config_from_reference(dict(ref_cfg, is_parity=True))
