"""Reading the reference's training inputs without ASE: ``structure_list`` files and VASP OUTCARs
-> per-structure labelled graphs for ``finetune.collate``; the neighbour lists come from the CUDA cell list
(``ops.radius_graph``), not from ``ase.neighborlist``.

Reference: ``sevenn/train/dataload.py:223-300`` (``structure_list_reader``: ``[label]`` headers, brace-expanded file
patterns, an index expression per line; every ionic step of an OUTCAR parsed by ASE's ``outcarchunks`` with the
``PositionsAndForces, Stress, Energy, Cell`` parsers) and ``:71-154`` (``atoms_to_graph``: labels taken from the ASE atoms).
ASE is not installed in this environment, so the conventions below restate what those ASE parsers deliver and what the
reference does with it (parity with ASE itself is unpinned; the shipped OUTCARs parse, see tests/test_dataio_cpu.py):
  * energy  = "free  energy   TOTEN" (``get_potential_energy(force_consistent=True)``, dataload.py:94-97), eV
  * forces  = columns 4-6 of the "POSITION ... TOTAL-FORCE (eV/Angst)" block, positions = columns 1-3
  * cell    = the "direct lattice vectors" block nearest before the positions of the step
  * stress  = the "in kB" line (XX YY ZZ XY YZ ZX, VASP's sign) x 0.1 GPa in eV/A^3: ASE stores minus that in Voigt order,
              the reference negates and reorders it back (dataload.py:99-106), so the label is VASP's own six numbers, in
              eV/A^3, order xx yy zz xy yz zx; NaN when the step prints no stress
  * species = "VRHFIN =El:" per POTCAR, repeated by "ions per type"
"""
import os
import re

import numpy as np
import torch

KB_TO_EV_A3 = 1.0 / 1602.1766208     # 0.1 GPa in eV/A^3 (ase.units.GPa, CODATA 2014, as sevenn/error_recorder.py:37)

CHEMICAL_SYMBOLS = ('X H He Li Be B C N O F Ne Na Mg Al Si P S Cl Ar K Ca Sc Ti V Cr Mn Fe Co Ni Cu Zn Ga Ge As Se Br Kr '
                    'Rb Sr Y Zr Nb Mo Tc Ru Rh Pd Ag Cd In Sn Sb Te I Xe Cs Ba La Ce Pr Nd Pm Sm Eu Gd Tb Dy Ho Er Tm Yb Lu '
                    'Hf Ta W Re Os Ir Pt Au Hg Tl Pb Bi Po At Rn Fr Ra Ac Th Pa U Np Pu Am Cm Bk Cf Es Fm Md No Lr Rf Db Sg '
                    'Bh Hs Mt Ds Rg Cn Nh Fl Mc Lv Ts Og').split()
ATOMIC_NUMBER = {s: z for z, s in enumerate(CHEMICAL_SYMBOLS)}


def braceexpand(expr):
    """The subset of shell brace expansion structure_list files use: ``{a..b}`` integer ranges (zero padding kept) and
    ``{x,y,z}`` lists, nested / repeated braces expanded left to right."""
    m = re.search(r'\{([^{}]*)\}', expr)
    if not m:
        return [expr]
    body, out = m.group(1), []
    rng = re.fullmatch(r'(-?\d+)\.\.(-?\d+)(?:\.\.(-?\d+))?', body)
    if rng:
        a, b = int(rng.group(1)), int(rng.group(2))
        step = abs(int(rng.group(3))) if rng.group(3) else 1
        width = max(len(rng.group(1)), len(rng.group(2))) if (rng.group(1).startswith('0') or rng.group(2).startswith('0')) else 0
        items = [str(v).zfill(width) for v in (range(a, b + 1, step) if a <= b else range(a, b - 1, -step))]
    else:
        items = body.split(',')
    for it in items:
        out.extend(braceexpand(expr[:m.start()] + it + expr[m.end():]))
    return out


def string2index(s):
    """ase.io.formats.string2index: 'i' -> int, 'a:b:c' -> slice."""
    s = s.strip()
    if ':' not in s:
        return int(s)
    parts = [int(p) if p.strip() else None for p in s.split(':')]
    return slice(*parts)


def read_outcar(path, index=slice(None)):
    """Every ionic step of a VASP OUTCAR as dict(numbers[N], pos[N,3], cell[3,3], energy, forces[N,3], stress[6], pbc)."""
    symbols, counts = [], None
    cell = stress = pos_force = None
    steps = []
    with open(path, 'r') as f:
        lines = f.readlines()
    i, n = 0, len(lines)
    while i < n:
        ln = lines[i]
        if 'VRHFIN =' in ln:
            symbols.append(ln.split('=')[1].split(':')[0].strip())
        elif 'ions per type =' in ln:
            counts = [int(t) for t in ln.split('=')[1].split()]
        elif 'direct lattice vectors' in ln:
            cell = np.array([[float(t) for t in lines[i + k].replace('-', ' -').split()[:3]] for k in (1, 2, 3)])
            i += 3
        elif ln.lstrip().startswith('in kB'):
            try:
                stress = np.array([float(t) for t in ln.split()[2:8]]) * KB_TO_EV_A3
            except ValueError:      # '******' overflow fields
                stress = np.full(6, np.nan)
        elif 'POSITION' in ln and 'TOTAL-FORCE' in ln:
            nat = sum(counts)
            rows = np.array([[float(t) for t in lines[i + 2 + k].split()[:6]] for k in range(nat)])
            pos_force = (rows[:, :3], rows[:, 3:6])
            i += 2 + nat
        elif 'free  energy   TOTEN' in ln:
            if pos_force is None or cell is None:
                raise RuntimeError(f'{path}: energy line {i + 1} without positions / cell before it')
            numbers = np.repeat([ATOMIC_NUMBER[s] for s in symbols[:len(counts)]], counts)
            steps.append({'numbers': numbers, 'pos': pos_force[0], 'cell': cell.copy(), 'forces': pos_force[1],
                          'energy': float(ln.split('=')[1].split()[0]),
                          'stress': stress if stress is not None else np.full(6, np.nan), 'pbc': (True, True, True),
                          'info': {'data_from': 'user_OUTCAR', 'file': os.path.abspath(path), 'ionic_step': len(steps)}})
            pos_force = stress = None
        i += 1
    return steps[index] if isinstance(index, slice) else [steps[index]]


def structure_list_reader(filename):
    """dataload.py:223-300: {label: [structures]}; relative file patterns are resolved against the list's directory as the
    reference resolves them against the working directory it is started in."""
    base = os.path.dirname(os.path.abspath(filename))
    out, label = {}, None
    with open(filename, 'r') as f:
        for raw in f:
            line = raw.strip()
            if not line:
                continue
            if line.startswith('['):
                if not line.endswith(']'):
                    raise ValueError('wrong structure_list title format')
                label = line[1:-1]
                out[label] = []
                continue
            if label is None:
                raise ValueError('wrong structure_list format')
            tok = line.split()
            if len(tok) == 1:
                tok.append(':')
            elif len(tok) != 2:
                raise ValueError('wrong structure_list format')
            index = string2index(tok[1])
            if isinstance(index, slice) and any(v is not None and v < 0 for v in (index.start, index.stop)):
                raise ValueError('Negative index is not supported yet')   # dataload.py:281-284
            for name in braceexpand(tok[0]):
                out[label].extend(read_outcar(name if os.path.isabs(name) else os.path.join(base, name), index))
    return out


def to_graphs(structures, cutoff, chemical_species, device):
    """``atoms_to_graph`` (dataload.py:71-154) + the one-hot index map of ``AtomGraphDataset.x_to_one_hot_idx``
    (dataset.py:203-217): per-structure dicts for ``finetune.collate``.  ``chemical_species``: the model's species list
    (symbols, in the model's order); the neighbour list is the CUDA cell list."""
    from . import ops
    type_of = {ATOMIC_NUMBER[s]: k for k, s in enumerate(chemical_species)}
    graphs = []
    for st in structures:
        try:
            types = np.array([type_of[int(z)] for z in st['numbers']])
        except KeyError as e:
            raise ValueError(f'atomic number {e} is not among the model species') from None
        g = ops.radius_graph(torch.tensor(st['pos'], dtype=torch.float64, device=device), st['cell'], st['pbc'], cutoff)
        graphs.append({'x': torch.tensor(types), 'edge_index': torch.stack([g['dst'], g['src']]).long().cpu(),
                       'edge_vec': g['edge_vec'].cpu(), 'cell_volume': abs(float(np.linalg.det(st['cell']))),
                       'num_atoms': len(types), 'total_energy': st['energy'],
                       'force_of_atoms': torch.tensor(st['forces'], dtype=torch.float32),
                       'stress': torch.tensor(st['stress'], dtype=torch.float32), 'info': st['info']})
    return graphs


# ---- extended XYZ (the reference's default data format, dataload.py:192-196 via ase.io.read) ---------------------------------
_KV = re.compile(r'(\w+)\s*=\s*("[^"]*"|\S+)')


def _parse_index(index, n):
    """ASE index strings: ':', 'a:b', 'a:b:c', '-1', or an int / slice."""
    if isinstance(index, slice):
        return range(n)[index]
    if isinstance(index, int):
        return [range(n)[index]]
    parts = str(index).split(':')
    if len(parts) == 1:
        return [range(n)[int(parts[0])]]
    vals = [int(p) if p.strip() else None for p in parts]
    return range(n)[slice(*vals)]


def read_extxyz(path, index=':'):
    """Frames of an extended-XYZ file as the dicts ``read_outcar`` returns.  Comment-line keys (case-insensitive): ``Lattice``
    (9 numbers, rows = lattice vectors), ``Properties`` (species:S:1:pos:R:3[:forces:R:3] ...), ``energy`` / ``free_energy``
    (the reference prefers the force-consistent free energy, dataload.py:94-97), ``stress`` (9 numbers, ASE's sign = minus
    VASP's; the label keeps VASP's sign and the order xx yy zz xy yz zx as dataload.py:99-106 ends up with), ``pbc``."""
    with open(path, 'r') as f:
        lines = f.readlines()
    frames, i = [], 0
    while i < len(lines):
        if not lines[i].strip():
            i += 1
            continue
        n = int(lines[i].split()[0])
        info = {k.lower(): v.strip('"') for k, v in _KV.findall(lines[i + 1])}
        props = info.get('properties', 'species:S:1:pos:R:3').split(':')
        cols, c = {}, 0
        for name, kind, cnt in zip(props[0::3], props[1::3], props[2::3]):
            cols[name.lower()] = (c, int(cnt), kind)
            c += int(cnt)
        rows = [ln.split() for ln in lines[i + 2:i + 2 + n]]
        sp = cols.get('species', cols.get('z'))
        if sp is None or 'pos' not in cols:
            raise ValueError(f'{path}: Properties without species / pos')
        if sp[2] == 'S':
            numbers = np.array([ATOMIC_NUMBER[r[sp[0]]] for r in rows])
        else:
            numbers = np.array([int(r[sp[0]]) for r in rows])
        take = lambda key: np.array([[float(v) for v in r[cols[key][0]:cols[key][0] + cols[key][1]]] for r in rows])  # noqa: E731
        fkey = 'forces' if 'forces' in cols else ('force' if 'force' in cols else None)
        cell = np.array([float(v) for v in info['lattice'].split()]).reshape(3, 3) if 'lattice' in info else np.zeros((3, 3))
        pbc = ([v.upper().startswith('T') for v in info['pbc'].split()] if 'pbc' in info else [bool(np.any(cell))] * 3)
        energy = float(info['free_energy']) if 'free_energy' in info else (float(info['energy']) if 'energy' in info else float('nan'))
        if 'stress' in info:
            s = -np.array([float(v) for v in info['stress'].split()]).reshape(3, 3)
            stress = np.array([s[0, 0], s[1, 1], s[2, 2], s[0, 1], s[1, 2], s[2, 0]])
        else:
            stress = np.full(6, np.nan)
        frames.append(dict(numbers=numbers, pos=take('pos'), cell=cell, energy=energy,
                           forces=take(fkey) if fkey else np.full((n, 3), np.nan), stress=stress, pbc=pbc,
                           info={'data_from': 'extxyz', 'file': os.path.abspath(path)}))
        i += 2 + n
    return [frames[k] for k in _parse_index(index, len(frames))]


def write_extxyz(path, structures):
    """The inverse of ``read_extxyz`` (ASE conventions: stress = minus VASP's, full 3x3)."""
    with open(path, 'w') as f:
        for st in structures:
            n = len(st['numbers'])
            lat = ' '.join(repr(float(v)) for v in np.asarray(st['cell']).reshape(-1))
            head = f'Lattice="{lat}" Properties=species:S:1:pos:R:3:forces:R:3 energy={float(st["energy"])!r} '
            if not np.isnan(np.asarray(st['stress'], float)).any():
                xx, yy, zz, xy, yz, zx = (-float(v) for v in st['stress'])
                head += 'stress="' + ' '.join(repr(v) for v in (xx, xy, zx, xy, yy, yz, zx, yz, zz)) + '" '
            head += 'pbc="' + ' '.join('T' if b else 'F' for b in st['pbc']) + '"'
            f.write(f'{n}\n{head}\n')
            for z, p, fr in zip(st['numbers'], st['pos'], st['forces']):
                f.write(f'{CHEMICAL_SYMBOLS[int(z)]} ' + ' '.join(repr(float(v)) for v in list(p) + list(fr)) + '\n')


# ---- checkpoints (trainer.py:98-107 get_checkpoint_dict, scripts/processing_epoch.py:34-40, util.py:186-231) -----------------
def config_from_reference(config):
    """The reference's checkpoint ``config`` (keys of sevenn/_keys.py) -> the keyword dict ``nn.build_E3_equivariant_model``
    reads.  Raises for model options the CUDA path does not implement instead of building something else silently."""
    c = dict(config)
    rb, cf = dict(c.get('radial_basis', {})), dict(c.get('cutoff_function', {}))
    if rb.get('radial_basis_name', 'bessel') != 'bessel':
        raise ValueError('B200 path: only the Bessel radial basis')
    if c.get('is_parity', False):
        raise ValueError('B200 path: is_parity=True (odd irreps) is not implemented')
    if c.get('interaction_type', 'nequip') != 'nequip':
        raise ValueError("B200 path: interaction_type 'nequip' only")
    name = cf.get('cutoff_function_name', 'poly_cut')
    out = dict(cutoff=float(c['cutoff']), num_species=int(c.get('_number_of_species', len(c.get('chemical_species', [])))),
               lmax=int(c['lmax']), num_convolution_layer=int(c['num_convolution_layer']),
               bessel_basis_num=int(rb.get('bessel_basis_num', 8)),
               cutoff_function=({'name': 'XPLOR', 'cutoff_on': float(cf['cutoff_on'])} if name.upper() == 'XPLOR'
                                else {'name': 'poly_cut', 'poly_cut_p_value': cf.get('poly_cut_p_value', 6)}),
               weight_nn_hidden_neurons=list(c.get('weight_nn_hidden_neurons', [64, 64])),
               self_connection_type=c.get('self_connection_type', 'nequip'),
               use_species_wise_shift_scale=bool(c.get('use_species_wise_shift_scale', False)),
               train_shift_scale=bool(c.get('train_shift_scale', False)), train_denominator=bool(c.get('train_denominator', False)),
               use_bias_in_linear=bool(c.get('use_bias_in_linear', False)), normalize_sph=bool(c.get('_normalize_sph', True)))
    tolist = lambda v: v.tolist() if hasattr(v, 'tolist') else v   # noqa: E731
    for k in ('conv_denominator', 'shift', 'scale'):
        if k in c:
            out[k] = tolist(c[k])
    if c.get('irreps_manual'):
        out['irreps_manual'] = list(c['irreps_manual'])
    else:   # model_build.py:300-334 with is_parity False: channel x (0e .. lmax e), scalars only after the last layer
        ch, lm, nl = int(c['channel']), int(c['lmax']), int(c['num_convolution_layer'])
        mid = '+'.join(f'{ch}x{l}e' for l in range(lm + 1))
        out['irreps_manual'] = [f'{ch}x0e'] + [mid] * (nl - 1) + [f'{ch}x0e']
    return out


def checkpoint_dict(model, optimizer=None, scheduler=None, config=None, epoch=None):
    """What the reference writes to checkpoint_{epoch}.pth: model / optimizer / scheduler state dicts + config + epoch."""
    m = model.module if hasattr(model, 'module') else model
    out = {'model_state_dict': m.state_dict()}
    if optimizer is not None:
        out['optimizer_state_dict'] = optimizer.state_dict()
    if scheduler is not None:
        out['scheduler_state_dict'] = scheduler.state_dict()
    if config is not None:
        out['config'] = config
    if epoch is not None:
        out['epoch'] = epoch
    return out


def model_from_checkpoint(checkpoint, device='cuda'):
    """``sevenn.util.model_from_checkpoint``: a checkpoint file or dict -> (model built from the B200 twins, config).  Every
    key of ``model_state_dict`` must have a twin parameter (strict)."""
    from . import nn as b200nn
    if isinstance(checkpoint, str):
        checkpoint = torch.load(checkpoint, map_location='cpu', weights_only=False)
    if not isinstance(checkpoint, dict) or 'model_state_dict' not in checkpoint or 'config' not in checkpoint:
        raise ValueError('checkpoint must be a path or a dict with model_state_dict and config')
    model = b200nn.build_E3_equivariant_model(config_from_reference(checkpoint['config']))
    missing = model.load_state_dict(checkpoint['model_state_dict'], strict=True)
    assert not missing.missing_keys and not missing.unexpected_keys
    return model.to(device), checkpoint['config']
