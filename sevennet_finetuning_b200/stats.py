"""Statistics either side of the fine-tuning loop, as device-side reductions.

* ``DatasetStatistics``: what the reference computes once at the start of a (continued) training to initialise
  ``conv_denominator``, ``shift`` and ``scale`` - ``AtomGraphDataset.get_avg_num_neigh``, ``get_force_rms``,
  ``get_species_wise_force_rms``, ``get_per_atom_energy_mean``, ``get_species_ref_energy_by_linear_comb``,
  ``get_statistics`` (``sevenn/train/dataset.py:247-356``; consumed by ``scripts/processing_dataset.py:38-143``).  The
  reference walks Python lists of per-structure tensors (``.tolist()`` of every force component, ``np.unique`` per
  structure); here the structures are collated once (``finetune.collate``) and every statistic is a bincount / scatter /
  small normal-equation solve on the device, with one host read per query.
* ``ErrorRecorder``: the per-epoch error metrics of ``sevenn/error_recorder.py:66-300`` (RMSE, ComponentRMSE, MAE,
  VectorMAE over Energy / Force / Stress, averaged over all values seen, unlabelled (NaN) references dropped).  The
  reference does ``values.sum().item()`` per metric and batch (a device sync each) and two all_reduce per metric at the
  epoch's end (``:70-76``); here the running sums stay on the device and an epoch ends with ONE all_reduce of one packed
  vector and one host read.
"""
import math

import torch
import torch.distributed as dist

from .finetune import TO_KB, collate

TO_GPA = 160.21766208   # error_recorder.py:44


class DatasetStatistics:
    def __init__(self, graphs, num_species, device=None):
        b = collate(graphs, device=None)
        dev = torch.device(device) if device is not None else torch.device('cpu')
        self.device, self.num_species = dev, int(num_species)
        self.x = b['x'].to(dev)
        self.batch = b['batch'].to(dev)
        self.centre = b['edge_index'][0].to(dev)
        self.num_atoms = b['num_atoms'].to(dev)
        self.n_struct = int(self.num_atoms.numel())
        self.energy = b['total_energy'].to(dev).double() if 'total_energy' in b else None
        self.force = b['force_of_atoms'].to(dev) if 'force_of_atoms' in b else None
        self.stress = b['stress'].to(dev) if 'stress' in b else None

    # dataset.py:344-354: np.unique(edge_index[0], return_counts=True) per structure, averaged over the atoms THAT HAVE
    # neighbours (np.unique never lists an atom with no edge)
    def get_avg_num_neigh(self):
        deg = torch.bincount(self.centre, minlength=int(self.x.numel()))
        has = deg > 0
        return float(deg[has].double().mean()) if bool(has.any()) else float('nan')

    def get_force_rms(self):                                    # dataset.py:311-322
        return float(self.force.double().square().mean().sqrt())

    def get_species_wise_force_rms(self):                       # dataset.py:324-342: per species, mean over atoms and x, y, z
        sq = torch.zeros(self.num_species, 3, dtype=torch.float64, device=self.device)
        sq.index_add_(0, self.x, self.force.double().square())
        cnt = torch.bincount(self.x, minlength=self.num_species).clamp(min=1).double()
        return (sq / cnt[:, None]).mean(1).sqrt().float()       # species absent from the data: 0 (scatter_reduce leaves zeros)

    def get_per_atom_mean(self, key='total_energy'):            # dataset.py:262-270
        v = {'total_energy': self.energy}[key]
        return float((v / self.num_atoms.double()).mean())

    def get_per_atom_energy_mean(self):                         # dataset.py:272-276
        return self.get_per_atom_mean('total_energy')

    def composition(self):
        """[n_structures, n_species] atom counts (the ``c`` matrix of dataset.py:291-295)."""
        c = torch.zeros(self.n_struct, self.num_species, dtype=torch.float64, device=self.device)
        c.index_put_((self.batch, self.x), torch.ones_like(self.x, dtype=torch.float64), accumulate=True)
        return c

    def get_species_ref_energy_by_linear_comb(self, alpha=0.1):
        """dataset.py:278-309: ``Ridge(alpha=0.1, fit_intercept=False)`` of the total energies on the compositions, species
        that never occur left at 0.  Ridge without intercept is the normal-equation solve (C^T C + alpha I) w = C^T y."""
        c = self.composition()
        present = c.sum(0) > 0
        cr = c[:, present]
        A = cr.T @ cr + alpha * torch.eye(cr.shape[1], dtype=torch.float64, device=self.device)
        w = torch.linalg.solve(A, cr.T @ self.energy)
        full = torch.zeros(self.num_species, dtype=torch.float64, device=self.device)
        full[present] = w
        return full.cpu().numpy()

    def get_statistics(self, key):
        """dataset.py:357-395 ('_total' entry): mean / std (unbiased, torch.std) / median / max / min of the flattened
        values with unlabelled (NaN) entries dropped (energy per structure as stored, i.e. total)."""
        v = {'total_energy': self.energy, 'force_of_atoms': self.force, 'stress': self.stress}[key].reshape(-1).double()
        v = v[~torch.isnan(v)]
        if v.numel() == 0:
            return {k: float('nan') for k in ('mean', 'std', 'median', 'max', 'min')}
        return {'mean': float(v.mean()), 'std': float(v.std()), 'median': float(v.median()),
                'max': float(v.max()), 'min': float(v.min())}

    def shift_scale(self, shift='per_atom_energy_mean', scale='force_rms'):
        """The initial ``shift`` / ``scale`` of ``scripts/processing_dataset.py:96-143`` for the two presets the shipped
        fine-tuning inputs use (per-atom energy mean or element-wise reference energies; global or element-wise force RMS)."""
        sh = (self.get_species_ref_energy_by_linear_comb().tolist() if shift == 'elemwise_reference_energies'
              else self.get_per_atom_energy_mean())
        sc = (self.get_species_wise_force_rms().tolist() if scale == 'elemwise_force_rms' else self.get_force_rms())
        return sh, sc


# ---- error metrics ------------------------------------------------------------------------------------------------------
_ERROR_TYPES = {   # error_recorder.py:11-63
    'TotalEnergy': dict(ref='total_energy', pred='inferred_total_energy', coeff=1.0, vdim=1, unit='eV'),
    'Energy': dict(ref='total_energy', pred='inferred_total_energy', coeff=1.0, vdim=1, unit='eV/atom', per_atom=True),
    'Force': dict(ref='force_of_atoms', pred='inferred_force', coeff=1.0, vdim=3, unit='eV/A'),
    'Stress': dict(ref='stress', pred='inferred_stress', coeff=TO_KB, vdim=6, unit='kbar'),
    'Stress_GPa': dict(ref='stress', pred='inferred_stress', coeff=TO_GPA, vdim=6, unit='GPa'),
}


def _errors(kind, y_ref, y_pred, vdim):
    """per-value error terms of one metric (error_recorder.py:146-232)"""
    if kind == 'RMSE':             # vector squared error, root at the end
        return (y_ref - y_pred).square().reshape(-1, vdim).sum(1)
    if kind == 'ComponentRMSE':
        return (y_ref - y_pred).square().reshape(-1)
    if kind == 'MAE':
        return (y_ref - y_pred).abs().reshape(-1)
    if kind == 'VectorMAE':
        return (y_ref - y_pred).square().reshape(-1, vdim).sum(1).sqrt()
    raise ValueError(f'unknown metric {kind}')


class ErrorRecorder:
    """``ErrorRecorder.from_config``-style list of (error type, metric) pairs, e.g. [('Energy', 'RMSE'), ('Force', 'RMSE'),
    ('Stress', 'RMSE')] (``_const.py`` DEFAULT error_record).  ``update(output)`` takes the model's output dict of a batch
    (predictions and labels side by side, as ``trainer.py:66-71`` passes it); ``epoch_forward()`` returns
    {'Energy_RMSE (eV/atom)': value, ...}, reduced over the ranks when torch.distributed is initialised, and resets."""

    def __init__(self, metrics=(('Energy', 'RMSE'), ('Force', 'RMSE'), ('Stress', 'RMSE')), device='cpu'):
        self.metrics = [(t, k) for t, k in metrics]
        for t, k in self.metrics:
            if t not in _ERROR_TYPES:
                raise ValueError(f'unknown error type {t}')
            _errors(k, torch.zeros(6), torch.zeros(6), 1)
        self.device = torch.device(device)
        self.reset()

    def reset(self):
        self._acc = torch.zeros(len(self.metrics), 2, dtype=torch.float64, device=self.device)   # [sum, count] per metric

    def update(self, output):
        row = []
        for t, k in self.metrics:
            d = _ERROR_TYPES[t]
            y_ref = output[d['ref']].detach().to(self.device, torch.float64) * d['coeff']
            y_pred = output[d['pred']].detach().to(self.device, torch.float64).reshape(y_ref.shape) * d['coeff']
            if d.get('per_atom'):
                n = output['num_atoms'].to(self.device, torch.float64).reshape(y_ref.shape)
                y_ref, y_pred = y_ref / n, y_pred / n
            e = _errors(k, y_ref, y_pred, d['vdim'])
            ok = ~torch.isnan(e)                                  # unlabelled references are NaN (error_recorder.py:118-121)
            row.append(torch.stack([torch.where(ok, e, torch.zeros_like(e)).sum(), ok.sum().double()]))
        self._acc += torch.stack(row)

    def key_str(self, t, k):
        return f'{t}_{k} ({_ERROR_TYPES[t]["unit"]})'

    def epoch_forward(self):
        acc = self._acc.clone()
        if dist.is_available() and dist.is_initialized():
            dist.all_reduce(acc, op=dist.ReduceOp.SUM)           # one collective for every metric (reference: two per metric)
        vals = acc.cpu()
        out = {}
        for i, (t, k) in enumerate(self.metrics):
            s, c = float(vals[i, 0]), float(vals[i, 1])
            v = float('nan') if c == 0 else s / c
            out[self.key_str(t, k)] = math.sqrt(v) if k in ('RMSE', 'ComponentRMSE') and c > 0 else v
        self.reset()
        return out
