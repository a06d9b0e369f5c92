"""Host side of the fine-tuning loop over the B200 modules: the reference's loss definitions and its
rehearsal + EWC step, kept as PyTorch (BASELINE north_star: "fine-tuning ... is otherwise left as PyTorch").

What is mirrored (names, argument meaning, arithmetic):

* ``LossDefinition`` / ``PerAtomEnergyLoss`` / ``ForceLoss`` / ``StressLoss`` / ``EWCLoss`` and
  ``get_loss_functions`` - ``sevenn/train/loss.py:18-93,96-132,135-175,178-214,217-265,268-307``;
* ``RehearsalTrainer.run_one_step`` - the body of ``run_one_epoch_rehearsal`` (``sevenn/train/trainer.py:173-210``):
  ``zero_grad`` -> train batch forward -> loss -> ``backward`` -> ``optimizer.step`` -> replay batch forward -> loss ->
  ``backward`` (no ``zero_grad`` in between: the replay gradients are added to the train-batch gradients, as in the
  reference) -> ``optimizer.step``;
* ``compute_fisher_matrix`` - ``sevenn/train/trainer.py:126-152``;
* ``collate`` - what torch_geometric's ``Batch.from_data_list`` does to ``AtomGraphData`` (``sevenn/train/collate.py``):
  node tensors concatenated, ``edge_index`` offset by the node count, per-structure ``cell_volume`` / ``num_atoms`` /
  labels stacked, ``batch`` = structure id of every atom;
* data parallelism (config 5): ``torch.nn.parallel.DistributedDataParallel`` over NCCL, one process per GPU
  (``sevenn/train/trainer.py:19-24``); ``shard_indices`` reproduces ``DistributedSampler``'s split
  (``sevenn/scripts/train.py:24-38``).

The model's forward / backward / double-backward arithmetic is in libsevennet_b200.so (``nn.py``, ``higher_order.py``).
"""
import math

import numpy as np
import torch

TO_KB = 1602.1766208   # eV/A^3 -> kbar (loss.py:199)


# ---- loss definitions (sevenn/train/loss.py) ---------------------------------------------------------------------------
class LossDefinition:
    """loss.py:10-93.  ``use_weight``: per-structure data weights (``load_dataset_with_weights``, dataset.py:430-450):
    ``batch_data['data_weight'][weight_key]`` holds one weight per structure, the criterion must have reduction 'none' and the
    loss is mean(criterion * weight).  (The reference's own weighted branch trips over ``assert isinstance(weights, list)``
    at loss.py:88 whenever unlabelled entries are filtered, i.e. always with its defaults; this is the computation it
    describes, with the weights filtered like the values.)"""

    def __init__(self, name=None, unit=None, criterion=None, ref_key=None, pred_key=None, vdim=1, use_weight=False,
                 weight_key=None, delete_unlabled=True):
        if criterion is not None and hasattr(criterion, 'reduction'):
            if use_weight:
                assert criterion.reduction == 'none' and weight_key is not None
            else:
                assert criterion.reduction != 'none'
        self.name, self.unit, self.criterion, self.vdim = name, unit, criterion, vdim
        self.ref_key, self.pred_key, self.use_weight, self.delete_unlabeled = ref_key, pred_key, use_weight, delete_unlabled
        self.weight_key = weight_key

    def _preprocess(self, batch_data, model=None):
        return batch_data[self.pred_key].reshape(-1), batch_data[self.ref_key].reshape(-1)

    def _get_data_weight(self, batch_data):   # loss.py:60-63: one weight per structure, repeated over the components
        return torch.repeat_interleave(batch_data['data_weight'][self.weight_key], self.vdim)

    def get_loss(self, batch_data, model=None):
        pred, ref = self._preprocess(batch_data, model)
        weights = self._get_data_weight(batch_data) if self.use_weight else None
        # a loader that saw the labels on the host can vouch that none is missing (collate): boolean-mask indexing has a
        # data-dependent size, i.e. a device->host synchronisation per loss term and another in its reverse pass
        if self.delete_unlabeled and self.ref_key not in batch_data.get('_b200_all_labeled', ()):
            unl = torch.isnan(ref)
            pred, ref = pred[~unl], ref[~unl]
            if len(pred) == 0:
                return torch.zeros(1, device=pred.device)
            if weights is not None:
                weights = weights[~unl]
        if weights is not None:
            return torch.mean(self.criterion(pred, ref) * weights)
        return self.criterion(pred, ref)


class PerAtomEnergyLoss(LossDefinition):
    def __init__(self, criterion=None, **kw):
        super().__init__('Energy', 'eV/atom', criterion, 'total_energy', 'inferred_total_energy', vdim=1,
                         weight_key='per_atom_energy', **kw)

    def _preprocess(self, batch_data, model=None):
        n = batch_data['num_atoms']
        return batch_data[self.pred_key] / n, batch_data[self.ref_key] / n


class ForceLoss(LossDefinition):
    def __init__(self, criterion=None, **kw):
        super().__init__('Force', 'eV/A', criterion, 'force_of_atoms', 'inferred_force', vdim=3, weight_key='force_of_atoms', **kw)

    def _get_data_weight(self, batch_data):   # loss.py:172-175: the structure's weight for each of its atoms
        w = batch_data['data_weight'][self.weight_key][batch_data['batch']]
        return torch.repeat_interleave(w, self.vdim)


class StressLoss(LossDefinition):
    def __init__(self, criterion=None, **kw):
        super().__init__('Stress', 'kbar', criterion, 'stress', 'inferred_stress', vdim=6, weight_key='stress', **kw)

    def _preprocess(self, batch_data, model=None):
        return (batch_data[self.pred_key] * TO_KB).reshape(-1), (batch_data[self.ref_key] * TO_KB).reshape(-1)


class EWCLoss(LossDefinition):
    """sum_k F_k (theta_k - theta*_k)^2 over the parameters present in both dictionaries (loss.py:250-265)."""

    def __init__(self, fisher_dict, opt_params_dict, device=None):
        super().__init__('EWC')
        self.fisher_dict, self.opt_params_dict, self.device = dict(fisher_dict), dict(opt_params_dict), device
        if device is not None:
            self.to(device)

    def to(self, device):
        self.fisher_dict = {k: v.to(device) for k, v in self.fisher_dict.items()}
        self.opt_params_dict = {k: v.to(device) for k, v in self.opt_params_dict.items()}
        self.device = device
        self._flat_key = None

    def get_loss(self, batch_data, model=None):
        if model is None:
            raise ValueError('EWC requires model to compute loss')
        # same sum as the reference's per-parameter loop (loss.py:259-264), evaluated on one flattened vector: a few
        # kernels per step instead of several per parameter tensor
        names, params = [], []
        for name, p in model.named_parameters():
            name = name[7:] if name.startswith('module.') else name   # DDP prefix
            if name in self.fisher_dict and name in self.opt_params_dict:
                names.append(name)
                params.append(p)
        if not params:
            return torch.zeros(1, device=self.device)
        key = tuple(names)
        if getattr(self, '_flat_key', None) != key:
            self._flat_key = key
            self._flat_fisher = torch.cat([self.fisher_dict[n].reshape(-1) for n in names])
            self._flat_opt = torch.cat([self.opt_params_dict[n].reshape(-1) for n in names])
        d = torch.cat([p.reshape(-1) for p in params]) - self._flat_opt
        return torch.sum(self._flat_fisher * d * d).reshape(1)


def get_loss_functions(loss='huber', loss_param=None, force_loss_weight=1.0, is_train_stress=True, stress_loss_weight=0.01,
                       fisher_dict=None, opt_params_dict=None, ewc_lambda=0.0, device=None, use_weight=False):
    """[(definition, weight)] as ``get_loss_functions_from_config`` builds it (loss.py:268-307); EWC enters as lambda/2;
    ``use_weight`` = the config carries ``load_dataset_with_weights`` (reduction 'none' + per-structure weights)."""
    crit = {'mse': torch.nn.MSELoss, 'huber': torch.nn.HuberLoss, 'l1': torch.nn.L1Loss}[loss.lower()](
        reduction='none' if use_weight else 'mean', **(loss_param or {}))
    common = dict(criterion=crit, use_weight=use_weight)
    out = [(PerAtomEnergyLoss(**common), 1.0), (ForceLoss(**common), float(force_loss_weight))]
    if is_train_stress:
        out.append((StressLoss(**common), float(stress_loss_weight)))
    if fisher_dict is not None and opt_params_dict is not None:
        out.append((EWCLoss(fisher_dict, opt_params_dict, device=device), float(ewc_lambda) / 2.0))
    return out


def total_loss(loss_functions, output, model, device):
    tot = torch.zeros(1, device=device)
    for ld, w in loss_functions:
        tot = tot + ld.get_loss(output, model) * w
    return tot


# ---- batching ---------------------------------------------------------------------------------------------------------------
def collate(graphs, device=None, pin=False):
    """List of per-structure dicts (x, edge_index[2,E], edge_vec[E,3], cell_volume, num_atoms and optionally
    total_energy, force_of_atoms, stress) -> one batched dict with ``batch``; host tensors unless ``device``."""
    off, xs, eis, vecs, batch = 0, [], [], [], []
    for b, g in enumerate(graphs):
        n = int(g['x'].shape[0])
        xs.append(torch.as_tensor(g['x'], dtype=torch.long))
        eis.append(torch.as_tensor(g['edge_index'], dtype=torch.long) + off)
        vecs.append(torch.as_tensor(g['edge_vec'], dtype=torch.float32))
        batch.append(torch.full((n,), b, dtype=torch.long))
        off += n
    out = {'x': torch.cat(xs), 'edge_index': torch.cat(eis, 1), 'edge_vec': torch.cat(vecs), 'batch': torch.cat(batch),
           'num_atoms': torch.tensor([int(g['x'].shape[0]) for g in graphs], dtype=torch.long),
           'cell_volume': torch.tensor([float(g['cell_volume']) for g in graphs], dtype=torch.float32)}
    if 'total_energy' in graphs[0]:
        out['total_energy'] = torch.tensor([float(g['total_energy']) for g in graphs], dtype=torch.float32)
        out['force_of_atoms'] = torch.cat([torch.as_tensor(g['force_of_atoms'], dtype=torch.float32) for g in graphs])
        out['stress'] = torch.stack([torch.as_tensor(g['stress'], dtype=torch.float32).reshape(6) for g in graphs])
    if 'data_weight' in graphs[0]:   # dataset.py:430-450: {'per_atom_energy': w, 'force_of_atoms': w, 'stress': w} per structure
        out['data_weight'] = {k: torch.tensor([float(g['data_weight'][k]) for g in graphs], dtype=torch.float32)
                              for k in graphs[0]['data_weight']}
    labeled = tuple(k for k in ('total_energy', 'force_of_atoms', 'stress') if k in out and not bool(torch.isnan(out[k]).any()))
    if pin:
        out = {k: ({kk: vv.pin_memory() for kk, vv in v.items()} if isinstance(v, dict) else v.pin_memory()) for k, v in out.items()}
    # edges of every structure sorted by their centre atom (what ase / the CUDA cell list emit) stay sorted under the
    # node offsets: let the convolution skip its sortedness check (nn._graph_of)
    if all(bool((torch.as_tensor(g['edge_index'])[0][1:] >= torch.as_tensor(g['edge_index'])[0][:-1]).all()) for g in graphs):
        out['_b200_dst_sorted'] = True
    out['_b200_all_labeled'] = labeled   # keys whose labels hold no NaN: the losses skip their unlabelled-entry filter
    if device is not None:
        out = to_device(out, device)
    return out


def to_device(batch, device):
    """``batch.to(device, non_blocking=True)`` (trainer.py:176) + the requires_grad the dataset sets on edge_vec
    (``sevenn/train/dataload.py`` via ``AtomGraphData``; ``util.py:243``)."""
    out = {k: (v.to(device, non_blocking=True) if torch.is_tensor(v) else
               {kk: vv.to(device, non_blocking=True) for kk, vv in v.items()} if isinstance(v, dict) else v) for k, v in batch.items()}
    out['edge_vec'].requires_grad_(True)
    return out


def batch_bytes(batch):
    return int(sum(v.numel() * v.element_size() for v in batch.values() if torch.is_tensor(v)))


def shard_indices(n_items, rank, world_size, seed=0, epoch=0, shuffle=True):
    """Indices ``DistributedSampler`` gives rank ``rank`` (padding by wrap-around so every rank gets the same count)."""
    if shuffle:
        g = torch.Generator()
        g.manual_seed(seed + epoch)
        idx = torch.randperm(n_items, generator=g).tolist()
    else:
        idx = list(range(n_items))
    per = math.ceil(n_items / world_size)
    total = per * world_size
    idx += idx[:total - len(idx)]
    return idx[rank:total:world_size]


# ---- trainer --------------------------------------------------------------------------------------------------------------
class RehearsalTrainer:
    """``sevenn/train/trainer.py:15-42,155-222`` over a model built from the ``nn.py`` twins."""

    def __init__(self, model, loss_functions, optimizer='adam', optim_param=None, device='cuda:0', distributed=False,
                 local_rank=0):
        self.distributed = distributed
        if distributed:
            from torch.nn.parallel import DistributedDataParallel as DDP
            cuda = torch.device(device).type == 'cuda'
            device = torch.device('cuda', local_rank) if cuda else torch.device('cpu')   # cpu: gloo tests of the host logic
            torch.distributed.barrier()
            self.model = DDP(model.to(device), device_ids=[device] if cuda else None)
            if hasattr(self.model.module, 'set_is_batch_data'):
                self.model.module.set_is_batch_data(True)
        else:
            self.model = model.to(device)
            if hasattr(self.model, 'set_is_batch_data'):
                self.model.set_is_batch_data(True)
        self.device = torch.device(device)
        params = [p for p in self.model.parameters() if p.requires_grad]
        opt = {'adam': torch.optim.Adam, 'adamw': torch.optim.AdamW, 'sgd': torch.optim.SGD}[optimizer.lower()]
        self.optimizer = opt(params, **(optim_param or {'lr': 1e-4}))
        self.loss_functions = loss_functions
        self._params = params
        for ld, _ in loss_functions:
            if isinstance(ld, EWCLoss):
                ld.to(self.device)

    def _fwd_loss(self, batch):
        output = self.model(batch)
        return output, total_loss(self.loss_functions, output, self.model, self.device)

    def _backward(self, loss):
        # the reference calls loss.backward() (trainer.py:185,209), which also fills edge_vec.grad (a leaf that requires
        # grad) and never reads it; restricting the reverse pass to the parameters gives identical parameter gradients
        # without that branch.  Under DDP the plain call is kept (its reducer hooks expect it).
        if self.distributed:
            loss.backward()
        else:
            loss.backward(inputs=self._params)

    def run_one_step(self, batch, batch_mem=None):
        """One iteration of the rehearsal loop (trainer.py:173-210).  ``batch`` / ``batch_mem`` are device dicts
        (``to_device``).  Returns (train loss, replay loss) as device tensors."""
        self.model.train()
        self.optimizer.zero_grad()
        _, loss = self._fwd_loss(batch)
        self._backward(loss)
        self.optimizer.step()
        mem_loss = None
        if batch_mem is not None:
            _, mem_loss = self._fwd_loss(batch_mem)
            self._backward(mem_loss)
            self.optimizer.step()
        return loss.detach(), (None if mem_loss is None else mem_loss.detach())

    def compute_fisher_matrix(self, batches, loss_thr=-1.0):
        """Fisher[name] = mean over the kept batches of grad^2; theta* = current parameters (trainer.py:126-152)."""
        # names without DDP's 'module.' prefix: EWCLoss strips it before looking a name up, so a Fisher dict computed under
        # DDP must not carry it (with the prefix no name matches and the penalty is silently zero)
        def bare(n):
            return n[7:] if n.startswith('module.') else n
        fisher = {bare(n): torch.zeros_like(p) for n, p in self.model.named_parameters()}
        self.model.train()
        cnt = 0
        for batch in batches:
            self.model.zero_grad()
            _, loss = self._fwd_loss(batch)
            if loss_thr < 0 or float(loss) < loss_thr:
                loss.backward()
                for n, p in self.model.named_parameters():
                    if p.grad is not None:
                        fisher[bare(n)] += p.grad.detach().clone() ** 2
                cnt += 1
        for n in fisher:
            fisher[n] /= max(cnt, 1)
        opt = {bare(k): v.data.detach().clone() for k, v in self.model.named_parameters()}
        return fisher, opt, cnt


# ---- synthetic labelled structures -------------------------------------------------------------------
def synthetic_graphs(structures, cutoff, device, label_model=None, noise=0.0, seed=0):
    """Radius graphs (the CUDA cell-list kernel) of synthetic structures as host dicts for ``collate``; labels come from
    ``label_model`` (a ``B200SevenNet`` engine evaluated on the same structure) plus Gaussian noise."""
    from . import ops
    from .synthetic import SPECIES_INDEX
    rng = np.random.default_rng(seed)
    out = []
    for pos, cell, pbc, sym in structures:
        types = np.array([SPECIES_INDEX[s] for s in sym], dtype=np.int64)
        g = ops.radius_graph(torch.tensor(pos, dtype=torch.float64, device=device), cell, pbc, cutoff)
        d = {'x': torch.tensor(types), 'edge_index': torch.stack([g['dst'], g['src']]).long().cpu(),
             'edge_vec': g['edge_vec'].cpu(), 'cell_volume': abs(float(np.linalg.det(cell))), 'num_atoms': len(types)}
        if label_model is not None:
            o = label_model.calculate(torch.tensor(pos, dtype=torch.float64, device=device), cell, pbc,
                                      torch.tensor(types, dtype=torch.int32, device=device))
            n = len(types)
            d['total_energy'] = float(o['energy'].item()) + noise * n * float(rng.normal())
            d['force_of_atoms'] = o['forces'].cpu() + torch.tensor(rng.normal(0, noise * 10, (n, 3)), dtype=torch.float32)
            d['stress'] = o['stress'].cpu().reshape(6) + torch.tensor(rng.normal(0, noise, 6), dtype=torch.float32)
        out.append(d)
    return out
