"""Whole-model energy/forces/stress through the C-ABI engine (b200_model_*).

This is the call ``SevenNetCalculator.calculate`` (sevenn/sevennet_calculator.py:119-157) and
``PairE3GNN::compute`` (sevenn/pair_e3gnn/pair_e3gnn.cpp:205-266) make into the model; the
reverse pass is hand-written kernels (ForceStressOutputFromEdge semantics,
sevenn/nn/force_output.py:158-215), so no autograd graph is built.
"""
import ctypes as C

import numpy as np
import torch

from . import ops
from ._lib import B200_CUTOFF_POLY, B200_CUTOFF_XPLOR, HALO_FN, EdgeParams, ModelConfig, check, lib

SEVENNET0_CONFIG = dict(
    cutoff=5.0, num_species=89, lmax=2, num_convolution_layer=5,
    irreps_manual=['128x0e', '128x0e+64x1e+32x2e', '128x0e+64x1e+32x2e', '128x0e+64x1e+32x2e',
                   '128x0e+64x1e+32x2e', '128x0e'],
    bessel_basis_num=8, cutoff_function=dict(name='XPLOR', cutoff_on=4.5),
    weight_nn_hidden_neurons=[64, 64], use_species_wise_shift_scale=True, normalize_sph=True,
)


def edge_params_from_config(cfg):
    cf = cfg['cutoff_function']
    xpl = cf['name'].upper() == 'XPLOR'
    return EdgeParams(int(cfg['bessel_basis_num']), B200_CUTOFF_XPLOR if xpl else B200_CUTOFF_POLY,
                      int(bool(cfg.get('normalize_sph', True))), float(cfg['cutoff']),
                      float(cf.get('cutoff_on', 0.0)), float(cf.get('poly_cut_p_value', 6)))


def _wrap_rows(ptr, n_rows, width, dev):
    """Zero-copy fp32 [n_rows, width] view of engine workspace memory handed to the halo hook."""
    class _Arr:
        __cuda_array_interface__ = dict(shape=(n_rows, width), typestr='<f4', data=(int(ptr), False), version=2)
    return torch.as_tensor(_Arr(), device=dev)


class B200SevenNet:
    """Device-resident SevenNet-style model (nequip interaction, linear self connection)."""

    def __init__(self, config=None, state_dict=None, device='cuda:0'):
        cfg = dict(SEVENNET0_CONFIG)
        cfg.update(config or {})
        self.config = cfg
        if cfg.get('self_connection_type', 'linear') != 'linear':
            raise RuntimeError("B200SevenNet: the fused engine implements the 'linear' self connection (SevenNet-0); "
                               "'nequip' models run through the module API (nn.build_E3_equivariant_model)")
        self.device = torch.device(device)
        mc = ModelConfig()
        nl = cfg['num_convolution_layer']
        mc.n_layers, mc.n_species, mc.lmax = nl, cfg['num_species'], cfg['lmax']
        for t, ir in enumerate(cfg['irreps_manual']):
            for l, m in enumerate(ops.muls_of(ir)):
                mc.mul[t][l] = m
        mc.hidden[0], mc.hidden[1] = cfg['weight_nn_hidden_neurons']
        mc.readout_hidden = ops.muls_of(cfg['irreps_manual'][-1])[0] // 2
        mc.edge = edge_params_from_config(cfg)
        mc.species_wise_rescale = int(bool(cfg['use_species_wise_shift_scale']))
        self._cfg_struct = mc
        self._h = C.c_void_p()
        with torch.cuda.device(self.device):
            check(lib().b200_model_create(C.byref(mc), C.byref(self._h)), 'model_create')
        self._ws = None
        if state_dict is not None:
            self.load_state_dict(state_dict)

    def load_state_dict(self, sd):
        """sd: reference parameter names (named_parameters of the sevenn model) -> array/tensor."""
        with torch.cuda.device(self.device):
            for k, v in sd.items():
                a = np.ascontiguousarray(v.detach().cpu().numpy() if torch.is_tensor(v) else v, dtype=np.float32).reshape(-1)
                check(lib().b200_model_set_param(self._h, k.encode(), a.ctypes.data_as(C.c_void_p), a.size),
                      f'model_set_param({k})')

    def __del__(self):
        try:
            if getattr(self, '_h', None) is not None and self._h.value:
                lib().b200_model_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def workspace_bytes(self, n_atoms, n_edges, n_ghost=0):
        return lib().b200_model_workspace_bytes_halo(self._h, int(n_atoms), int(n_ghost), int(n_edges))

    def compute(self, types, rowptr, src, dst, edge_vec, n_ghost=0, halo=None):
        """types int32[N+n_ghost]; CSR sorted by dst: rowptr int32[N+n_ghost+1], src/dst int32[E]; edge_vec f32[E,3].
        halo(phase, layer, rows) is the exchange hook of the spatially decomposed call (parallel.py).
        Returns energy[1], atomic_energy[N], force[N+n_ghost,3], virial[6] (device tensors)."""
        NA, E = types.shape[0], src.shape[0]
        N = NA - int(n_ghost)
        need = self.workspace_bytes(N, E, n_ghost)
        if self._ws is None or self._ws.numel() < need:
            self._ws = None
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        dev = self.device
        energy = torch.empty(1, dtype=torch.float32, device=dev)
        ae = torch.empty(N, dtype=torch.float32, device=dev)
        force = torch.empty(NA, 3, dtype=torch.float32, device=dev)
        virial = torch.empty(6, dtype=torch.float32, device=dev)
        p = ops._p
        err = []
        cb = None
        if halo is not None:
            def _hook(user, phase, layer, rows_ptr, width):
                try:
                    rows = _wrap_rows(rows_ptr, NA, width, dev)
                    halo(phase, layer, rows)
                    return 0
                except Exception as ex:  # never unwind through the C frame
                    err.append(ex)
                    return 1
            cb = HALO_FN(_hook)
        rc = lib().b200_model_compute_halo(self._h, N, int(n_ghost), E, p(types), p(rowptr), p(src), p(dst), p(edge_vec),
                                           p(self._ws), self._ws.numel(), p(energy), p(ae), p(force), p(virial),
                                           C.cast(cb, C.c_void_p) if cb is not None else None, None, ops._stream())
        if err:
            raise err[0]
        check(rc, 'model_compute')
        return energy, ae, force, virial

    SORT_MIN_ATOMS = 4096

    @staticmethod
    def cell_order(pos, cell, cutoff, pbc=None):
        """Permutation that puts the atoms in cell-list order (bins of about one cutoff in fractional coordinates):
        neighbours of consecutive atoms overlap, so the x[src] gathers and the d x[src] RED.ADD scatter of the
        convolution kernels stay in L2.  What LAMMPS' `atom_modify sort` does for the reference's pair styles.
        Fully periodic cells on the device: the binning kernel of the brick decomposition with ONE brick
        (`b200_partition_keys`; the inverse cell is formed on the host) and one sort of its keys."""
        if pos.is_cuda and pbc is not None and all(bool(b) for b in np.asarray(pbc).reshape(3)):
            from .parallel import cell_heights
            hc = np.asarray(cell, np.float64).reshape(3, 3)
            hts = cell_heights(hc)
            nbins = [max(1, int(hts[a] / cutoff)) for a in range(3)]
            key = ops.partition_keys(pos, np.linalg.inv(hc), (1, 1, 1), (0, 0, 0), cutoff / hts, nbins)
            return torch.sort(key, stable=True)[1]
        h = torch.as_tensor(np.asarray(cell, np.float64).reshape(3, 3), device=pos.device)
        frac = torch.linalg.solve(h.T, pos.T).T
        frac = frac - torch.floor(frac)
        vol = abs(float(torch.linalg.det(h)))
        nb = []
        for k in range(3):
            area = float(torch.linalg.norm(torch.linalg.cross(h[(k + 1) % 3], h[(k + 2) % 3])))
            nb.append(max(1, int(vol / area / cutoff)))
        c = [torch.clamp((frac[:, k] * nb[k]).long(), 0, nb[k] - 1) for k in range(3)]
        return torch.argsort((c[0] * nb[1] + c[1]) * nb[2] + c[2], stable=True)

    def calculate(self, pos, cell, pbc, types, sort_atoms=True):
        """One MD force call: radius graph + E/F/S.  pos float64 [N,3] (device), cell 3x3 (host),
        types int32 [N] (device).  stress = -virial/volume in the model's order [xx,yy,zz,xy,yz,zx]
        (sevenn/nn/force_output.py:186-213).  Large systems are evaluated in cell-list atom order (`cell_order`);
        per-atom results come back in the caller's order, `graph` is in the internal order and carries `perm`
        (internal index -> caller's index)."""
        perm = None
        if sort_atoms and pos.shape[0] >= self.SORT_MIN_ATOMS:
            perm = self.cell_order(pos, cell, self.config['cutoff'], pbc)
            pos, types = pos[perm].contiguous(), types[perm].contiguous()
        g = ops.radius_graph(pos, cell, pbc, self.config['cutoff'])
        e, ae, f, vir = self.compute(types, g['rowptr'], g['src'], g['dst'], g['edge_vec'])
        if perm is not None:
            inv = torch.empty_like(perm)
            inv[perm] = torch.arange(perm.shape[0], device=perm.device)
            ae, f = ae[inv], f[inv]
            g = dict(g, perm=perm)
        vol = abs(float(np.linalg.det(np.asarray(cell, np.float64).reshape(3, 3))))
        stress = -vir / vol if vol > 0 else None
        return dict(energy=e, atomic_energy=ae, forces=f, virial=vir, stress=stress, graph=g)
