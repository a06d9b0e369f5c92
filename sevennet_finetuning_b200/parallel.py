"""Spatially decomposed MD force call (BASELINE config 4), one process per GPU.

Mirrors sevenn/pair_e3gnn/pair_e3gnn_parallel.cpp: every rank owns the atoms of one brick of the cell
plus a ghost shell of one cutoff (:277-314); edges live on the rank of their centre atom; before each
interaction layer t >= 1 the features of the ghosts are fetched from their owners (:397-400, LAMMPS
forward_comm) and in the reverse pass the gradients w.r.t. ghost rows are returned and added (:446-453,
:907-933); ghost force contributions go back the same way (LAMMPS reverse_comm, newton on, :685).
Transport here: ``torch.distributed.all_to_all_single`` (NCCL over NVLink on the GPUs, gloo in the CPU
tests of the bookkeeping).  The per-rank math is the C-ABI engine (``b200_model_compute_halo``).

A ghost is a remote *atom*, not a periodic image: features are translation invariant and the edge vectors
(which carry the image shift) are built by the radius-graph kernel on the rank's sub-list of atoms in the
global periodic cell.
"""
import numpy as np
import torch
import torch.distributed as dist

from ._lib import B200_HALO_BWD_X, B200_HALO_FORCE, B200_HALO_FWD_X


class DistComm:
    """Transport = torch.distributed (NCCL on GPUs, gloo on CPU)."""

    def __init__(self, group=None):
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)

    def all_to_all(self, out, inp, out_splits, in_splits):
        dist.all_to_all_single(out, inp, output_split_sizes=out_splits, input_split_sizes=in_splits, group=self.group)

    def all_reduce_sum(self, t):
        dist.all_reduce(t, group=self.group)

    def all_gather(self, t):
        out = torch.empty((self.world,) + tuple(t.shape), dtype=t.dtype, device=t.device)
        dist.all_gather_into_tensor(out, t, group=self.group)
        return out


class ThreadComm:
    """In-process transport: the ranks are threads of one process sharing one device and one stream.
    Used to run the complete decomposed path on a single GPU (tests, debugging); not a product transport."""

    class _Shared:
        def __init__(self, world):
            import threading
            self.barrier = threading.Barrier(world)
            self.slots = [None] * world

    @classmethod
    def create(cls, world):
        sh = cls._Shared(world)
        return [cls(sh, r, world) for r in range(world)]

    def __init__(self, shared, rank, world):
        self.sh, self.rank, self.world = shared, rank, world

    def all_to_all(self, out, inp, out_splits, in_splits):
        off = np.concatenate([[0], np.cumsum(in_splits)]).astype(int)
        self.sh.slots[self.rank] = (inp, off)
        self.sh.barrier.wait()
        o = 0
        for r in range(self.world):
            src, soff = self.sh.slots[r]
            n = int(soff[self.rank + 1] - soff[self.rank])
            assert n == out_splits[r], 'all_to_all: split sizes disagree between ranks'
            out[o:o + n].copy_(src[soff[self.rank]:soff[self.rank + 1]])
            o += n
        self.sh.barrier.wait()

    def all_reduce_sum(self, t):
        self.sh.slots[self.rank] = t.clone()
        self.sh.barrier.wait()
        tot = sum(self.sh.slots[r] for r in range(self.world))
        self.sh.barrier.wait()
        t.copy_(tot)

    def all_gather(self, t):
        self.sh.slots[self.rank] = t.clone()
        self.sh.barrier.wait()
        out = torch.stack([self.sh.slots[r] for r in range(self.world)])
        self.sh.barrier.wait()
        return out


def brick_grid(world):
    """2 -> (2,1,1), 4 -> (2,2,1), 8 -> (2,2,2), otherwise the most cubic factorisation."""
    best = (world, 1, 1)
    for a in range(1, world + 1):
        if world % a:
            continue
        for b in range(1, world // a + 1):
            if (world // a) % b:
                continue
            c = world // a // b
            cand = tuple(sorted((a, b, c), reverse=True))
            if max(cand) - min(cand) < max(best) - min(best):
                best = cand
    return best


def cell_heights(cell):
    """Perpendicular widths of the cell along each lattice direction."""
    c = np.asarray(cell, np.float64).reshape(3, 3)
    vol = abs(np.linalg.det(c))
    return np.array([vol / np.linalg.norm(np.cross(c[(a + 1) % 3], c[(a + 2) % 3])) for a in range(3)])


class HaloPlan:
    """Ownership and exchange maps of one rank for one set of positions."""

    def __init__(self, pos, cell, pbc, cutoff, grid, rank, comm=None):
        """pos: [N,3] float64 (any device), identical on every rank."""
        world = int(np.prod(grid))
        self.comm, self.rank, self.world, self.grid = comm, rank, world, tuple(int(g) for g in grid)
        dev = pos.device
        cell = np.asarray(cell, np.float64).reshape(3, 3)
        pbc = [bool(b) for b in np.asarray(pbc).reshape(3)]
        h = cell_heights(cell)
        for a in range(3):
            if pbc[a] and h[a] <= cutoff:
                raise ValueError('parallel: periodic cell thinner than the cutoff is not supported')
        self._n_global = pos.shape[0]
        self.send_rows = None
        self.send_counts = None
        if pos.is_cuda and all(pbc):
            # device path (fully periodic cell): ONE kernel classifies every atom (owned / ghost of which owner / not needed)
            # and bins it into the cell list, one stable sort of the keys orders the rank's atoms (cell-list order) and its
            # ghosts (grouped by owner), one histogram gives all counts: no per-atom eager tensor algebra, one host read
            from . import ops
            mine = np.array(np.unravel_index(rank, self.grid))
            nb = [max(1, int(h[a] / cutoff)) for a in range(3)]
            key = ops.partition_keys(pos, np.linalg.inv(cell), self.grid, mine, cutoff / h, nb)
            skey, order = torch.sort(key, stable=True)
            nbins = nb[0] * nb[1] * nb[2]
            self._class_counts = torch.bincount(torch.div(skey, nbins, rounding_mode='floor'), minlength=world + 2)
            self._order = order
            self.presorted = True        # local_ids / ghost_ids already follow the cell list
            return
        self.presorted = False
        inv = torch.tensor(np.linalg.inv(cell), dtype=torch.float64, device=dev)
        frac = pos @ inv
        pb = torch.tensor(pbc, device=dev)
        frac = torch.where(pb, frac - torch.floor(frac), frac)
        g = torch.tensor(self.grid, device=dev)
        lo_np = np.zeros(3)
        ext_np = np.ones(3)
        if not all(pbc):  # non-periodic axes: bricks span the bounding interval of the atoms
            fmin, fmax = frac.min(0).values.cpu().numpy(), frac.max(0).values.cpu().numpy()
            for a in range(3):
                if not pbc[a]:
                    lo_np[a], ext_np[a] = fmin[a], max(fmax[a] - fmin[a], 1e-9) * (1 + 1e-9)
        lo_t = torch.tensor(lo_np, device=dev)
        ext_t = torch.tensor(ext_np, device=dev)
        u = (frac - lo_t) / ext_t                                     # brick coordinate in [0, 1)
        b = torch.clamp(torch.floor(u * g).long(), min=torch.zeros(3, dtype=torch.long, device=dev), max=g - 1)
        owner = (b[:, 0] * self.grid[1] + b[:, 1]) * self.grid[2] + b[:, 2]
        mine = np.array(np.unravel_index(rank, self.grid))
        blo = torch.tensor(mine / np.array(self.grid), device=dev)
        bhi = torch.tensor((mine + 1) / np.array(self.grid), device=dev)
        margin = torch.tensor(cutoff / (h * ext_np), device=dev)
        is_local = owner == rank
        near = torch.ones(pos.shape[0], dtype=torch.bool, device=dev)
        for a in range(3):
            ua = u[:, a]
            d = torch.clamp(torch.maximum(blo[a] - ua, ua - bhi[a]), min=0.0)
            if pbc[a]:
                for k in (-1.0, 1.0):
                    d = torch.minimum(d, torch.clamp(torch.maximum(blo[a] - (ua + k), (ua + k) - bhi[a]), min=0.0))
            near &= d <= margin[a]
        self.local_ids = torch.nonzero(is_local).flatten()
        gh = torch.nonzero(near & ~is_local).flatten()
        gown = owner[gh]
        order = torch.argsort(gown, stable=True)
        self.ghost_ids = gh[order]                                     # grouped by owner rank
        self.n_local, self.n_ghost = int(self.local_ids.numel()), int(self.ghost_ids.numel())
        self.recv_counts = torch.bincount(gown, minlength=world).cpu().tolist()
        self.subset = torch.cat([self.local_ids, self.ghost_ids])
        self._class_counts = None

    def connect(self):
        """Tell every owner which of its atoms this rank needs (two small all-to-alls per graph build)."""
        if self._class_counts is not None:
            # device path: every rank learns every (needer, owner) count from one all-gather of the class histograms; one host
            # read per graph build (the split sizes of the exchanges are host arguments of the collective)
            dev = self._order.device
            allc = self.comm.all_gather(self._class_counts[:self.world + 1].contiguous()).cpu()   # [needer, class]
            me = self.rank
            self.n_local = int(allc[me, 0])
            self.recv_counts = allc[me, 1:].tolist()           # ghosts of mine owned by r
            self.send_counts = allc[:, 1 + me].tolist()        # atoms of mine that rank r needs
            self.n_ghost = int(sum(self.recv_counts))
            self.subset = self._order[:self.n_local + self.n_ghost]
            self.local_ids, self.ghost_ids = self.subset[:self.n_local], self.subset[self.n_local:]
            want = torch.empty(sum(self.send_counts), dtype=torch.int64, device=dev)
            self.comm.all_to_all(want, self.ghost_ids.contiguous(), self.send_counts, self.recv_counts)
            g2l = torch.full((self._n_global,), -1, dtype=torch.int32, device=dev)
            g2l[self.local_ids] = torch.arange(self.n_local, device=dev, dtype=torch.int32)
            self.send_rows32 = g2l[want]
            self.send_rows = self.send_rows32.long()
            self._bufs = {}
            return self
        dev = self.subset.device
        rc = torch.tensor(self.recv_counts, dtype=torch.int64, device=dev)
        sc = torch.empty_like(rc)
        self.comm.all_to_all(sc, rc, [1] * self.world, [1] * self.world)
        self.send_counts = sc.cpu().tolist()
        want = torch.empty(sum(self.send_counts), dtype=torch.int64, device=dev)
        self.comm.all_to_all(want, self.ghost_ids.contiguous(), self.send_counts, self.recv_counts)
        g2l = torch.full((self._n_global,), -1, dtype=torch.int64, device=dev)
        g2l[self.local_ids] = torch.arange(self.n_local, device=dev)
        self.send_rows = g2l[want]
        if self.send_rows.numel() and int(self.send_rows.min()) < 0:
            raise RuntimeError('parallel: a peer asked for an atom this rank does not own')
        self.send_rows32 = self.send_rows.to(torch.int32)
        self._bufs = {}
        return self

    # ---- the three exchanges (rows: [n_local + n_ghost, width], ghosts after the locals) ----
    def _buf(self, width, dtype, device):
        # one packed send / receive buffer per row width, reused by every exchange of the step
        key = (width, dtype)
        if key not in self._bufs:
            self._bufs[key] = torch.empty(self.send_rows.numel(), width, dtype=dtype, device=device)
        return self._bufs[key]

    def fetch_ghost_rows(self, rows):
        """pack_forward_init + forward_comm + unpack_forward_init (pair_e3gnn_parallel.cpp:803-866): the owners pack the rows
        their peers asked for (CUDA kernel b200_halo_pack); the received rows ARE the ghost rows, no unpack copy."""
        from . import ops
        send = self._buf(rows.shape[1], rows.dtype, rows.device)
        if rows.is_cuda:
            ops.halo_pack(rows, self.send_rows32, send)
        else:   # gloo tests of the bookkeeping (CPU tensors never reach the engine)
            torch.index_select(rows[:self.n_local], 0, self.send_rows, out=send)
        self.comm.all_to_all(rows[self.n_local:], send, self.recv_counts, self.send_counts)

    def return_ghost_rows(self, rows):
        """pack_reverse + reverse_comm + unpack_reverse (:867-933): ghost rows go back to their owners and are added
        (b200_halo_unpack_add, RED.ADD like the reference's index_add_)."""
        from . import ops
        back = self._buf(rows.shape[1], rows.dtype, rows.device)
        self.comm.all_to_all(back, rows[self.n_local:], self.send_counts, self.recv_counts)
        if rows.is_cuda:
            ops.halo_unpack_add(rows, self.send_rows32, back)
        else:
            rows[:self.n_local].index_add_(0, self.send_rows, back)

    def hook(self, phase, layer, rows):
        if phase == B200_HALO_FWD_X:
            self.fetch_ghost_rows(rows)
        elif phase in (B200_HALO_BWD_X, B200_HALO_FORCE):
            self.return_ghost_rows(rows)
        else:
            raise ValueError(f'unknown halo phase {phase}')


class ParallelSevenNet:
    """pair_e3gnn_parallel equivalent: one instance per rank, all ranks call ``calculate`` together."""

    def __init__(self, model, grid=None, comm=None):
        self.model = model
        self.comm = comm if comm is not None else DistComm()
        self.world, self.rank = self.comm.world, self.comm.rank
        self.grid = tuple(grid) if grid is not None else brick_grid(self.world)
        self.last_plan = None

    def calculate(self, pos, cell, pbc, types):
        """pos float64 [N,3] and types int32 [N] on this rank's device, identical on all ranks.
        Returns energy/virial/stress of the whole cell and the forces of this rank's atoms
        (``local_ids`` = their global indices)."""
        from . import ops
        cutoff = self.model.config['cutoff']
        # cell-list atom order (as in B200SevenNet.calculate): local and ghost rows of a brick follow it, so the
        # gathers and the RED scatter of the convolution kernels stay L2-local; local_ids are mapped back below
        perm = None
        device_path = pos.is_cuda and all(bool(b) for b in np.asarray(pbc).reshape(3))
        if not device_path and pos.shape[0] >= self.model.SORT_MIN_ATOMS:
            perm = self.model.cell_order(pos, cell, cutoff)
            pos, types = pos[perm].contiguous(), types[perm].contiguous()
        plan = HaloPlan(pos, cell, pbc, cutoff, self.grid, self.rank, self.comm).connect()
        self.last_plan = plan
        nl, ng = plan.n_local, plan.n_ghost
        g = ops.radius_graph(pos.index_select(0, plan.subset).contiguous(), cell, pbc, cutoff)
        e_loc = int(g['rowptr'][nl].item())                 # edges whose centre is a local atom
        rowptr = torch.clamp(g['rowptr'], max=e_loc)
        src, dst, vec = g['src'][:e_loc], g['dst'][:e_loc], g['edge_vec'][:e_loc]
        t_sub = types.index_select(0, plan.subset).contiguous()
        e, ae, f, vir = self.model.compute(t_sub, rowptr, src, dst, vec, n_ghost=ng, halo=plan.hook)
        tot = torch.cat([e, vir])
        self.comm.all_reduce_sum(tot)
        vol = abs(float(np.linalg.det(np.asarray(cell, np.float64).reshape(3, 3))))
        local_ids = plan.local_ids if perm is None else perm[plan.local_ids.long()].to(plan.local_ids.dtype)
        return dict(energy=tot[:1], virial=tot[1:], stress=-tot[1:] / vol if vol > 0 else None,
                    atomic_energy=ae, forces=f[:nl], local_ids=local_ids, n_local=nl, n_ghost=ng, n_edges=e_loc)
