"""Seeded synthetic structure generators for the benchmark configurations.

Pure numpy; returns (pos[N,3] f64, cell[3,3] f64, pbc[3] bool, symbols list[str]).
"""
import numpy as np

# 89-species list of SevenNet-0 ('chemical_symbols_to_index' extra file of the deployed model)
SEVENNET0_SPECIES = (
    'Ac Ag Al Ar As Au B Ba Be Bi Br C Ca Cd Ce Cl Co Cr Cs Cu Dy Er Eu F Fe Ga Gd Ge H He Hf '
    'Hg Ho I In Ir K Kr La Li Lu Mg Mn Mo N Na Nb Nd Ne Ni Np O Os P Pa Pb Pd Pm Pr Pt Pu Rb Re '
    'Rh Ru S Sb Sc Se Si Sm Sn Sr Ta Tb Tc Te Th Ti Tl Tm U V W Xe Y Yb Zn Zr').split()
SPECIES_INDEX = {s: i for i, s in enumerate(SEVENNET0_SPECIES)}


def si_diamond(reps=3, a=5.43, sigma=0.05, seed=0):
    """C1: Si diamond, reps^3 conventional cells (216 atoms for reps=3), Gaussian noise."""
    basis = np.array([[0, 0, 0], [0, .5, .5], [.5, 0, .5], [.5, .5, 0],
                      [.25, .25, .25], [.25, .75, .75], [.75, .25, .75], [.75, .75, .25]])
    g = np.stack(np.meshgrid(*[np.arange(reps)] * 3, indexing='ij'), -1).reshape(-1, 3)
    pos = ((g[:, None, :] + basis[None]) * a).reshape(-1, 3)
    rng = np.random.default_rng(seed)
    pos = pos + rng.normal(0.0, sigma, pos.shape)
    return pos, np.eye(3) * a * reps, np.array([True] * 3), ['Si'] * len(pos)


_PROTOS = {
    'fcc': (np.array([[0, 0, 0], [0, .5, .5], [.5, 0, .5], [.5, .5, 0]]), [0, 0, 0, 0]),
    'bcc': (np.array([[0, 0, 0], [.5, .5, .5]]), [0, 1]),
    'rocksalt': (np.array([[0, 0, 0], [0, .5, .5], [.5, 0, .5], [.5, .5, 0],
                           [.5, .5, .5], [.5, 0, 0], [0, .5, 0], [0, 0, .5]]), [0, 0, 0, 0, 1, 1, 1, 1]),
    'diamond': (np.array([[0, 0, 0], [0, .5, .5], [.5, 0, .5], [.5, .5, 0], [.25, .25, .25],
                          [.25, .75, .75], [.75, .25, .75], [.75, .75, .25]]), [0, 0, 0, 0, 1, 1, 1, 1]),
}


def bulk_crystal(rng, target_atoms=100):
    """C2: one ~100-atom bulk crystal: random prototype, 2-3 species, 12-20 A^3/atom,
    sigma = 0.08 A displacement, random 3 % strain."""
    name = rng.choice(sorted(_PROTOS))
    basis, sub = _PROTOS[name]
    nb = len(basis)
    reps = max(1, int(round((target_atoms / nb) ** (1 / 3))))
    shape = [reps, reps, reps]
    # grow one axis at a time toward the target
    while nb * np.prod(shape) < target_atoms - 4:
        shape[int(np.argmin(shape))] += 1
    vol_per_atom = rng.uniform(12.0, 20.0)
    a = (vol_per_atom * nb) ** (1 / 3)
    g = np.stack(np.meshgrid(*[np.arange(s) for s in shape], indexing='ij'), -1).reshape(-1, 3)
    frac = (g[:, None, :] + basis[None]).reshape(-1, 3)
    cell = np.diag(np.array(shape, float) * a)
    strain = rng.uniform(-0.03, 0.03, (3, 3))
    strain = 0.5 * (strain + strain.T)
    cell = cell @ (np.eye(3) + strain)
    pos = (frac * a) @ (np.eye(3) + strain) + rng.normal(0, 0.08, frac.shape)
    nsp = int(rng.integers(2, 4))
    sp = rng.choice(len(SEVENNET0_SPECIES), nsp, replace=False)
    sites = np.tile(np.array(sub), len(g))
    types = sp[sites % nsp]
    # third species substitutes at random
    if nsp == 3:
        m = rng.random(len(types)) < 0.2
        types = np.where(m, sp[2], types)
    return pos, cell, np.array([True] * 3), [SEVENNET0_SPECIES[t] for t in types]


def bulk_batch(n_struct=32, seed=0, target_atoms=100):
    rng = np.random.default_rng(seed)
    return [bulk_crystal(rng, target_atoms) for _ in range(n_struct)]


def amorphous_cell(n_atoms=100_000, density=0.0928, dmin=1.4, species=('Si', 'O', 'O'), seed=0):
    """C3/C4: amorphous cubic cell, jittered-grid placement with a minimum distance of
    ~dmin (seeded), sized so that edges/atom ~ 50 at r_c = 5 A (rho = 0.0955 / A^3).
    Random sequential insertion is emulated by a perturbed simple-cubic grid, which keeps
    generation O(N) and guarantees the minimum distance by construction."""
    rng = np.random.default_rng(seed)
    L = (n_atoms / density) ** (1 / 3)
    m = int(np.ceil(n_atoms ** (1 / 3)))
    h = L / m
    amp = max(0.0, 0.5 * (h - dmin))  # max jitter keeping neighbours >= dmin apart
    sites = rng.permutation(m ** 3)[:n_atoms]
    g = np.stack(np.unravel_index(sites, (m, m, m)), -1).astype(float)
    pos = (g + 0.5) * h + rng.uniform(-amp, amp, (n_atoms, 3))
    types = [species[i % len(species)] for i in rng.permutation(n_atoms)]
    return pos, np.eye(3) * L, np.array([True] * 3), types
