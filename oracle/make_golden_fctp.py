"""Golden vectors for the 'nequip' self connection (FullyConnectedTensorProduct of the node features with the one-hot
species vector, sevenn/nn/self_connection.py:11-38).  Test infrastructure: run in the build container only.

e3nn is not vendored, so the reference's arithmetic is taken from the frozen TorchScript graph of its own example model
(example_inputs/md_serial_example/deployed_serial.pt, written by sevenn/scripts/deploy.py).  The script pulls the two
weight constants of layer 1's ``1_self_connection_intro`` out of that file and evaluates the graph's own lines 331-356
(mul by 1/sqrt(mul_in n_species), einsum / tensordot with the constants) on seeded inputs.

    python oracle/make_golden_fctp.py     ->  tests/golden/fctp_hfo.npz
"""
import os

import numpy as np
import torch

REF = '/root/reference/example_inputs/md_serial_example/deployed_serial.pt'
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', 'tests', 'golden', 'fctp_hfo.npz')


def main():
    m = torch.jit.load(REF, map_location='cpu')
    consts = m.code_with_constants[1].const_mapping
    c12, c13 = consts['c12'].detach(), consts['c13'].detach()      # [4, 2, 12] (0e -> 0e), [4, 2, 4] (1 -> 1)
    g = torch.Generator().manual_seed(7)
    n = 37
    x = torch.randn(n, 16, generator=g)                             # 4x0e + 4x1
    types = torch.randint(0, 2, (n,), generator=g)
    onehot = torch.nn.functional.one_hot(types, 2).to(torch.float32)
    # the frozen graph's operations, forward() lines 338-352
    x0, x1 = x[:, 0:4], x[:, 4:16].reshape(n, 4, 3)
    mul3 = onehot * 0.35355339059327379
    out0 = torch.tensordot(torch.einsum('cb,ca->cba', mul3, x0), c12, [[1, 2], [1, 0]]).reshape(n, 12)
    out1 = torch.einsum('dbc,dba->dca', torch.tensordot(mul3, c13, [[1], [1]]), x1).reshape(n, 12)
    np.savez(OUT, x=x.numpy(), types=types.numpy().astype(np.int64), w0=c12.numpy(), w1=c13.numpy(),
             out=torch.cat([out0, out1], 1).numpy())
    print('wrote', OUT)


if __name__ == '__main__':
    main()
