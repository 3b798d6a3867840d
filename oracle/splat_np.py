"""TEST INFRASTRUCTURE ONLY (imported by tests/; never by simubox_b200/).

numpy restatement of ``XmlTransformer.transform`` (SimuBox/toolkit/xmltrans.py:117-153), pinned against the
live reference by tests/golden/splat_*.npz (oracle/make_golden.py::gen_splat)."""
from itertools import product

import numpy as np


def transform(xyz, index, n_types, NxNyNz, lxlylz, r_cut=2.0):
    NxNyNz = np.array(NxNyNz, dtype=int)
    lxlylz = np.asarray(lxlylz, dtype=np.float64)
    h = lxlylz / NxNyNz                                              # :102
    phi = np.zeros([n_types, *NxNyNz])                               # :128
    idx_lst = np.array(r_cut * 2 / h + 2, dtype=int)                 # :130
    res = np.array(list(product(*[range(i) for i in idx_lst])), dtype=int)
    xyz = np.asarray(xyz, dtype=np.float64)
    base = np.array((np.floor((xyz - r_cut) / h) * h + lxlylz / 2.0) / h, dtype=int)  # :136-143
    for i in range(len(base)):                                       # :145-151
        grids = (res + base[i]) % NxNyNz
        coord = grids * h - lxlylz / 2
        dis = np.sqrt(np.sum((coord - xyz[i]) ** 2, axis=1))
        g = grids[dis <= r_cut]
        np.add.at(phi[index[i]], (g[:, 0], g[:, 1], g[:, 2]), 1)
    return phi
