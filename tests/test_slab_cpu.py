"""Host-side logic of the slab decomposition on CPU: layout bookkeeping, and the halo exchange over torch.distributed
(gloo, world_size 2 and 3) with the CPU oracle standing in for the rank-local operator."""
import os
import socket

import numpy as np
import pytest

from conftest import MATERIAL, ROOT


def test_layout_covers_grid_exactly_once():
    from ndr_b200.slab import SlabLayout
    for E0, levels, world in ((128, 3, 1), (128, 3, 2), (256, 3, 8), (512, 3, 8), (48, 3, 3), (40, 2, 4), (32, 1, 5)):
        gE = [E0, 8, 4]
        w = 1 << levels
        node_owner = np.zeros(E0 + 1, int)
        elem_owner = np.zeros(E0, int)
        for r in range(world):
            L = SlabLayout(gE, levels, r, world)
            assert L.layer_lo % w == 0 and L.layer_hi % w == 0 and L.local_elems[0] % w == 0
            assert L.local_planes == L.local_elems[0] + 1
            assert (L.own0 == 0) == (r == 0) and L.own0 in (0, w)
            # interior faces: first owned plane even, last owned plane odd (the half-sweep exchange schedule relies on it)
            if r > 0:
                assert (L.layer_lo + L.own0) % 2 == 0
            if r < world - 1:
                assert (L.layer_lo + L.own1) % 2 == 1 and L.own1 == L.local_planes - 2
            node_owner[L.layer_lo + L.own0: L.layer_lo + L.own1 + 1] += 1
            elem_owner[L.layer_lo + L.elem_own0: L.layer_lo + L.elem_own1] += 1
            # every element touching an owned node is local
            assert L.layer_lo <= max(L.layer_lo + L.own0 - 1, 0) and L.layer_lo + L.own1 <= L.layer_hi
            ns, es = L.node_slice(), L.elem_slice()
            assert (ns.stop - ns.start) == L.local_planes * L.plane_nodes and (es.stop - es.start) == L.local_elems[0] * L.plane_elems
        assert np.all(node_owner == 1) and np.all(elem_owner == 1)


def test_layout_rejects_bad_splits():
    from ndr_b200.slab import SlabLayout
    with pytest.raises(RuntimeError):
        SlabLayout([36, 8, 8], 3, 0, 2)   # not a multiple of 2^levels
    with pytest.raises(RuntimeError):
        SlabLayout([32, 8, 8], 3, 0, 8)   # fewer coarsest layers than ranks


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, grid, levels, q):
    import sys
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from ndr_b200.slab import SlabLayout, exchange_ghost_planes
    from oracle import oracle as O
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        O.set_num_threads(1)
        dom = np.array([2.0, 1.0, 1.0])
        og = O.OracleSim([0, 0, 0], dom, grid); og.read_material(MATERIAL)
        rho = np.random.default_rng(88).uniform(0, 1, og.ne); og.set_densities(rho); og.set_simp(1.0, 1e-4, 3.0)
        u = np.random.default_rng(89).standard_normal((og.nn, 3))
        Ku = og.apply_k(u)
        L = SlabLayout(grid, levels, rank, world)
        h0 = dom[0] / grid[0]
        ol = O.OracleSim([L.layer_lo * h0, 0, 0], [L.layer_hi * h0, dom[1], dom[2]], L.local_elems); ol.read_material(MATERIAL)
        ol.set_densities(rho[L.elem_slice()]); ol.set_simp(1.0, 1e-4, 3.0)
        ul = torch.from_numpy(u[L.node_slice()].copy())
        planes = ul.view(L.local_planes, -1)
        planes[: L.own0] = 7.0e3            # every ghost plane starts with garbage
        planes[L.own1 + 1:] = -3.0e3
        exchange_ghost_planes(ul, L, dist)
        out = ol.apply_k(ul.numpy())
        err = np.max(np.abs(out[L.owned_node_rows()] - Ku[L.owned_global_node_rows()])) / np.max(np.abs(Ku))
        # global dot product assembled from the owned rows
        part = torch.tensor([float(np.sum(ul.numpy()[L.owned_node_rows()] * out[L.owned_node_rows()]))], dtype=torch.float64)
        dist.all_reduce(part)
        derr = abs(part.item() - float(np.sum(u * Ku))) / abs(float(np.sum(u * Ku)))
        q.put((rank, float(err), float(derr)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("grid,levels,world", [([16, 4, 4], 2, 2), ([24, 4, 4], 1, 3)])
def test_halo_exchange_gloo(oracle_mod, grid, levels, world):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, world, port, grid, levels, q)) for r in range(world)]
    for p in ps:
        p.start()
    res = [q.get(timeout=180) for _ in range(world)]
    for p in ps:
        p.join(60)
        assert p.exitcode == 0
    for rank, err, derr in res:
        assert err < 1e-13, (rank, err)
        assert derr < 1e-12, (rank, derr)
