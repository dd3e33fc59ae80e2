"""World-size-2 (and 3) run of the multi-GPU host logic on CPU over gloo: sample sharding + film reduce
(path_tracing_b200/multi.py).  The per-rank renderer is the TEST-ONLY host build of the device stage bodies
(tests/hostsim), so the whole N>1 data path except the CUDA launch itself is exercised: shard films are raw sums of
disjoint sample index sets, and their reduction equals the oracle's single-process film."""
import os
import socket
import sys
import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CFG = dict(W=20, H=16, spp=5, spl=4, depth=4, seed=13)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_path, integrator=0):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from hostsim_py import HostSim
    from path_tracing_b200 import multi
    from path_tracing_b200.scene import reference_scene
    hs = HostSim(reference_scene("cornell"))
    c = CFG

    def render(off, stride):
        if integrator == 4:                                     # B2_STRATEGIES_MIS: eye samples AND their light subpaths shard by sample index (csrc/b2_mis.cuh)
            return hs.render_strategies(4, c["W"], c["H"], spp=c["spp"], eye_depth=c["depth"], light_depth=c["depth"] + 1, seed=c["seed"],
                                        sample_offset=off, sample_stride=stride, film_mode=1)[0]
        if integrator == 1:                                     # the pt mode shards its samples the same way (csrc/b2_integrators.cu)
            return hs.render_other(1, c["W"], c["H"], spp=c["spp"], eye_depth=c["depth"], seed=c["seed"], sample_offset=off, sample_stride=stride, film_mode=1)[0]
        film, _ = hs.render(c["W"], c["H"], c["spp"], c["spl"], c["depth"], c["depth"], seed=c["seed"], sample_offset=off,
                            sample_stride=stride, film_mode=1)
        return film

    film, n_local = multi.render_sharded(render, c["spp"], torch.zeros(1))
    counts = torch.tensor([n_local])
    dist.all_reduce(counts)
    assert int(counts) == c["spp"]
    if rank == 0:
        np.save(out_path, film.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_render_and_reduce_equal_the_single_process_film(tmp_path, world):
    from hostsim_py import build
    build()
    out = str(tmp_path / "film.npy")
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    got = np.load(out)
    from oracle import oracle_py as O
    from path_tracing_b200.scene import reference_scene
    f = reference_scene("cornell").to_flat()
    d = O.SceneDesc(f["kind"], f["group"], f["geom"], f["mtl"], f["lights"], f["cam"], f["W"], f["H"])
    P = O.Port(); h = P.scene(d); c = CFG
    want = P.render(h, d.cam, c["depth"], c["depth"], c["W"], c["H"], c["spp"], c["spl"], P.opts(rng_mode=P.PHILOX, seed=c["seed"], max_delta=256))
    assert got.shape == want.shape and want.mean() > 0
    assert np.allclose(got, want, rtol=1e-4, atol=1e-6)


def test_sharded_pt_render_equals_the_single_process_film(tmp_path):
    from hostsim_py import build
    build()
    out = str(tmp_path / "film_pt.npy")
    mp.spawn(_worker, args=(2, _free_port(), out, 1), nprocs=2, join=True)
    got = np.load(out)
    from oracle import oracle_py as O
    from path_tracing_b200.scene import reference_scene
    f = reference_scene("cornell").to_flat()
    d = O.SceneDesc(f["kind"], f["group"], f["geom"], f["mtl"], f["lights"], f["cam"], f["W"], f["H"])
    P = O.Port(); h = P.scene(d); c = CFG
    want = P.render_pt(h, d.cam, c["depth"], c["W"], c["H"], c["spp"], P.opts(rng_mode=P.PHILOX, seed=c["seed"], max_delta=256))
    assert got.shape == want.shape and want.mean() > 0
    assert np.allclose(got.mean(axis=(0, 1)), want.mean(axis=(0, 1)), rtol=2e-3)
    assert (np.abs(got - want).max(axis=2) <= 1e-3 * want.mean()).mean() >= 0.99


def test_sharded_mis_render_equals_the_single_process_film(tmp_path):
    """The t = 1 splats of a shard land in pixels of every other shard's samples: the film reduce has to be a SUM of raw films."""
    from hostsim_py import build, HostSim
    from path_tracing_b200.scene import reference_scene
    build()
    out = str(tmp_path / "film_mis.npy")
    mp.spawn(_worker, args=(2, _free_port(), out, 4), nprocs=2, join=True)
    got = np.load(out)
    hs = HostSim(reference_scene("cornell")); c = CFG
    want, _ = hs.render_strategies(4, c["W"], c["H"], spp=c["spp"], eye_depth=c["depth"], light_depth=c["depth"] + 1, seed=c["seed"])
    hs.close()
    assert got.shape == want.shape and want.mean() > 0
    assert np.allclose(got, want, rtol=1e-4, atol=1e-5 * want.mean())


def test_shard_plan():
    from path_tracing_b200 import multi
    for world in (1, 2, 3, 8):
        for spp in (0, 1, 5, 64):
            seen = []
            for r in range(world):
                off, stride, n = multi.shard(r, world, spp)
                idx = list(range(off, spp, stride))
                assert len(idx) == n
                seen += idx
            assert sorted(seen) == list(range(spp))
    with pytest.raises(ValueError):
        multi.shard(2, 2, 4)
