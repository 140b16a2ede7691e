"""Host-side data-parallel logic on CPU with the gloo backend, world_size 2 (the N>1 path of acai_omr_b200.train):
batch sharding, the flat parameter / gradient layout, the single-buffer gradient all-reduce and the global loss
normaliser.  The CUDA kernels are not involved (and FlatTrainer.step must refuse to run without them)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from acai_omr_b200 import ops, train


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _make_model():
    torch.manual_seed(0)
    m = torch.nn.Sequential(torch.nn.Linear(7, 5), torch.nn.LayerNorm(5), torch.nn.Linear(5, 3))
    m[0].bias.requires_grad_(False)          # a frozen parameter must stay out of the flat buffers
    return m


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        m = _make_model()
        before = {k: v.clone() for k, v in m.state_dict().items()}
        tr = train.FlatTrainer(m, lr=1e-3)
        ok = all(torch.equal(v, before[k]) for k, v in m.state_dict().items())              # flattening keeps the values
        ok &= all(p.data_ptr() == tr.flat_param.data_ptr() + 4 * o for p, o in zip(tr.params, tr.offsets))
        ok &= all(o % train.ALIGN == 0 for o in tr.offsets) and m[0].bias.grad is None
        x = torch.full((4, 7), float(rank + 1))
        tr.zero_grad()
        m(x).sum().backward()                                                                    # torch autograd on CPU
        ok &= all(p.grad.data_ptr() == tr.flat_grad.data_ptr() + 4 * o for p, o in zip(tr.params, tr.offsets))
        local = tr.flat_grad.clone()
        tr.allreduce_gradients()
        gathered = [torch.zeros_like(local) for _ in range(world)]
        dist.all_gather(gathered, local)
        ok &= torch.allclose(tr.flat_grad, sum(gathered))
        # overlapped variant: the last layer's gradients are reduced from a tensor hook while the backward still runs, the rest
        # at the end - same sums, every element reduced exactly once
        tr.zero_grad()
        h = m[1](m[0](x))
        tr.grad_ready_hook(h, [m[2]])                                                            # fires once d loss / d h exists
        m[2](h).sum().backward()
        local2 = local.clone()
        ok &= tr._reduced == [(next(o for p, o in zip(tr.params, tr.offsets) if p is m[2].weight), tr.numel)] and len(tr._works) == 1
        tr.allreduce_gradients()
        ok &= tr._works == [] and tr._reduced == []
        ok &= torch.allclose(tr.flat_grad, sum(gathered))
        ok &= torch.equal(local2, local)
        try:
            tr.allreduce_ready(m[2].parameters()); tr.allreduce_ready(m[2].parameters())
            ok = False                                                                           # a range must not be reduced twice
        except RuntimeError:
            tr.allreduce_gradients()
        norm = tr.mean_count(torch.tensor(10 * (rank + 1)))
        ok &= float(norm) == sum(10 * (r + 1) for r in range(world)) / world
        try:
            tr.step()
            ok = False
        except ops.AcaiError:
            pass
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_flat_allreduce():
    world, port = 2, _free_port()
    out = mp.get_context("spawn").Manager().dict()
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    assert dict(out) == {0: True, 1: True}


def test_complement_of_reduced_ranges():
    assert train._complement([], 10) == [(0, 10)]
    assert train._complement([(4, 10)], 10) == [(0, 4)]
    assert train._complement([(6, 8), (0, 2)], 10) == [(2, 6), (8, 10)]
    assert train._complement([(0, 10)], 10) == []


def test_shard_partitions_the_batch():
    items = list(range(13))
    for world in (1, 2, 4, 8):
        parts = [train.shard(items, r, world) for r in range(world)]
        assert sorted(sum(parts, [])) == items
        assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1


def test_cost_balanced_sharding():
    lens = [1024, 968, 972, 1000, 1015, 1008, 980, 1024, 120, 64, 500, 333, 90, 800, 640, 77]
    for world in (2, 4, 8):
        parts = [train.shard(lens, r, world, cost=lambda v: v) for r in range(world)]
        assert sorted(sum(parts, [])) == sorted(lens)
        sums = [sum(p) for p in parts]
        naive = [sum(lens[r::world]) for r in range(world)]
        assert max(sums) - min(sums) <= max(naive) - min(naive) and max(sums) <= max(naive), (sums, naive)


def test_global_normaliser_gives_single_process_gradient():
    """sum_r (local_sum_r / (G / W)) / W == global_sum / G: the identity behind FlatTrainer.mean_count."""
    sums, counts = [3.0, 11.0, 0.5], [4.0, 9.0, 1.0]
    W, G = len(sums), sum(counts)
    assert abs(sum(s / (G / W) for s in sums) / W - sum(sums) / G) < 1e-12


def test_trainer_state_roundtrip():
    m = _make_model()
    tr = train.FlatTrainer(m, param_groups=[{"params": m[0].parameters(), "lr": 1e-3}, {"params": m[2].parameters(), "lr": 1e-4}])
    assert [g["lr"] for g in tr.param_groups] == [1e-3, 1e-4]
    assert tr.param_groups[0]["end"] == tr.param_groups[1]["start"]
    tr.exp_avg.fill_(2.0)
    tr.step_count = 7
    sd = tr.state_dict()
    tr2 = train.FlatTrainer(_make_model(), param_groups=None)
    with pytest.raises(ValueError):
        tr2.load_state_dict(sd)                    # different layout (LayerNorm parameters included)
    m3 = _make_model()
    tr3 = train.FlatTrainer(m3, param_groups=[{"params": m3[0].parameters(), "lr": 5.0}, {"params": m3[2].parameters(), "lr": 5.0}])
    tr3.load_state_dict(sd)
    assert tr3.step_count == 7 and float(tr3.exp_avg[0]) == 2.0 and [g["lr"] for g in tr3.param_groups] == [1e-3, 1e-4]
