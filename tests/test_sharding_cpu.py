"""World-size-2 gloo test of the data-parallel path (host logic; CUDA launchers emulated):
gradients after the flat-bucket all-reduce == single-process gradients on the concatenated batch."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, name, out_dir):
    for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
        sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import emu_kernels
    import cmf_b200.kernels as K
    for n in emu_kernels.ALL:
        setattr(K, n, getattr(emu_kernels, n))
    from conftest import load_golden
    from cases import CASES, make_input
    from helpers import build_product, strip_wrappers
    from cmf_b200.non_square_helpers import get_non_square_train_metrics
    from cmf_b200.parallel import GradientBucket, allreduce_gradients, shard_batch
    g = load_golden(name)
    density = build_product(g)
    bucket = GradientBucket(density.parameters())
    x = shard_batch(make_input(CASES[name]))
    loss = get_non_square_train_metrics(g["config"])[0](strip_wrappers(density), x, g["case"]["epoch"])["loss"]
    loss.backward()
    loss = allreduce_gradients(bucket, loss)
    first = bucket.flat.clone()
    # a second step the way the reference's trainer runs it (`cmf/trainer.py:211-222`): a real optimizer
    # whose zero_grad() drops the .grad tensors (set_to_none=True is torch's default) -- the bucket must
    # pick the fresh gradients up again, not reduce its stale buffer
    opt = torch.optim.SGD(density.parameters(), lr=0.0)
    opt.zero_grad()
    assert all(p.grad is None for p in density.parameters())
    loss2 = get_non_square_train_metrics(g["config"])[0](strip_wrappers(density), x, g["case"]["epoch"])["loss"]
    loss2.backward()
    assert bucket.sync_aliases() > 0 and bucket.sync_aliases() == 0
    allreduce_gradients(bucket, loss2)
    opt.step()
    assert torch.allclose(bucket.flat, first, rtol=1e-6, atol=1e-12), "second step (after zero_grad) reduced stale gradients"
    assert all(p.grad is not None and p.grad.data_ptr() >= bucket.flat.data_ptr() for p in bucket.params)
    if rank == 0:
        torch.save({"loss": loss, "flat": bucket.flat.clone(),
                    "names": [n for n, p in density.named_parameters() if p.requires_grad]},
                   os.path.join(out_dir, "ddp.pt"))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("name", ["sphere_gij_d3", "img8_gij_d8"])
def test_two_rank_gradients_equal_single_process(tmp_path, name):
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    mp.spawn(_worker, args=(2, port, name, str(tmp_path)), nprocs=2, join=True)
    res = torch.load(tmp_path / "ddp.pt")
    from conftest import load_golden
    g = load_golden(name)
    assert abs(float(res["loss"]) - float(g["loss"])) <= 1e-4 * abs(float(g["loss"]))
    # flat bucket order == named_parameters order; compare against the reference's gradient probes
    from cases import grad_probe
    import emu_kernels, cmf_b200.kernels as K
    from helpers import build_product
    density = build_product(g)
    off = 0
    for n, p in density.named_parameters():
        k = n[7:] if n.startswith("module.") else n
        grad = res["flat"][off:off + p.numel()].view_as(p)
        off += p.numel()
        if k in g["grads"]:
            probe, mine = g["grads"][k], grad_probe(k, grad)
            assert abs(mine["norm"] - probe["norm"]) <= 1e-4 * max(probe["norm"], 1e-12), k
        else:
            assert float(grad.abs().max()) == 0.0
    assert off == res["flat"].numel()
