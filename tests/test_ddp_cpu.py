"""DDP wrapper on CPU/gloo, 2 processes: replicas stay identical and match single-process training on the
concatenated batch; no_sync() defers the reduction (gradient accumulation)."""
import torch

from dist_utils import run_distributed


def _model():
    from llm_in_practise_b200.models.gpt import GPTLike
    torch.manual_seed(0)
    return GPTLike(vocab_size=128, block_size=16, n_layer=2, n_head=2, d_model=32, dropout=0.0)


def _worker(rank, world, tmp):
    from llm_in_practise_b200.parallel.dist import setup_distributed
    from llm_in_practise_b200.parallel.ddp import DistributedDataParallel as DDP
    setup_distributed()
    torch.manual_seed(100 + rank)            # different init per rank: the ctor broadcast must fix it
    from llm_in_practise_b200.models.gpt import GPTLike
    m = GPTLike(vocab_size=128, block_size=16, n_layer=2, n_head=2, d_model=32, dropout=0.0)
    if rank == 0:
        m.load_state_dict(_model().state_dict())
    ddp = DDP(m, bucket_cap_mb=0.02)
    assert len(ddp.buckets) > 1
    opt = torch.optim.SGD(ddp.parameters(), lr=0.1)
    g = torch.Generator().manual_seed(3)
    data = torch.randint(0, 128, (8, 4, 16), generator=g)
    for step in range(2):
        ddp.zero_grad()
        with ddp.no_sync():
            x = data[step * 4 + rank]
            (ddp(x, x) / 2).backward()
        x = data[step * 4 + 2 + rank]
        (ddp(x, x) / 2).backward()
        opt.step()
    torch.save(ddp.arena.data.clone(), f"{tmp}/r{rank}.pt")


def test_ddp_matches_single_process(tmp_path):
    run_distributed(_worker, 2, (str(tmp_path),))
    a, b = torch.load(tmp_path / "r0.pt"), torch.load(tmp_path / "r1.pt")
    assert torch.equal(a, b)
    from llm_in_practise_b200.utils.arena import flatten_module
    m = _model()
    arena = flatten_module(m, pad_to=8)
    opt = torch.optim.SGD(m.parameters(), lr=0.1)
    g = torch.Generator().manual_seed(3)
    data = torch.randint(0, 128, (8, 4, 16), generator=g)
    for step in range(2):
        arena.zero_grad()
        for j in range(4):
            x = data[step * 4 + j]
            (m(x, x) / 4).backward()
        opt.step()
    assert torch.allclose(a[: arena.numel], arena.data, atol=1e-5), (a[: arena.numel] - arena.data).abs().max()
