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


def _worker_mixed(rank, world, tmp):
    from llm_in_practise_b200.parallel.dist import setup_distributed
    from llm_in_practise_b200.parallel.ddp import DistributedDataParallel as DDP
    setup_distributed()
    m = _model().to(torch.bfloat16)
    ddp = DDP(m, bucket_cap_mb=0.02, mixed_precision=torch.bfloat16)
    assert sum(len(b.params) for b in ddp.buckets) == len(list(m.parameters()))   # every parameter owns a bucket slot
    opt = ddp.make_optimizer(lr=1e-2, weight_decay=0.0)
    g = torch.Generator().manual_seed(3)
    data = torch.randint(0, 128, (8, 4, 16), generator=g)
    for step in range(2):
        x = data[2 * step + rank]            # different data per rank: replicas stay equal only if gradients are reduced
        ddp(x, x).backward()
        opt.step()
        ddp.zero_grad()
    torch.save(ddp.arena.data.clone(), f"{tmp}/m{rank}.pt")


def test_ddp_mixed_precision_reduces_gradients(tmp_path):
    """bf16 working weights + fp32 masters: parameters are views of the low-precision twin, and their gradients must
    still find their buckets (the replicas see different data and must end up with identical masters)."""
    run_distributed(_worker_mixed, 2, (str(tmp_path),))
    a, b = torch.load(tmp_path / "m0.pt"), torch.load(tmp_path / "m1.pt")
    assert torch.equal(a, b)
    assert a.isfinite().all() and a.abs().sum() > 0
