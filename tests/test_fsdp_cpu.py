"""FSDP wrapper on CPU/gloo (2 processes): FULL_SHARD / SHARD_GRAD_OP / root-only fully_shard all reproduce
single-process SGD on the concatenated batch, and the full state dict gathers every parameter."""
import pytest
import torch

from dist_utils import run_distributed


def _model():
    from llm_in_practise_b200.models.gpt import GPTLike
    torch.manual_seed(0)
    return GPTLike(vocab_size=128, block_size=16, n_layer=2, n_head=2, d_model=32, dropout=0.0)


def _worker(rank, world, mode, tmp):
    from llm_in_practise_b200.models.transformer import TransformerBlock
    from llm_in_practise_b200.parallel import fsdp as F
    from llm_in_practise_b200.parallel.dist import setup_distributed
    setup_distributed()
    m = _model()
    if mode == "fsdp2":
        model = F.fully_shard(m, mesh=F.DeviceMesh("cpu", list(range(world))))
    else:
        strat = F.ShardingStrategy.FULL_SHARD if mode == "full" else F.ShardingStrategy.SHARD_GRAD_OP
        model = F.FSDP(m, auto_wrap_policy=F.transformer_auto_wrap_policy(transformer_layer_cls={TransformerBlock}),
                       sharding_strategy=strat, sync_module_states=True)
        assert len(model.units) == 3
    opt = torch.optim.SGD(model.parameters(), lr=0.1)
    g = torch.Generator().manual_seed(3)
    data = torch.randint(0, 128, (4, 4, 16), generator=g)
    for step in range(2):
        model.zero_grad()
        x = data[step * 2 + rank]
        model(x, x).backward()
        opt.step()
    sd, _ = F.get_state_dict(model, options=F.StateDictOptions(full_state_dict=True))
    if rank == 0:
        torch.save(sd, f"{tmp}/sd.pt")


@pytest.mark.parametrize("mode", ["full", "grad_op", "fsdp2"])
def test_fsdp_matches_single_process(mode, tmp_path):
    run_distributed(_worker, 2, (mode, str(tmp_path)))
    sd = torch.load(tmp_path / "sd.pt")
    m = _model()
    opt = torch.optim.SGD(m.parameters(), lr=0.1)
    g = torch.Generator().manual_seed(3)
    data = torch.randint(0, 128, (4, 4, 16), generator=g)
    for step in range(2):
        opt.zero_grad()
        for r in range(2):
            x = data[step * 2 + r]
            (m(x, x) / 2).backward()
        opt.step()
    ref = m.state_dict()
    assert set(ref) == set(sd)
    for k in ref:
        assert torch.allclose(ref[k].float(), sd[k].float(), atol=2e-5), (k, (ref[k].float() - sd[k].float()).abs().max())
