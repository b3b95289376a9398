"""Model definitions on CPU: shapes, parameter counts, autograd path, DDP wrapping of GPT-2."""
import torch

from distributed_training_b200.models.gpt2 import GPT2, GPT2Config
from distributed_training_b200.models.mlp import MLP, wide_mlp
from distributed_training_b200.parallel.comm import SingleProcessComm
from distributed_training_b200.parallel.ddp import ReplicatedModule
from distributed_training_b200.parallel.sharded import ShardedModule, default_units


def test_gpt2_parameter_counts():
    small = GPT2(GPT2Config.small(), dtype=torch.bfloat16)
    # 124M + 47 padding rows of the embedding table (vocab padded 50257 -> 50304)
    assert small.num_parameters() == 124_439_808 + (50304 - 50257) * 768
    with torch.device("meta"):
        xl = GPT2(GPT2Config.xl(), dtype=torch.bfloat16)
    assert abs(xl.num_parameters() - 1_557_611_200) < 2_000_000


def test_gpt2_forward_backward_cpu_tiny():
    cfg = GPT2Config(vocab_size=100, n_positions=32, n_embd=64, n_layer=2, n_head=1)
    model = GPT2(cfg)
    tok = torch.randint(0, 100, (2, 32))
    logits = model(tok)
    assert logits.shape == (2, 32, 100)
    loss = model.loss(tok, tok)
    loss.backward()
    assert all(p.grad is not None for p in model.parameters())
    assert not model.native_ready()
    # causal: changing a later token must not change earlier logits
    tok2 = tok.clone()
    tok2[:, -1] = (tok2[:, -1] + 1) % 100
    assert torch.allclose(model(tok2)[:, :-1], logits[:, :-1], atol=1e-5)


def test_mlp_shapes_and_units():
    m = MLP(20, 64, 1, num_layers=3)
    assert m(torch.randn(5, 20)).shape == (5, 1)
    assert len(m.fsdp_units()) == 3 and not m.native_ready()
    with torch.device("meta"):
        w = wide_mlp()
    assert sum(p.numel() for p in w.parameters()) == 8 * (8192 * 8192 + 8192)


def test_gpt2_under_replicated_and_sharded_wrappers():
    cfg = GPT2Config(vocab_size=50, n_positions=16, n_embd=64, n_layer=2, n_head=1)
    comm = SingleProcessComm(torch.device("cpu"))
    torch.manual_seed(0)
    a, b = GPT2(cfg), GPT2(cfg)
    b.load_state_dict(a.state_dict())
    ra = ReplicatedModule(a, comm)
    sb = ShardedModule(b, comm)
    assert len(default_units(b)) == 2 + cfg.n_layer + 1
    tok = torch.randint(0, 50, (2, 16))
    la = torch.nn.functional.cross_entropy(ra(tok).view(-1, 50), tok.view(-1))
    lb = torch.nn.functional.cross_entropy(sb(tok).view(-1, 50), tok.view(-1))
    la.backward()
    lb.backward()
    assert torch.allclose(la, lb, atol=1e-6)
    for (n, p), (_, q) in zip(a.named_parameters(), b.named_parameters()):
        assert torch.allclose(p.grad, q.grad, atol=1e-5), n
    assert list(sb.state_dict().keys()) == list(a.state_dict().keys())


def test_step_params_equals_whole_model_step():
    """Stepping parameters bucket by bucket (what the data-parallel wrapper does behind each all-reduce) must give
    bit-identical results to one step() over the whole model, for both fused optimizers."""
    import torch
    from distributed_training_b200.optim import FusedAdamW, FusedSGD

    def run(split, cls, **kw):
        torch.manual_seed(0)
        ps = [torch.nn.Parameter(torch.randn(5, 3)) for _ in range(4)]
        opt = cls(ps, lr=1e-2, **kw)
        for _ in range(3):
            for p in ps:
                p.grad = torch.randn_like(p)
            if split:
                opt.step_params(ps[:2])
                opt.step_params(ps[1:3])        # overlapping call: already-updated parameters are skipped
                assert [len(opt._pending(g)) for g in opt.param_groups] == [1]
            opt.step()
            assert not opt._already_updated
        return [p.detach().clone() for p in ps]

    for cls, kw in ((FusedAdamW, dict(master_weights=False)), (FusedSGD, dict(momentum=0.9, weight_decay=0.1))):
        a, b = run(False, cls, **kw), run(True, cls, **kw)
        assert all(torch.equal(x, y) for x, y in zip(a, b)), cls.__name__
