"""Model family used by the DDP benchmark: shapes, parameter count, a CPU forward/backward."""
import torch

from akka_allreduce_b200.models.gpt2 import GPT2, GPT2Config


def test_gpt2_small_parameter_count():
    cfg = GPT2Config.small()
    with torch.device("meta"):
        m = GPT2(cfg)
    assert m.num_parameters() == 124_439_808  # GPT-2 small with tied output embedding


def test_gpt2_tiny_forward_backward_cpu():
    torch.manual_seed(0)
    cfg = GPT2Config.tiny()
    m = GPT2(cfg)
    x = torch.randint(0, cfg.vocab_size, (2, 17))
    loss = m(x[:, :-1], x[:, 1:])
    loss.backward()
    assert torch.isfinite(loss) and abs(float(loss) - torch.log(torch.tensor(float(cfg.vocab_size)))) < 1.0
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in m.parameters())
    assert m(x[:, :-1]).shape == (2, 16, cfg.vocab_size)
