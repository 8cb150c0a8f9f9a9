"""Host logic of the rollout prompt-token cache (lipvq_vae_b200/cache.py) with a stand-in tokenizer:
the cache never computes tokens itself, so a counting module shows exactly when the wrapped module runs."""
import torch

from lipvq_vae_b200.cache import PromptTokenCache


class _CountingTokenizer(torch.nn.Module):
    """Same call contract as LLFQVAE_V4.forward (v5:70-84): x[N,F] -> (z_latent detached, 0-dim loss)."""

    def __init__(self):
        super().__init__()
        self.lin = torch.nn.Linear(12, 8)
        self.calls = 0

    def forward(self, x):
        self.calls += 1
        z = self.lin(x)
        self.last_indices = z.argmax(dim=1)
        return z.detach(), (z ** 2).mean()


def test_same_prompt_is_tokenized_once_under_no_grad():
    m = _CountingTokenizer()
    c = PromptTokenCache(m, capacity=4)
    g = torch.Generator().manual_seed(0)
    x = torch.randn(80, 12, generator=g)
    with torch.no_grad():
        z0, l0 = c(x)
        z1, l1 = c(x.clone())            # a NEW tensor object with the same content: host key is the content hash
        z2, _ = c(x + 1.0)               # different prompt
    assert m.calls == 2 and c.hits == 1 and c.misses == 2
    assert torch.equal(z0, z1) and torch.equal(l0, l1) and not torch.equal(z0, z2)
    assert torch.equal(c.last_indices, z2.argmax(dim=1))


def test_parameter_update_invalidates():
    m = _CountingTokenizer()
    c = PromptTokenCache(m)
    x = torch.randn(10, 12)
    opt = torch.optim.AdamW(c.parameters(), lr=1e-3, weight_decay=1e-4)     # icl.py:886-889 builds it the same way
    assert len(list(c.parameters())) == len(list(m.parameters()))
    with torch.no_grad():
        z0, _ = c(x)
        c(x)
    assert m.calls == 1
    _, loss = c(x)                        # autograd on: straight to the module, graph intact
    assert m.calls == 2 and loss.requires_grad
    loss.backward()
    opt.step()                            # in-place update bumps the parameter versions
    with torch.no_grad():
        z1, _ = c(x)
    assert m.calls == 3 and not torch.equal(z0, z1)
    sd = {k: v.clone() for k, v in m.state_dict().items()}
    m.load_state_dict(sd)                 # strict reload (algo.py:337) also invalidates
    with torch.no_grad():
        c(x)
    assert m.calls == 4


def test_lru_capacity_and_passthrough_with_grad():
    m = _CountingTokenizer()
    c = PromptTokenCache(m, capacity=2)
    xs = [torch.full((4, 12), float(i)) for i in range(3)]
    with torch.no_grad():
        for x in xs:
            c(x)
        c(xs[2]); c(xs[1])               # both still cached
        assert m.calls == 3
        c(xs[0])                          # evicted by capacity 2
    assert m.calls == 4
    before = m.calls
    c(xs[0]); c(xs[0])                    # grad enabled: never cached
    assert m.calls == before + 2
    c.clear()
    with torch.no_grad():
        c(xs[0])
    assert m.calls == before + 3


def test_wrapper_is_state_dict_transparent():
    """wrapping must not change the checkpoint keys (strict load/save at robomimic/algo/algo.py:323-337)."""
    m = _CountingTokenizer()
    c = PromptTokenCache(m)
    assert list(c.state_dict().keys()) == list(m.state_dict().keys()) == ["lin.weight", "lin.bias"]
    assert [id(p) for p in c.parameters()] == [id(p) for p in m.parameters()]
    # as a registered submodule of a parent, the keys carry the parent's prefix only
    parent = torch.nn.Module()
    parent.action_network = c
    ref_parent = torch.nn.Module()
    ref_parent.action_network = _CountingTokenizer()
    assert list(parent.state_dict().keys()) == list(ref_parent.state_dict().keys())
    parent.load_state_dict(ref_parent.state_dict(), strict=True)
    assert torch.equal(m.lin.weight, ref_parent.action_network.lin.weight)
    ref_parent.load_state_dict(parent.state_dict(), strict=True)
    # attributes of the wrapped tokenizer stay reachable through the wrapper
    assert c.calls == m.calls and c.model is m


def test_raw_pointer_update_invalidates_through_param_epoch():
    """FlatAdamW / ema_update / TokenizerTrainer write parameters through raw pointers or ``.data`` --
    no autograd version bump -- and advance ``model._param_epoch`` instead; the cache must miss afterwards."""
    from lipvq_vae_b200.dist import _bump_param_epoch
    m = _CountingTokenizer()
    c = PromptTokenCache(m)
    x = torch.randn(10, 12)
    with torch.no_grad():
        z0, _ = c(x)
        c(x)
        assert m.calls == 1
        m.lin.weight.data.add_(1.0)          # what a raw-pointer kernel write looks like to autograd: nothing
        versions_only = tuple((p.data_ptr(), p._version) for p in m.parameters())
        assert versions_only == c._param_state[1:]
        _bump_param_epoch(m)                  # what FlatAdamW.step / ema_update / TokenizerTrainer.step do
        z1, _ = c(x)
    assert m.calls == 2 and not torch.equal(z0, z1)
